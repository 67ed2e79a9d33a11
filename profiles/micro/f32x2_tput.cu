// Issue-rate probe: scalar FFMA/FADD vs packed FFMA2/FADD2 on sm_100a.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o f32x2_tput f32x2_tput.cu && ./f32x2_tput
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE> __global__ void __launch_bounds__(512) k(float *out, float a, float b, int iters) {
    float2 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
    const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) { v[i].x = fmaf(v[i].x, a, b); v[i].y = fmaf(v[i].y, a, b); }
            if (MODE == 1) v[i] = __ffma2_rn(v[i], a2, b2);
            if (MODE == 2) { v[i].x = v[i].x + a; v[i].y = v[i].y + a; }
            if (MODE == 3) v[i] = __fadd2_rn(v[i], a2);
            if (MODE == 4) { v[i].x = fmaf(v[i].x, a, v[(i + 1) & 7].y); v[i].y = fmaf(v[i].y, b, v[(i + 3) & 7].x); }   // 3-register FFMA
            if (MODE == 5) v[i] = __ffma2_rn(v[i], a2, v[(i + 1) & 7]);
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += v[i].x + v[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char *name, float *d) {
    int iters = 4096;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * 2, 512>>>(d, 1.0001f, 0.5f, 16);
    cudaEventRecord(e0);
    k<MODE><<<148 * 2, 512>>>(d, 1.0001f, 0.5f, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double lane_ops = 148.0 * 2 * 512 * (double)iters * 16;   // scalar-equivalent ops
    printf("%-28s %.3f ms  %.1f G lane-ops/s  (%.1f lane-ops/clk/SM at 1.965 GHz)\n", name, ms,
           lane_ops / ms / 1e6, lane_ops / (ms * 1e-3) / 148 / 1.965e9);
}
int main() {
    float *d; cudaMalloc(&d, 148 * 2 * 512 * 4);
    run<0>("FFMA (reg,imm/const)", d); run<1>("FFMA2", d); run<2>("FADD", d); run<3>("FADD2", d);
    run<4>("FFMA 3-reg", d); run<5>("FFMA2 3-reg", d);
    return 0;
}
