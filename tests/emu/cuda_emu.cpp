// Fiber scheduler for tests/emu/cuda_emu.h (test infrastructure only).
#include "cuda_emu.h"

namespace emu {
State S;
static const size_t STACK = 256 * 1024;
static std::vector<ucontext_t> ctxs;
static std::vector<unsigned char *> stacks;
static std::vector<char> done;
static const std::function<void()> *g_body;
static std::vector<unsigned char> g_smem;

void yield() { swapcontext(&ctxs[S.cur], &S.main_ctx); }

void block_barrier() {
    int gen = S.block_gen;
    if (++S.block_arrived == S.alive) {
        S.block_arrived = 0;
        ++S.block_gen;
        return;
    }
    while (S.block_gen == gen) yield();
}

void warp_barrier() {
    int w = S.cur / 32;
    int lanes = std::min(32, S.nthreads - w * 32);
    int gen = S.warp_gen[w];
    if (++S.warp_arrived[w] == lanes) {
        S.warp_arrived[w] = 0;
        ++S.warp_gen[w];
        return;
    }
    while (S.warp_gen[w] == gen) yield();
}

static void trampoline() {
    (*g_body)();
    done[S.cur] = 1;
    --S.alive;
    // a thread that exits releases a pending block barrier (CUDA semantics
    // for exited threads on sm_70+: they no longer participate)
    if (S.alive > 0 && S.block_arrived == S.alive) {
        S.block_arrived = 0;
        ++S.block_gen;
    }
    swapcontext(&ctxs[S.cur], &S.main_ctx);
}

static void set_tid(int t) {
    S.cur = t;
    S.tIdx.x = t % S.bDim.x;
    S.tIdx.y = (t / S.bDim.x) % S.bDim.y;
    S.tIdx.z = t / (S.bDim.x * S.bDim.y);
}

void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body) {
    int nt = block.x * block.y * block.z;
    if (nt <= 0 || nt > 1024) { std::fprintf(stderr, "emu: bad block size %d\n", nt); std::abort(); }
    if (smem > 232448) { std::fprintf(stderr, "emu: dynamic smem %zu > 227 KB\n", smem); std::abort(); }
    while ((int)stacks.size() < nt) stacks.push_back((unsigned char *)std::malloc(STACK));
    ctxs.resize(nt);
    done.assign(nt, 0);
    g_smem.assign(smem + 16, 0);
    S.dyn_smem = (unsigned char *)(((uintptr_t)g_smem.data() + 15) & ~(uintptr_t)15);
    S.bDim = block;
    S.gDim = grid;
    S.nthreads = nt;
    int nw = (nt + 31) / 32;
    g_body = &body;
    for (unsigned bz = 0; bz < grid.z; ++bz)
        for (unsigned by = 0; by < grid.y; ++by)
            for (unsigned bx = 0; bx < grid.x; ++bx) {
                S.bIdx = dim3(bx, by, bz);
                S.block_gen = S.block_arrived = 0;
                S.alive = nt;
                S.warp_gen.assign(nw, 0);
                S.warp_arrived.assign(nw, 0);
                S.warp_slots.assign((size_t)nw * 32 * 16, 0);
                std::fill(done.begin(), done.end(), 0);
                for (int t = 0; t < nt; ++t) {
                    getcontext(&ctxs[t]);
                    ctxs[t].uc_stack.ss_sp = stacks[t];
                    ctxs[t].uc_stack.ss_size = STACK;
                    ctxs[t].uc_link = &S.main_ctx;
                    makecontext(&ctxs[t], trampoline, 0);
                }
                long spins = 0;
                while (S.alive > 0) {
                    for (int t = 0; t < nt; ++t) {
                        if (done[t]) continue;
                        set_tid(t);
                        swapcontext(&S.main_ctx, &ctxs[t]);
                    }
                    if (++spins > 100000000L) { std::fprintf(stderr, "emu: deadlock\n"); std::abort(); }
                }
            }
}
}  // namespace emu
