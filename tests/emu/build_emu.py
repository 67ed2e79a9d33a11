"""Build the kernel sources with g++ against the CUDA execution-model emulator
(tests/emu/cuda_emu.h).  TEST INFRASTRUCTURE ONLY: the resulting
libpfb_b200_emu.so exports the same C-ABI as the product library with
"device" pointers being host pointers, so the unchanged kernels can be checked
against the oracle in the GPU-less build container.  The product package never
loads it."""
import glob
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "pfb-clean_b200", "csrc")
SO = os.path.join(HERE, "libpfb_b200_emu.so")


def build(force=False):
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    deps = srcs + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) + \
        [os.path.join(HERE, "cuda_emu.h"), os.path.join(HERE, "cuda_emu.cpp"),
         os.path.join(ROOT, "include", "pfb_b200.h")]
    if not force and os.path.exists(SO) and all(
            os.path.getmtime(SO) >= os.path.getmtime(d) for d in deps):
        return SO
    objs = []
    bdir = os.path.join(HERE, "build")
    os.makedirs(bdir, exist_ok=True)
    procs = []
    for s in srcs + [os.path.join(HERE, "cuda_emu.cpp")]:
        o = os.path.join(bdir, os.path.basename(s) + ".o")
        objs.append(o)
        if (not force and os.path.exists(o) and
                all(os.path.getmtime(o) >= os.path.getmtime(d) for d in deps)):
            continue
        cmd = ["g++", "-O2", "-g", "-std=c++17", "-fPIC", "-DPFB_EMU", "-Wno-unknown-pragmas",
               "-include", os.path.join(HERE, "cuda_emu.h"), "-x", "c++", "-c", s, "-o", o]
        procs.append((cmd, subprocess.Popen(cmd)))
    for cmd, p in procs:
        if p.wait() != 0:
            raise RuntimeError("emu build failed: " + " ".join(cmd))
    subprocess.check_call(["g++", "-shared", "-o", SO] + objs)
    return SO


if __name__ == "__main__":
    print(build(force=True))
