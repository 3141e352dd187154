"""In-tree build of the native libraries (no JIT cache: the built .so files travel with the tree).

  lib/libb200spmv.so : CUDA kernels + C-ABI, nvcc, sm_100a only
  lib/libb200amg.so  : host-side synthetic matrix source, gcc
  lib/libb200mtx.so  : host-side Matrix Market reader, gcc
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "lib")
CU_SOURCES = ["capi.cu", "csr_kernel.cu", "sell_kernel.cu", "convert.cu", "partition.cu", "pcg.cu", "panels.cu", "amg_device.cu", "coo.cu", "f32.cu", "power.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared", "--cudart", "static", "-ldl"]      # -ldl: power.cu binds NCCL with dlopen (no link-time NCCL)


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_spmv(force=False, verbose=False):
    os.makedirs(LIB, exist_ok=True)
    so = os.path.join(LIB, "libb200spmv.so")
    srcs = [os.path.join(CSRC, f) for f in CU_SOURCES]
    deps = srcs + [os.path.join(CSRC, f) for f in ("common.cuh", "amg_fill.h", "amg_fill_plan.h", "coo_sort.h", "f32_tile.h", "csr_perm.h")] + \
           [os.path.join(HERE, "..", "include", f) for f in ("b200spmv.h", "b200amg.h")]
    if force or _stale(so, deps):
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", so] + srcs
        subprocess.check_call(cmd)
    return so


# experimental builds of the same library (kernel variants behind compile-time switches of csr_kernel.cu), timed against
# the shipped one by tools/sweep_exp.py on the GPU box; never loaded unless B200SPMV_LIB points at one
EXPERIMENTS = {
    "nolds128": ["-DCSR_EXP_LDS128=0"],                            # round 1's shipped row sum (one LDS.64 per product)
    "lpr2": ["-DCSR_EXP_MIN_LPR=2"],
    "lpr4": ["-DCSR_EXP_MIN_LPR=4"],
    "lpr8": ["-DCSR_EXP_MIN_LPR=8"],
    "perm": ["-DCSR_EXP_PERM=1"],                                  # conflict-free row -> thread table (csr_perm.h); y bit-identical
    "perm_nolds128": ["-DCSR_EXP_PERM=1", "-DCSR_EXP_LDS128=0"],
    "rot_perm": ["-DCSR_EXP_ROT=1", "-DCSR_EXP_PERM=1"],           # rotation in the thread-per-row bin + table in the lanes-per-row bins
    "rot": ["-DCSR_EXP_ROT=1"],
    "t128": ["-DCSR_EXP_THREADS=128"],                             # 128-thread CTAs, 1 024-nnz tiles, 12 resident per SM
    "xna": ["-DCSR_EXP_XNA=1"],                                    # x gathers without L1 allocation                                    # per-row start rotation (csr_perm.h, csr_rot_build): no holes, y coalesced
}


def build_experiments(tags=None, force=False):
    out = {}
    exp_dir = os.path.join(LIB, "exp")
    os.makedirs(exp_dir, exist_ok=True)
    srcs = [os.path.join(CSRC, f) for f in CU_SOURCES]
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    for tag in (tags or sorted(EXPERIMENTS)):
        so = os.path.join(exp_dir, "libb200spmv_%s.so" % tag)
        if force or _stale(so, srcs + [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "csr_perm.h")]):
            subprocess.check_call([nvcc] + NVCC_FLAGS + EXPERIMENTS[tag] + ["-o", so] + srcs)
        out[tag] = so
    return out


def build_amg(force=False):
    from . import amg
    return amg.build(force)


def build_mtx(force=False):
    from . import mtx
    return mtx.build(force)


def build_all(force=False, verbose=False):
    return build_spmv(force, verbose), build_amg(force), build_mtx(force)


if __name__ == "__main__":
    import sys
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
