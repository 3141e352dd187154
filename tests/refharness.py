"""Runs the reference's OWN benchmark harness executable (benchmark_code/BENCH/src/bench.cpp + bench_spmv.cpp,
built unmodified by oracle/Makefile into oracle/_ref/) with a format plugin, and parses what it prints.

  spmv_b200_csr_d.exe / spmv_b200_sell_d.exe : the harness linked with host/b200_plugin.cpp over libb200spmv.so
  spmv_csr_d.exe                              : the harness with the reference's CPU CSR plugin (K/csr.cpp)
  spmv_csr_f.exe / spmv_b200_csr_f.exe        : the `_f` build axis (make.sh:215-216) of the two, float vectors

Environment variables are the ones run.sh exports.  The harness prints the CSV
row on stderr (bench_spmv.cpp:484-487) and timing / accuracy lines on stdout.
"""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def exe_path(name):
    return os.path.join(REF_DIR, name)


def available(name):
    return os.path.exists(exe_path(name))


def write_matrix_market(path, rp, col, val, n):
    """general real coordinate file, 1-based, one entry per line (lib/storage_formats/matrix_market)."""
    m = len(rp) - 1
    with open(path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real general\n")
        f.write("%d %d %d\n" % (m, n, int(rp[-1])))
        for i in range(m):
            for j in range(int(rp[i]), int(rp[i + 1])):
                f.write("%d %d %.17g\n" % (i + 1, int(col[j]) + 1, float(val[j])))


def run(name, args, gpu_kernel, threads=None, extra_env=None, timeout=900):
    """args: [mtx path] or the 11 generator parameters.  Returns a dict with the parsed output."""
    env = dict(os.environ)
    env.update(USE_ARTIFICIAL_MATRICES="0" if len(args) == 1 else "1", USE_PROCESSES="0", GPU_KERNEL="1" if gpu_kernel else "0",
               CLEAR_CACHES="0", USE_RCM_REORDERING="0", COOLDOWN="0", RAPL_REGISTERS="0", PROGG=name, OMP_DYNAMIC="false")
    if threads:
        env["OMP_NUM_THREADS"] = str(threads)
    if extra_env:
        env.update(extra_env)
    p = subprocess.run([exe_path(name)] + [str(a) for a in args], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=timeout)
    out = dict(returncode=p.returncode, stdout=p.stdout, stderr=p.stderr, test_failed=("Test failed" in p.stdout))
    cg = re.findall(r"k = (\d+)\s+error = (\S+)\s+error_explicit = (\S+)\s+error_best = (\S+)", p.stdout)
    if cg:
        out["cg_trace"] = [(int(k), float(e), float(ee), float(eb)) for k, e, ee, eb in cg]
        mf = re.search(r"^error = (\S+)", p.stdout, re.M)
        out["cg_final_error"] = float(mf.group(1)) if mf else None
    mt = re.search(r"time iter: min=([0-9.eE+-]+), median=([0-9.eE+-]+), max=([0-9.eE+-]+)", p.stdout)
    if mt:
        out.update(time_min=float(mt.group(1)), time_median=float(mt.group(2)), time_max=float(mt.group(3)))
    ml = re.search(r"number of loops = (\d+)", p.stdout)
    if ml:
        out["loops"] = int(ml.group(1))
    me = re.search(r"errors spmv: (.*)", p.stdout)
    if me:
        out["errors"] = {k: float(v) for k, v in (kv.split("=") for kv in me.group(1).split(", "))}
    mf = re.search(r"Test failed! \(([0-9.eE+-]+)\)", p.stdout)
    if mf:
        out["max_rel_diff"] = float(mf.group(1))          # bench_spmv.cpp:198-199: printed only above epsilon (1e-10 `_d`, 1e-7 `_f`)
    rows = [ln for ln in p.stderr.splitlines() if ln.count(",") > 10]
    if rows:
        out["csv"] = rows[-1].split(",")
    return out
