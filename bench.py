#!/usr/bin/env python
"""bench.py -- headline measurement of the B200 SpMV engine (contract: see DESIGN.md "Measurement").

  python bench.py [--gpus 1] [--steps K] [--warmup W]            N=1: stratified sample of the medium dataset
  torchrun --nproc-per-node N bench.py --gpus N ...              N>1: row-sharded power iteration (config 5 series)
  python bench.py --impl reference ...                           the reference's CPU CSR kernel on the host cores

One JSON line on stdout (rank 0).  A "step" is one pass of the hot path over the workload:
one SpMV of every sample matrix (N=1) or one power iteration of the sharded matrix (N>1).
"""
import argparse
import json
import os

# torchrun sets OMP_NUM_THREADS=1; the host-side generator (OpenMP) should use this rank's share of the cores.
# The reference arm (--impl reference) runs on rank 0 alone and gets ALL host cores: both arms then do the same work.
import sys
if int(os.environ.get("WORLD_SIZE", "1")) > 1 and os.environ.get("B200_KEEP_OMP") != "1":
    _ref_arm = any(a == "reference" or a == "--impl=reference" for a in sys.argv[1:])
    _share = 1 if _ref_arm else int(os.environ.get("LOCAL_WORLD_SIZE", os.environ["WORLD_SIZE"]))
    os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // _share))

import subprocess
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "SpMV GFLOP/s (2*nnz/t), fp64; effective HBM GB/s over compulsory bytes"
NOMINAL_HBM_GBS = 8000.0


def ncu_traffic(kernel, sub=None):
    """DRAM bytes of one launch of `kernel` from the committed ncu --set full captures (profiles/ncu_traffic.json); `sub` selects
    a variant of an entry (the config-5 shard with 1 / 2 column panels)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        k = t[kernel]
        matrix, bc = k.get("matrix", t["matrix"]), k.get("compulsory_bytes", t["compulsory_bytes"])
        if sub is not None:
            k = k[sub]
        tot = int(k["dram_bytes_read"]) + int(k["dram_bytes_write"])
        return tot, "one product on '%s' (compulsory bytes: %d, ratio %.3f); %s" % (matrix, bc, tot / bc, t["source"])
    except Exception:
        return None, None


def measured_peak():
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(pk["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the GPU runs the benchmark's steps.  The timed region of a
    default run lasts tens of milliseconds, shorter than one nvidia-smi period, so the caller keeps the SAME steps
    running for a further `soak` seconds straight after the timed region (untimed) and the samples taken between
    mark_start() and the end of that soak are summarised; `window` in the summary says so."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0, soak=0.6):
        self.rows, self.proc, self.index, self.soak_s = [], None, index, soak
        self.t0 = self.t1 = None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in ln.split(",")]))

    def mark_start(self):
        self.t0 = time.time()

    def soak(self, step_fn, sync_fn):
        """Keeps the load of the timed region on the GPU for soak_s more seconds (not timed), then closes the window."""
        end = time.time() + self.soak_s
        while time.time() < end:
            step_fn()
            sync_fn()
        self.t1 = time.time()

    def __exit__(self, *a):
        if self.t1 is None:
            self.t1 = time.time()
        if self.proc:
            time.sleep(0.1)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        t0 = (self.t0 or 0.0) + 0.06            # a sample reports the state of the preceding period
        for ts, r in self.rows:
            if ts < t0 or ts > (self.t1 or ts) + 0.06:
                continue
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                pw.append(float(r[2]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)), "reasons": sorted(reasons),
                "samples": len(sm), "window": "timed region + %.1f s of the same steps run straight after it (untimed)" % self.soak_s}


def cpu_csr_runner(rp, col, val, n):
    """The timed CPU arm: the reference's own OpenMP CSR plugin when oracle/_ref travelled with the
    tree (kind "reference"), else the oracle port of the same kernel (kind "port")."""
    from oracle import oracle as O
    threads = O.max_threads()
    if O.ref_available():
        try:
            R = O.RefCSR(rp, col, val, n)
            return (lambda x, y: R.spmv(x, y)), "reference", threads
        except OSError:
            pass
    bounds = np.full(2 * threads, -1, np.int64)
    return (lambda x, y: O.csr_spmv_omp(rp, col, val, x, y, bounds, threads)), "port", threads


# C code under oracle/_ref printf()s to stdout ("balance time = ..."); keep the real stdout for the one JSON line.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(d):
    os.write(_REAL_STDOUT, (json.dumps(d) + "\n").encode())


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "medium_sample", "config1", "g1", "g2", "power"])
    ap.add_argument("--format", default=None, choices=["auto", "csr", "sell"],
                    help="default: auto at N=1 (both formats built, the faster kept per matrix / per column panel), csr at N>1")
    ap.add_argument("--rows-per-gpu", type=int, default=4194304)
    ap.add_argument("--exchange", default="p2p", choices=["nccl", "p2p", "mc"],   # --power-impl python only (mc implies it)
                    help="N>1: p2p = fused unicast peer stores (default; measured fastest at N=8, profiles/r01g_bench_n8*.json), "
                         "mc = fused stores to an NVSwitch multicast address (falls back to p2p), nccl = all-gather")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sell", action="store_true")
    ap.add_argument("--full-exchange", action="store_true", help="N>1: send every y row to the peers, not only the referenced column band")
    ap.add_argument("--no-nccl-leg", action="store_true", help="N>1: skip the plain-NCCL comparison leg")
    ap.add_argument("--no-graph", action="store_true", help="N>1: time an eager step loop instead of CUDA-graph replays")
    ap.add_argument("--x", default="ones", choices=["ones", "random"])
    ap.add_argument("--host-pipeline", action="store_true",
                    help="N=1 e2e leg: opts.host_pipeline = 1 (x uploaded range by range so H2D overlaps D2H; off by default until verified on a B200)")
    ap.add_argument("--device-generator", action="store_true",
                    help="power workload: generate each rank's non-zeros on the GPU (b200spmv_amg_fill_device; bit-identical, "
                         "not yet verified on a B200 -> off by default)")
    ap.add_argument("--line", default=None, help="one generator parameter line to use as the N=1 workload (profiling)")
    ap.add_argument("--power-impl", default="c", choices=["c", "python"],
                    help="power workload: c = b200spmv_power_* (C host loop, default); python = round 1's loop over torch.distributed")
    ap.add_argument("--variant", default="auto", help="power workload, N>1: auto = verify + trial every entry of POWER_VARIANTS and run the fastest, or one name")
    ap.add_argument("--host-generator", action="store_true", help="power workload: generate the shard on the host (default: pass 2 on the GPU)")
    ap.add_argument("--no-power-parity", action="store_true", help="power workload: skip the per-step oracle check on the small member")
    ap.add_argument("--no-n1-legs", action="store_true", help="power workload, N>1: skip the two single-GPU scaling references")
    return ap.parse_args()


def workload_params(args, world):
    from spmv_research_b200 import harness as H
    wl = args.workload
    if args.line:
        from spmv_research_b200 import amg
        return "line", [amg.MatrixParams.parse(args.line)]
    if wl == "auto":
        wl = "power" if world > 1 else "medium_sample"
    if wl == "medium_sample":
        return wl, H.medium_sample()
    if wl == "config1":
        return wl, [H.CONFIG1]
    if wl == "g1":
        return wl, [H.CONFIG4_G1]
    if wl == "g2":
        return wl, [H.CONFIG4_G2]
    return wl, [H.config5(world, args.rows_per_gpu)]


L2_NOTE = "inputs larger than the 126 MB L2 (every matrix / shard >= 320 MB, a step streams GBs); no explicit flush"


def single_config(wl, plist, x):
    """`config` of the N = 1 line: the same dict on the b200 arm and on the reference arm."""
    return {"workload": wl, "matrices": len(plist), "lines": [p.line() for p in plist], "x": x, "l2": L2_NOTE}


def power_config(p, rows_per_gpu, world):
    """`config` of the N > 1 line (row-sharded power iteration): the same dict on both arms."""
    return {"workload": "power_iteration config5-series", "line": p.line(), "rows_per_gpu": int(rows_per_gpu), "shards": int(world),
            "x": "power iteration from x = 1", "l2": L2_NOTE}


# ------------------------------------------------------------------------------------------------
# N = 1: one SpMV of every matrix of the workload per step, everything resident in HBM
# ------------------------------------------------------------------------------------------------
def run_single(args):
    import torch
    import spmv_research_b200 as S
    from spmv_research_b200 import amg, capi, harness as H
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(0)
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream().cuda_stream
    wl, plist = workload_params(args, 1)
    FM = {"auto": S.AUTO, "csr": S.CSR_ADAPTIVE, "sell": S.SELL_C_SIGMA}
    fmt = FM[args.format or "auto"]
    others = [] if args.no_sell else [f for f in (S.CSR_ADAPTIVE, S.SELL_C_SIGMA) if f != fmt]
    peak, peak_src = measured_peak()

    mats, t_gen = [], 0.0
    cpu = {"t": 0.0, "flops": 0.0, "kind": None, "cores": 0, "iters": 3, "max_err_over_bound": 0.0}
    for p in plist:
        t0 = time.time()
        rp, col, val = amg.generate(p)
        t_gen += time.time() - t0
        m, n, nnz = p.rows, p.cols, int(rp[-1])
        t0 = time.time()
        MF = S.csr_to_format(rp, col, val, m, n, nnz, 0, 0, format=fmt, host_pipeline=(1 if args.host_pipeline else 0))
        t_create = time.time() - t0
        MFo = [S.csr_to_format(rp, col, val, m, n, nnz, 0, 0, format=f) for f in others]
        x_h = np.ones(n) if args.x == "ones" else np.random.default_rng(14).uniform(-1, 1, n)
        dx = torch.from_numpy(x_h).to(dev)
        dy = torch.full((m,), 1.0, dtype=torch.float64, device=dev)
        rec = dict(p=p, m=m, n=n, nnz=nnz, MF=MF, MFo=MFo, dx=dx, dy=dy, x_h=x_h, bc=H.compulsory_bytes(m, n, nnz), create_s=t_create)
        # CPU baseline leg (rank 0, bounded: 1 warm-up + 3 timed calls per matrix) and a parity spot check
        if not args.no_cpu_baseline:
            from oracle import oracle as O
            run, kind, cores = cpu_csr_runner(rp, col, val, n)
            y_cpu = np.full(m, 1.0)
            run(x_h, y_cpu)
            t0 = time.time()
            for _ in range(cpu["iters"]):
                run(x_h, y_cpu)
            cpu["t"] += time.time() - t0
            cpu["flops"] += 2.0 * nnz * cpu["iters"]
            cpu["kind"], cpu["cores"] = kind, cores
            # Parity leg: x ~ U[-1, 1) (seed 14) whatever the timed x is.  With x = 1 and the generator's dyadic 31-bit values
            # every row sum is exact in any order and every gathered x equal, so that check could not fail.
            x_par = np.random.default_rng(14).uniform(-1.0, 1.0, n)
            y_ser = O.csr_spmv(rp, col, val, x_par)
            bound = O.row_abs_bound(rp, col, val, x_par)
            for F in [MF] + MFo:
                y_gpu = np.full(m, 1.0)
                F.spmv(x_par, y_gpu)
                ratio = float(np.max(np.abs(y_gpu - y_ser) / (1e-12 * np.maximum(bound, 1e-300))))
                cpu["max_err_over_bound"] = max(cpu["max_err_over_bound"], ratio)
                if not ratio <= 1.0:
                    raise SystemExit("bench.py: parity check failed on %s, %s (err / (1e-12 bound) = %.3g)" % (p.line(), F.format_name, ratio))
            del x_par, y_ser, bound
        mats.append(rec)
        del rp, col, val

    def one_step(which=-1, events=None):
        for i, r in enumerate(mats):
            if events is not None:
                events[i][0].record()
            (r["MF"] if which < 0 else r["MFo"][which]).spmv_device(r["dx"], r["dy"], stream)
            if events is not None:
                events[i][1].record()

    K, W = args.steps, max(args.warmup, 3)
    for _ in range(W):
        one_step()
    torch.cuda.synchronize()
    ev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in mats] for _ in range(K)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(0) as clk:
        torch.cuda.synchronize()
        clk.mark_start()
        e0.record()
        for k in range(K):
            one_step(events=ev[k])
        e1.record()
        torch.cuda.synchronize()
        clk.soak(one_step, torch.cuda.synchronize)
    ms_total = e0.elapsed_time(e1)
    ms_step = ms_total / K
    tot_nnz = sum(r["nnz"] for r in mats)
    tot_bc = sum(r["bc"] for r in mats)
    gflops = 2.0 * tot_nnz / (ms_step * 1e-3) * 1e-9
    eff = tot_bc / (ms_step * 1e-3) * 1e-9
    per = []
    kern_ms = 0.0
    for i, r in enumerate(mats):
        t = float(np.mean([ev[k][i][0].elapsed_time(ev[k][i][1]) for k in range(K)]))
        kern_ms += t
        st = r["MF"].handle.stats()
        per.append({"matrix": r["p"].line(), "format": r["MF"].format_name, "trial_ms": r["MF"].trial_ms, "nnz": r["nnz"], "ms": round(t, 5), "gflops": round(2.0 * r["nnz"] / t * 1e-6, 1),
                    "eff_GBs": round(r["bc"] / t * 1e-6, 1), "frac_of_measured": round(r["bc"] / t * 1e-6 / peak, 4),
                    "launches": r["MF"].gpu_launches_per_spmv, "split_rows": int(st["nsplit"]),
                    "convert_ms": round(float(st["convert_ms"]), 3), "csr_to_format_s": round(r["create_s"], 3)})
    launches = K * sum(r["MF"].gpu_launches_per_spmv for r in mats)

    # the other format, same protocol, reported beside (not the headline)
    other = []
    for oi in range(len(others)):
        for _ in range(W):
            one_step(oi)
        torch.cuda.synchronize()
        e0.record()
        for k in range(K):
            one_step(oi, events=ev[k])
        e1.record()
        torch.cuda.synchronize()
        ms2 = e0.elapsed_time(e1) / K
        other.append({"format": mats[0]["MFo"][oi].format_name, "ms_per_step": round(ms2, 5), "gflops": round(2.0 * tot_nnz / ms2 * 1e-6, 1),
                      "eff_GBs": round(tot_bc / ms2 * 1e-6, 1),
                      "per_matrix_ms": [round(float(np.mean([ev[k][i][0].elapsed_time(ev[k][i][1]) for k in range(K)])), 5) for i in range(len(mats))]})

    # end to end through the reference-facing call with host buffers: H2D x + kernel + D2H y + sync per call
    hx, hy = [], []
    for r in mats:
        tx = torch.from_numpy(r["x_h"]).pin_memory()
        ty = torch.full((r["m"] + 64,), 1.0, dtype=torch.float64).pin_memory()
        hx.append(tx.numpy())
        hy.append(ty.numpy())
        r["_pin"] = (tx, ty)
    for r, x, y in zip(mats, hx, hy):
        r["MF"].spmv(x, y)
    Ke = max(3, min(K, 10))
    t0 = time.perf_counter()
    for _ in range(Ke):
        for r, x, y in zip(mats, hx, hy):
            r["MF"].spmv(x, y)
    te = (time.perf_counter() - t0) / Ke
    e2e = {"value": round(2.0 * tot_nnz / te * 1e-9, 2), "unit": "GFLOP/s", "h2d_bytes_per_step": int(sum(8 * (r["MF"].handle.col_range()[1] - r["MF"].handle.col_range()[0]) for r in mats)),
           "h2d_note": "x is uploaded for the referenced column band [col_lo, col_hi) only (8 n = %d bytes would be the full vectors)" % int(sum(8 * r["n"] for r in mats)),
           "d2h_bytes_per_step": int(sum(8 * r["m"] for r in mats)), "ms_per_step": round(te * 1e3, 4),
           "call": "Matrix_Format::spmv(x, y) via b200spmv_spmv_host, pinned host x / y" + (", host_pipeline = 1" if args.host_pipeline else "")}

    n_csr = sum(1 for r in mats if r["MF"].format == S.CSR_ADAPTIVE)
    traffic, traffic_note = ncu_traffic("csr_tile_kernel" if 2 * n_csr >= len(mats) else "sell_kernel")
    out = {
        "metric": METRIC, "value": round(gflops, 2), "unit": "GFLOP/s", "n_gpus": 1, "steps": K, "warmup": W,
        "ms_per_step": round(ms_step, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (reference artificial-matrix-generator restated bit-exactly, seed 14), x = %s" % args.x,
        "config": single_config(wl, plist, args.x),
        "config_detail": {"format": {S.AUTO: "B200_AUTO (per matrix: CSR-adaptive or SELL-C-sigma, picked by a device-timed trial inside csr_to_format)",
                                     S.CSR_ADAPTIVE: "B200_CSR_ADAPTIVE", S.SELL_C_SIGMA: "B200_SELL_C_SIGMA"}[fmt], "total_nnz": int(tot_nnz),
                          "compulsory_bytes_per_step": int(tot_bc), "generate_s": round(t_gen, 1),
                          "csr_to_format_s_total": round(sum(r["create_s"] for r in mats), 2),
                          "csr_to_format_note": "format construction (upload + plan / conversion; AUTO adds two SELL conversions and 3 x 5 timed "
                                                "trial products) is outside GFLOP/s, as in the reference harness (bench.cpp:600-633); per matrix in per_matrix[]"},
        "effective_GBs": round(eff, 1), "roofline_frac_nominal_8TBs": round(eff / NOMINAL_HBM_GBS, 4),
        "roofline": {"bound": "hbm", "kernel": {S.AUTO: "csr_tile_kernel / sell_kernel (per matrix, see per_matrix[].format)",
                                                S.CSR_ADAPTIVE: "csr_tile_kernel", S.SELL_C_SIGMA: "sell_kernel"}[fmt],
                     "achieved": round(tot_bc / kern_ms * 1e-6, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(tot_bc / kern_ms * 1e-6 / peak, 4), "traffic": traffic, "traffic_note": traffic_note, "peak_source": peak_src,
                     "achieved_def": "compulsory bytes B_c = 12 nnz + 12 m + 8 n + 4 per SpMV / CUDA-event time of that SpMV, summed over the sample"},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clk.summary(), "per_matrix": per, "other_formats": other,
    }
    if not args.no_cpu_baseline and cpu["t"] > 0:
        out["cpu_baseline"] = {"value": round(cpu["flops"] / cpu["t"] * 1e-9, 3), "unit": "GFLOP/s", "cores": cpu["cores"], "kind": cpu["kind"],
                               "sample": "every matrix of the workload, 1 warm-up + %d timed calls each of the OpenMP CSR kernel" % cpu["iters"],
                               "parity_max_err_over_1e-12_bound": cpu["max_err_over_bound"],
                               "parity_note": "x ~ U[-1,1) seed 14, every format built (AUTO pick, CSR, SELL) vs the serial oracle, must be <= 1"}
    emit(out)


# ------------------------------------------------------------------------------------------------
# N >= 1 under torchrun: row-sharded power iteration, one process per GPU
# ------------------------------------------------------------------------------------------------
def run_power_python(args):
    """Round 1's implementation of the sharded power iteration: the loop in Python over torch.distributed (power.ShardedPowerIteration).
    Kept for --power-impl python (cross-check of the C layer, and the NVSwitch-multicast experiment --exchange mc)."""
    import torch
    import torch.distributed as dist
    import spmv_research_b200 as S
    from spmv_research_b200 import amg, capi, harness as H, power
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.current_stream().cuda_stream
    # CSR for the sharded run: the fused exchange stores y rows to the peers from the epilogue, and only the CSR kernel
    # writes rows in order (coalesced NVLink stores); SELL-C-sigma scatters them through row_order -- measured at N=8:
    # 2.61 ms / iteration with AUTO -> SELL against 1.10 ms with CSR (profiles/r01s_bench_n8_auto_sell_panels.json)
    fmt = {"auto": S.AUTO, "csr": S.CSR_ADAPTIVE, "sell": S.SELL_C_SIGMA}[args.format or "csr"]
    peak, peak_src = measured_peak()
    p = H.config5(world, args.rows_per_gpu)

    # every rank runs pass 1 (row lengths; cheap, deterministic), then generates only its own rows
    t0 = time.time()
    rp, bw = amg.row_ptr(p)
    m, n, nnz = p.rows, p.cols, int(rp[-1])
    capi.check(capi.lib().b200spmv_set_device(local))
    bounds = power.shard_bounds(rp, m, nnz, world, capi.partition)        # reference partitioner, on the GPU
    i_s, i_e = int(bounds[rank, 0]), int(bounds[rank, 1])
    rp_loc = (rp[i_s:i_e + 1] - rp[i_s]).astype(np.int32)
    nnz_loc = int(rp_loc[-1])
    if args.device_generator:
        # pass 2 on the GPU: the shard's non-zeros never exist on the host
        d_rp = torch.from_numpy(rp).to(dev)
        col = torch.empty(nnz_loc, dtype=torch.int32, device=dev)
        val = torch.empty(nnz_loc, dtype=torch.float64, device=dev)
        amg.fill_device(p, d_rp, bw, i_s, i_e, col, val, stream)
        del d_rp
    else:
        col, val = amg.fill(p, rp, bw, i_s, i_e)
    t_gen = time.time() - t0
    # Column panels are off under the fused exchange: the rows become final (and can be stored to the peers) only in the
    # LAST panel's launch, which squeezes the NVLink traffic into half of the iteration -- measured at N=8: 1.28 ms / iteration
    # with 2 panels against 1.10 ms without, although the kernels alone are 10 % faster (profiles/r01s_bench_n8_csr_panels.json;
    # the two runs were on different boxes with different power caps, see DESIGN.md 7 -- to be re-measured on one box).
    fused = world > 1 and args.exchange in ("p2p", "mc")
    if args.device_generator:
        from spmv_research_b200 import formats
        MF = formats.csr_to_format_device(torch.from_numpy(rp_loc).to(dev), col, val, i_e - i_s, n, nnz_loc, format=fmt, device=local,
                                          col_panels=(1 if fused else 0))
    else:
        MF = S.csr_to_format(rp_loc, col, val, i_e - i_s, n, nnz_loc, 0, 0, format=fmt, device=local, col_panels=(1 if fused else 0))
    del col, val

    exchange = args.exchange if world > 1 else "nccl"
    pv, mc_note = None, None
    if exchange == "mc":
        # fused exchange through NVSwitch multicast: every finished row is stored once, the switch replicates it.
        # Every rank must agree before and after the (collective) symmetric-memory rendezvous, else fall back to p2p.
        ok = torch.tensor([1 if power.multicast_supported(local) else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()):
            try:
                pv = power.MulticastVectors(n, rank, world, dev)
            except Exception as ex:
                mc_note = "multicast setup failed: %s" % str(ex).splitlines()[0][:160]
                pv = None
            ok = torch.tensor([1 if pv is not None else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if not int(ok.item()):
                pv = None
                mc_note = mc_note or "multicast setup failed on another rank"
        else:
            mc_note = "CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED is 0 on some rank"
        if pv is None:
            exchange = "p2p"
    if exchange == "p2p":
        # fused exchange: the SpMV epilogue stores every finished row into all peers' next-x buffers over NVLink
        pv = power.PeerVectors(n, rank, world, dev)

    # rows of the next x that ANY rank reads: the union of the shards' referenced column bands.  Only those rows are sent.
    peer_rows = None
    band = (0, n)
    if world > 1 and not args.full_exchange:
        band, peer_rows = power.exchange_rows(MF.handle.col_range(), bounds, rank)

    def local_spmv(cur, slot, scale, norm2, parity):
        peers = ()
        if exchange == "mc":
            peers = [pv.mc[1 - parity] + 8 * i_s]
        elif pv is not None:
            b = 1 - parity
            peers = [pv.peer[b][q] + 8 * i_s for q in range(world) if q != rank]
        MF.spmv_device(cur, slot, torch.cuda.current_stream().cuda_stream, scale=scale, norm2=norm2, peers=peers,
                       peer_rows=(peer_rows if peers else None))

    it = power.ShardedPowerIteration(n, bounds, rank, world, local_spmv, dev, exchange=exchange,
                                     buffers=(pv.tensors if pv is not None else None))
    it.set_start(np.ones(n))
    K, W = args.steps, max(args.warmup, 3)
    for _ in range(W):
        it.step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # kernel-only time of the local shard (no exchange), for the roofline entry
    torch.cuda.synchronize()
    k0.record()
    for _ in range(5):
        MF.spmv_device(it.cur, it.nxt[i_s:i_e], stream)
    k1.record()
    torch.cuda.synchronize()
    kern_ms = k0.elapsed_time(k1) / 5
    it.set_start(np.ones(n))
    for _ in range(W):
        it.step()
    # two iterations (one per buffer parity) captured into a CUDA graph: K timed iterations = ceil(K/2) replays
    launch_mode = "eager"
    if not args.no_graph:
        try:
            it.capture_pair()
            launch_mode = "cuda_graph(2 iterations per replay)"
        except Exception as ex:                      # capture unsupported for this collective: keep the eager loop, say so
            launch_mode = "eager (graph capture failed: %s)" % str(ex).splitlines()[0][:120]
            torch.cuda.synchronize()
    K_run = K
    with ClockSampler(local) as clk:
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        clk.mark_start()
        e0.record()
        if launch_mode.startswith("cuda_graph"):
            it.replay(K // 2)
            if K % 2:
                it.step()
        else:
            for _ in range(K):
                it.step()
        e1.record()
        torch.cuda.synchronize()
        # keep the same load running for the clock sampler: a fixed number of further iterations (identical on every rank)
        n_soak = int(0.6 / max(e0.elapsed_time(e1) / K * 1e-3, 1e-5))
        if world > 1:
            tn = torch.tensor([n_soak], dtype=torch.int64, device=dev)
            dist.all_reduce(tn, op=dist.ReduceOp.MAX)
            n_soak = int(tn.item())
        n_soak = min(n_soak, 4000)
        if launch_mode.startswith("cuda_graph"):
            it.replay(n_soak // 2 + 1)
        else:
            for _ in range(n_soak):
                it.step()
        torch.cuda.synchronize()
        clk.t1 = time.time()
        if world > 1:
            dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1) * K / K_run], dtype=torch.float64, device=dev)
    kms = torch.tensor([kern_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(kms, op=dist.ReduceOp.MAX)
    ms_step = float(ms.item()) / K
    eig = it.eigenvalue()

    # the plain-NCCL exchange (all-gather of y slots after the kernel), same protocol, reported beside the fused one
    nccl_cmp = None
    if world > 1 and exchange in ("p2p", "mc") and not args.no_nccl_leg:
        def plain_spmv(cur, slot, scale, norm2, parity):
            MF.spmv_device(cur, slot, stream, scale=scale, norm2=norm2)
        it2 = power.ShardedPowerIteration(n, bounds, rank, world, plain_spmv, dev, exchange="nccl")
        it2.set_start(np.ones(n))
        for _ in range(W):
            it2.step()
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(K):
            it2.step()
        e1.record()
        torch.cuda.synchronize()
        ms2 = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
        nccl_cmp = {"ms_per_step": round(float(ms2.item()) / K, 5), "gflops": round(2.0 * nnz / (float(ms2.item()) / K * 1e-3) * 1e-9, 2),
                    "eigenvalue_estimate": it2.eigenvalue(),
                    "exchange": "NCCL in-place all-gather of y slots (one broadcast per owner) + scalar all-reduce"}
        # both legs ran W + K (+ graph warm-up) iterations from the same start on different exchange paths: the eigenvalue
        # estimates converge to the same value; a fused exchange that dropped or misplaced rows would not
        rel = abs(nccl_cmp["eigenvalue_estimate"] - eig) / abs(nccl_cmp["eigenvalue_estimate"])
        nccl_cmp["fused_vs_nccl_eigenvalue_rel_diff"] = rel
        if rank == 0 and not rel < 1e-6:
            raise SystemExit("bench.py: fused exchange disagrees with the NCCL all-gather (rel diff %.3g)" % rel)
        del it2

    # the N = 1 member of the same weak-scaling series (rows_per_gpu rows), timed on rank 0 alone while the others wait:
    # lets the scaling of THIS workload be read off one line (the default N = 1 bench line is the medium-dataset sample)
    n1_same = None
    if world > 1:
        if rank == 0:
            p1 = H.config5(1, args.rows_per_gpu)
            rp1, col1, val1 = amg.generate(p1)
            MF1 = S.csr_to_format(rp1, col1, val1, p1.rows, p1.cols, int(rp1[-1]), 0, 0, format=fmt, device=local)
            it1 = power.ShardedPowerIteration(p1.cols, np.array([[0, p1.rows]], np.int64), 0, 1,
                                              lambda cur, slot, scale, norm2, parity: MF1.spmv_device(cur, slot, stream, scale=scale, norm2=norm2), dev)
            it1.set_start(np.ones(p1.cols))
            for _ in range(W):
                it1.step()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(K):
                it1.step()
            e1.record()
            torch.cuda.synchronize()
            t1 = e0.elapsed_time(e1) / K
            n1_same = {"matrix": p1.line(), "ms_per_step": round(t1, 5), "gflops": round(2.0 * int(rp1[-1]) / t1 * 1e-6, 2)}
            MF1.close()
            del it1, rp1, col1, val1
        dist.barrier()

    # end to end through the host-vector call: this rank's shard product with x from pinned host memory, y back to the host
    tx = torch.ones(n, dtype=torch.float64).pin_memory()
    ty = torch.ones(i_e - i_s + 64, dtype=torch.float64).pin_memory()
    MF.spmv(tx.numpy(), ty.numpy())
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    Ke = 5
    for _ in range(Ke):
        MF.spmv(tx.numpy(), ty.numpy())
    te = torch.tensor([(time.perf_counter() - t0) / Ke], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    te = float(te.item())

    if rank == 0:
        bc = H.compulsory_bytes(m, n, nnz)
        bc_loc = H.compulsory_bytes(i_e - i_s, n, nnz_loc)
        gflops = 2.0 * nnz / (ms_step * 1e-3) * 1e-9
        out = {
            "metric": METRIC, "value": round(gflops, 2), "unit": "GFLOP/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms_step, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (reference artificial-matrix-generator restated bit-exactly, seed 14), power iteration from x = 1",
            "config": power_config(p, args.rows_per_gpu, world),
            "config_detail": {"format": MF.format_name, "col_panels": MF.col_panels,
                       "total_nnz": int(nnz), "shard_rows": [int(b[1] - b[0]) for b in bounds], "exchange": {"mc": "fused: SpMV epilogue stores each y row once to an NVSwitch multicast address (NVLS) mapped over every rank's next-x buffer + NCCL scalar all-reduce",
                                    "p2p": "fused: SpMV epilogue stores y rows into every peer's next-x buffer over NVLink (CUDA IPC) + NCCL scalar all-reduce",
                                    "nccl": "NCCL in-place all-gather of y slots (one broadcast per owner) + scalar all-reduce"}[exchange],
                       "exchange_note": mc_note,
                       "exchange_rows": "rows [%d, %d) of %d (the union of the shards' referenced column bands)" % (band[0], band[1], n),
                       "l2_note": "each shard is %.1f GB (> 126 MB L2); no explicit flush" % (bc_loc * 1e-9), "generate_s": round(t_gen, 1),
                       "generator": "device pass 2" if args.device_generator else "host"},
            "effective_GBs": round(bc / (ms_step * 1e-3) * 1e-9, 1),
            "roofline_frac_nominal_8TBs": round(bc / (ms_step * 1e-3) * 1e-9 / (NOMINAL_HBM_GBS * world), 4),
            "roofline": {"bound": "hbm", "kernel": "csr_tile_kernel" if MF.format == S.CSR_ADAPTIVE else "sell_kernel",
                         "achieved": round(bc_loc / float(kms.item()) * 1e-6, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(bc_loc / float(kms.item()) * 1e-6 / peak, 4), "traffic": None, "peak_source": peak_src,
                         "achieved_def": "rank-0 shard compulsory bytes / slowest rank's kernel-only SpMV time (CUDA events, 5 launches)"},
            "e2e": {"value": round(2.0 * nnz / te * 1e-9, 2), "unit": "GFLOP/s", "h2d_bytes_per_step": int(8 * (MF.handle.col_range()[1] - MF.handle.col_range()[0]) * world),
                    "d2h_bytes_per_step": int(8 * m), "ms_per_step": round(te * 1e3, 4),
                    "call": "per rank Matrix_Format::spmv(x, y) on its row shard, pinned host x / y, max over ranks"},
            "gpu_launches": int(K * (MF.gpu_launches_per_spmv + 1)), "clocks": clk.summary(), "eigenvalue_estimate": eig, "launch": launch_mode, "nccl_exchange": nccl_cmp, "n1_same_series": n1_same,
        }
        emit(out)
    # Teardown.  A captured graph holding NCCL kernels must be released before the communicator, and a communicator
    # teardown that stalls must not keep the GPUs leased: a watchdog ends the process 30 s after the line is out.
    def _bail():
        os._exit(0)
    wd = threading.Timer(30.0, _bail)
    wd.daemon = True
    wd.start()
    torch.cuda.synchronize()
    if getattr(it, "_graph", None) is not None:
        it._graph.reset()
        it._graph = None
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    wd.cancel()


# ------------------------------------------------------------------------------------------------
# N >= 1 under torchrun: row-sharded power iteration through the C-ABI (b200spmv_power_*), one process per GPU
# ------------------------------------------------------------------------------------------------
# What is tried per run (all of them verified per step against the oracle on a small member first, then timed for a few
# iterations; the fastest runs the timed K steps and every trial time is reported):
POWER_VARIANTS = {
    "fused_allreduce":   dict(exchange="fused", scalar_sync=0, panels=1, chunks=0),     # round 1's default, now in C
    "fused_mailbox":     dict(exchange="fused", scalar_sync=1, panels=1, chunks=0),     # no NCCL call inside the iteration
    "fused_mailbox_p2c2": dict(exchange="fused", scalar_sync=1, panels=2, chunks=2),    # 2 column panels interleaved over 2 row chunks
    "fused_mailbox_p2c3": dict(exchange="fused", scalar_sync=1, panels=2, chunks=3),
    "fused_mailbox_p2c4": dict(exchange="fused", scalar_sync=1, panels=2, chunks=4),    # ... over 4 (each chunk re-reads the panel's x)
    "nccl_c4":           dict(exchange="nccl", scalar_sync=0, panels=1, chunks=4),      # NCCL all-gather, chunk c under kernel c + 1
    "nccl_p2c2":         dict(exchange="nccl", scalar_sync=0, panels=2, chunks=2),
}


def _power_shard(p, rp, bw, i_s, i_e, fmt, local, dev, panels_wanted, device_generator, keep_host=False):
    """This rank's rows [i_s, i_e) as format handles, one per column-panel setting in panels_wanted (0 = auto)."""
    import torch
    import spmv_research_b200 as S
    from spmv_research_b200 import amg, formats
    n = p.cols
    rp_loc = (rp[i_s:i_e + 1] - rp[i_s]).astype(np.int32)
    nnz_loc = int(rp_loc[-1])
    host = None
    out = {}
    if device_generator and not keep_host:
        d_rp = torch.from_numpy(rp).to(dev)
        col = torch.empty(max(nnz_loc, 1), dtype=torch.int32, device=dev)
        val = torch.empty(max(nnz_loc, 1), dtype=torch.float64, device=dev)
        amg.fill_device(p, d_rp, bw, i_s, i_e, col, val, torch.cuda.current_stream().cuda_stream)
        d_rp_loc = torch.from_numpy(rp_loc).to(dev)
        for cp in panels_wanted:
            out[cp] = formats.csr_to_format_device(d_rp_loc, col, val, i_e - i_s, n, nnz_loc, format=fmt, device=local, col_panels=cp)
        del d_rp, d_rp_loc, col, val
    else:
        col, val = amg.fill(p, rp, bw, i_s, i_e)
        for cp in panels_wanted:
            out[cp] = S.csr_to_format(rp_loc, col, val, i_e - i_s, n, nnz_loc, 0, 0, format=fmt, device=local, col_panels=cp)
        if keep_host:
            host = (rp_loc, col, val)
    return out, nnz_loc, host


def _power_parity(world, rank, local, dev, fmt, names, rows_per_gpu=262144, steps=3):
    """SURVEY 8e 'parity for iterated runs', on hardware, for every variant about to be timed: on a small member of the series
    copy x_k to the host, apply the oracle's serial CSR kernel (K/csr.cpp:334-350) to this rank's shard, require the GPU's slot
    of x_{k+1} within 1e-12 * sum |a_ij x_j| per row, and the exchanged band rows bit-equal on every rank (what the peers
    received).  Returns {variant: worst err / (1e-12 bound) over ranks and steps}; raises SystemExit on any failure."""
    import hashlib
    import torch
    import torch.distributed as dist
    from spmv_research_b200 import amg, capi, harness as H, power
    from oracle import oracle as O          # checker only
    p = H.config5(world, rows_per_gpu)
    rp, bw = amg.row_ptr(p)
    m, n, nnz = p.rows, p.cols, int(rp[-1])
    bounds = power.shard_bounds(rp, m, nnz, world, capi.partition)
    i_s, i_e = int(bounds[rank, 0]), int(bounds[rank, 1])
    need = sorted({POWER_VARIANTS[v]["panels"] for v in names})
    shards, _, (rp_loc, col, val) = _power_shard(p, rp, bw, i_s, i_e, fmt, local, dev, need, False, keep_host=True)
    x0 = np.random.default_rng(5).uniform(0.5, 1.5, n)
    res = {}
    for name in names:
        v = POWER_VARIANTS[name]
        MF = shards[v["panels"]]
        it = power.CPowerIteration(MF, bounds, n, rank, world, local, exchange=v["exchange"], chunks=v["chunks"], use_graph=False,
                                   scalar_sync=v["scalar_sync"])
        info = it.info()
        it.set_start(x0)
        c_lo, c_hi = MF.handle.col_range()
        worst, digests = 0.0, []
        for k in range(steps):
            xk, s = it.get_x()
            xr = np.zeros(n)
            xr[c_lo:c_hi] = xk[c_lo:c_hi]          # columns this shard never reads may hold stale rows: not part of the product
            y_ref = s * O.csr_spmv(rp_loc, col, val, xr)
            bound = abs(s) * O.row_abs_bound(rp_loc, col, val, xr)
            it.iterate(1)
            xn, _ = it.get_x()
            worst = max(worst, float(np.max(np.abs(xn[i_s:i_e] - y_ref) / (1e-12 * np.maximum(bound, 1e-300)))))
            digests.append(hashlib.sha1(xn[info["band_lo"]:info["band_hi"]].tobytes()).hexdigest())
        it.close()
        ok_bits = True
        if world > 1:
            every = [None] * world
            dist.all_gather_object(every, digests)
            ok_bits = all(e == every[0] for e in every)
            t = torch.tensor([worst], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            worst = float(t.item())
        if not (worst <= 1.0 and ok_bits):
            raise SystemExit("bench.py: multi-GPU per-step parity failed for variant %s: worst err / (1e-12 bound) = %.3g, band rows "
                             "bit-equal on every rank: %s" % (name, worst, ok_bits))
        res[name] = round(worst, 4)
    for MF in shards.values():
        MF.close()
    return {"matrix": p.line(), "steps": steps, "worst_err_over_1e-12_bound": res, "band_rows_bit_equal_on_every_rank": True,
            "check": "x_k copied to the host, oracle serial CSR on the rank's shard, GPU slot of x_{k+1} under the per-row bound"}


def run_power(args):
    import torch
    import torch.distributed as dist
    import spmv_research_b200 as S
    from spmv_research_b200 import amg, capi, formats, harness as H, power
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)      # plumbing: id hand-over, barriers, max-over-ranks of the timings
    stream = torch.cuda.current_stream().cuda_stream
    capi.check(capi.lib().b200spmv_set_device(local))
    # CSR for the sharded run: the fused exchange stores y rows to the peers from the epilogue, and only the CSR kernel writes
    # rows in order (coalesced NVLink stores); SELL-C-sigma scatters them through row_order (2.61 vs 1.10 ms at N=8, r01s)
    fmt = {"auto": S.AUTO, "csr": S.CSR_ADAPTIVE, "sell": S.SELL_C_SIGMA}[args.format or "csr"]
    peak, peak_src = measured_peak()
    p = H.config5(world, args.rows_per_gpu)
    K, W = args.steps, max(args.warmup, 3)
    use_graph = not args.no_graph

    def max_over_ranks(v):
        if world == 1:
            return float(v)
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if world == 1:
        names = ["n1"]
        POWER_VARIANTS["n1"] = dict(exchange="fused", scalar_sync=0, panels=0, chunks=0)        # panels: the auto rule
    elif args.variant == "auto":
        names = list(POWER_VARIANTS)
    else:
        names = [args.variant]
    parity = None
    if not args.no_power_parity:
        parity = _power_parity(world, rank, local, dev, fmt, names)

    # the matrix: every rank runs pass 1 (row lengths; cheap, deterministic), then generates only its own rows -- on its GPU
    t0 = time.time()
    rp, bw = amg.row_ptr(p)
    m, n, nnz = p.rows, p.cols, int(rp[-1])
    bounds = power.shard_bounds(rp, m, nnz, world, capi.partition)        # reference partitioner, on the GPU
    i_s, i_e = int(bounds[rank, 0]), int(bounds[rank, 1])
    need = sorted({POWER_VARIANTS[v]["panels"] for v in names})
    shards, nnz_loc, _ = _power_shard(p, rp, bw, i_s, i_e, fmt, local, dev, need, not args.host_generator)
    torch.cuda.synchronize()
    t_gen = time.time() - t0

    def make(name, graph):
        v = POWER_VARIANTS[name]
        it = power.CPowerIteration(shards[v["panels"]], bounds, n, rank, world, local, exchange=v["exchange"], chunks=v["chunks"],
                                   use_graph=graph, scalar_sync=v["scalar_sync"])
        it.set_start(np.ones(n))
        return it

    # trial: every variant, a few iterations each (eager launches; the graph is for the timed run)
    trial = {}
    for name in names:
        it = make(name, False)
        it.iterate(W)
        barrier()
        trial[name] = round(max_over_ranks(it.iterate(6)) / 6, 5)
        it.close()
    best = min(names, key=lambda k: trial[k])
    vbest = POWER_VARIANTS[best]
    MF = shards[vbest["panels"]]

    # kernel-only time of the local shard (no exchange, no epilogue), for the roofline entry
    xk = torch.ones(n, dtype=torch.float64, device=dev)
    yk = torch.zeros(max(i_e - i_s, 1), dtype=torch.float64, device=dev)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(2):
        MF.spmv_device(xk, yk, stream)
    torch.cuda.synchronize()
    k0.record()
    for _ in range(5):
        MF.spmv_device(xk, yk, stream)
    k1.record()
    torch.cuda.synchronize()
    kern_ms = max_over_ranks(k0.elapsed_time(k1) / 5)
    del xk, yk

    # the timed run: W warm-up iterations (the CUDA graph of two iterations is captured inside them), then exactly K
    it = make(best, use_graph)
    it.iterate(max(W, 4) if use_graph else W)
    with ClockSampler(local) as clk:
        barrier()
        clk.mark_start()
        ms = it.iterate(K)
        torch.cuda.synchronize()
        barrier()
        ms_step = max_over_ranks(ms) / K
        n_soak = min(int(max_over_ranks(0.6 / max(ms_step * 1e-3, 1e-5))), 4000)
        it.iterate(n_soak + (n_soak & 1))
        clk.t1 = time.time()
        barrier()
    eig = it.eigenvalue()
    info = it.info()
    it.close()

    # scaling references, timed on rank 0 alone while the others wait (scaling = t(1 GPU) / t(P)):
    #   n1_same_series: the N = 1 member of the weak-scaling series (rows_per_gpu rows on one GPU)
    #   n1_same_matrix: THIS N-GPU matrix on one GPU with the best single-GPU settings (AUTO format, automatic column panels)
    n1_series = n1_matrix = None
    if world > 1 and not args.no_n1_legs:
        if rank == 0:
            def one_gpu(pp, f):
                rp1, bw1 = amg.row_ptr(pp)
                sh, _, _ = _power_shard(pp, rp1, bw1, 0, pp.rows, f, local, dev, [0], True)
                b1 = np.array([[0, pp.rows]], np.int64)
                it1 = power.CPowerIteration(sh[0], b1, pp.cols, 0, 1, local, use_graph=use_graph)
                it1.set_start(np.ones(pp.cols))
                it1.iterate(max(W, 4))
                t1 = it1.iterate(K) / K
                r = {"matrix": pp.line(), "format": sh[0].format_name, "col_panels": sh[0].col_panels, "ms_per_step": round(t1, 5),
                     "gflops": round(2.0 * int(rp1[-1]) / t1 * 1e-6, 2)}
                it1.close()
                sh[0].close()
                return r
            try:
                n1_series = one_gpu(H.config5(1, args.rows_per_gpu), fmt)
            except Exception as ex:
                n1_series = {"error": str(ex).splitlines()[0][:200]}
            for MFs in shards.values():
                if MFs is not MF:
                    MFs.close()
            torch.cuda.empty_cache()
            try:
                n1_matrix = one_gpu(p, S.AUTO)
            except Exception as ex:
                n1_matrix = {"error": str(ex).splitlines()[0][:200]}
        barrier()

    # end to end through the host-vector call: this rank's shard product with x from pinned host memory, y back to the host
    tx = torch.ones(n, dtype=torch.float64).pin_memory()
    ty = torch.ones(i_e - i_s + 64, dtype=torch.float64).pin_memory()
    MF.spmv(tx.numpy(), ty.numpy())
    barrier()
    t0 = time.perf_counter()
    Ke = 5
    for _ in range(Ke):
        MF.spmv(tx.numpy(), ty.numpy())
    te = max_over_ranks((time.perf_counter() - t0) / Ke)

    if rank == 0:
        bc = H.compulsory_bytes(m, n, nnz)
        bc_loc = H.compulsory_bytes(i_e - i_s, n, nnz_loc)
        gflops = 2.0 * nnz / (ms_step * 1e-3) * 1e-9
        # DRAM traffic of the shard kernel: captured for the rank-0 shard of config 5 at N = 8 (the default series member)
        shard_traffic = shard_traffic_note = None
        if world == 8 and args.rows_per_gpu == 4194304 and MF.format == S.CSR_ADAPTIVE and MF.col_panels in (1, 2):
            shard_traffic, shard_traffic_note = ncu_traffic("c5shard_n8", "panels_%d" % MF.col_panels)
        xdesc = {"fused": "fused: the SpMV epilogue stores each finished y row of the band into every peer's next-x buffer over NVLink (CUDA IPC)",
                 "nccl": "NCCL: y slots all-gathered in place (grouped ncclSend / ncclRecv between every pair, band rows only) per row chunk on a second stream"}[vbest["exchange"]]
        sdesc = "mailbox barrier over peer memory (one-warp kernel, no NCCL call in the iteration)" if vbest["scalar_sync"] else "ncclAllReduce of the norm"
        out = {
            "metric": METRIC, "value": round(gflops, 2), "unit": "GFLOP/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms_step, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (reference artificial-matrix-generator restated bit-exactly, seed 14), power iteration from x = 1",
            "config": power_config(p, args.rows_per_gpu, world),
            "config_detail": {"format": MF.format_name, "col_panels": MF.col_panels, "total_nnz": int(nnz),
                              "shard_rows": [int(b[1] - b[0]) for b in bounds], "variant": best, "exchange": xdesc, "scalar_sync": sdesc,
                              "row_chunks": info["chunks"], "exchange_rows": "rows [%d, %d) of %d (the union of the shards' referenced column bands)"
                              % (info["band_lo"], info["band_hi"], n),
                              "loop": "b200spmv_power_iterate (C host code, csrc/power.cu); NCCL bound inside the library",
                              "launch": "cuda_graph(2 iterations per replay)" if info["graph"] else "eager",
                              "generate_s": round(t_gen, 1), "generator": "host" if args.host_generator else "device pass 2 (b200spmv_amg_fill_device)"},
            "effective_GBs": round(bc / (ms_step * 1e-3) * 1e-9, 1),
            "roofline_frac_nominal_8TBs": round(bc / (ms_step * 1e-3) * 1e-9 / (NOMINAL_HBM_GBS * world), 4),
            "roofline": {"bound": "hbm", "kernel": "csr_tile_kernel" if MF.format == S.CSR_ADAPTIVE else "sell_kernel",
                         "achieved": round(bc_loc / kern_ms * 1e-6, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(bc_loc / kern_ms * 1e-6 / peak, 4), "traffic": shard_traffic, "traffic_note": shard_traffic_note,
                         "peak_source": peak_src, "kernel_only_ms": round(kern_ms, 5),
                         "achieved_def": "rank-0 shard compulsory bytes / slowest rank's kernel-only SpMV time (CUDA events, 5 launches, plain product)"},
            "limiter": {"kernel_only_ms": round(kern_ms, 5), "iteration_ms": round(ms_step, 5),
                        "exchange_and_sync_ms": round(ms_step - kern_ms, 5)},
            "e2e": {"value": round(2.0 * nnz / te * 1e-9, 2), "unit": "GFLOP/s", "h2d_bytes_per_step": int(8 * (MF.handle.col_range()[1] - MF.handle.col_range()[0]) * world),
                    "d2h_bytes_per_step": int(8 * m), "ms_per_step": round(te * 1e3, 4),
                    "call": "per rank Matrix_Format::spmv(x, y) on its row shard, pinned host x / y, max over ranks"},
            "gpu_launches": int(K * info["launches_per_iteration"]), "nccl_calls_per_iteration": info["nccl_calls_per_iteration"],
            "clocks": clk.summary(), "eigenvalue_estimate": eig,
            "variant_trials_ms_per_iteration": trial, "parity_per_step": parity,
            "n1_same_series": n1_series, "n1_same_matrix": n1_matrix,
        }
        if n1_series and "ms_per_step" in n1_series:
            out["scaling_weak_vs_same_series"] = round(world * n1_series["ms_per_step"] / ms_step, 3)
        if n1_matrix and "ms_per_step" in n1_matrix:
            out["scaling_strong_vs_same_matrix"] = round(n1_matrix["ms_per_step"] / ms_step, 3)
        emit(out)
    for MFs in shards.values():
        MFs.close()
    # a communicator teardown that stalls must not keep the GPUs leased: a watchdog ends the process 30 s after the line is out
    def _bail():
        os._exit(0)
    wd = threading.Timer(30.0, _bail)
    wd.daemon = True
    wd.start()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    wd.cancel()


# ------------------------------------------------------------------------------------------------
# reference arm: the reference's CPU CSR kernel on the host cores, same workload definition
# ------------------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    from spmv_research_b200 import amg, harness as H
    wl, plist = workload_params(args, world)
    K, W = args.steps, args.warmup
    from oracle import oracle as O

    def time_items(items_iter):
        """One matrix (or row shard) at a time: the reference's CSR plugin keeps its per-thread row ranges in file-scope arrays
        (K/csr.cpp thread_i_s / thread_i_e), so a second csr_to_format() re-partitions EVERY live object; timing several
        objects in turn would run all of them on the last matrix's row ranges (an incomplete, far too fast product)."""
        cores, kind, t, nnzs, ok = 0, "port", 0.0, [], True
        for rp, col, val, n in items_iter:
            run, kind, cores = cpu_csr_runner(rp, col, val, n)
            x, y = np.ones(n), np.full(len(rp) - 1, 1.0)
            for _ in range(max(W, 1)):
                run(x, y)
            t0 = time.perf_counter()
            for _ in range(K):
                run(x, y)
            t += (time.perf_counter() - t0) / K
            nnzs.append(int(rp[-1]))
            xr = np.random.default_rng(14).uniform(-1.0, 1.0, n)          # verification with a random x (the timed x is ones)
            run(xr, y)
            y_ser = O.csr_spmv(rp, col, val, xr)
            ok = ok and bool(np.all(np.abs(y - y_ser) <= 1e-12 * O.row_abs_bound(rp, col, val, xr)))
            del run
        return cores, kind, t, nnzs, ok

    if wl == "power":
        # The WHOLE N-GPU matrix on ALL host cores (the b200 arm's value is the aggregate over all shards): the row shards of
        # the reference partitioner are generated and timed one after the other, their times summed -- one step = one product
        # of the whole matrix.  Bounded: shards are taken in order until `budget_s` of wall time is spent (at least one, all of
        # them on a box with >= 32 cores); value = 2 * (nnz of the shards timed) / (their summed time).
        p = plist[0]
        cfg = power_config(p, args.rows_per_gpu, world)
        rp, bw = amg.row_ptr(p)
        b = O.partition(world, rp, p.rows, int(rp[-1]))
        budget_s, t_begin, done = float(os.environ.get("B200_REF_BUDGET_S", "240")), time.time(), []

        def shards():
            for q in range(world):
                if q > 0 and time.time() - t_begin > budget_s:
                    return
                i_s, i_e = int(b[q, 0]), int(b[q, 1])
                col, val = amg.fill(p, rp, bw, i_s, i_e)
                done.append(q)
                yield (rp[i_s:i_e + 1] - rp[i_s]).astype(np.int32), col, val, p.cols
        cores, kind, t, nnzs, verified = time_items(shards())
        t_scale = float(rp[-1]) / max(1.0, float(sum(nnzs)))          # 1.0 when every shard was timed
        sample = "row shards %s of %d (reference partitioner, nnz-balanced) of %s, each timed on all %d host cores in turn, times summed; " \
                 "one product per step" % (done, world, p.line(), cores)
    else:
        cfg = single_config(wl, plist, "ones")
        t_scale = 1.0

        def mats():
            for p in plist:
                rp, col, val = amg.generate(p)
                yield rp, col, val, p.cols
        cores, kind, t, nnzs, verified = time_items(mats())
        sample = "every matrix of the %s workload, one SpMV each per step (timed matrix by matrix, K calls each)" % wl
    if not verified:
        raise SystemExit("bench.py --impl reference: the CPU kernel's y does not match the serial oracle")
    flops = sum(2.0 * z for z in nnzs)
    val_ = round(flops / t * 1e-9, 3)
    emit({"impl": "reference", "metric": METRIC, "value": val_, "unit": "GFLOP/s", "n_gpus": world, "steps": K, "warmup": W,
          "ms_per_step": round(t * t_scale * 1e3, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
          "data": "synthetic (same generator and parameter lines as the b200 arm)",
          "config": cfg, "config_detail": {"format": "Custom_CSR_B (reference OpenMP CSR, K/csr.cpp)", "total_nnz_timed": int(sum(nnzs))},
          "cpu_baseline": {"value": val_, "unit": "GFLOP/s", "cores": cores, "kind": kind, "sample": sample,
                           "verified_against_serial_oracle": verified},
          "e2e": {"value": val_, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 or args.workload == "power":
        return run_power_python(args) if (args.power_impl == "python" or args.exchange == "mc") else run_power(args)
    return run_single(args)


if __name__ == "__main__":
    main()
