"""BASELINE.json config 2 -- the synthetic_matrices_medium_dataset.txt sweep, fp64 CSR vs SELL-C-sigma on one B200 -- made
feasible by generating on the GPU: pass 1 of the generator (row lengths) runs once per (rows, avg, skew) group on the host
(the 16 190 lines share 360 of them), the corrected band width per line on a host thread pool while the GPU works, pass 2
(the non-zeros) on the GPU straight into HBM (b200spmv_amg_fill_device), then b200spmv_create_device and device-timed
products.  Round 1 measured stratified samples only (the host generator needs ~13 h for the full file).

  python tools/run_medium_sweep.py [--stride K] [--offset O] [--max-seconds S] [--check-every N] [--generator device|host]
                                   > gpurun_out/medium_sweep.jsonl

Per matrix (one JSON line): cold device time (L2 flushed before every product) of the CSR and the SELL-C-sigma handle,
effective GB/s over compulsory bytes, fraction of the measured HBM peak; y(CSR) against y(SELL) on the device under the
per-row bound for EVERY matrix (x = 1 and values >= 0, so sum |a_ij x_j| is y itself), and against the CPU oracle's serial
CSR kernel with random x every --check-every-th matrix (the matrix is copied back to the host for that).  Last line: the
summary per CSR-footprint class of SURVEY 8d.  Measurement tooling: one of the places allowed to load oracle/ (checker).
"""
import argparse
import collections
import concurrent.futures as cf
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from spmv_research_b200 import amg, capi, formats, harness as H  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stride", type=int, default=1, help="take every K-th line of the dataset (systematic sample)")
    ap.add_argument("--offset", type=int, default=0)
    ap.add_argument("--max-seconds", type=float, default=1e9, help="stop after this much wall time (the summary covers what ran)")
    ap.add_argument("--check-every", type=int, default=50, help="oracle parity with random x on every N-th matrix (0 = never)")
    ap.add_argument("--generator", default="device", choices=["device", "host"])
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--dataset", default="medium", choices=["medium", "twins_range"],
                    help="medium: synthetic_matrices_medium_dataset.txt (config 2); twins_range: the 3 637 lines of "
                         "validation_matrices_10_samples_30_range_twins.txt (config 3), harness.validation_twins_range()")
    ap.add_argument("--max-nnz", type=float, default=1.6e9, help="skip lines whose nominal rows x avg exceeds this (int32 index space / HBM)")
    ap.add_argument("--max-skew", type=float, default=0.0,
                    help="skip lines whose skew parameter exceeds this (0 = no limit): one row of skew x avg entries is ONE sequential random "
                         "stream for the device generator (the mawi twin's 8e6-skew variants: 100 s each); the summary counts what was skipped")
    ap.add_argument("--min-skew", type=float, default=0.0, help="only lines whose skew parameter is at least this (re-measuring the giant-row lines)")
    ap.add_argument("--limit", type=int, default=0, help="run at most this many lines (0 = all selected)")
    ap.add_argument("--resume", action="append", default=[],
                    help="JSONL of an earlier (interrupted) run of the same sweep: lines already measured there are not run again and "
                         "count in the summary (the reference's sweep script retries / skips finished work too, benchmark_code/BENCH/run.sh:672-745); repeatable")
    args = ap.parse_args()
    from oracle import oracle as O          # checker only

    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    dataset = H.medium_dataset() if args.dataset == "medium" else [p for p, _ in H.validation_twins_range()]
    lines = [p for i, p in enumerate(dataset) if i % args.stride == args.offset % args.stride and p.rows * p.avg <= args.max_nnz]
    if args.min_skew > 0:
        lines = [p for p in lines if p.skew >= args.min_skew]
    n_skew_skipped = sum(1 for p in lines if args.max_skew > 0 and p.skew > args.max_skew)
    if args.max_skew > 0:
        lines = [p for p in lines if p.skew <= args.max_skew]
    # resume: records of earlier runs (kept for the summary), their lines skipped
    prior = {}
    for path in args.resume:
        if os.path.exists(path):
            for ln in open(path):
                try:
                    r = json.loads(ln)
                except ValueError:
                    continue                     # a line cut off when the earlier run was killed
                if "matrix" in r and "best_ms" in r:
                    prior[r["matrix"]] = r
    if args.limit > 0:
        lines = lines[:args.limit]
    wanted = {p.line() for p in lines}
    resumed = [r for k, r in prior.items() if k in wanted]
    lines = [p for p in lines if p.line() not in prior]
    # pass 1 depends on (rows, cols, avg, std, distribution, skew, seed) only: lines that share those share row_ptr
    groups = collections.OrderedDict()
    for p in lines:
        groups.setdefault((p.rows, p.cols, p.avg, p.std, p.distribution, p.skew, p.seed), []).append(p)
    pool = cf.ThreadPoolExecutor(max_workers=max(1, (os.cpu_count() or 2) - 1))
    flush = 256 << 20
    t_start = time.time()
    recs, done, stopped = list(resumed), 0, False
    t_pass1 = t_fill = 0.0
    for key, plist in groups.items():
        if stopped:
            break
        t0 = time.time()
        rp, _ = amg.row_ptr(plist[0])                       # pass 1: shared by the whole group
        t_pass1 += time.time() - t0
        m, n = plist[0].rows, plist[0].cols
        nnz = int(rp[-1])
        bws = [pool.submit(amg.bw_corrected, p, nnz) for p in plist]      # ctypes releases the GIL
        d_rp = torch.from_numpy(rp).to(dev)
        d_col = torch.empty(nnz, dtype=torch.int32, device=dev)
        d_val = torch.empty(nnz, dtype=torch.float64, device=dev)
        dx = torch.ones(n, dtype=torch.float64, device=dev)
        bc = H.compulsory_bytes(m, n, nnz)
        for p, fut in zip(plist, bws):
            if time.time() - t_start > args.max_seconds:
                stopped = True
                break
            bw = fut.result()
            t0 = time.time()
            if args.generator == "device":
                amg.fill_device(p, d_rp, bw, 0, m, d_col, d_val, st)
            else:
                c, v = amg.fill(p, rp, bw, 0, m)
                d_col.copy_(torch.from_numpy(c))
                d_val.copy_(torch.from_numpy(v))
                torch.cuda.synchronize()
            t_f = time.time() - t0
            t_fill += t_f
            rec = {"matrix": p.line(), "class": H.footprint_class(p), "m": m, "nnz": nnz, "compulsory_bytes": int(bc),
                   "fill_s": round(t_f, 4), "formats": {}}
            ys = {}
            for name, fmt in (("csr", capi.CSR_ADAPTIVE), ("sell", capi.SELL_C_SIGMA)):
                try:
                    MF = formats.csr_to_format_device(d_rp, d_col, d_val, m, n, nnz, format=fmt)
                except capi.B200Error as ex:
                    rec["formats"][name] = {"error": str(ex)[:160]}
                    continue
                dy = torch.full((m,), 1.0, dtype=torch.float64, device=dev)
                MF.handle.time_device(dx, dy, 2, flush)
                ms = float(np.median(MF.handle.time_device(dx, dy, args.iters, flush)))
                ys[name] = dy
                rec["formats"][name] = {"ms_cold": round(ms, 5), "gflops": round(2.0 * nnz / ms * 1e-6, 1),
                                        "eff_GBs": round(bc / ms * 1e-6, 1), "frac_of_measured_peak": round(bc / ms * 1e-6 / peak, 4),
                                        "convert_ms": round(MF.convert_ms, 2)}
                if args.check_every and done % args.check_every == 0:
                    # oracle parity, random x, through the host-vector entry point
                    rp_h, col_h, val_h = rp, d_col.cpu().numpy(), d_val.cpu().numpy()
                    x = np.random.default_rng(14).uniform(-1.0, 1.0, n)
                    y = np.full(m, 1.0)
                    MF.spmv(x, y)
                    y_ref = O.csr_spmv(rp_h, col_h, val_h, x)
                    bound = O.row_abs_bound(rp_h, col_h, val_h, x)
                    rec["formats"][name]["oracle_parity_ok"] = bool(np.all(np.abs(y - y_ref) <= 1e-12 * bound))
                MF.close()
            if len(ys) == 2:
                # x = 1, values in [0, 1): sum_j |a_ij x_j| = y_i, so the per-row bound needs no extra pass
                rec["csr_vs_sell_ok"] = bool(torch.all((ys["csr"] - ys["sell"]).abs() <= 2e-12 * ys["csr"].abs()).item())
            ok = [f for f in rec["formats"].values() if "ms_cold" in f]
            if ok:
                best = min(ok, key=lambda f: f["ms_cold"])
                rec["best_ms"] = best["ms_cold"]
                rec["best_frac"] = best["frac_of_measured_peak"]
            recs.append(rec)
            print(json.dumps(rec), flush=True)
            done += 1
        del d_rp, d_col, d_val, dx
        torch.cuda.empty_cache()
    # summary per footprint class
    summ = {"summary": True, "lines_run": done, "lines_resumed": len(resumed), "lines_selected": len(lines) + len(resumed), "dataset": args.dataset, "dataset_lines": len(dataset), "stride": args.stride,
            "generator": args.generator, "skipped_for_skew": n_skew_skipped, "wall_s": round(time.time() - t_start, 1), "pass1_host_s": round(t_pass1, 1),
            "fill_s": round(t_fill, 1), "hbm_peak_GBs": peak, "stopped_early": stopped, "classes": {}}
    for cls in ("<32MB", "32-128MB", "128-512MB", ">=512MB", "all"):
        rs = [r for r in recs if "best_ms" in r and (cls == "all" or r["class"] == cls)]
        if not rs:
            continue
        fr = np.array([r["best_frac"] for r in rs])
        d = {"matrices": len(rs), "frac_median": round(float(np.median(fr)), 4), "frac_p10": round(float(np.percentile(fr, 10)), 4),
             "frac_p90": round(float(np.percentile(fr, 90)), 4), "share_at_or_above_0.75": round(float(np.mean(fr >= 0.75)), 4),
             "aggregate_gflops": round(2.0 * sum(r["nnz"] for r in rs) / sum(r["best_ms"] for r in rs) * 1e-6, 1),
             "aggregate_eff_GBs": round(sum(r["compulsory_bytes"] for r in rs) / sum(r["best_ms"] for r in rs) * 1e-6, 1)}
        for name in ("csr", "sell"):
            f = [r["formats"][name]["frac_of_measured_peak"] for r in rs if "ms_cold" in r["formats"].get(name, {})]
            if f:
                d[name + "_frac_median"] = round(float(np.median(f)), 4)
        d["sell_faster_share"] = round(float(np.mean([r["formats"].get("sell", {}).get("ms_cold", 1e30) < r["formats"].get("csr", {}).get("ms_cold", 1e30)
                                                      for r in rs])), 4)
        summ["classes"][cls] = d
    summ["csr_vs_sell_failures"] = sum(1 for r in recs if r.get("csr_vs_sell_ok") is False)
    summ["oracle_parity_checked"] = sum(1 for r in recs for f in r["formats"].values() if "oracle_parity_ok" in f)
    summ["oracle_parity_failures"] = sum(1 for r in recs for f in r["formats"].values() if f.get("oracle_parity_ok") is False)
    print(json.dumps(summ), flush=True)


if __name__ == "__main__":
    main()
