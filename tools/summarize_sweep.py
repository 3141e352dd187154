"""Per-class tables of a tools/run_medium_sweep.py JSONL (config 2 / config 3): count, median / p10 / p90 of the fraction of the
measured HBM peak (best of CSR and SELL = what AUTO's device-timed trial would pick), CSR vs SELL win counts, share of
matrices at or above 75 % of the NOMINAL 8 TB/s, parity failures.  python tools/summarize_sweep.py run.jsonl [more.jsonl] > table.txt"""
import collections
import json
import sys

import numpy as np


def main():
    recs, peak = {}, 6535.1
    for path in sys.argv[1:]:
        for ln in open(path):
            try:
                r = json.loads(ln)
            except ValueError:
                continue
            if r.get("summary"):
                peak = r.get("hbm_peak_GBs", peak)
            elif "matrix" in r and "best_ms" in r:
                recs[r["matrix"]] = r
    rs = list(recs.values())
    nominal = 8000.0
    print("# %d matrices; fraction = compulsory bytes / cold device time / %.1f GB/s (measured copy peak); 'nominal' = / 8000 GB/s" % (len(rs), peak))
    print("# best = min(CSR, SELL-C-sigma sigma 16384) per matrix (AUTO additionally tries sigma 256 and keeps the fastest by a device-timed trial)")

    def table(title, keyfn, order=None):
        groups = collections.OrderedDict()
        for r in rs:
            groups.setdefault(keyfn(r), []).append(r)
        keys = order if order else sorted(groups)
        print("\n%s" % title)
        print("%-22s %6s | %7s %7s %7s | %7s %7s | %9s %9s | %10s %8s" % ("group", "count", "median", "p10", "p90", "csr med", "sell med", "csr wins", "sell wins", ">=.75 nom", "GF/s agg"))
        for k in keys:
            g = groups.get(k)
            if not g:
                continue
            fr = np.array([r["best_frac"] for r in g])
            c = [r["formats"]["csr"]["frac_of_measured_peak"] for r in g if "ms_cold" in r["formats"].get("csr", {})]
            s = [r["formats"]["sell"]["frac_of_measured_peak"] for r in g if "ms_cold" in r["formats"].get("sell", {})]
            sw = sum(1 for r in g if r["formats"].get("sell", {}).get("ms_cold", 1e30) < r["formats"].get("csr", {}).get("ms_cold", 1e30))
            nom = float(np.mean(fr * peak / nominal >= 0.75))
            agg = 2.0 * sum(r["nnz"] for r in g) / sum(r["best_ms"] for r in g) * 1e-6
            print("%-22s %6d | %7.3f %7.3f %7.3f | %7.3f %7.3f | %9d %9d | %10.3f %8.0f" % (str(k), len(g), np.median(fr), np.percentile(fr, 10), np.percentile(fr, 90),
                  np.median(c) if c else float("nan"), np.median(s) if s else float("nan"), len(g) - sw, sw, nom, agg))

    table("by CSR footprint class (SURVEY 8d)", lambda r: r["class"], ["<32MB", "32-128MB", "128-512MB", ">=512MB"])
    big = [r for r in rs if r["class"] in ("128-512MB", ">=512MB")]
    rs_all, rs[:] = list(rs), big
    if big:
        table("matrices >= 128 MB, by average non-zeros per row", lambda r: int(round(float(r["matrix"].split()[2]))))
        table("matrices >= 128 MB, by band width x similarity (bw, crs)", lambda r: (float(r["matrix"].split()[6]), float(r["matrix"].split()[9])))
        table("matrices >= 128 MB, by x band in MB (8 * bw * cols): <= 63 fits one L2 half", lambda r: ("<=63MB" if 8 * float(r["matrix"].split()[6]) * int(r["matrix"].split()[1]) / 2 ** 20 <= 63 else ("63-126MB" if 8 * float(r["matrix"].split()[6]) * int(r["matrix"].split()[1]) / 2 ** 20 <= 126 else ">126MB")), ["<=63MB", "63-126MB", ">126MB"])
    rs[:] = rs_all
    fails = sum(1 for r in rs if r.get("csr_vs_sell_ok") is False)
    chk = sum(1 for r in rs for f in r["formats"].values() if "oracle_parity_ok" in f)
    bad = sum(1 for r in rs for f in r["formats"].values() if f.get("oracle_parity_ok") is False)
    print("\nparity: CSR vs SELL on the device under the per-row bound for every matrix: %d failures of %d; oracle (serial CSR, random x): %d checked, %d failures"
          % (fails, len(rs), chk, bad))
    worst = sorted(rs, key=lambda r: r["best_frac"])[:8]
    print("\nslowest (fraction of measured peak):")
    for r in worst:
        print("  %.4f  %s  %s" % (r["best_frac"], r["matrix"], {k: v.get("ms_cold") for k, v in r["formats"].items()}))


if __name__ == "__main__":
    main()
