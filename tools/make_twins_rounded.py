"""Writes spmv_research_b200/params/validation_twins_rounded.txt: for each of the 52 validation twins
(params/validation_twins.txt, benchmark_code/BENCH/config.sh:402-453) the 5-decimal base features the reference's
matrix_generation_parameters/validation_matrices_10_samples_30_range_twins.txt (BASELINE config 3) was expanded from
(create_validation_twins_param_file.py rounds validation_matrices_features.csv, whose last digits differ from config.sh's).
Run in the build container (needs /root/reference); harness.validation_twins_range() regenerates the 3 637 lines from it and
tests/test_amg.py checks them against the reference's file line by line."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from spmv_research_b200 import harness as H  # noqa: E402

REF = "/root/reference/matrix_generation_parameters/validation_matrices_10_samples_30_range_twins.txt"
PAR = os.path.join(ROOT, "spmv_research_b200", "params")


def main():
    src = os.path.join(PAR, "validation_twins.txt")
    twins = H.load_param_file(src)
    names = [ln.split()[-1] for ln in open(src) if ln.strip() and not ln.startswith("#")]
    ref = [ln.split() for ln in open(REF) if ln.strip()]
    S = np.linspace(0.7, 1.3, 10)
    ptr, out = 0, []
    for p, name in zip(twins, names):
        L = ref[ptr]
        assert int(L[0]) == int(p.rows * S[0]), (name, L)
        base = [L[2], L[3], L[6], L[7], L[8], L[9]]
        out.append("%d %d %s %s normal random %s %s %s %s 14 %s" % (p.rows, p.cols, base[0], base[1], base[2], base[3], base[4], base[5], name))
        scaled = {int(p.rows * x) for x in S}
        while ptr < len(ref):                                 # skip this twin's block (<= 70 lines, duplicates were dropped)
            L = ref[ptr]
            f = [L[2], L[3], L[6], L[7], L[8], L[9]]
            if (f == base and int(L[0]) in scaled) or (int(L[0]) == p.rows and sum(a == b for a, b in zip(f, base)) >= 5):
                ptr += 1
            else:
                break
    assert ptr == len(ref) and len(out) == 52
    with open(os.path.join(PAR, "validation_twins_rounded.txt"), "w") as f:
        f.write("# base features (5 decimals) of the 52 validation twins as the reference's\n"
                "# matrix_generation_parameters/validation_matrices_10_samples_30_range_twins.txt uses them; made by\n"
                "# tools/make_twins_rounded.py.  rows cols avg std distribution placement bw skew neigh crs seed name\n")
        f.write("\n".join(out) + "\n")
    print("wrote", len(out), "twins")


if __name__ == "__main__":
    main()
