"""Picks the stratified sample of matrix_generation_parameters/synthetic_matrices_medium_dataset.txt
(16 190 lines; a full sweep is ~13 h of CPU generation) that bench.py times at N=1.
Run in the authoring container only; the chosen lines are committed in
spmv_research_b200/params/medium_sample.txt (parameter lines are data, not code).

Strata: one matrix per avg-nnz class {5,10,10,20,50,50,100,500} from the 320-512 MB footprint
classes (above the 126 MB L2, where an HBM roofline is meaningful), cycling through the file's
bw / skew / neighbour / similarity grid values so that easy and hard gather patterns both appear.
"""
import os
import sys

REF = "/root/reference/matrix_generation_parameters/synthetic_matrices_medium_dataset.txt"
WANT = [  # rows, bw, skew, neigh, crs
    (5242879, "0.3", "0", "0.5", "0.5"),
    (4329604, "0.05", "0", "0.05", "0.05"),
    (4329604, "0.6", "100", "0.95", "0.5"),
    (2200290, "0.3", "1000", "1.4", "0.95"),
    (888859, "0.6", "0", "0.5", "0.5"),
    (888859, "0.3", "10000", "1.9", "0.05"),
    (445906, "0.05", "0", "0.95", "0.95"),
    (89418, "0.3", "100", "0.5", "0.5"),
]

lines = [l.strip() for l in open(REF)]
index = {}
for l in lines:
    t = l.split()
    index[(int(t[0]), t[6], t[7], t[8], t[9])] = l
out = []
for w in WANT:
    if w not in index:
        sys.exit("not in the reference file: %r" % (w,))
    out.append(index[w])
dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "spmv_research_b200", "params", "medium_sample.txt")
with open(dst, "w") as f:
    f.write("# stratified sample of the reference's synthetic_matrices_medium_dataset.txt (tools/make_medium_sample.py)\n")
    f.write("# rows cols avg std distribution placement bw skew neigh crs seed\n")
    f.write("\n".join(out) + "\n")
print("\n".join(out))
