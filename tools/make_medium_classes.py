"""Writes spmv_research_b200/params/medium_classes.txt: 12 parameter lines per CSR-footprint class (<32 MB, 32-128,
128-512, >=512 MB) of the reference's synthetic_matrices_medium_dataset.txt, drawn with
numpy default_rng(14) (a fixed stride would resonate with the period of the parameter grid).  Authoring container only (reads /root/reference); the chosen lines are data and committed."""
import collections

import numpy as np

REF = "/root/reference/matrix_generation_parameters/synthetic_matrices_medium_dataset.txt"
lines = [l.strip() for l in open(REF) if l.strip()]
cls = collections.OrderedDict([("<32MB", []), ("32-128MB", []), ("128-512MB", []), (">=512MB", [])])
for l in lines:
    t = l.split()
    rows, avg = int(t[0]), float(t[2])
    fp = (12 * rows * avg + 4 * (rows + 1)) / 2 ** 20
    cls["<32MB" if fp < 32 else "32-128MB" if fp < 128 else "128-512MB" if fp < 512 else ">=512MB"].append(l)
out = ["# stratified sample of the reference's synthetic_matrices_medium_dataset.txt by CSR footprint class (SURVEY 8d):",
       "# 12 lines per class drawn with numpy default_rng(14) (tools/make_medium_classes.py)",
       "# rows cols avg std distribution placement bw skew neigh crs seed class"]
rng = np.random.default_rng(14)
for k, v in cls.items():
    out += [v[i] + " " + k for i in sorted(rng.choice(len(v), 12, replace=False))]
open("spmv_research_b200/params/medium_classes.txt", "w").write("\n".join(out) + "\n")
