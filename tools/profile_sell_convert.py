import sys
sys.path.insert(0, '/root/repo')
import torch
import spmv_research_b200 as S
from spmv_research_b200 import amg, formats
p = amg.MatrixParams.parse("888859 888859 50 16.6667 normal random 0.3 10000 1.9 0.05 14")
rp, d_rp, d_col, d_val = amg.generate_device(p)
for _ in range(2):
    MF = formats.csr_to_format_device(d_rp, d_col, d_val, p.rows, p.cols, int(rp[-1]), format=S.SELL_C_SIGMA)
    print("convert_ms", MF.convert_ms)
    MF.close()
