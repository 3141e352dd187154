"""CPU model of the shared-memory wavefronts of csr_tile_kernel's thread-per-row reads (no GPU needed): for sampled tiles of a
generator matrix, count per warp and per step j the wavefronts of `LDS.64 s_prod[a_l + j]` -- a 64-bit warp load is served
per half-warp, one wavefront per distinct multiplicity level of a bank pair ((a_l + j) mod 16) among its active lanes --
for (a) the shipped order (row = thread) and (b) the CSR_EXP_PERM table (csr_perm.h via tests/emul/libperm_emul.so).
Prints wavefronts per 32 non-zeros; the ncu figure to compare with is 79 500 shared wavefronts per SM on the captured
matrix = 9.5 per 32 non-zeros, 2 of which are the product stores (profiles/r01e_ncu_full_csr_tile_kernel_key_metrics.txt).
  python tools/model_lds.py ["generator line" ...]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
T, EXT = 2048, 256


def wavefronts(starts, lens, phi=None):
    """starts / lens of the rows given to consecutive threads (-1 start = idle lane), optional start rotation of each row
    (CSR_EXP_ROT); returns LDS wavefronts"""
    n = len(starts)
    pad = (-n) % 32
    a = np.r_[starts, np.full(pad, -1)].reshape(-1, 2, 16)            # warp, half-warp, lane
    ln = np.r_[lens, np.zeros(pad, np.int64)].reshape(-1, 2, 16)
    ph = np.r_[phi if phi is not None else np.zeros(n, np.int64), np.zeros(pad, np.int64)].reshape(-1, 2, 16)
    total = 0
    steps = ln.reshape(len(a), -1).max(axis=1)                        # a warp runs until its longest row ends
    for w in range(len(a)):
        for j in range(int(steps[w])):
            for h in range(2):
                act = (ln[w, h] > j) & (a[w, h] >= 0)
                if act.any():
                    total += np.bincount((a[w, h][act] + (j + ph[w, h][act]) % ln[w, h][act]) & 15, minlength=16).max()
    return int(total)


def wavefronts_lpr(starts, lens, order=None):
    """the lanes-per-row bins (rows_full <= 128): LPR adjacent lanes stride one row's products; `order` = the tile's table of
    group slots (row or -1), default the natural order; returns (wavefronts, LPR, y store lines)"""
    rows = len(starts)
    lpr = 2
    while lpr < 32 and rows * (lpr * 2) <= 256:
        lpr *= 2
    order = np.arange(rows) if order is None else order
    gw, gh = 32 // lpr, max(16 // lpr, 1)
    total = ylines = 0
    for w in range(0, len(order), gw):
        grp = order[w:w + gw]
        act_rows = grp[grp >= 0]
        if act_rows.size == 0:
            continue
        ylines += np.unique((act_rows + 0) >> 4).size
        steps = int(np.ceil(lens[act_rows].max() / lpr)) if lens[act_rows].max() > 0 else 0
        for k in range(steps):
            for h in range(0, len(grp), gh):
                cnt = np.zeros(16, np.int64)
                for r in grp[h:h + gh]:
                    if r < 0:
                        continue
                    pos = np.arange(lpr) + k * lpr
                    pos = pos[pos < lens[r]]
                    np.add.at(cnt, (starts[r] + pos) & 15, 1)
                total += int(cnt.max())
    return int(total), lpr, int(ylines)


def y_lines(rows_global):
    """distinct 128-byte lines of y one warp-wide store touches, summed over the warps (rows_global: -1 = idle lane)"""
    n = len(rows_global)
    r = np.r_[rows_global, np.full((-n) % 32, -1)].reshape(-1, 32)
    return int(sum(np.unique(w[w >= 0] >> 4).size for w in r))


def main():
    from spmv_research_b200 import amg
    from test_perm_emul import load, tile_info
    L = load()
    lines = sys.argv[1:] or ["1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14", "1082401 1082401 10 3.3333 normal random 0.6 100 0.95 0.5 14",
                             "1310719 1310719 5 1.6667 normal random 0.3 0 0.5 0.5 14", "550072 550072 20 6.6667 normal random 0.3 0 0.5 0.5 14"]
    rng = np.random.default_rng(0)
    for line in lines:
        p = amg.MatrixParams.parse(line)
        rp, _ = amg.row_ptr(p)
        rp = np.ascontiguousarray(rp, np.int32)
        ntiles, ti = tile_info(rp, T)
        off = np.zeros(ntiles + 1, np.int64)
        perm = np.zeros(4 * len(rp) + 64, np.uint16)
        st = np.zeros(4, np.int64)
        L.perm_emul_build(rp.ctypes.data, ti.ctypes.data, ntiles, T, EXT, off.ctypes.data, perm.ctypes.data, perm.size, st.ctypes.data)
        L.rot_emul_build.restype = None
        rot = np.zeros(len(rp), np.uint8)
        st2 = np.zeros(2, np.int64)
        L.rot_emul_build(C.c_void_p(rp.ctypes.data), C.c_void_p(ti.ctypes.data), C.c_int64(ntiles), C.c_int64(len(rp) - 1), C.c_int(T), C.c_int(EXT),
                         C.c_void_p(rot.ctypes.data), C.c_void_p(st2.ctypes.data))
        w2 = 0
        w0 = w1 = nnz = tpr = slots = rows = y0 = y1 = lw = lw1 = ly0 = ly1 = lnnz = 0
        lprs = []
        for t in rng.choice(ntiles - 1, size=min(200, ntiles - 1), replace=False):
            R0, R1 = int(ti[t, 0]), int(ti[t + 1, 0])
            rows_full = R1 - R0
            if rows_full <= 128:
                if rows_full > 0 and int(ti[t + 1, 1]) - (t * T + T) <= EXT:
                    a = rp[R0:R1].astype(np.int64) - t * T
                    ln = np.diff(rp[R0:R1 + 1]).astype(np.int64)
                    wl, lpr, yl = wavefronts_lpr(a, ln)
                    lw += wl
                    ly0 += yl
                    tab = perm[off[t]:off[t + 1]].astype(np.int64)
                    if tab.size:
                        wl, _, yl = wavefronts_lpr(a, ln, np.where(tab == 0xFFFF, -1, tab))
                    lw1 += wl
                    ly1 += yl
                    lnnz += int(ln.sum())
                    lprs.append(lpr)
                continue
            tpr += 1
            a = rp[R0:R1].astype(np.int64) - t * T
            ln = np.diff(rp[R0:R1 + 1]).astype(np.int64)
            nnz += int(ln.sum())
            w0 += wavefronts(a, ln)
            w2 += wavefronts(a, ln, rot[R0:R1].astype(np.int64))
            y0 += y_lines(np.arange(R0, R1))
            tab = perm[off[t]:off[t + 1]].astype(np.int64)
            if tab.size == 0:
                w1 += wavefronts(a, ln)
                y1 += y_lines(np.arange(R0, R1))
                slots += rows_full
            else:
                hole = tab == 0xFFFF
                r = np.where(hole, 0, tab)
                w1 += wavefronts(np.where(hole, -1, a[r]), np.where(hole, 0, ln[r]))
                y1 += y_lines(np.where(hole, -1, R0 + r))
                slots += tab.size
            rows += rows_full
        if lnnz:
            print("%-64s lanes-per-row tiles %3d (LPR median %d): LDS wavefronts per 32 nnz: shipped %.2f  perm %.2f (floor 2.00); y store lines: shipped %.2f  perm %.2f"
                  % (line, len(lprs), int(np.median(lprs)), 32.0 * lw / lnnz, 32.0 * lw1 / lnnz, 32.0 * ly0 / lnnz, 32.0 * ly1 / lnnz))
        if not tpr:
            continue
        print("%-64s tiles %3d  rows/tile %5.1f  slots/row %.2f  LDS wavefronts per 32 nnz: shipped %.2f  perm %.2f  rot %.2f (floor 2.00); y store lines per 32 nnz: shipped %.2f  perm %.2f  rot %.2f; tables on %d of %d tiles"
              % (line, tpr, rows / tpr, slots / rows, 32.0 * w0 / nnz, 32.0 * w1 / nnz, 32.0 * w2 / nnz, 32.0 * y0 / nnz, 32.0 * y1 / nnz, 32.0 * y0 / nnz, st[1], st[0]))


if __name__ == "__main__":
    main()
