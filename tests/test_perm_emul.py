"""csrc/csr_perm.h -- the row -> thread table of the CSR_EXP_PERM experiment (not in the shipped library) --
checked on the CPU: for every tile given a table each completed row appears in exactly one group slot, and the rows of every
half-warp (16 / LPR group slots, LPR = lanes per row of the tile's bin) start at the same offset mod LPR and in different
bank windows, which is what makes each step of the row sums free of shared-memory bank conflicts (replayed for a sample).  tile_info is recomputed here the way tile_info_kernel (convert.cu) does."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import matrices as M

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "emul", "perm_emul.cpp")
HDR = os.path.join(HERE, "..", "spmv_research_b200", "csrc", "csr_perm.h")
SO = os.path.join(HERE, "emul", "libperm_emul.so")
T, EXT = 2048, 256


def load():
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in (SRC, HDR)):
        subprocess.check_call(["g++", "-O2", "-fPIC", "-shared", "-Wall", "-Wextra", "-o", SO, SRC])
    L = C.CDLL(SO)
    L.perm_emul_build.restype = C.c_int64
    L.perm_emul_build.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    return L


def tile_info(rp, T):
    m, nnz = len(rp) - 1, int(rp[-1])
    ntiles = max(1, (nnz + T - 1) // T)
    ti = np.zeros((ntiles + 1, 2), np.int32)
    for t in range(ntiles):
        R0 = 0 if t == 0 else int(np.searchsorted(rp[:m], t * T, side="left"))       # first row with row_ptr >= base (row_ptr[m] = nnz)
        ti[t] = (R0, rp[R0])
    ti[ntiles] = (m, nnz)
    return ntiles, ti


def cases():
    rng = np.random.default_rng(11)
    out = [("avg 10", M.csr_from_lengths(np.clip(rng.normal(10, 3.3, 60000).round(), 0, None).astype(np.int64), 60000, 1)[0]),
           ("avg 5 + empties", M.csr_from_lengths(rng.integers(0, 11, 80000), 80000, 2)[0]),
           ("all length 8 (every start = 0 mod 8)", M.csr_from_lengths(np.full(30000, 8), 30000, 3)[0]),
           ("length 1 rows (2048 rows per tile)", M.csr_from_lengths(np.ones(20000, np.int64), 20000, 4)[0]),
           ("mostly empty", M.csr_from_lengths((rng.random(200000) < 0.05).astype(np.int64) * 7, 200000, 5)[0]),
           ("long rows (no thread-per-row tiles)", M.csr_from_lengths(np.full(500, 300), 5000, 6)[0]),
           ("avg 20 (2 lanes per row)", M.csr_from_lengths(np.clip(rng.normal(20, 6.7, 30000).round(), 0, None).astype(np.int64), 30000, 8)[0]),
           ("avg 50 (4 lanes per row)", M.csr_from_lengths(np.clip(rng.normal(50, 16.7, 12000).round(), 0, None).astype(np.int64), 12000, 9)[0]),
           ("avg 100 (8 lanes per row)", M.csr_from_lengths(np.clip(rng.normal(100, 33, 6000).round(), 0, None).astype(np.int64), 6000, 10)[0]),
           ("giant row inside short ones", M.csr_from_lengths(np.r_[rng.integers(1, 9, 3000), 9000, rng.integers(1, 9, 3000)], 20000, 7)[0])]
    return out


def lpr_of(rows_full):
    if rows_full > 128:
        return 1
    lpr = 2
    while lpr < 32 and rows_full * (lpr * 2) <= 256:
        lpr *= 2
    return lpr


def bank_conflicts(rp, R0, base, groups, lpr):
    """highest multiplicity of a bank pair over all steps of one half-warp (groups: global rows of its G group slots)"""
    worst = 0
    steps = max([-(-int(rp[r + 1] - rp[r]) // lpr) for r in groups] + [0])
    for k in range(steps):
        cnt = np.zeros(16, np.int64)
        for r in groups:
            a, ln = int(rp[r]) - base, int(rp[r + 1] - rp[r])
            for l in range(lpr):
                if l + k * lpr < ln:
                    cnt[(a + l + k * lpr) & 15] += 1
        worst = max(worst, int(cnt.max()))
    return worst


def test_perm_table_covers_rows_once_and_is_conflict_free():
    L = load()
    total = 0
    stats = {}
    seen_lpr = set()
    for name, rp in cases():
        tables = skipped = 0
        rp = np.ascontiguousarray(rp, np.int32)
        ntiles, ti = tile_info(rp, T)
        off = np.zeros(ntiles + 1, np.int64)
        cap = 4 * len(rp) + 64
        perm = np.zeros(cap, np.uint16)
        st = np.zeros(4, np.int64)
        n = L.perm_emul_build(rp.ctypes.data, ti.ctypes.data, ntiles, T, EXT, off.ctypes.data, perm.ctypes.data, cap, st.ctypes.data)
        assert 0 <= n <= cap and off[0] == 0 and off[-1] == n and np.all(np.diff(off) >= 0), name
        for t in range(ntiles):
            R0, R1 = int(ti[t, 0]), int(ti[t + 1, 0])
            nrows = R1 - R0
            over = int(ti[t + 1, 1]) - (t * T + T)
            rows_full = nrows - 1 if (nrows > 0 and over > EXT) else nrows
            tab = perm[off[t]:off[t + 1]]
            lpr = lpr_of(rows_full) if rows_full > 0 else 32
            if rows_full < 1 or lpr > 8:
                assert tab.size == 0, (name, t)               # nothing to order / conflict-free as it is
                continue
            total += 1
            if tab.size == 0:                                  # the wavefront model kept the natural order for this tile
                skipped += 1
                continue
            tables += 1
            seen_lpr.add(lpr)
            G = 16 // lpr
            assert tab.size % G == 0 and tab.size <= 4 * rows_full, (name, t)                     # whole half-warps, bounded size
            rows = tab[tab != 0xFFFF].astype(np.int64)
            assert np.array_equal(np.sort(rows), np.arange(rows_full)), (name, t)                  # every completed row exactly once
            for g in range(0, tab.size, G):
                grp = tab[g:g + G]
                r = grp[grp != 0xFFFF].astype(np.int64) + R0
                ne = r[rp[r + 1] > rp[r]]
                a = rp[ne].astype(np.int64) - t * T
                assert np.unique(a % lpr).size <= 1, (name, t, g)                                  # one offset class per half-warp
                assert np.unique((a // lpr) % G).size == a.size, (name, t, g)                      # G different bank windows
                if t % 7 == 0 and g % (5 * G) == 0:
                    assert bank_conflicts(rp, R0, t * T, list(ne), lpr) <= 1, (name, t, g)         # ... hence no conflict at any step
        stats[name] = (tables, skipped)
        assert st[0] == tables + skipped and st[1] == tables and st[3] <= st[2], (name, st)       # never worse than natural, by the model
        if name in ("avg 10", "avg 5 + empties", "avg 20 (2 lanes per row)"):
            assert st[3] < 0.9 * st[2], (name, st)                                               # modelled wavefronts (sums + y stores) drop by > 10 %
    assert total > 300 and {1, 2, 4} <= seen_lpr
    assert stats["all length 8 (every start = 0 mod 8)"][0] == 0                  # two residues only: never worth a table
    assert stats["avg 10"][1] <= 0.05 * stats["avg 10"][0] and stats["avg 5 + empties"][1] <= 0.05 * stats["avg 5 + empties"][0]


def test_rotation_table_is_valid_and_lowers_the_modelled_conflicts():
    """CSR_EXP_ROT (csr_rot_build): every rotation is smaller than its row's length (0 for rows outside the thread-per-row
    bin), and the wavefronts of the rotated reads -- recomputed here from the bank arithmetic -- are what the builder reports
    and fewer than the natural order's; rows of one even length (stride-10 starts: a 2-way conflict) become conflict-free."""
    L = load()
    L.rot_emul_build.restype = None
    L.rot_emul_build.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    extra = [("all length 10", M.csr_from_lengths(np.full(30000, 10), 30000, 12)[0])]
    for name, rp in cases()[:5] + extra:
        rp = np.ascontiguousarray(rp, np.int32)
        m = len(rp) - 1
        ntiles, ti = tile_info(rp, T)
        rot = np.zeros(max(m, 1), np.uint8)
        st = np.zeros(2, np.int64)
        L.rot_emul_build(rp.ctypes.data, ti.ctypes.data, ntiles, m, T, EXT, rot.ctypes.data, st.ctypes.data)
        ln = np.diff(rp)
        assert np.all(rot[:m] < np.maximum(ln, 1)), name
        nat = got = 0
        in_bin = np.zeros(m, bool)
        for t in range(ntiles):
            R0, R1 = int(ti[t, 0]), int(ti[t + 1, 0])
            over = int(ti[t + 1, 1]) - (t * T + T)
            rows_full = R1 - R0 - 1 if (R1 > R0 and over > EXT) else R1 - R0
            if rows_full <= 128:
                continue
            in_bin[R0:R0 + rows_full] = True
            if t % 5:
                continue
            for h in range(R0, R0 + rows_full, 16):
                rows = np.arange(h, min(h + 16, R0 + rows_full))
                a, l, ph = rp[rows].astype(np.int64) - t * T, ln[rows].astype(np.int64), rot[rows].astype(np.int64)
                for j in range(int(l.max()) if l.size else 0):
                    act = l > j
                    nat += np.bincount((a[act] + j) & 15, minlength=16).max()
                    got += np.bincount((a[act] + (j + ph[act]) % l[act]) & 15, minlength=16).max()
        assert not rot[:m][~in_bin].any(), name
        assert st[1] <= st[0], (name, st)
        if nat:
            assert got <= nat and abs(got / nat - st[1] / st[0]) < 0.05, (name, got, nat, st)      # the sample agrees with the builder's totals
        if name in ("avg 10", "avg 5 + empties"):
            assert st[1] < 0.7 * st[0], (name, st)
        if name == "all length 10":
            assert st[0] == 2 * st[1], (name, st)                                                 # 2-way conflict -> none
