"""The multi-GPU layer under the C-ABI (csrc/power.cu: b200spmv_power_*, NCCL bound inside the library): per-step parity
against the oracle's serial CSR kernel on every rank's slot, and the rows the peers received bit-equal on every rank
(SURVEY.md 8e "parity for iterated runs").  World size 1 runs on any GPU box; world size 2 needs two GPUs (skipped otherwise;
run by tools/gpu_n2.sh, log under profiles/).  Each rank is its own process (tests/power_worker.py), as in production."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(tmp_path, world, **kw):
    uid = str(tmp_path / "uid")
    procs, outs = [], []
    for r in range(world):
        out = str(tmp_path / ("rank%d.npz" % r))
        outs.append(out)
        cmd = [sys.executable, os.path.join(ROOT, "tests", "power_worker.py"), "--rank", str(r), "--world", str(world), "--uid-file", uid, "--out", out]
        for k, v in kw.items():
            cmd += ["--" + k.replace("_", "-"), str(v)]
        procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    logs = []
    for p in procs:
        try:
            o, _ = p.communicate(timeout=300)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        logs.append(o)
    for p, o in zip(procs, logs):
        assert p.returncode == 0, o[-3000:]
    res = [np.load(o) for o in outs]
    for r in res:
        assert float(r["worst"]) <= 1.0, float(r["worst"])
    # what every rank holds of the exchanged band after each step is the same bits (own rows and received rows alike)
    for r in res[1:]:
        assert np.array_equal(r["bands"], res[0]["bands"])
        assert np.array_equal(r["eigs"], res[0]["eigs"])
    return res


def _info(r):
    return dict(zip([str(k) for k in r["info_keys"]], [int(v) for v in r["info"]]))


@pytest.mark.parametrize("kw", [
    dict(exchange="fused"),
    dict(exchange="nccl"),
    dict(exchange="fused", panels=2, chunks=4),
    dict(exchange="fused", format="sell"),
    dict(exchange="fused", panels=2, chunks=3, graph=1),
], ids=["fused", "nccl", "panels_chunks", "sell", "panels_graph"])
def test_power_c_world1(native, oracle, tmp_path, kw):
    res = _run(tmp_path, 1, **kw)
    inf = _info(res[0])
    if kw.get("panels", 1) > 1:
        assert inf["panels"] == kw["panels"]
    if kw.get("graph"):
        assert inf["graph"] == 1


def _gpus():
    from spmv_research_b200 import capi
    return capi.device_count()


@pytest.mark.parametrize("kw", [
    dict(exchange="nccl", chunks=4),
    dict(exchange="fused", scalar_sync=0),
    dict(exchange="fused", scalar_sync=1),
    dict(exchange="fused", scalar_sync=1, panels=2, chunks=4, graph=1),
    dict(exchange="nccl", chunks=4, panels=2, graph=1),
    dict(exchange="nccl", format="sell"),
], ids=["nccl_chunks", "fused_allreduce", "fused_mailbox", "fused_mailbox_panels_graph", "nccl_panels_graph", "nccl_sell"])
def test_power_c_world2(native, oracle, tmp_path, kw):
    if _gpus() < 2:
        pytest.skip("needs two GPUs")
    res = _run(tmp_path, 2, **kw)
    inf = _info(res[0])
    if kw["exchange"] == "nccl" and kw.get("format", "csr") == "csr":
        assert inf["chunks"] == 4 and inf["nccl_calls_per_iteration"] == 5
    if kw.get("scalar_sync") == 1:
        assert inf["nccl_calls_per_iteration"] == 0


def test_fused_and_nccl_agree_world2(native, oracle, tmp_path):
    """Two exchanges, same arithmetic: the iterates are bit-equal (same kernels, same scale bits by the fixed-order sum)."""
    if _gpus() < 2:
        pytest.skip("needs two GPUs")
    (tmp_path / "a").mkdir()
    (tmp_path / "b").mkdir()
    a = _run(tmp_path / "a", 2, exchange="nccl", chunks=1)
    b = _run(tmp_path / "b", 2, exchange="fused", scalar_sync=1)
    np.testing.assert_allclose(a[0]["eigs"], b[0]["eigs"], rtol=1e-13)
    np.testing.assert_allclose(a[0]["bands"], b[0]["bands"], rtol=1e-12)


def test_c_host_program(native, oracle):
    """tests/c/power_test.c: a plain C program over the C-ABI alone (no Python / torch in the ranks), one forked process per
    GPU, every step of four exchange modes checked against the oracle's serial kernel.  Two ranks when the box has two GPUs."""
    cdir = os.path.join(ROOT, "tests", "c")
    subprocess.check_call(["make", "-C", cdir, "-s"])
    world = 2 if _gpus() >= 2 else 1
    r = subprocess.run([os.path.join(cdir, "power_test"), str(world), "300000"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert "all ranks passed" in r.stdout
