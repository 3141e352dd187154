"""The edge-case set through EVERY product path in one process (tools/run_sanitize_cases.py: CSR plain / compressed rows, SELL
with several C / sigma, column panels, the fused epilogue with a peer destination, fp32, COO -> CSR, the device generator, PCG),
each result checked against the oracle and the guard region behind y[m] required untouched.

That tool was written to run under compute-sanitizer (memcheck / racecheck).  compute-sanitizer is closed on
this GPU pool ("runs under it have left GPUs needing a reset" -- gpurun refuses it, profiles/r02d_sanitizer_closed.txt), so
what is left is what the pool's message recommends: small cases, result comparison with the CPU reference, guard regions."""
import importlib.util
import os

import pytest

pytestmark = [pytest.mark.gpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_edge_case_set_through_every_path(native, oracle, capsys):
    spec = importlib.util.spec_from_file_location("run_sanitize_cases", os.path.join(ROOT, "tools", "run_sanitize_cases.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.main() == 0, capsys.readouterr().out[-3000:]
