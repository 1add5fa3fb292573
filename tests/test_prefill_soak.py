"""Randomised soak of the prompt kernels (tools/prefill_soak.py) as a test: ragged batches of random shapes through
the plain and the paged operator, every mix of item types in both grid orders, with and without context splits.
The failure it guards against is a hang (the output-store hand-over once dead-locked on one particular order of
item types), so it runs in a child process under a timeout."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [11, 12])
def test_random_batches_finish_and_match(seed):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "prefill_soak.py"), "40", str(seed)],
                       capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert "soak ok: 40 iterations" in r.stdout
