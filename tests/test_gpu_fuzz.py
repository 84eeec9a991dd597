"""A short soak of tools/fuzz_gpu.py: random batch shapes, mixed content, runs and damaged streams through the C ABI,
every result compared with the CPU checker (the long runs are recorded in profiles/r2_fuzz_gpu.txt)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [3, 4])
def test_fuzz_soak(seed):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_gpu.py"), "--seconds", "6", "--seed", str(seed)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "all equal to the reference" in r.stdout
