"""Memory-safety sweep (the substitute for compute-sanitizer, which is closed on this GPU pool):
every kernel family on ragged shapes with operands and results in canary-guarded caller buffers,
results pre-filled with garbage, each case repeated; tools/sanitize_sweep.py has the details."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_guarded_shape_sweep_has_no_stray_writes_and_is_repeatable():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sanitize_sweep.py"), "--guard", "--repeat", "3"],
                         capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "0 failures" in out.stdout
