"""GPU: the bounds evidence that stands in for compute-sanitizer (closed on this pool, profiles/r2_sanitizer_closed.txt):
tiny-shape launches of the three hand-synchronised kernels with guard bands around every output."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_outputs_stay_inside_their_buffers():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "sanitize_kernels.py")], capture_output=True, text=True,
                       timeout=600)
    print(r.stdout[-2000:], r.stderr[-2000:])
    assert r.returncode == 0 and "sanitize_kernels: ok" in r.stdout
