"""L2 rerank stage on the device (drm_l2_rerank): FIRST HARDWARE RUN.  The kernels were written after the
round's GPU budget was spent; host planning and the selection order are tested on the CPU (tests/test_oracle_l2.py), the
device path has not run yet.  The check runs in a child process (tests/l2_gpu_check.py) so a fault cannot poison this
session's CUDA context, and is marked xfail(strict=False): the outcome is recorded (XPASS / xfail) without deciding the
suite, until a round with GPU time confirms it and removes the marker."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.xfail(strict=False, reason="first hardware run of the L2 rerank kernels (not yet validated on a GPU)")
def test_l2_rerank_stage_against_oracle_first_hardware_run():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "l2_gpu_check.py")], capture_output=True, text=True, timeout=300)
    print(r.stdout[-2000:])
    print(r.stderr[-4000:])
    assert r.returncode == 0, r.stderr[-2000:]
