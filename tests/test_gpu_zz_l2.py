"""L2 rerank stage on the device (drm_l2_rerank) against oracle/l2_oracle.py.

The kernels were written after round 1's GPU budget was spent, so their first execution on hardware is this test.  Before
that, the same check (tests/l2_gpu_check.py) passed against the host-thread emulation of the very same sources - api.cu,
l2_plan.cpp, embed_kernel.cu, l2_kernel.cu - including under AddressSanitizer (tests/test_emu_kernels.py).  The check runs
in a child process so that a device fault, should one happen, cannot take this session's CUDA context with it."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cache", ["default", "always", "off"])
def test_l2_rerank_stage_against_oracle(cache):
    """With the window-embedding cache forced (DRM_L2_CACHE_SHARE=0: every call builds or reuses it; the check switches between
    references and between decorated / bare windows on one embedder, so the cache key is exercised), under the default rule,
    and without it (DRM_L2_CACHE_GB=0: every batch embeds its own windows)."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ)
    if cache == "always":
        env["DRM_L2_CACHE_SHARE"] = "0"
    elif cache == "off":
        env["DRM_L2_CACHE_GB"] = "0"
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "l2_gpu_check.py")], capture_output=True, text=True, timeout=300, env=env)
    print(r.stdout[-2000:])
    print(r.stderr[-4000:])
    assert r.returncode == 0 and "device == oracle" in r.stdout, r.stderr[-2000:]


def test_pipeline_l2_stage_host_code_against_oracle():
    """post_process_l2_flat (what `pipeline` runs under DRM_RERANK=l2) through `formats_tool l2flat`: dynamic over a FASTA
    genome, static over the bare lookup-mode window table (device mode 2).  Same child-process arrangement."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "l2_pipeline_check.py")], capture_output=True, text=True, timeout=300)
    print(r.stdout[-2000:])
    print(r.stderr[-4000:])
    assert r.returncode == 0 and "host code == oracle" in r.stdout, r.stderr[-2000:]
