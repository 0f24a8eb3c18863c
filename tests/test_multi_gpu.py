# SPDX-License-Identifier: WTFPL
"""N-GPU image == 1-GPU image (SURVEY.md 8e "determinism note"; the invariant is the one store per pixel of
shader.comp:56-57, split over ranks): launches tests/multi_gpu_worker.py under torch.distributed.run on
min(2, device_count) GPUs (VCRT_TEST_GPUS overrides).  bench.py repeats the check at the driver's N = 2 / 4 / 8 and puts the
outcome into its JSON line (`parity`)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sharded_render_equals_single_gpu_render():
    import torch
    have = torch.cuda.device_count()
    want = int(os.environ.get("VCRT_TEST_GPUS", "2"))
    n = min(want, have)
    if n < 2:
        pytest.skip(f"needs 2 GPUs, this box has {have} (run `gpurun --gpus 2 -- python -m pytest tests/test_multi_gpu.py -m gpu`; "
                    "bench.py carries the same check at N = 2 / 4 / 8)")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    sys.stdout.write(out.stdout[-6000:])
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert f"MULTI_GPU_CHECK PASS ranks {n}" in out.stdout
