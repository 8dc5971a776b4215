"""Multi-GPU tests (need >= 2 GPUs on the box; skipped otherwise): the peer-memory framebuffer (kernels of rank 1 store into
rank 0's framebuffer over NVLink, vx_b200_export_buffer / vx_b200_set_remote_framebuffer) must deliver the same frame as
the NCCL gather, which tests/test_host_cpu.py checks against the single-rank frame with gloo."""
import json
import os
import subprocess
import sys

import pytest

from rt_common import ROOT

pytestmark = pytest.mark.gpu


def test_p2p_framebuffer_equals_nccl_gather():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "2", "--warmup", "1",
           "--workload", "cfg1_1080p_primary+shadow_100k", "--no-cpu", "--scene-upload", "broadcast"]
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["n_gpus"] == 2
    assert line["config"]["p2p_frame_equals_nccl_gather_frame"] is True
    assert "peer memory" in line["config"]["framebuffer_transport"]
    # the end-to-end leg replicates the scene by NCCL broadcast from rank 0 (+ vx_b200_mark_written; `auto` picks it from four GPUs
    # up, so it is requested here): same frame
    assert line["config"]["e2e_frame_equals_checked_frame"] is True
    assert "broadcast" in line["config"]["scene_replication"]
