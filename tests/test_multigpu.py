"""Multi-GPU tests.  GPU part: slab mode on >= 2 B200 (skipped on one GPU).
CPU part (gloo, world_size 2, runs here): the slab decomposition arithmetic --
which rank owns which pass-A lines and which layers, and that scattering the
pruned pass-A output to the layer owners followed by local passes B and C
reproduces the single-rank field -- restated with numpy FFTs."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def torchrun(script, nproc, extra_env=None, timeout=600):
    env = dict(os.environ)
    env.update(extra_env or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1", "--master-port", str(free_port()), script]
    return subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout, env=env, cwd=ROOT)


@pytest.mark.gpu
def test_slab_mode_equals_single_gpu():
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    nproc = 8 if n >= 8 else (4 if n >= 4 else 2)
    out = torchrun(os.path.join(ROOT, "tests", "mgpu_slab_check.py"), nproc)
    print(out.stdout[-3000:])
    assert out.returncode == 0, out.stdout[-3000:]
    assert "SLAB PARITY" in out.stdout and "OK" in out.stdout


@pytest.mark.gpu
def test_slab_peer_wait_is_bounded_and_fatal():
    """one rank sleeps past PF_SLAB_TIMEOUT_S: the waiting rank's call fails
    within the limit and its plan stays failed (no hang, no silent result)"""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    out = torchrun(os.path.join(ROOT, "tests", "mgpu_slab_timeout.py"), 2, extra_env={"PF_SLAB_TIMEOUT_S": "2"}, timeout=180)
    print(out.stdout[-3000:])
    assert out.returncode == 0, out.stdout[-3000:]
    assert "SLAB TIMEOUT OK" in out.stdout


def test_slab_decomposition_gloo_world2():
    out = torchrun(os.path.join(ROOT, "tests", "gloo_slab_emulation.py"), 2, timeout=300)
    print(out.stdout[-3000:])
    assert out.returncode == 0, out.stdout[-3000:]
    assert "GLOO SLAB OK" in out.stdout
