"""Multi-GPU tests: one process per GPU under torchrun, exercising the symmetric heap (IPC and VMM /
multicast), the sharded centre, device collectives, the mailbox and asynchronous EAMSGD."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
HERE = os.path.dirname(os.path.abspath(__file__))


def _torchrun(cases, nproc, env=None, timeout=600):
    e = dict(os.environ)
    e.update(env or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(HERE, "mp_gpu_cases.py"), *cases]
    r = subprocess.run(cmd, env=e, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


@pytest.mark.parametrize("mode", ["ipc", "vmm"])
def test_center_collectives_mailbox(mode):
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["center", "collectives", "mailbox"], n, {"MPIT_SYMM_MODE": mode})
    assert "case mailbox: ok" in out


def test_async_eamsgd_training():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["eamsgd_train"], n)
    assert "case eamsgd_train: ok" in out


def test_server_side_rules_on_device_path():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["server_rules"], n)
    assert "case server_rules: ok" in out


def test_async_parameter_server_under_rank_skew():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["ps_stress"], n)
    assert "case ps_stress: ok" in out


def test_sync_easgd_allreduce_fused_with_elastic_step():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["sync_easgd"], n)
    assert "case sync_easgd: ok" in out


def test_pull_fused_with_consumer_gemm():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["pulled_linear"], n)
    assert "case pulled_linear: ok" in out


def test_one_sided_window():
    """MPI_Put / Accumulate / Get / Win_fence on the symmetric heap."""
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["window"], n)
    assert "case window: ok" in out


def test_ring_smoke_script_on_device():
    """/root/reference/test.lua over the NVLink mailbox (one float around the ring of GPUs)."""
    n = min(torch.cuda.device_count(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", "29534", "-m", "mpit_b200.launch.test", "--device", "cuda"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=os.path.dirname(HERE))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "after: rbuf" in r.stdout
