"""Multi-GPU tests: one process per GPU under torchrun, exercising the symmetric heap (IPC and VMM /
multicast), the sharded centre, device collectives, the device transport of the message layer, the asynchronous
parameter server and the EAMSGD / sync-EASGD trainers."""
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
def test_center_and_collectives(mode):
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["center", "collectives"], n, {"MPIT_SYMM_MODE": mode})
    assert "case collectives: ok" in out


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


def test_message_api_on_device_tensors():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["msg_device"], n)
    assert "case msg_device: ok" in out


def test_ps_protocol_gpu_to_gpu_over_device_transport():
    """The reference's ptest (pServer / pClient, aio_send / aio_recv) with every vector in device memory."""
    n = min(torch.cuda.device_count(), 8) // 2 * 2
    cmd = [sys.executable, "-m", "mpit_b200.launch.run", "-np", str(n), "-m", "mpit_b200.launch.ptest", "--device", "cuda",
           "--protocol", "ps", "--ssize", str(3 * 1024 * 1024 + 5), "--rounds", "5"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=os.path.dirname(HERE))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "(device transport)" in r.stdout and "'rendezvous': 0" not in r.stdout.split("device transport:")[-1]


def test_pull_fused_with_consumer_gemm():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["pulled_linear"], n)
    assert "case pulled_linear: ok" in out


def test_model_trained_through_pull_fused_gemms():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["pulled_mlp"], n)
    assert "case pulled_mlp: ok" in out


def test_one_sided_window():
    """MPI_Put / Accumulate / Get / Win_fence on the symmetric heap."""
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun(["window"], n)
    assert "case window: ok" in out


def test_ring_smoke_script_on_device():
    """/root/reference/test.lua with CUDA tensors: mpiT.Send / Recv over the device transport (one float around the ring of GPUs)."""
    n = min(torch.cuda.device_count(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", "29534", "-m", "mpit_b200.launch.test", "--device", "cuda"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=os.path.dirname(HERE))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "after: rbuf" in r.stdout
