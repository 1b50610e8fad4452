"""GPU: the end-to-end step through the public API -- native step executor (csrc/kernels/stepper.cu) against the
torch stream/event path it replaces, and the native batch loader feeding it."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(native: bool, steps: int, B: int = 32):
    from mpit_b200.models.cifar_cnn import random_crop_offsets
    from mpit_b200.workloads.cifar import CifarOpts, CifarTrainer, synthetic_cifar
    os.environ["MPIT_NATIVE_STEPPER"] = "1" if native else "0"
    try:
        opts = CifarOpts(name="eamsgd", lr=1e-2, mom=0.99, mb=B, su=3, mva=0.9, impl="umma", use_graphs=True, seed=5)
        tr = CifarTrainer(opts, torch.device("cuda:0"))
    finally:
        os.environ.pop("MPIT_NATIVE_STEPPER", None)
    assert (tr._stepper is not None) == native
    x, y = synthetic_cifar(steps * B, seed=3)
    g = torch.Generator().manual_seed(9)
    handed_back = []
    for i in range(steps):
        crop = random_crop_offsets(B, g).pin_memory()
        handed_back.append(tr.train_step(x[i * B:(i + 1) * B], y[i * B:(i + 1) * B], crop))
    newest = tr.last_loss()
    rest = tr.drain_losses()
    theta = tr.params().clone()
    torch.cuda.synchronize()
    return handed_back, newest, rest, theta


def test_native_stepper_matches_the_torch_path_bit_for_bit():
    steps = 11
    a_back, a_new, a_rest, a_theta = _run(False, steps)
    b_back, b_new, b_rest, b_theta = _run(True, steps)
    assert a_back[:2] == [None, None] and b_back[:2] == [None, None]            # LAG = 2
    assert a_back == b_back and a_new == b_new and a_rest == b_rest
    assert len(b_rest) == 2 and b_rest[-1] == b_new
    assert all(l is None or 0.5 < l < 10 for l in b_back)
    assert torch.equal(a_theta, b_theta)


def test_loader_feeds_train_step():
    from mpit_b200.data.loader import BatchLoader
    from mpit_b200.workloads.cifar import CifarOpts, CifarTrainer, synthetic_cifar
    B = 32
    opts = CifarOpts(name="sgd", lr=1e-2, mom=0.9, mb=B, impl="umma", use_graphs=True, seed=2)
    tr = CifarTrainer(opts, torch.device("cuda:0"))
    x, y = synthetic_cifar(8 * B, seed=4)
    ld = BatchLoader(x, y, B, nslots=3, crop_range=5, seed=1)
    for _ in range(20):
        tr.train_step(*ld.next())
    losses = tr.drain_losses()
    ld.close()
    assert ld.epoch == 0 or ld.epoch >= 1
    assert all(0.5 < l < 10 for l in losses) and torch.isfinite(tr.params()).all()


@pytest.mark.parametrize("su,mom", [(3, 0.99), (10, 0.9), (2, 0.0), (1, 0.9)])
def test_step_finished_inside_the_reduction_is_bit_identical(su, mom):
    """On steps without an exchange the weight-gradient reduction applies the gradients, scales the momentum and
    refreshes the operands itself and the boundary kernel is skipped (optim/device.py DeviceEAMSGD.fuse_reduce).  The
    parameters after every step -- also with params() / test() interleaved, which re-pack the operands -- must equal
    those of the unfused schedule bit for bit, and the fused schedule must launch fewer kernels."""
    from mpit_b200._native import launches
    from mpit_b200.models.cifar_cnn import random_crop_offsets
    from mpit_b200.workloads.cifar import CifarOpts, CifarTrainer, synthetic_cifar
    B, steps = 32, 27
    x, y = synthetic_cifar(steps * B, seed=3, pinned=False)
    out = {}
    for fuse in ("0", "1"):
        os.environ["MPIT_FUSE_REDUCE"] = fuse
        try:
            opts = CifarOpts(name="eamsgd", lr=1e-2, mom=mom, mb=B, su=su, mva=0.3, impl="umma", use_graphs=True, seed=5)
            tr = CifarTrainer(opts, torch.device("cuda:0"))
        finally:
            os.environ.pop("MPIT_FUSE_REDUCE", None)
        assert tr.opt.fuse_reduce == (fuse == "1")
        g = torch.Generator().manual_seed(9)
        snaps, l0 = [], launches.count
        for i in range(steps):
            tr.train_step(x[i * B:(i + 1) * B].cuda(), y[i * B:(i + 1) * B].cuda(), random_crop_offsets(B, g).cuda())
            if i in (4, 5, 11, 20):
                snaps.append(tr.params().clone())          # finalises the pending step mid-run
            if i == 13:
                tr.test(x[:4 * B].cuda(), y[:4 * B].cuda())  # re-packs the operands from the parameters
        snaps.append(tr.params().clone())
        torch.cuda.synchronize()
        tr.engine._umma.check()
        out[fuse] = (snaps, launches.count - l0, tr.center.gather_full().clone())
    for a, b in zip(out["0"][0], out["1"][0]):
        assert torch.equal(a, b)
    assert torch.equal(out["0"][2], out["1"][2]), "the centre variable took the same pushes"
    if su > 1:
        assert out["1"][1] < out["0"][1], (out["0"][1], out["1"][1])
