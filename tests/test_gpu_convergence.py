"""Convergence on a learnable synthetic set: the tensor-core path (TF32 products), its fp32-accurate 3xTF32 variant, the exact-fp32
CUDA-core path and a plain PyTorch fp32 model are trained from ONE initialisation on the SAME mini-batches with the same plain-SGD rule
(/root/reference/asyncsgd/optim-msgd.lua:30-36 with mom = 0) for 300 steps.  What is checked is what a wrong kernel
would break: every path learns (loss far below ln 10, accuracy far above chance), and the loss curves of the three
paths stay together -- TF32 does not change what is learned.  Dropout is the one source of divergence between the
paths (three different mask streams), so the curves are compared smoothed over 20-step windows."""
import math

import pytest
import torch

from mpit_b200.models.cifar_cnn import TorchCifarCNN, flat_from_torch, torch7_reset, CROP
from mpit_b200.workloads.cifar import CifarTrainer, CifarOpts, learnable_cifar

pytestmark = pytest.mark.gpu

B, STEPS, LR = 64, 360, 0.04


def _batches():
    x, y = learnable_cifar(32 * B, seed=5, pinned=False)
    g = torch.Generator().manual_seed(11)
    out = []
    for i in range(STEPS):
        j = (i % 32) * B
        crop = torch.randint(0, 5, (B, 2), dtype=torch.int32, generator=g)
        out.append((x[j:j + B].contiguous(), y[j:j + B].contiguous(), crop))
    return out, learnable_cifar(8 * B, seed=99, pinned=False)


def _theta0():
    m = TorchCifarCNN()
    torch7_reset(m, torch.Generator().manual_seed(3))
    return m, flat_from_torch(m)


def _train_engine(impl, batches, theta0):
    opts = CifarOpts(name="sgd", lr=LR, mom=0.0, mb=B, impl=impl, use_graphs=(impl != "ref"), seed=21)
    tr = CifarTrainer(opts, torch.device("cuda:0"), theta0=theta0.clone())
    losses = []
    for x, y, c in batches:
        tr.train_step(x.cuda(), y.cuda(), c.cuda())
        losses.append(float(tr.engine.loss_sum.item()) / B)
    return tr, losses


def _train_torch(batches, model):
    model = model.cuda()
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    opt = torch.optim.SGD(model.parameters(), lr=LR)
    gen = torch.Generator(device="cuda").manual_seed(21)
    losses = []
    for x, y, c in batches:
        xf = x.cuda().float() / 255.0
        xs = torch.stack([xf[i, :, int(c[i, 0]):int(c[i, 0]) + CROP, int(c[i, 1]):int(c[i, 1]) + CROP] for i in range(B)])
        mask = (torch.rand(B, 256, device="cuda", generator=gen) < 0.5).float() * 2.0
        opt.zero_grad()
        loss = torch.nn.functional.nll_loss(model(xs, drop_mask=mask), y.cuda().long())
        loss.backward()
        opt.step()
        losses.append(float(loss))
    return model, losses


def _smooth(v, w=20):
    return [sum(v[i:i + w]) / w for i in range(0, len(v) - w + 1, w)]


def test_three_paths_learn_the_same_thing():
    batches, (tx, ty) = _batches()
    model, theta0 = _theta0()
    curves = {}
    acc = {}
    for impl in ("umma", "umma3x", "ref"):
        tr, curves[impl] = _train_engine(impl, batches, theta0)
        acc[impl] = tr.test(tx.cuda(), ty.cuda())["acc"]
        if impl != "ref":
            tr.engine._umma.check()
    tm, curves["torch"] = _train_torch(batches, model)
    with torch.no_grad():
        xs = tx.cuda().float()[:, :, 2:2 + CROP, 2:2 + CROP] / 255.0
        acc["torch"] = float((tm(xs, drop_mask=torch.ones(xs.shape[0], 256, device="cuda")).argmax(1) == ty.cuda()).float().mean())
    sm = {k: _smooth(v) for k, v in curves.items()}
    # "learns" is judged on the best of the late windows: plain SGD with dropout at this step size throws an occasional
    # loss spike after convergence (and the exact-fp32 path accumulates weight gradients with float atomics, so where
    # its spikes fall differs from run to run); a path with a broken kernel never gets near these levels
    for k, v in sm.items():
        assert v[0] > 1.5, (k, v[0])                                   # starts near ln 10
        assert min(v[-6:]) < 0.25 * v[0], (k, v)                        # and every path learns
        assert acc[k] > 0.5 or min(v[-3:]) < 0.1, (k, acc, v)           # chance is 0.1
    print("smoothed loss curves:", {k: [round(x, 3) for x in v] for k, v in sm.items()}, "accuracy:", acc)
    # the curves stay together: the paths differ in their dropout streams (and, for TF32, in rounding), which shifts
    # the steep part of the descent by a few steps -- hence a tolerance that is relative to the loss level
    # Compared over the descent (windows in which the fp32 torch model is still above 0.5): below that a path's loss is
    # dominated by rare SGD spikes of its own dropout stream, which land in different windows on different paths (seen:
    # 0.44 against 0.16 in one window, 0.14 against 0.42 in the next).  After the descent the paths are compared on
    # the level they settle at (median of the late windows), not window by window.
    descent = [i for i, b in enumerate(sm["torch"]) if b > 0.5 and all(x > 0.5 for x in sm["torch"][:i])]
    assert len(descent) >= 6, sm["torch"]
    late = {k: sorted(v[-6:])[2] for k, v in sm.items()}
    assert max(late.values()) < 0.3, late
    for k in ("umma", "umma3x", "ref"):
        for i in descent:
            a, b = sm[k][i], sm["torch"][i]
            assert abs(a - b) < 0.2 + 0.35 * b, (k, i, sm[k], sm["torch"])
        assert min(sm[k][-6:]) < 0.15, (k, sm[k])
    for k in ("umma", "umma3x"):
        for i in descent:
            a, b = sm[k][i], sm["ref"][i]
            assert abs(a - b) < 0.2 + 0.35 * b, (k, i, sm[k], sm["ref"])
    assert math.isfinite(curves["umma"][-1])
