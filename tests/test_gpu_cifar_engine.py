"""GPU numerics: CIFAR CNN engine (fp32 CUDA-core path) vs a PyTorch CPU fp32 oracle."""
import pytest
import torch

from mpit_b200.models.cifar_cnn import (CifarEngine, TorchCifarCNN, flat_from_torch, random_crop_offsets,
                                        center_crop_offsets, NUM_CLASSES, CROP)

pytestmark = pytest.mark.gpu


def _oracle(model, images_u8, labels, crop, drop_mask):
    x = images_u8.float() / 255.0
    xs = torch.stack([x[i, :, crop[i, 0]:crop[i, 0] + CROP, crop[i, 1]:crop[i, 1] + CROP]
                      for i in range(x.shape[0])])
    model.zero_grad()
    logp = model(xs, drop_mask=drop_mask)
    loss = torch.nn.functional.nll_loss(logp, labels.long(), reduction="mean")
    return logp, loss


def _run(impl, B=16, train=True, tol=2e-4):
    torch.manual_seed(0)
    model = TorchCifarCNN()
    theta = flat_from_torch(model)
    images = torch.randint(0, 256, (B, 3, 32, 32), dtype=torch.uint8)
    labels = torch.randint(0, NUM_CLASSES, (B,), dtype=torch.int32)
    crop = random_crop_offsets(B) if train else center_crop_offsets(B)
    eng = CifarEngine(B, impl=impl, seed=7)
    eng.load_flat(theta.cuda())
    eng.set_batch(images, labels, crop)
    if train:
        eng.fwd_bwd()
    else:
        eng.evaluate()
    torch.cuda.synchronize()
    a4 = eng.act[3].view(B, 256).cpu()
    mask = (a4 > 0).float() * 2.0 if train else torch.ones(B, 256)
    logp, loss = _oracle(model, images, labels, crop, mask)
    got_logits = eng.logits().cpu()
    got_logp = torch.log_softmax(got_logits, dim=1)
    assert torch.allclose(got_logp, logp.detach(), atol=tol, rtol=tol)
    assert abs(float(eng.loss_sum.item()) / B - float(loss)) < tol * 10
    if train:
        loss.backward()
        ref_g = flat_from_torch(model, grads=True)
        got_g = eng.flat_grad().cpu()
        denom = ref_g.abs().max().item()
        assert (got_g - ref_g).abs().max().item() <= tol * max(1.0, denom) + tol
        # about half of the active units are dropped
        frac = float((a4 > 0).float().mean())
        assert 0.05 < frac < 0.6
    return eng


def test_engine_ref_train_matches_torch_cpu():
    _run("ref", train=True)


def test_engine_ref_eval_matches_torch_cpu():
    eng = _run("ref", train=False)
    assert int(eng.confusion.sum().item()) == eng.B


def test_dropout_counter_advances():
    eng = _run("ref", train=True)
    assert int(eng.rng[1].item()) == 1
