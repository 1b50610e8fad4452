"""CIFAR-10 Torch7 loader (goot.lua:62-84) and the reference's custom nn modules (dropout.lua, DivideConstant.lua,
Normalize.lua) restated in PyTorch."""
import torch

from mpit_b200 import t7
from mpit_b200.data import load_cifar10, write_cifar10_t7
from mpit_b200.models.modules import DivideConstant, Dropout, Normalize


def test_t7_ascii_round_trip(tmp_path):
    obj = {"data": torch.randint(0, 256, (7, 5), dtype=torch.uint8), "labels": torch.tensor([[1.0, 2.0, 3.5]]),
           "name": "hello world", "n": 5, "nested": {"a": [1, 2, 3], "f": torch.randn(3, 2)}, "flag": True}
    p = tmp_path / "x.t7"
    t7.save(str(p), obj, "ascii")
    o = t7.load(str(p), "ascii")
    assert torch.equal(o["data"], obj["data"]) and torch.equal(o["labels"], obj["labels"])
    assert o["name"] == "hello world" and o["n"] == 5 and o["flag"] is True
    assert torch.allclose(o["nested"]["f"], obj["nested"]["f"])
    assert o["nested"]["a"] == {1: 1, 2: 2, 3: 3}


def test_t7_ascii_reads_torch7_line_layout():
    """Sizes/strides on ONE line, as THDiskFile writes them."""
    text = (b"4\n1\n3\nV 1\n16\ntorch.ByteTensor\n2\n2 3\n3 1\n1\n4\n2\n3\nV 1\n17\ntorch.ByteStorage\n6\n"
            b"\x01\x02\x03\x04\x05\x06\n")
    out = t7.loads(text) if False else t7._AsciiReader(__import__("io").BytesIO(text)).obj()
    assert out.tolist() == [[1, 2, 3], [4, 5, 6]]


def test_cifar10_loader_round_trip(tmp_path):
    g = torch.Generator().manual_seed(3)
    tr = (torch.randint(0, 256, (50, 3, 32, 32), dtype=torch.uint8, generator=g),
          torch.randint(0, 10, (50,), dtype=torch.int32, generator=g))
    te = (torch.randint(0, 256, (12, 3, 32, 32), dtype=torch.uint8, generator=g),
          torch.randint(0, 10, (12,), dtype=torch.int32, generator=g))
    for mode in ("ascii", "binary"):
        root = tmp_path / mode
        write_cifar10_t7(str(root), tr, te, mode)
        (x, y), (tx, ty), real = load_cifar10(str(root), trsize=45, tesize=10, synthetic_ok=False)
        assert real and x.shape == (45, 3, 32, 32) and tx.shape == (10, 3, 32, 32)
        assert torch.equal(x, tr[0][:45]) and torch.equal(y, tr[1][:45])
        assert torch.equal(tx, te[0][:10]) and torch.equal(ty, te[1][:10])
    (x, y), _, real = load_cifar10(str(tmp_path / "missing"), trsize=20, tesize=5)
    assert not real and x.shape == (20, 3, 32, 32) and x.dtype == torch.uint8


def test_dropout_module_matches_reference_semantics():
    torch.manual_seed(0)
    d = Dropout(0.5)
    x = torch.randn(64, 32, requires_grad=True)
    y = d(x)
    assert torch.equal(y.detach(), x.detach() * d.noise / 0.5)
    y.sum().backward()
    assert torch.equal(x.grad, d.noise / 0.5)                       # dropout.lua:24-29
    assert 0.3 < float(d.noise.mean()) < 0.7
    assert torch.equal(Dropout(0.0)(x).detach(), x.detach())


def test_divide_constant_and_normalize():
    x = torch.rand(5, 7, dtype=torch.float64) + 0.5
    x.requires_grad_(True)
    assert torch.autograd.gradcheck(DivideConstant(3.0), (x,))
    assert torch.allclose(DivideConstant(3.0)(x), 3.0 / x)
    z = torch.randn(4, 9, dtype=torch.float64, requires_grad=True)
    for p in (2.0, 3.0):
        n = Normalize(p)
        assert torch.allclose(n(z).abs().pow(p).sum(-1), torch.ones(4, dtype=torch.float64), atol=1e-8)
        assert torch.autograd.gradcheck(n, (z,))


def test_checkpoint_resume_round_trip(tmp_path):
    """A ``model-epoch-N.net`` written by train_epochs (goot.lua:246-254) restores the exact flat parameters;
    a flat tensor file (the BiCNN style, bicnn.lua:590-594) is accepted too."""
    from mpit_b200.models.cifar_cnn import TorchCifarCNN, flat_from_torch
    from mpit_b200.workloads.cifar import load_checkpoint
    theta = flat_from_torch(TorchCifarCNN())
    p = tmp_path / "model-epoch-1.net"
    t7.save(str(p), t7.cifar_model_to_t7(theta))
    assert torch.equal(load_checkpoint(str(p)), theta)
    q = tmp_path / "flat_model"
    t7.save(str(q), theta)
    assert torch.equal(load_checkpoint(str(q)), theta)


def test_native_batch_loader_epochs_crops_and_determinism():
    """csrc/io/loader.cpp: every sample exactly once per epoch, labels travel with their images, crop offsets in
    range, same seed -> same stream, sequential order without shuffling."""
    from mpit_b200.data.loader import BatchLoader
    n, B = 40, 8
    x = torch.arange(n, dtype=torch.uint8).view(n, 1, 1, 1).expand(n, 3, 4, 4).contiguous()
    y = (torch.arange(n, dtype=torch.int32) * 7) % 10
    ld = BatchLoader(x, y, B, nslots=3, crop_range=5, seed=11)
    seen, crops = [], []
    for step in range(10):                                   # two epochs
        bx, by, bc = ld.next()
        ids = bx[:, 0, 0, 0].to(torch.int64)
        assert torch.equal(bx, x[ids]) and torch.equal(by, y[ids])
        assert int(bc.min()) >= 0 and int(bc.max()) <= 4
        seen.append(ids.clone())
        crops.append(bc.clone())
        assert ld.epoch == step // 5
    e0, e1 = torch.cat(seen[:5]), torch.cat(seen[5:])
    assert sorted(e0.tolist()) == list(range(n)) and sorted(e1.tolist()) == list(range(n))
    assert not torch.equal(e0, e1) and not torch.equal(e0, torch.arange(n))
    assert len({tuple(c.view(-1).tolist()) for c in crops}) > 1
    ld2 = BatchLoader(x, y, B, nslots=3, crop_range=5, seed=11)
    for step in range(10):
        bx, by, bc = ld2.next()
        assert torch.equal(bx[:, 0, 0, 0].to(torch.int64), seen[step]) and torch.equal(bc, crops[step])
    ld.close()
    ld2.close()
    seq = BatchLoader(x, y, B, nslots=2, crop_range=0, seed=1, shuffle=False)
    for step in range(7):                                    # 40 // 8 = 5 batches per epoch, then wraps
        bx, by, bc = seq.next()
        lo = (step % 5) * B
        assert bx[:, 0, 0, 0].tolist() == list(range(lo, lo + B)) and int(bc.abs().sum()) == 0
    seq.close()
