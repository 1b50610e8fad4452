"""Torch7 binary format: reader on the reference's shipped blobs, writer round trips, model graphs."""
import io
import os

import pytest
import torch

from mpit_b200 import t7

REF = "/root/reference/BiCNN"
needs_ref = pytest.mark.skipif(not os.path.exists(f"{REF}/binary_mapWordIdx2WordStr"), reason="reference blobs absent")


@needs_ref
def test_reads_reference_vocabulary_tables():
    idx2str = t7.load(f"{REF}/binary_mapWordIdx2WordStr")
    str2idx = t7.load(f"{REF}/binary_mapWordStr2WordIdx")
    assert len(idx2str) == 22354 and len(str2idx) == 22354     # V of the BiCNN embedding
    assert all(isinstance(k, int) for k in list(idx2str)[:100])
    for i in (1, 2, 100, 22354):
        assert str2idx[idx2str[i]] == i


@needs_ref
def test_rewrites_reference_blob_identically():
    raw = open(f"{REF}/binary_mapWordIdx2WordStr", "rb").read()
    obj = t7.loads(raw)
    again = t7.dumps(obj)
    assert t7.loads(again) == obj
    assert len(again) == len(raw)      # same grammar, same sizes


def test_roundtrip_scalars_tables_tensors():
    obj = {"n": 3.25, "i": 7, "s": "héllo".encode("latin-1").decode("latin-1"), "b": True, "nil": None,
           "list": [1, 2, 3], "nested": {"x": {"y": "z"}},
           "f": torch.arange(12, dtype=torch.float32).view(3, 4),
           "d": torch.randn(5, dtype=torch.float64), "l": torch.tensor([1, -2, 3]),
           "u8": torch.tensor([0, 255], dtype=torch.uint8), "empty": torch.empty(0)}
    back = t7.loads(t7.dumps(obj))
    assert back["n"] == 3.25 and back["i"] == 7 and back["b"] is True and back["nil"] is None
    assert t7.table_to_list(back["list"]) == [1, 2, 3] and back["nested"]["x"]["y"] == "z"
    for k in ("f", "d", "l", "u8"):
        assert torch.equal(back[k], obj[k]) and back[k].dtype == obj[k].dtype
    assert back["empty"].numel() == 0


def test_non_contiguous_and_shared_references():
    base = torch.arange(20, dtype=torch.float32).view(4, 5)
    tr = base.t()
    shared = {"a": base, "b": base, "t": tr}
    back = t7.loads(t7.dumps(shared))
    assert torch.equal(back["t"], tr)
    assert back["a"] is back["b"]          # reference semantics survive (memoised object index)


def test_float_tensor_layout_bytes():
    """The on-disk layout the reference's loaders rely on."""
    b = t7.dumps(torch.tensor([1.0, 2.0], dtype=torch.float32))
    import struct
    assert struct.unpack_from("<i", b, 0)[0] == 4               # torch object
    assert b[8:8 + 7] == struct.pack("<i", 3) + b"V 1"
    assert b"torch.FloatTensor" in b and b"torch.FloatStorage" in b
    assert b.endswith(struct.pack("<q", 2) + struct.pack("<ff", 1.0, 2.0))


def test_cifar_model_graph_roundtrip(tmp_path):
    theta = torch.randn(351946)
    model = t7.cifar_model_to_t7(theta)
    path = tmp_path / "model-epoch-10.net"
    t7.save(str(path), model)
    back = t7.load(str(path))
    assert back.classname == "nn.Sequential"
    mods = t7.table_to_list(back.fields["modules"])
    assert [m.classname for m in mods][:3] == ["nn.SpatialConvolution", "nn.ReLU", "nn.SpatialMaxPooling"]
    assert mods[0].fields["weight"].shape == (64, 3, 5, 5)
    assert torch.equal(t7.cifar_theta_from_t7(back), theta)


def test_flat_parameter_checkpoint_like_bicnn(tmp_path):
    """BiCNN saves/loads the flat parameter tensor (bicnn.lua:259-261,590-594)."""
    p = torch.randn(1000)
    t7.save(str(tmp_path / "x_12.5_model"), p)
    assert torch.equal(t7.load(str(tmp_path / "x_12.5_model")), p)
