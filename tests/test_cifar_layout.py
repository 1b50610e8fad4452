"""CPU tests of the flat <-> packed parameter layout of the CIFAR CNN."""
import torch

from mpit_b200.models.cifar_cnn import (ParamLayout, TorchCifarCNN, flat_from_torch, flat_to_torch,
                                        cifar_layers)


def test_param_count_matches_reference():
    # asyncsgd/models/7-layers-cnn.lua: 4864+204928+73792+65792+2570
    assert ParamLayout().n_flat == 351946
    m = TorchCifarCNN()
    assert flat_from_torch(m).numel() == 351946


def test_pack_unpack_roundtrip():
    lay = ParamLayout()
    theta = torch.randn(lay.n_flat)
    pk = lay.pack(theta)
    assert torch.equal(lay.unpack(pk), theta)
    # padded entries stay zero (forward copy + data-gradient copy of the weights)
    assert int((pk != 0).sum()) <= 2 * lay.n_flat


def test_dgrad_operand_copy_is_transposed_weight():
    lay = ParamLayout()
    m = TorchCifarCNN()
    pk = lay.pack(flat_from_torch(m))
    L = lay.layers[2]   # conv3: Wd[c][kh][kw][co]
    wd = pk[L.d_off:L.d_off + L.C * L.k * L.k * L.CO].view(L.C, L.k, L.k, L.CO)
    ref = m.features[6].weight.detach().permute(1, 2, 3, 0)
    assert torch.equal(wd, ref)
    assert lay.layers[0].d_off == -1 and int((lay.dmap >= 0).sum()) > 0


def test_packed_conv_layout_matches_torch_weight():
    lay = ParamLayout()
    m = TorchCifarCNN()
    theta = flat_from_torch(m)
    pk = lay.pack(theta)
    L = lay.layers[1]
    w = pk[L.w_off:L.w_off + L.wsize].view(L.CO, L.k, L.KWP, L.C)
    ref = m.features[3].weight.detach().permute(0, 2, 3, 1)  # co,kh,kw,ci
    assert torch.equal(w[:, :, :L.k, :], ref)
    # conv1: pair-expanded layout, 8 four-channel slots per kernel row; kw -> slot [0,1,3,5,7]
    L = lay.layers[0]
    w = pk[L.w_off:L.w_off + L.wsize].view(L.CO, L.k, 8, 4)
    ref = m.features[0].weight.detach().permute(0, 2, 3, 1)
    assert torch.equal(w[:, :, [0, 1, 3, 5, 7], :3], ref)
    assert float(w[:, :, [2, 4, 6], :].abs().sum()) == 0.0 and float(w[..., 3].abs().sum()) == 0.0


def test_fc1_hwc_permutation():
    lay = ParamLayout()
    m = TorchCifarCNN()
    pk = lay.pack(flat_from_torch(m))
    L = lay.layers[3]
    w = pk[L.w_off:L.w_off + L.wsize].view(256, 256)
    feat_chw = torch.randn(64, 2, 2)
    y_ref = m.fc1.weight.detach() @ feat_chw.reshape(-1)
    y_pk = w @ feat_chw.permute(1, 2, 0).reshape(-1)
    assert torch.allclose(y_ref, y_pk, atol=1e-5)


def test_flat_roundtrip_through_torch_model():
    m = TorchCifarCNN()
    theta = torch.randn(351946)
    flat_to_torch(m, theta)
    assert torch.equal(flat_from_torch(m), theta)
