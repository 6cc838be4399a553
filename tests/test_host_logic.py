"""Host-side logic that needs no GPU: integer push bookkeeping against the oracle, the PPNet drop-in's
construction / pruning / pickling, and the 'no CPU fallback' contract."""
import io
import pickle

import numpy as np
import pytest
import torch
import torch.nn as nn

from oracle import head_oracle as orc
from adlm_inherently_interpretable_segmentation_b200 import _lib, head, model, push


class MSC(nn.Module):
    """stub backbone whose repr starts with 'MSC' and whose .base holds >= 2 convs (model.py:88-93)"""

    def __init__(self, k=64):
        super().__init__()
        self.base = nn.Sequential(nn.Conv2d(3, k, 1), nn.Conv2d(k, k, 1))

    def forward(self, x):
        return self.base(x)


def make_net(N=190, K=64, C=19, **kw):
    return model.PPNet(features=MSC(K), img_size=64, prototype_shape=(N, K, 1, 1), proto_layer_rf_info=None,
                       num_classes=C, init_weights=True, prototype_activation_function='log',
                       add_on_layers_type='deeplab_simple', patch_classification=True, **kw)


@pytest.mark.parametrize("H,W,h,w", [(1024, 2048, 129, 257), (513, 513, 65, 65), (64, 96, 9, 13), (50, 70, 7, 9)])
def test_patch_luts_and_rf_boxes_match_oracle(H, W, h, w):
    r, c = push.patch_luts(H, W, h, w)
    ro, co = orc.patch_luts(H, W, h, w)
    assert np.array_equal(r, ro) and np.array_equal(c, co)
    for pi, pj in [(0, 0), (h - 1, w - 1), (h // 2, w // 3)]:
        assert push.rf_box(pi, pj, H, W, h, w) == orc.rf_box(pi, pj, H, W, h, w)


def test_unpack_packed_words():
    d = np.array([0.0, 1.5, 3.25e-7], dtype=np.float32)
    ids = np.array([0, 12345, 2 ** 32 - 1], dtype=np.int64)
    packed = (d.view(np.uint32).astype(np.int64) << 32) | ids
    packed = np.concatenate([packed, [_lib.PUSH_NONE]])
    dist, got = push.unpack(packed)
    assert np.array_equal(dist[:3], d.astype(np.float64)) and np.isinf(dist[3])
    assert np.array_equal(got[:3], ids) and got[3] == -1
    # ordering of the packed words == ordering by (distance, id)
    order = np.argsort(packed[:3])
    assert list(order) == sorted(range(3), key=lambda i: (d[i], ids[i]))


def test_dedup_matches_oracle(golden):
    g = golden("push_small")
    ui, dup = push.dedup_prototypes(g["gpatch"])
    assert np.array_equal(np.sort(ui), g["unique_index"]) and np.array_equal(dup, g["duplicate_idx"])


def test_grow_box_stays_inside_and_contains_seed():
    rng = np.random.default_rng(0)
    act = rng.random((40, 60))
    box = push.grow_high_activation_box(act, (10, 14, 20, 25), threshold=0.5)
    assert 0 <= box[0] <= 10 and 14 < box[1] <= 40 and 0 <= box[2] <= 20 and 25 < box[3] <= 60
    # nothing above threshold around the seed -> only the margin is added
    box = push.grow_high_activation_box(np.zeros((40, 60)), (10, 14, 20, 25), threshold=0.5)
    assert box == (5, 20, 15, 31)


def test_ppnet_structure_matches_reference_contract():
    net = make_net()
    assert tuple(net.prototype_shape) == (190, 64, 1, 1) and net.num_prototypes == 190 and net.num_classes == 19
    assert net.num_prototypes_per_class == 10 and net.epsilon == 1e-4
    assert isinstance(net.add_on_layers, nn.Sequential) and isinstance(net.add_on_layers[0], nn.Sigmoid)
    assert net.ones.shape == net.prototype_vectors.shape and not net.ones.requires_grad
    assert net.last_layer.bias is None and tuple(net.last_layer.weight.shape) == (19, 190)
    ident = net.prototype_class_identity.numpy()
    assert np.array_equal(ident, orc.class_identity(190, 19).astype(np.float32))
    W = net.last_layer.weight.detach().numpy()
    assert np.array_equal(W, np.where(ident.T == 1, 1.0, -0.5).astype(np.float32))
    assert set(dict(net.named_parameters())) >= {"prototype_vectors", "ones", "last_layer.weight"}
    assert np.array_equal(net.proto_class().numpy(), np.arange(190) // 10)


def test_ppnet_prune_and_pickle_roundtrip():
    net = make_net(N=30, C=3)
    net.prune_prototypes([0, 7, 29])
    assert net.num_prototypes == 27 and tuple(net.ones.shape) == (27, 64, 1, 1)
    assert tuple(net.last_layer.weight.shape) == (3, 27) and net.last_layer.in_features == 27
    assert tuple(net.prototype_class_identity.shape) == (27, 3)
    assert list(net.proto_class().numpy()[:9]) == [0] * 8 + [1]
    buf = io.BytesIO()
    torch.save(net, buf)          # the reference checkpoints whole module objects (module.py:292-297)
    buf.seek(0)
    back = torch.load(buf, weights_only=False)
    assert isinstance(back, model.PPNet) and back.num_prototypes == 27
    assert torch.equal(back.prototype_vectors, net.prototype_vectors)


def test_add_on_layer_variants():
    for kind, n_layers in (("bottleneck", 4), ("regular", 4)):
        net = model.PPNet(features=MSC(64), img_size=64, prototype_shape=(20, 32, 1, 1), proto_layer_rf_info=None,
                          num_classes=2, add_on_layers_type=kind)
        assert len(net.add_on_layers) == n_layers and isinstance(net.add_on_layers[-1], nn.Sigmoid)
    with pytest.raises(Exception):
        model.PPNet(features=nn.Sequential(nn.Linear(2, 2)), img_size=64, prototype_shape=(20, 32, 1, 1),
                    proto_layer_rf_info=None, num_classes=2)


def test_cpu_tensors_are_refused_not_emulated():
    net = make_net(N=30, C=3)
    z = torch.rand(1, 64, 5, 5)
    with pytest.raises(ValueError, match="CUDA"):
        net.forward_from_conv_features(z)
    with pytest.raises(ValueError, match="CUDA"):
        head.fused_step(z, net.prototype_vectors, net.last_layer.weight, torch.zeros(1, 5, 5), net.proto_class())
    net.prototype_activation_function = lambda d: -d
    with pytest.raises(NotImplementedError):
        net.forward_from_conv_features(z)
    with pytest.raises(ValueError):
        head._prec("bf16")


def test_non_onehot_identity_is_rejected():
    ident = torch.zeros(4, 2)
    ident[:, 0] = 1
    ident[0, 1] = 1
    with pytest.raises(NotImplementedError):
        head.proto_class_from_identity(ident)


def test_eval_source_tables_match_oracle_and_tile_extent():
    """host tables of the evaluation epilogue (torch's align_corners=False source index, float32)"""
    from adlm_inherently_interpretable_segmentation_b200 import evalops, _lib
    for in_size, out_size in [(129, 1024), (257, 2048), (65, 513), (9, 70), (13, 100), (7, 7), (5, 3)]:
        idx, lam = evalops.source_tables(in_size, out_size)
        i0, i1, l1 = orc.bilinear_tables(in_size, out_size)
        assert np.array_equal(idx, i0) and np.array_equal(lam, l1)
        assert idx.min() >= 0 and idx.max() <= in_size - 1 and (lam >= 0).all() and (lam < 1).all()
        ext = evalops.tile_patch_extent(idx, in_size, _lib.EVAL_TILE_Y)
        # brute force: rows touched by every tile (second tap = min(idx + 1, in - 1))
        worst = 0
        for a in range(0, out_size, _lib.EVAL_TILE_Y):
            rows = set()
            for y in range(a, min(a + _lib.EVAL_TILE_Y, out_size)):
                rows.update((int(idx[y]), min(int(idx[y]) + 1, in_size - 1)))
            worst = max(worst, max(rows) - min(rows) + 1)
        assert ext == worst


def test_nearest_patch_label_rule_and_pruning():
    """find_nearest.py:196-205 label rule and prune.py:31-42 on hand-made cases"""
    from adlm_inherently_interpretable_segmentation_b200 import nearest
    lab = np.zeros((8, 12), np.uint8)                     # 2 x 3 patches of 4 x 4 pixels, 0 = void
    lab[:4, :4] = 3                                       # patch (0,0): class 2 everywhere
    lab[:4, 4:8] = 1; lab[0, 4] = 5                       # patch (0,1): class 0 with one pixel of class 4
    lab[4:, 8:] = 0                                       # patch (1,2): void only
    assert nearest.patch_label(lab, 0, 0, 2, 3, target_class=7) == 2       # most common
    assert nearest.patch_label(lab, 0, 1, 2, 3, target_class=4) == 4       # any pixel of the prototype's class wins
    assert nearest.patch_label(lab, 0, 1, 2, 3, target_class=9) == 0
    assert nearest.patch_label(lab, 1, 2, 2, 3, target_class=1) == -1      # void
    labels = np.array([[0, 0, 1], [1, 1, 1], [2, 0, -1], [0, 3, 3]])
    pc = np.array([0, 1, 2, 3])
    assert nearest.prototypes_to_prune(labels, pc, 2) == orc.prototypes_to_prune(labels, pc, 2) == [2]
    assert nearest.prototypes_to_prune(labels, pc, 3) == [0, 2, 3]


@pytest.mark.parametrize("name", ["push_small", "push_mds"])
def test_activation_bound_boxes_equal_the_reference_fixture(golden, name):
    """push.activation_bound_box / grow_high_activation_box (host code of push_prototypes) reproduce the boxes the
    reference stored for every winner (push.py:309-350, helpers.py:48-82), exactly"""
    pytest.importorskip("cv2")
    from oracle import head_oracle as orc
    g = golden(name)
    n_img, K, h, w, N, C, H, Wimg, seed = [int(v) for v in g["meta"]]
    pc = np.arange(N) // (N // C)
    rf = g["rf_boxes"]
    n = 0
    for j in range(N):
        if rf[j, 0] < 0:
            assert np.all(g["bound_boxes"][j] == -1)
            continue
        idx = int(rf[j, 0])
        d = orc.l2_distances(g["z"][idx:idx + 1], g["P"])[0]
        bb = push.activation_bound_box(d[j], g["labels"][idx], int(pc[j]), tuple(int(v) for v in rf[j, 1:5]), 'log', 1e-4)
        assert (idx,) + tuple(int(v) for v in bb) == tuple(int(v) for v in g["bound_boxes"][j][:5])
        n += 1
    assert n > 0
