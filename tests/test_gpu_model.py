"""The PPNet drop-in on the GPU: reference API surface (model.py:39-386) with the fused head underneath,
autograd connectivity, list-in/list-out, pruning, and push_prototypes end to end on a synthetic dataset."""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

from oracle import head_oracle as orc
import parity
from adlm_inherently_interpretable_segmentation_b200 import model, push, synth
from test_host_logic import MSC, make_net

pytestmark = pytest.mark.gpu


class Backbone(nn.Module):
    """tiny stand-in for MSC(DeepLabV2): stride-8 features with K channels"""

    def __init__(self, k=64):
        super().__init__()
        self.base = nn.Sequential(nn.Conv2d(3, k, 3, padding=1), nn.AvgPool2d(8, ceil_mode=True), nn.Conv2d(k, k, 1))

    def forward(self, x):
        return self.base(x)

    def __repr__(self):
        return "MSC(stub)"


def test_forward_api_and_autograd(golden):
    g = golden("head_fwd_small_B")
    N, K = g["P"].shape
    C = g["W"].shape[0]
    net = make_net(N, K, C).cuda()
    with torch.no_grad():
        net.prototype_vectors.copy_(torch.from_numpy(g["P"]).reshape(N, K, 1, 1))
        net.last_layer.weight.copy_(torch.from_numpy(g["W"]))
    z = torch.from_numpy(g["z"]).cuda().requires_grad_(True)
    logits, dist = net.forward_from_conv_features(z)
    assert tuple(logits.shape) == g["logits"].shape and tuple(dist.shape) == g["dist"].shape
    assert logits.is_contiguous() and dist.is_contiguous() and logits.dtype == torch.float32
    assert np.abs(dist.detach().cpu().numpy() - g["dist"]).max() <= 4 * K * 2.0 ** -21
    logits2, act = net.forward_from_conv_features(z, return_activations=True)
    assert tuple(act.shape) == g["act"].shape and torch.equal(logits2, logits)
    # autograd reaches features, prototypes and the last layer through both outputs
    rng = np.random.default_rng(0)
    dl = (rng.standard_normal(g["logits"].shape) / 100).astype(np.float32)
    dd = (rng.standard_normal(g["dist"].shape) / 100).astype(np.float32)
    ((logits * torch.from_numpy(dl).cuda()).sum() + (dist * torch.from_numpy(dd).cuda()).sum()).backward()
    ref = parity.torch_head_reference(g["z"], g["P"], g["W"], dl, dd)
    assert parity.rel_err(z.grad.cpu().numpy(), ref["dz"]) <= 2e-5
    assert parity.rel_err(net.prototype_vectors.grad.cpu().numpy().reshape(N, K), ref["dP"]) <= 2e-5
    assert parity.rel_err(net.last_layer.weight.grad.cpu().numpy(), ref["dW"]) <= 2e-5
    assert net.ones.grad is None
    # list in -> list out (multi-scale training, model.py:243-244,260-261)
    with torch.no_grad():
        outs = net.forward_from_conv_features([z, z[:, :, :5, :7].contiguous()])
    assert isinstance(outs, list) and len(outs) == 2 and tuple(outs[1][0].shape) == (z.shape[0], 5, 7, C)
    # the other entry points
    with torch.no_grad():
        d2 = net._l2_convolution(z)
        lg, am = net.predict(z)
    assert torch.equal(d2, dist.detach()) and torch.equal(lg, logits.detach())
    assert torch.equal(am.long(), logits.detach().argmax(-1))
    net.materialize_distances = False
    with torch.no_grad():
        lg3, none = net.forward_from_conv_features(z)
    assert none is None and torch.equal(lg3, logits.detach())


def test_full_forward_through_backbone_and_prune():
    torch.manual_seed(0)
    net = model.PPNet(features=Backbone(16), img_size=64, prototype_shape=(30, 16, 1, 1), proto_layer_rf_info=None,
                      num_classes=3, add_on_layers_type='deeplab_simple', patch_classification=True).cuda()
    x = torch.randn(2, 3, 64, 96, device="cuda")
    logits, dist = net(x)
    assert tuple(logits.shape) == (2, 8, 12, 3) and tuple(dist.shape) == (2, 30, 8, 12)
    feats, d2 = net.push_forward(x)
    assert tuple(feats.shape) == (2, 16, 8, 12) and torch.allclose(d2, dist)
    lg, dd, cf = net.forward_with_features(x)
    assert torch.allclose(cf, feats)
    (logits.sum() + dist.mean()).backward()
    assert net.features.base[0].weight.grad is not None and net.prototype_vectors.grad is not None
    net.prune_prototypes([3, 4, 25])
    logits2, dist2 = net(x)
    assert tuple(dist2.shape) == (2, 27, 8, 12)
    keep = [i for i in range(30) if i not in (3, 4, 25)]
    assert torch.allclose(dist2, dist[:, keep].detach(), atol=1e-6)
    # legacy min-pool classification branch (model.py:284-299)
    net.patch_classification = False
    lg, mind = net(x)
    assert tuple(lg.shape) == (2, 3) and tuple(mind.shape) == (2, 27)
    assert torch.allclose(mind, dist2.flatten(2).min(dim=2).values)


def test_fused_step_method_trains_the_head(golden):
    g = golden("step_small_A")
    N, K = g["P"].shape
    C = g["W"].shape[0]
    net = make_net(N, K, C).cuda()
    with torch.no_grad():
        net.prototype_vectors.copy_(torch.from_numpy(g["P"]).reshape(N, K, 1, 1))
        net.last_layer.weight.copy_(torch.from_numpy(g["W"]))
    z = torch.from_numpy(g["z"]).cuda().requires_grad_(True)
    loss, scalars = net.fused_step(z, torch.from_numpy(g["target"]).cuda(), ignore_void_class=True,
                                   loss_weights=tuple(float(v) for v in g["weights"]))
    loss.backward()
    assert float(loss) == pytest.approx(float(g["loss"]), rel=2e-5)
    assert parity.rel_err(z.grad.cpu().numpy(), g["dz"]) <= 2e-4
    assert parity.rel_err(net.prototype_vectors.grad.cpu().numpy().reshape(N, K), g["dP"]) <= 2e-4
    assert parity.rel_err(net.last_layer.weight.grad.cpu().numpy(), g["dW"]) <= 2e-4


def test_fused_step_from_backbone_fuses_the_sigmoid_add_on():
    """features(x) -> [Sigmoid add-on + head step] in the kernels == add_on_layers(features(x)) -> head step"""
    N, K, C, B, h, w = 190, 64, 19, 2, 17, 21
    net = make_net(N, K, C).cuda()
    net.head_precision = 'tf32'
    assert net.add_on_is_sigmoid()
    target = torch.from_numpy(synth.block_labels(B, h, w, C, block=4, seed=77).astype(np.int32)).cuda()
    feats = (torch.randn(B, K, h, w, device="cuda") * 1.5)
    grads = []
    losses = []
    for fused in (True, False):
        f = feats.clone().requires_grad_(True)
        net.zero_grad()
        if fused:
            loss, scalars = net.fused_step_from_backbone(f, target)
        else:
            loss, scalars = net.fused_step(net.add_on_layers(f), target)
        (loss / 5).backward()
        losses.append(float(loss))
        grads.append((f.grad.clone(), net.prototype_vectors.grad.clone(), net.last_layer.weight.grad.clone()))
    assert losses[0] == pytest.approx(losses[1], rel=1e-4)
    for a, b in zip(grads[0], grads[1]):
        assert parity.rel_err(a.cpu().numpy(), b.cpu().numpy()) <= 1e-3
    # any other add-on stack or precision runs the add-on layers as stock torch first
    net.head_precision = 'tf32x3'
    f = feats.clone().requires_grad_(True)
    loss, _ = net.fused_step_from_backbone(f, target)
    assert float(loss) == pytest.approx(losses[1], rel=1e-3)


class _DiskDataset:
    """the slice of PatchClassificationDataset that push_prototypes touches (push.py:101-115)"""
    image_margin_size = 0
    convert_targets = None

    def __init__(self, root, imgs, labels):
        from PIL import Image
        self.img_dir = os.path.join(root, "img")
        self.annotations_dir = os.path.join(root, "ann")
        os.makedirs(self.img_dir)
        os.makedirs(self.annotations_dir)
        self.img_ids = [f"s{i:03d}" for i in range(len(imgs))]
        for i, (im, lab) in enumerate(zip(imgs, labels)):
            Image.fromarray(im).save(os.path.join(self.img_dir, self.img_ids[i] + ".png"))
            np.save(os.path.join(self.annotations_dir, self.img_ids[i] + ".npy"), lab)

    def __len__(self):
        return len(self.img_ids)

    def get_img_path(self, img_id):
        return os.path.join(self.img_dir, img_id + ".png")


@pytest.mark.parametrize("head_precision", [None, "tf32"])
def test_push_prototypes_end_to_end(tmp_path, head_precision):
    """head_precision 'tf32' (the pre-push training mode INTEGRATION.md recommends) must not leak into the search: the
    winners are the fp32 arg-min either way (push.py:259-265)"""
    torch.manual_seed(1)
    rng = np.random.default_rng(5)
    n_img, H, W, K, N, C = 3, 64, 96, 16, 30, 3
    imgs = rng.integers(0, 255, size=(n_img, H, W, 3), dtype=np.uint8)
    labels = synth.block_labels(n_img, H, W, C, block=8, seed=9)
    labels[labels == 3] = 0                              # class 2 never appears -> zero vectors -> de-dup
    ds = _DiskDataset(str(tmp_path), imgs, labels)
    net = model.PPNet(features=Backbone(K), img_size=64, prototype_shape=(N, K, 1, 1), proto_layer_rf_info=None,
                      num_classes=C, add_on_layers_type='deeplab_simple', patch_classification=True).cuda()
    net.head_precision = head_precision
    P0 = net.prototype_vectors.detach().cpu().numpy().reshape(N, K).copy()
    # expected result from the oracle on the same features
    feats = []
    with torch.no_grad():
        for i in range(n_img):
            x = push._to_normalized_tensor(imgs[i]).unsqueeze(0).cuda()
            feats.append(net.conv_features(x)[0].cpu().numpy())
    feats = np.stack(feats)
    og, opatch, oidx, orf = orc.push_search(feats, P0, labels, C)
    out_dir = os.path.join(str(tmp_path), "protos")
    logs = []
    push.push_prototypes(ds, net, root_dir_for_saving_prototypes=out_dir, epoch_number=7,
                         proto_bound_boxes_filename_prefix="bb", log=logs.append)
    ui, dup = orc.push_finalize(opatch)
    # 10 zero vectors (absent class) collapse to 1; same-class prototypes that picked the same patch collapse too
    assert net.num_prototypes == len(ui) == N - len(dup) and len(dup) >= 9
    got = net.prototype_vectors.detach().cpu().numpy().reshape(-1, K)
    assert np.allclose(got, opatch[ui].astype(np.float32), atol=0)
    assert json.load(open(os.path.join(out_dir, "unique_prototypes.json"))) == [int(i) for i in ui]
    rf = np.load(os.path.join(out_dir, "epoch-7", "bb-receptive_field7.npy"))
    assert rf.shape == (N, 6) and np.array_equal(rf, orf)
    bb = np.load(os.path.join(out_dir, "epoch-7", "bb7.npy"))
    has = orf[:, 0] >= 0
    assert np.all(bb[has, 0] == orf[has, 0]) and np.all(bb[~has] == -1)
    assert np.all(bb[has, 1] <= orf[has, 1]) and np.all(bb[has, 2] >= orf[has, 2])
    assert any("push time" in str(l) for l in logs)
    # the pushed prototypes now sit on feature pixels: distance ~0 at the winner patch
    net.head_precision = None
    with torch.no_grad():
        d = net._l2_convolution(torch.from_numpy(feats).cuda())
    dmin = d.flatten(2).min(dim=2).values.min(dim=0).values.cpu().numpy()
    pushed = np.abs(got).sum(1) > 0
    assert pushed.sum() == len(ui) - 1 and dmin[pushed].max() <= 16 * 2.0 ** -19


def test_find_nearest_and_prune_drop_ins(tmp_path):
    """find_nearest.find_k_nearest_patches_to_prototypes (find_nearest.py:66-73) and prune.prune_prototypes
    (prune.py:11-20) with the reference's own signatures, as segmentation/run_pruning.py:60-69 calls them: a dataset on
    disk, a DataParallel-style wrapper exposing `.module`"""
    from adlm_inherently_interpretable_segmentation_b200 import nearest
    torch.manual_seed(3)
    rng = np.random.default_rng(11)
    n_img, H, W, K, N, C, k = 7, 64, 96, 16, 30, 3, 3
    imgs = rng.integers(0, 255, size=(n_img, H, W, 3), dtype=np.uint8)
    labels = synth.block_labels(n_img, H, W, C, block=8, seed=13)
    ds = _DiskDataset(str(tmp_path), imgs, labels)
    net = model.PPNet(features=Backbone(K), img_size=64, prototype_shape=(N, K, 1, 1), proto_layer_rf_info=None,
                      num_classes=C, add_on_layers_type='deeplab_simple', patch_classification=True).cuda()

    class Wrapper(nn.Module):
        def __init__(self, m):
            super().__init__()
            self.module = m
    wrapped = Wrapper(net)
    with torch.no_grad():
        feats = np.stack([net.conv_features(push._to_normalized_tensor(imgs[i]).unsqueeze(0).cuda())[0].cpu().numpy()
                          for i in range(n_img)])
    P0 = net.prototype_vectors.detach().cpu().numpy().reshape(N, K)
    pc = np.arange(N) // (N // C)
    ol, od, oi, op = orc.find_k_nearest(feats, P0, labels, pc, k)
    logs = []
    got = nearest.find_k_nearest_patches_to_prototypes(ds, wrapped, k=k, preprocess_input_function=None,
                                                       full_save=False, log=logs.append, chunk=3)
    assert got.shape == (N, k) and any("find nearest patches time" in str(l) for l in logs)
    # a different image may win only on a distance tie inside the fp32 tolerance
    same = np.all(got == ol, axis=1)
    assert same.mean() >= 0.9
    out_dir = os.path.join(str(tmp_path), "run")
    os.makedirs(out_dir)
    expect_prune = orc.prototypes_to_prune(got, pc, 2)
    info = nearest.prune_prototypes(dataset=ds, prototype_network_parallel=wrapped, k=k, prune_threshold=2,
                                    preprocess_input_function=None, original_model_dir=out_dir, epoch_number=0,
                                    log=logs.append, copy_prototype_imgs=False)
    assert [int(v) for v in info[:, 0]] == expect_prune and np.array_equal(info[:, 1], pc[expect_prune])
    assert net.num_prototypes == N - len(expect_prune)
    keep = json.load(open(os.path.join(out_dir, "prototypes_to_keep.json")))
    assert keep == sorted(set(range(N)) - set(expect_prune))
    assert np.array_equal(np.load(os.path.join(out_dir, "pruned_prototypes_epoch0_k3_pt2", "prune_info.npy")), info)
    assert np.array_equal(np.load(os.path.join(out_dir, "img", "full_class_id.npy")), got)
    assert np.array_equal(np.load(os.path.join(out_dir, "img", "5", "class_id.npy")), got[5])
