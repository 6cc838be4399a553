"""GPU parity of the per-image scan (adlm_push_scan_images) and of the legacy min-pool classification branch
(adlm_min_pool_fwd / adlm_min_pool_bwd, model.py:284-299) against fixtures from the unmodified reference
(oracle/make_golden_minpool.py) and the oracle."""
from ctypes import c_int, c_longlong

import numpy as np
import pytest
import torch

from oracle import head_oracle as orc
from adlm_inherently_interpretable_segmentation_b200 import _lib, head, push, synth
from test_gpu_head import available, cuda

pytestmark = pytest.mark.gpu
PRECS = ["fp32", "tf32x3"]


def class_specific_loss(logits, mind, label, ident, W, K):
    """train_and_test.py:38-63,92 on the device (a (B,N)-sized tail of stock torch ops, as in the reference)"""
    ce = torch.nn.functional.cross_entropy(logits, label)
    correct = ident[:, label].t()
    cluster = (K - ((K - mind) * correct).max(dim=1).values).mean()
    separation = (K - ((K - mind) * (1 - correct)).max(dim=1).values).mean()
    l1 = (W * (1 - ident.t())).norm(p=1)
    return ce + 0.8 * cluster - 0.08 * separation + 1e-4 * l1, ce, cluster, separation


@pytest.mark.parametrize("prec", PRECS)
@pytest.mark.parametrize("name", ["minpool_small_A", "minpool_small_B"])
def test_min_pool_forward_matches_reference_fixture(golden, name, prec):
    if not available(prec):
        pytest.skip(f"{prec} kernels not built")
    g = golden(name)
    B, K, h, w, N, C, seed = [int(v) for v in g["meta"]]
    z, P = cuda(g["z"], g["P"])
    mind, amin, packed = head.min_pool_forward(z, P, precision=prec)
    tol = 6 * K * 2.0 ** -21
    assert np.abs(mind.cpu().numpy() - g["min_distances"]).max() <= tol
    got = amin.cpu().numpy().astype(np.int64)
    diff = got != g["argmin"]
    # another patch is acceptable only as a near-tie of the reference's own distances
    d_ref = orc.l2_distances(g["z"], g["P"]).reshape(B, N, h * w)
    for b, n in zip(*np.nonzero(diff)):
        assert d_ref[b, n, got[b, n]] - g["min_distances"][b, n] <= 2 * tol, (b, n)
    assert diff.mean() <= 0.02


@pytest.mark.parametrize("prec", PRECS)
def test_min_pool_model_step_matches_reference_fixture(golden, prec):
    """PPNet(patch_classification=False) forward + the reference's class-specific loss + backward"""
    if not available(prec):
        pytest.skip(f"{prec} kernels not built")
    from adlm_inherently_interpretable_segmentation_b200 import model as m
    from test_host_logic import MSC
    g = golden("minpool_small_A")
    B, K, h, w, N, C, seed = [int(v) for v in g["meta"]]
    net = m.PPNet(features=MSC(K), img_size=64, prototype_shape=(N, K, 1, 1), proto_layer_rf_info=None,
                  num_classes=C, init_weights=True, prototype_activation_function='log',
                  add_on_layers_type='deeplab_simple', patch_classification=False).cuda()
    net.head_precision = prec
    with torch.no_grad():
        net.prototype_vectors.copy_(torch.from_numpy(g["P"]).reshape(N, K, 1, 1))
        net.last_layer.weight.copy_(torch.from_numpy(g["W"]))
    z = torch.from_numpy(g["z"]).cuda().requires_grad_(True)
    logits, mind = net.forward_from_conv_features(z)
    assert logits.shape == (B, C) and mind.shape == (B, N)
    assert np.allclose(logits.detach().cpu().numpy(), g["logits"], rtol=1e-3, atol=1e-3)
    label = torch.from_numpy(g["label"]).cuda()
    ident = net.prototype_class_identity.cuda()
    loss, ce, cl, sep = class_specific_loss(logits, mind, label, ident, net.last_layer.weight, K)
    loss.backward()
    assert abs(ce.item() - float(g["cross_entropy"])) <= 1e-4 * max(1.0, float(g["cross_entropy"]))
    assert abs(cl.item() - float(g["cluster"])) <= 1e-4 and abs(sep.item() - float(g["separation"])) <= 1e-4
    ref = orc.min_pool_step(g["z"], g["P"], g["W"], g["label"])
    for got, key in ((z.grad, "dz"), (net.prototype_vectors.grad.reshape(N, K), "dP"), (net.last_layer.weight.grad, "dW")):
        a = got.cpu().numpy()
        scale = np.abs(g[key]).max()
        assert np.abs(a - g[key]).max() <= 2e-4 * scale, key
        assert np.abs(a - ref[key]).max() <= 2e-4 * scale, key
    # the gradient touches only the arg-min patches
    nz = (z.grad.reshape(B, K, h * w).abs().sum(1) != 0).sum().item()
    assert nz <= B * N
    # return_distances=True materialises the map and yields the same minima
    with torch.no_grad():
        l2, m2, dist = net.forward_from_conv_features(z.detach(), return_distances=True)
    assert dist.shape == (B, N, h, w)
    assert torch.allclose(m2, mind.detach(), atol=6 * K * 2.0 ** -21)


def test_min_pool_backward_relu_gate_and_accumulation():
    """several prototypes share one arg-min patch (atomics accumulate); a prototype equal to a pixel has d = 0
    and passes no gradient"""
    B, K, h, w, N = 3, 64, 9, 11, 24
    zn, Pn, _ = synth.head_inputs(B, K, h, w, N, 4, seed=411)
    Pn[5] = zn[1, :, 4, 7]
    Pn[6] = Pn[7]                                   # two identical prototypes -> same arg-min everywhere
    z, P = cuda(zn, Pn)
    mind, amin, packed = head.min_pool_forward(z, P, precision="fp32")
    assert mind[1, 5].item() == 0.0 and amin[1, 5].item() == 4 * w + 7
    assert torch.equal(amin[:, 6], amin[:, 7])
    gmin = torch.randn(B, N, device="cuda")
    dz = torch.zeros_like(z)
    dP = torch.zeros_like(P)
    lib = _lib.load()
    rc = lib.adlm_min_pool_bwd(_lib.ptr(z), B, K, h * w, _lib.ptr(P), N, _lib.ptr(packed), _lib.ptr(gmin),
                               _lib.ptr(dz), _lib.ptr(dP), _lib.stream_ptr())
    _lib.check(rc, "adlm_min_pool_bwd")
    zf = zn.reshape(B, K, h * w).astype(np.float64)
    am = amin.cpu().numpy()
    md = mind.cpu().numpy()
    gn = gmin.cpu().numpy().astype(np.float64)
    edz, edP = np.zeros_like(zf), np.zeros((N, K))
    for b in range(B):
        for n in range(N):
            if md[b, n] > 0:
                diff = zf[b, :, am[b, n]] - Pn[n]
                edz[b, :, am[b, n]] += 2 * gn[b, n] * diff
                edP[n] -= 2 * gn[b, n] * diff
    assert np.abs(dz.cpu().numpy().reshape(B, K, h * w) - edz).max() <= 1e-5 * np.abs(edz).max()
    assert np.abs(dP.cpu().numpy() - edP).max() <= 1e-5 * np.abs(edP).max()


@pytest.mark.parametrize("prec", PRECS + ["tf32"])
@pytest.mark.parametrize("n,h,w,N", [(5, 9, 13, 190), (40, 17, 19, 190), (300, 12, 11, 30), (3, 129, 257, 190)])
def test_per_image_scan_equals_image_by_image_launches(prec, n, h, w, N):
    """adlm_push_scan_images (one launch, tiles that never straddle images, contiguous tile ranges per CTA) must give
    bit for bit what n single-image adlm_push_scan launches give — with a class mask and without any mask"""
    if not available(prec):
        pytest.skip(f"{prec} kernels not built")
    K, C = 64, 19 if N == 190 else 3
    zn, Pn, _ = synth.head_inputs(n, K, h, w, N, C, seed=500 + n)
    z, P = cuda(zn, Pn)
    pc = torch.arange(N, dtype=torch.int32, device="cuda") // (N // C)
    rng = np.random.default_rng(n)
    mask = torch.from_numpy(rng.integers(0, 2 ** C, size=(n, h * w), dtype=np.int64).astype(np.int32)).cuda()
    mask[n // 2, : (h * w) // 2] = 0
    ones = torch.full((n, h * w), -1, dtype=torch.int32, device="cuda")
    zero_class = torch.zeros(N, dtype=torch.int32, device="cuda")
    lib = _lib.load()
    p = _lib.PREC[prec]
    for mk, cls, mk_ref, cls_ref in ((mask, pc, mask, pc), (None, None, ones, zero_class)):
        one = torch.full((n, N), _lib.PUSH_NONE, dtype=torch.int64, device="cuda")
        rc = lib.adlm_push_scan_images(_lib.ptr(z), n, K, h * w, _lib.ptr(mk), _lib.ptr(P), N, _lib.ptr(cls),
                                       c_int(p), _lib.ptr(one), _lib.stream_ptr())
        _lib.check(rc, "adlm_push_scan_images")
        ref = torch.full((n, N), _lib.PUSH_NONE, dtype=torch.int64, device="cuda")
        for i in range(n):
            rc = lib.adlm_push_scan(_lib.ptr(z[i]), 1, K, h * w, _lib.ptr(mk_ref[i]), _lib.ptr(P), N,
                                    _lib.ptr(cls_ref), c_longlong(0), c_int(p), _lib.ptr(ref[i]), _lib.stream_ptr())
            _lib.check(rc, "adlm_push_scan")
        # distances must agree bit for bit.  The patch may differ only in a rounding tie: the running minimum inside
        # a CTA compares t = |z|^2 - 2 z.p (before |p|^2 is added), CTAs merge on the rounded d = t + |p|^2, so two
        # patches whose d round to the same float resolve by visiting order in one case and by patch id in the other
        assert torch.equal(one >> 32, ref >> 32), (prec, n, h, w, N, mk is None)
        differ = (one != ref)
        assert differ.float().mean().item() <= 1e-3, (prec, n, h, w, N, mk is None, int(differ.sum()))
        if differ.any():
            print(f"per-image scan: {int(differ.sum())} rounding ties of {one.numel()} resolved to another patch "
                  f"({prec}, n={n}, {h}x{w}, N={N})")
    assert (one != _lib.PUSH_NONE).all()


def test_min_pool_large_shapes_take_the_fp32_scan():
    """N > 224 or K > 64 (ProtoPNet's 2000 x 128 prototypes, main.py-style models) run on the CUDA-core scan"""
    B, K, h, w, N = 2, 128, 7, 7, 400
    zn, Pn, _ = synth.head_inputs(B, K, h, w, N, 4, seed=77)
    z, P = cuda(zn, Pn)
    mind, amin, _ = head.min_pool_forward(z, P)
    d = orc.l2_distances(zn, Pn).reshape(B, N, h * w)
    assert np.abs(mind.cpu().numpy() - d.min(axis=2)).max() <= 6 * K * 2.0 ** -21
    assert (amin.cpu().numpy() == d.argmin(axis=2)).mean() >= 0.99
