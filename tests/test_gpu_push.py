"""GPU parity of the push search through the C ABI (adlm_patch_class_mask / adlm_push_scan /
adlm_push_gather) against fixtures from the reference's own update_prototypes_on_image and the oracle."""
import numpy as np
import pytest
import torch

from oracle import head_oracle as orc
import parity
from adlm_inherently_interpretable_segmentation_b200 import _lib, push, synth
from test_gpu_head import available, cuda

pytestmark = pytest.mark.gpu
PRECS = ["fp32", "tf32x3"]


def proto_class(N, C):
    return torch.arange(N, dtype=torch.int32, device="cuda") // (N // C)


@pytest.mark.parametrize("kernel", ["thread_per_patch", "cta_per_patch_row"])
@pytest.mark.parametrize("H,W,h,w,block", [(64, 96, 9, 13, 5), (50, 70, 7, 9, 5), (1024, 2048, 129, 257, 32),
                                           (1024, 2048, 129, 257, 3), (37, 53, 5, 7, 5), (513, 520, 65, 65, 7),
                                           (96, 64, 96, 64, 2), (40, 80, 3, 2, 9)])
def test_patch_class_mask_bit_exact(H, W, h, w, block, kernel, monkeypatch):
    """both label -> patch-mask kernels (one thread per patch for widths that are a multiple of 8; one CTA per patch row
    otherwise and with ADLM_MASK_CTA_KERNEL set) against push.py:201-230 as restated by the oracle: coarse and
    fine-grained label maps, odd sizes (no vector path), one pixel per patch, patches wider than two 8-byte words"""
    if kernel == "cta_per_patch_row":
        monkeypatch.setenv("ADLM_MASK_CTA_KERNEL", "1")
    else:
        monkeypatch.delenv("ADLM_MASK_CTA_KERNEL", raising=False)
    labels = synth.block_labels(2, H, W, 19, block=block, seed=3)
    labels[1, : H // 3] = 0
    s = push.PushSearch(torch.rand(19, 8, device="cuda"), proto_class(19, 19))
    mask = s.class_mask(torch.from_numpy(labels).cuda(), h, w).cpu().numpy().view(np.uint32)
    for i in range(2):
        assert np.array_equal(mask[i].reshape(h, w), orc.patch_class_mask(labels[i], h, w))


@pytest.mark.parametrize("prec", PRECS)
@pytest.mark.parametrize("name", ["push_small", "push_mds"])
def test_push_matches_reference_fixture(golden, name, prec):
    if not available(prec):
        pytest.skip(f"{prec} kernels not built")
    g = golden(name)
    n_img, K, h, w, N, C, H, Wimg, seed = [int(v) for v in g["meta"]]
    z, P = cuda(g["z"], g["P"])
    labels = torch.from_numpy(g["labels"]).cuda()
    s, new_protos = push.push_search_tensors(z, labels, P, proto_class(N, C), precision=prec, chunk=2)
    dmin, w_img, w_pi, w_pj = s.winners()
    fin = np.isfinite(g["gmin"])
    assert np.array_equal(np.isfinite(dmin), fin)
    assert np.allclose(dmin[fin], g["gmin"][fin], rtol=0, atol=6 * K * 2.0 ** -21)
    got = new_protos.cpu().numpy()
    same = np.all(got == g["gpatch"].astype(np.float32), axis=1)
    # a different winner is acceptable only as a near-tie: its reference distance is within tolerance of the min
    o_gmin, o_patch, o_idx, o_rf = orc.push_search(g["z"], g["P"], g["labels"], C)
    for j in np.nonzero(~same)[0]:
        d_ref_at_ours = orc.l2_distances(g["z"][w_img[j]:w_img[j] + 1], g["P"])[0, j, w_pi[j], w_pj[j]]
        assert d_ref_at_ours - g["gmin"][j] <= 8 * K * 2.0 ** -21, f"prototype {j}: not a near-tie"
    assert same.mean() >= 0.98
    assert np.all(got[~fin] == 0)                   # classes that never appear keep the zero vector
    for j in np.nonzero(same & fin)[0]:
        box = push.rf_box(int(w_pi[j]), int(w_pj[j]), H, Wimg, h, w)
        assert (int(w_img[j]),) + box == tuple(int(v) for v in g["rf_boxes"][j][:5])
    ui, dup = push.dedup_prototypes(got.astype(np.float64))
    if same.all():
        assert np.array_equal(np.sort(ui), g["unique_index"]) and np.array_equal(dup, g["duplicate_idx"])
    # activation-crop boxes (push.py:309-350) of the winners, EXACTLY as the reference stored them: similarity map
    # of the winning prototype on the winner image -> cubic upsample -> 95th percentile -> crop growth on the
    # ground-truth pixels of the prototype's class
    try:
        import cv2  # noqa: F401
    except ImportError:
        return
    from adlm_inherently_interpretable_segmentation_b200 import head
    W_last = torch.zeros(C, N, device="cuda")
    pcn = proto_class(N, C).cpu().numpy()
    n_checked = 0
    for idx in sorted(set(int(i) for i in w_img[same & fin])):
        dist_map = head.head_forward(z[idx:idx + 1], P, W_last, precision="fp32" if prec == "tf32" else prec,
                                     want_dist=True)["dist"][0].cpu().numpy()
        for j in np.nonzero((w_img == idx) & same & fin)[0]:
            box = push.rf_box(int(w_pi[j]), int(w_pj[j]), H, Wimg, h, w)
            bb = push.activation_bound_box(dist_map[j], g["labels"][idx], int(pcn[j]), box, 'log', 1e-4)
            assert (idx,) + tuple(int(v) for v in bb) == tuple(int(v) for v in g["bound_boxes"][j][:5]), j
            assert int(g["bound_boxes"][j][5]) == int(pcn[j])
            n_checked += 1
    assert n_checked == int((same & fin).sum()) > 0


@pytest.mark.parametrize("prec", PRECS)
def test_push_shard_merge_and_order_independence(prec):
    """sharding the images (the multi-GPU split), merging with an element-wise min and scanning in any
    order give the same packed words; exact ties resolve to the lowest global patch id."""
    if not available(prec):
        pytest.skip(f"{prec} kernels not built")
    n_img, K, h, w, N, C, H, W = 6, 64, 17, 19, 190, 19, 136, 152
    zn, Pn, _ = synth.head_inputs(n_img, K, h, w, N, C, seed=71)
    zn[4] = zn[1]                                   # duplicate image: every winner in it ties with image 1
    labels = synth.block_labels(n_img, H, W, C, block=8, seed=72)
    labels[4] = labels[1]
    z, P = cuda(zn, Pn)
    lab = torch.from_numpy(labels).cuda()
    pc = proto_class(N, C)
    whole = push.PushSearch(P, pc, precision=prec)
    whole.scan(z, lab, 0)
    parts = []
    for lo, hi in ((3, 6), (0, 3)):                 # reversed order on purpose
        s = push.PushSearch(P, pc, precision=prec)
        for i in range(hi - 1, lo - 1, -1):
            s.scan(z[i:i + 1], lab[i:i + 1], i)
        parts.append(s.packed)
    merged = torch.minimum(parts[0], parts[1])
    assert torch.equal(merged, whole.packed)
    _, img, _, _ = whole.winners()
    assert not np.any(img == 4)                     # strict '<' keeps the earliest image (push.py:265)
    og, _, oidx, _ = orc.push_search(zn, Pn, labels, C)
    d, ids = push.unpack(whole.packed.cpu().numpy())
    agree = ids == oidx
    assert agree.mean() >= 0.98 and np.allclose(d, og, atol=6 * K * 2.0 ** -21)


def test_push_none_and_masking():
    """no labelled pixel -> every word stays ADLM_PUSH_NONE; a class present in one patch only wins there"""
    K, h, w, N, C, H, W = 16, 4, 5, 8, 4, 32, 40
    z = torch.rand(1, K, h, w, device="cuda")
    P = torch.rand(N, K, device="cuda")
    lab = torch.zeros(1, H, W, dtype=torch.uint8, device="cuda")
    s = push.PushSearch(P, proto_class(N, C))
    s.scan(z, lab, 0)
    assert bool((s.packed == _lib.PUSH_NONE).all())
    lab[0, 17, 23] = 3                               # class 2, patch (2, 2)
    s.scan(z, lab, 5)
    d, ids = push.unpack(s.packed.cpu().numpy())
    assert list(np.nonzero(ids >= 0)[0]) == [4, 5]
    assert np.all(ids[[4, 5]] == 5 * h * w + 2 * w + 2)
    ref = ((z[0, :, 2, 2][None] - P[4:6]) ** 2).sum(1).cpu().numpy()
    assert np.allclose(d[[4, 5]], ref, atol=1e-5)


@pytest.mark.parametrize("prec", PRECS)
def test_push_full_resolution_image(prec):
    """one Cityscapes-sized sample (1024x2048 labels, 129x257 feature map) against the oracle"""
    if not available(prec):
        pytest.skip(f"{prec} kernels not built")
    K, h, w, N, C, H, W = 64, 129, 257, 190, 19, 1024, 2048
    zn, Pn, _ = synth.head_inputs(1, K, h, w, N, C, seed=81)
    labels = synth.block_labels(1, H, W, C, block=32, seed=82)
    z, P = cuda(zn, Pn)
    s, new_protos = push.push_search_tensors(z, torch.from_numpy(labels).cuda(), P, proto_class(N, C), precision=prec)
    og, opatch, oidx, _ = orc.push_search(zn, Pn, labels, C)
    d, ids = push.unpack(s.packed.cpu().numpy())
    assert np.allclose(d, og, atol=6 * K * 2.0 ** -21)
    assert (ids == oidx).mean() >= 0.98
    same = ids == oidx
    assert np.array_equal(new_protos.cpu().numpy()[same], opatch[same].astype(np.float32))


def test_push_single_pass_tf32_only_differs_in_near_ties():
    """precision='tf32' (one TF32 pass) is NOT the default for push: its winners may differ from the fp32 argmin,
    but only where the fp32 distances of the two candidates are within the TF32 error of each other."""
    n_img, K, h, w, N, C, H, W = 3, 64, 33, 41, 190, 19, 264, 328
    zn, Pn, _ = synth.head_inputs(n_img, K, h, w, N, C, seed=91)
    labels = synth.block_labels(n_img, H, W, C, block=8, seed=92)
    z, P = cuda(zn, Pn)
    lab = torch.from_numpy(labels).cuda()
    pc = proto_class(N, C)
    exact = push.PushSearch(P, pc, precision="tf32x3")
    exact.scan(z, lab, 0)
    fast = push.PushSearch(P, pc, precision="tf32")
    fast.scan(z, lab, 0)
    d3, id3 = push.unpack(exact.packed.cpu().numpy())
    d1, id1 = push.unpack(fast.packed.cpu().numpy())
    assert np.array_equal(id3 >= 0, id1 >= 0)
    found = id3 >= 0
    assert np.allclose(d1[found], d3[found], atol=0.05)
    dist = orc.l2_distances(zn, Pn)                                  # (n_img, N, h, w) fp32-class reference
    flat = dist.transpose(1, 0, 2, 3).reshape(N, -1)
    for j in np.nonzero(found & (id1 != id3))[0]:
        assert flat[j, id1[j]] - flat[j, id3[j]] <= 0.05, f"prototype {j}: not a near-tie"
    assert (id1 == id3)[found].mean() >= 0.8


@pytest.mark.parametrize("N,C", [(190, 19), (30, 3), (128, 4)])
@pytest.mark.parametrize("prec", PRECS)
def test_push_many_tiles_per_cta_equals_image_by_image(prec, N, C):
    """8 feature maps of 129 x 257 in ONE scan launch are 2073 tiles — 14 per persistent CTA, enough to wrap the
    operand ring, the TMEM double buffer and the metadata slots several times; one image per launch is at most two
    tiles per CTA (the regime checked against the oracle above).  Per-patch arithmetic does not depend on the tile a
    patch lands in, so both ways must agree bit for bit.
    N <= 128 (one prototype block, mds_new.gin's 30 prototypes): the epilogue warps without a block once ran ahead on
    the metadata barrier and let the producer overwrite a class table in use from the third tile of a CTA on."""
    if not available(prec):
        pytest.skip(f"{prec} kernels not built")
    n, K, h, w, H, W = 8, 64, 129, 257, 1024, 2048
    g = torch.Generator(device="cuda").manual_seed(17)
    z = torch.sigmoid(torch.randn(n, K, h, w, device="cuda", generator=g))
    labels = torch.from_numpy(synth.block_labels(2, H, W, C, block=32, seed=18)).cuda().repeat(n // 2, 1, 1).contiguous()
    P = torch.rand(N, K, device="cuda", generator=g)
    pc = proto_class(N, C)
    for rep in range(2):
        a, pa = push.push_search_tensors(z, labels, P, pc, precision=prec, chunk=n)
        b, pb = push.push_search_tensors(z, labels, P, pc, precision=prec, chunk=1)
        assert torch.equal(a.packed >> 32, b.packed >> 32), rep          # distances: bit for bit
        same = a.packed == b.packed                                      # patches: up to rounding ties of t + |p|^2
        assert same.float().mean().item() >= 0.99, rep
        assert torch.equal(pa[same], pb[same]), rep


# ------------------------------------------------------------------ k nearest patches / pruning
@pytest.mark.parametrize("prec", PRECS)
def test_k_nearest_matches_reference_fixture(golden, prec):
    if not available(prec):
        pytest.skip(f"{prec} kernels not built")
    from adlm_inherently_interpretable_segmentation_b200 import nearest
    g = golden("nearest_small")
    n_img, K, h, w, N, C, H, W, k, seed = [int(v) for v in g["meta"]]
    z, P = cuda(g["z"], g["P"])
    pc = proto_class(N, C)
    res = nearest.find_k_nearest_tensors(z, torch.from_numpy(g["labels"]).cuda(), P, pc, k=k, precision=prec)
    ol, od, oi, op = orc.find_k_nearest(g["z"], g["P"], g["labels"], pc.cpu().numpy(), k)
    assert np.allclose(res["distances"], od, atol=6 * K * 2.0 ** -21)
    same = (res["images"] == oi) & (res["patch_i"] * w + res["patch_j"] == op)
    assert same.mean() >= 0.97                                # order flips only between near-equal distances
    assert np.array_equal(res["labels"][same], g["nearest_labels"][same])
    if same.all():
        assert nearest.prototypes_to_prune(res["labels"], pc, 2) == orc.prototypes_to_prune(g["nearest_labels"], pc.cpu().numpy(), 2)


def test_k_nearest_void_and_short_datasets():
    from adlm_inherently_interpretable_segmentation_b200 import nearest
    n, K, h, w, N, C, H, W = 3, 64, 17, 19, 190, 19, 136, 152
    zn, Pn, _ = synth.head_inputs(n, K, h, w, N, C, seed=171)
    labels = synth.block_labels(n, H, W, C, block=8, seed=172)
    labels[2] = 0                                             # an all-void image never yields a candidate
    z, P = cuda(zn, Pn)
    pc = proto_class(N, C)
    res = nearest.find_k_nearest_tensors(z, torch.from_numpy(labels).cuda(), P, pc, k=4)
    assert (res["images"][:, :2] >= 0).all() and (res["images"][:, :2] < 2).all()
    assert (res["images"][:, 2:] == -1).all() and (res["labels"][:, 2:] == -2).all()
    assert np.all(np.diff(res["distances"][:, :2], axis=1) >= 0)
    # the winning patch of every prototype is never a void one (NEAREST-resized label)
    rl, cl = nearest.pil_nearest_lut(H, h), nearest.pil_nearest_lut(W, w)
    for j in range(0, N, 17):
        for t in range(2):
            i, pi, pj = res["images"][j, t], res["patch_i"][j, t], res["patch_j"][j, t]
            assert labels[i][rl[pi], cl[pj]] != 0
