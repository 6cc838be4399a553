"""GPU parity of the fused evaluation epilogue (adlm_eval_upsample: bilinear upsampling + argmax / argmin + counters,
segmentation/eval_valid.py:165-199) against the torch-generated fixture, the oracle and torch itself at image size."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import head_oracle as orc
from adlm_inherently_interpretable_segmentation_b200 import evalops, synth

pytestmark = pytest.mark.gpu


def _near_tie_only(up, got, want, axis_max):
    """positions where got != want must be near-ties of the upsampled stack (|v[got] - v[want]| tiny)"""
    bad = np.argwhere(got != want)
    for idx in bad:
        col = up[(slice(None),) + tuple(idx)]
        assert abs(float(col[got[tuple(idx)]]) - float(col[want[tuple(idx)]])) <= 1e-5 * max(1.0, float(np.abs(col).max()))
    return len(bad)


def test_eval_upsample_matches_fixture(golden):
    g = golden("eval_small")
    B, h, w, C = g["logits"].shape
    N = g["dist"].shape[1]
    H, W = g["ann"].shape[1:]
    res = evalops.upsample_argmax(torch.from_numpy(g["logits"]).cuda(), (H, W),
                                  distances=torch.from_numpy(g["dist"]).cuda(),
                                  labels=torch.from_numpy(g["ann"]).cuda(),
                                  proto_class=torch.from_numpy(g["proto_class"]).cuda())
    pred, near = res["pred"].cpu().numpy().astype(np.int64), res["nearest"].cpu().numpy().astype(np.int64)
    flips = 0
    for i in range(B):
        flips += _near_tie_only(g["up_logits"][i], pred[i], g["pred"][i].astype(np.int64), True)
        flips += _near_tie_only(g["up_dist"][i], near[i], g["nearest"][i].astype(np.int64), False)
    assert flips <= 2
    if flips == 0:                                    # counters are a function of (pred, nearest, ann)
        c = evalops.counter_views(res["counters"].cpu().numpy(), C, N)
        assert int(c["correct_pixels"]) == int(g["correct_pixels"]) and int(c["total_pixels"]) == int(g["total_pixels"])
        assert np.array_equal(c["CLS_I"], g["CLS_I"]) and np.array_equal(c["CLS_U"], g["CLS_U"])
        assert np.array_equal(c["proto_counts"], g["proto_counts"])


def test_eval_counters_follow_their_definition_and_accumulate():
    """whatever the argmax decides on near-ties, the counters must be consistent with the returned planes"""
    B, h, w, C, N, H, W = 3, 17, 23, 19, 190, 130, 181
    rng = np.random.default_rng(5)
    logits = torch.from_numpy(rng.normal(size=(B, h, w, C)).astype(np.float32)).cuda()
    dist = torch.from_numpy(rng.uniform(0, 9, size=(B, N, h, w)).astype(np.float32)).cuda()
    ann = synth.block_labels(B, H, W, C, block=11, seed=6).astype(np.uint8)
    pc = (np.arange(N) // (N // C)).astype(np.int32)
    res = evalops.upsample_argmax(logits, (H, W), distances=dist, labels=torch.from_numpy(ann).cuda(),
                                  proto_class=torch.from_numpy(pc).cuda())
    pred, near = res["pred"].cpu().numpy().astype(np.int64), res["nearest"].cpu().numpy().astype(np.int64)
    c = evalops.counter_views(res["counters"].cpu().numpy(), C, N)
    assert int(c["correct_pixels"]) == int(np.sum(((pred + 1) == ann) & (ann != 0)))
    assert int(c["total_pixels"]) == int(np.sum(ann != 0))
    for k in range(C):
        pr, gt = pred == k, ann == k + 1
        assert int(c["CLS_I"][k]) == int(np.sum(pr & gt)) and int(c["CLS_U"][k]) == int(np.sum((pr | gt) & (ann != 0)))
    counts = np.zeros(N, np.int64)
    ok = pc[near] == pred
    np.add.at(counts, near[ok], 1)
    assert np.array_equal(c["proto_counts"], counts)
    # a second call with the same counter tensor accumulates
    res2 = evalops.upsample_argmax(logits, (H, W), distances=dist, labels=torch.from_numpy(ann).cuda(),
                                   proto_class=torch.from_numpy(pc).cuda(), counters=res["counters"])
    assert int(res2["counters"][1]) == 2 * int(c["total_pixels"])
    # oracle on one image
    o = orc.eval_image(logits[0].cpu().numpy(), dist[0].cpu().numpy(), ann[0], pc, C)
    assert float(np.mean(o["pred"] != pred[0])) <= 1e-4 and float(np.mean(o["nearest"] != near[0])) <= 1e-4


def test_eval_upsample_at_cityscapes_size_against_torch():
    """129 x 257 maps -> 1024 x 2048 (configs[2] / cfg4 eval shape): same planes as torch's own interpolate + argmax
    except near-ties; logits-only mode (eval_test.py:105-106) returns no nearest plane"""
    B, h, w, C, N, H, W = 1, 129, 257, 19, 190, 1024, 2048
    g = torch.Generator(device="cuda").manual_seed(11)
    logits = torch.randn(B, h, w, C, device="cuda", generator=g)
    dist = torch.rand(B, N, h, w, device="cuda", generator=g) * 10
    res = evalops.upsample_argmax(logits, (H, W), distances=dist)
    up_l = F.interpolate(logits.permute(0, 3, 1, 2), size=(H, W), mode="bilinear", align_corners=False)
    up_d = F.interpolate(dist, size=(H, W), mode="bilinear", align_corners=False)
    want_p, want_n = torch.argmax(up_l, 1), torch.argmin(up_d, 1)
    dp = res["pred"].long() != want_p
    dn = res["nearest"].long() != want_n
    assert float(dp.float().mean()) <= 1e-4 and float(dn.float().mean()) <= 1e-4
    # every disagreement is a near-tie of the upsampled stack
    if int(dp.sum()):
        a = up_l.gather(1, res["pred"].long()[:, None])[:, 0][dp]
        b = up_l.gather(1, want_p[:, None])[:, 0][dp]
        assert float((a - b).abs().max()) <= 1e-5
    if int(dn.sum()):
        a = up_d.gather(1, res["nearest"].long()[:, None])[:, 0][dn]
        b = up_d.gather(1, want_n[:, None])[:, 0][dn]
        assert float((a - b).abs().max()) <= 1e-4
    only = evalops.upsample_argmax(logits, (H, W))
    assert only["nearest"] is None and only["counters"] is None and torch.equal(only["pred"], res["pred"])
    with pytest.raises(ValueError):
        evalops.upsample_argmax(logits.cpu(), (H, W))


@pytest.mark.parametrize("hw,HW", [((17, 23), (40, 61)), ((17, 23), (70, 92)), ((9, 13), (33, 53)), ((12, 16), (12, 16))])
def test_eval_upsample_both_kernels_against_torch(hw, HW):
    """x factors below 4 take the per-pixel kernel (torch's association), factors >= 4 the strip kernel; both must
    agree with torch's interpolate + argmax / argmin except at near-ties, for widths that are no multiple of 4"""
    (h, w), (H, W) = hw, HW
    B, C, N = 2, 19, 190
    g = torch.Generator(device="cuda").manual_seed(h * 100 + W)
    logits = torch.randn(B, h, w, C, device="cuda", generator=g)
    dist = torch.rand(B, N, h, w, device="cuda", generator=g) * 10
    res = evalops.upsample_argmax(logits, (H, W), distances=dist)
    up_l = F.interpolate(logits.permute(0, 3, 1, 2), size=(H, W), mode="bilinear", align_corners=False)
    up_d = F.interpolate(dist, size=(H, W), mode="bilinear", align_corners=False)
    got_p, got_n = res["pred"].long(), res["nearest"].long()
    a = up_l.gather(1, got_p[:, None])[:, 0]
    assert float((a - up_l.max(1).values).abs().max()) <= 1e-5
    b = up_d.gather(1, got_n[:, None])[:, 0]
    assert float((b - up_d.min(1).values).abs().max()) <= 1e-4
    assert float((got_p != torch.argmax(up_l, 1)).float().mean()) <= 2e-3
    assert float((got_n != torch.argmin(up_d, 1)).float().mean()) <= 2e-3
