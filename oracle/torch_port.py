"""CPU port of the reference's hot path in the reference's own operator sequence (stock torch ops).

TEST / BENCH INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this; the product package never does.

The reference is Python on top of torch, so its CPU cost is the cost of THESE library calls in THIS
order: two F.conv2d (one against `ones`), three element-wise passes, permute+contiguous, log, nn.Linear,
cross_entropy, the Python triple loop of masked_select / log_softmax / kl_div, and autograd through all
of it.  The numpy oracle (head_oracle.py) restates the maths; this file restates the execution, so that
the timed CPU baseline on the GPU box (where /root/reference does not exist) costs what the reference
costs.  Pinned against the same reference-generated fixtures (tests/test_oracle_golden.py).

Reference lines followed:
  head_forward   model.py:203-221, 231-237, 259-283
  step           segmentation/module.py:143-222, 236
  push_image     segmentation/push.py:209-280 (the per-image search, dumps off)
"""
import numpy as np
import torch
import torch.nn.functional as F

EPS = 1e-4


def head_forward(z, P4, W, ones=None):
    """z (B,K,h,w), P4 (N,K,1,1), W (C,N) -> logits (B,h,w,C), distances (B,N,h,w)"""
    if ones is None:
        ones = torch.ones_like(P4)
    x2_patch_sum = F.conv2d(z ** 2, ones)                       # model.py:207-208
    p2 = torch.sum(P4 ** 2, dim=(1, 2, 3)).view(-1, 1, 1)       # model.py:210-214
    xp = F.conv2d(z, P4)                                        # model.py:216
    dist = F.relu(x2_patch_sum + (-2 * xp + p2))                # model.py:217-219
    B, N, h, w = dist.shape
    flat = dist.permute(0, 2, 3, 1).contiguous().reshape(-1, N)  # model.py:271-272
    act = torch.log((flat + 1) / (flat + EPS))                  # model.py:233
    logits = F.linear(act, W).reshape(B, h, w, -1)              # model.py:161-162,275
    return logits, dist


def step(z, P4, W, target, ident, ignore_void=True, weights=(1.0, 0.25, 1e-4), backward=True):
    """One reference training step of the head on given features.  target (B,h,w) long, ident (N,C).
    Returns dict(loss, cross_entropy, kld_loss, n_correct, n_patches) (+ grads in z.grad, P4.grad, W.grad)."""
    w_ce, w_kld, w_l1 = weights
    logits, dist = head_forward(z, P4, W)
    C = logits.shape[-1]
    output = logits.reshape(-1, C)
    target_img = target.reshape(target.shape[0], -1)
    tflat = target.flatten()
    acts = dist.permute(0, 2, 3, 1)
    acts_img = acts.reshape(acts.shape[0], -1, acts.shape[-1])
    if ignore_void:                                             # module.py:156-160
        keep = (tflat != 0).nonzero().squeeze()
        tflat = tflat[keep] - 1
        output = output[keep]
    ce = F.cross_entropy(output, tflat.long())
    klds = []                                                   # module.py:167-200
    for b in range(len(target_img)):
        for cls in torch.unique(target_img[b]).cpu().numpy():
            if cls < 0 or cls >= ident.shape[1]:
                continue
            protos = torch.nonzero(ident[:, cls]).flatten().cpu().numpy()
            if len(protos) == 0:
                continue
            mask = target_img[b] == cls
            las = [F.log_softmax(torch.masked_select(acts_img[b, :, i], mask), dim=0) for i in protos]
            if len(protos) < 2 or len(las[0]) < 2:
                continue
            for i in range(len(protos)):
                for j in range(i + 1, len(protos)):
                    k1 = F.kl_div(las[i], las[j], log_target=True, reduction='sum')
                    k2 = F.kl_div(las[j], las[i], log_target=True, reduction='sum')
                    klds.append((k1 + k2) / 2.0)
    kld = torch.exp(-torch.stack(klds)).mean() if klds else 0.0  # module.py:202-208
    is_correct = torch.argmax(output, dim=-1) == tflat
    l1 = (W * (1 - torch.t(ident))).norm(p=1)                   # module.py:213-218
    loss = w_ce * ce + w_kld * kld + w_l1 * l1
    if backward:
        loss.backward()
    return dict(loss=float(loss), cross_entropy=float(ce), kld_loss=float(kld), l1=float(l1),
                n_correct=int(is_correct.sum()), n_patches=int(output.shape[0]))


def class_identity(N, C):
    ident = torch.zeros(N, C)
    per = N // C
    for c in range(C):
        ident[c * per:(c + 1) * per, c] = 1
    return ident


def push_image(z_img, P4, labels, num_classes, gmin, gpatch, img_index=0):
    """The reference's per-image search in the reference's own style (push.py:209-280): distances to the host,
    a Python loop over every label pixel building class -> patch sets, a list comprehension + argmin per
    prototype.  z_img (1,K,h,w) tensor, labels (H,W) integer array.  Updates gmin / gpatch in place."""
    with torch.no_grad():
        _, dist = head_forward(z_img, P4, torch.zeros(num_classes, P4.shape[0]))
    feat = z_img[0].numpy()
    d = dist[0].permute(1, 2, 0).numpy()
    h, w = feat.shape[1:]
    H, Wimg = labels.shape
    ph, pw = H / h, Wimg / w
    lab = torch.LongTensor(np.asarray(labels, dtype=np.int64))
    sets = {c: set() for c in range(num_classes)}
    for i in range(H):                                          # push.py:216-223
        pi = int(i / ph)
        for j in range(Wimg):
            c = int(lab[i, j].item())
            if c > 0:
                sets[c - 1].add((pi, int(j / pw)))
    sets = {c: list(v) for c, v in sets.items()}
    N = P4.shape[0]
    per = N // num_classes
    for n in range(N):                                          # push.py:249-280
        cand = sets[n // per]
        if not cand:
            continue
        all_d = np.asarray([d[pi, pj, n] for pi, pj in cand])
        k = int(np.argmin(all_d))
        if all_d[k] < gmin[n]:
            pi, pj = cand[k]
            gmin[n] = all_d[k]
            gpatch[n] = feat[:, pi, pj]
    return gmin, gpatch
