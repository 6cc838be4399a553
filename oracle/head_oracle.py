"""CPU oracle for the ProtoSeg prototype head: a numpy restatement of the reference algorithm.

TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this; the product package never does.

Parity status: PINNED — every function here is checked in tests/test_oracle_golden.py against
fixtures produced by running the unmodified reference (oracle/make_golden.py -> tests/golden/).

Reference lines restated:
  l2_distances           model.py:203-221   (PPNet._l2_convolution, expanded ||z-p||^2, relu)
  distance_2_similarity  model.py:231-237   ('log': log((d+1)/(d+eps)), 'linear': -d)
  head_forward           model.py:259-283   (patch branch: NCHW->(M,N), similarity, last layer)
  step_loss              segmentation/module.py:143-222 (CE on non-void, KLD on distances, masked L1)
  step_grads             autograd of the above (module.py:236), written out analytically
  patch_class_mask       segmentation/push.py:201-230 (pixel -> patch via int(i / (H/h)))
  push_search_image      segmentation/push.py:249-307 (class-filtered per-prototype argmin + RF box)
  push_finalize          segmentation/push.py:142-152 (de-dup with np.unique(axis=0))
  min_pool_forward       model.py:284-299   (legacy classification branch: global min-pool, similarity, last layer)
  min_pool_step          train_and_test.py:35-101 (cross entropy + cluster / separation costs + masked L1, backward)
"""
import numpy as np

EPS = 1e-4  # model.py:50


def l2_distances(z, P, dtype=np.float32):
    """z (B,K,h,w), P (N,K) -> d (B,N,h,w) = relu(sum z^2 - 2 z.p + sum p^2)   [model.py:207-219]"""
    z = z.astype(dtype)
    P = P.astype(dtype)
    B, K, h, w = z.shape
    zf = z.reshape(B, K, h * w)
    x2 = (zf * zf).sum(axis=1, dtype=dtype)                     # (B,hw)   conv2d(x**2, ones)
    p2 = (P * P).sum(axis=1, dtype=dtype)                       # (N,)
    xp = np.einsum('nk,bkm->bnm', P, zf).astype(dtype)          # conv2d(x, prototypes)
    inter = (-2 * xp + p2[None, :, None]).astype(dtype)
    d = np.maximum(x2[:, None, :] + inter, 0).astype(dtype)
    return d.reshape(B, P.shape[0], h, w)


def distance_2_similarity(d, mode='log', eps=EPS):
    if mode == 'log':
        return np.log((d + d.dtype.type(1)) / (d + d.dtype.type(eps)))
    if mode == 'linear':
        return -d
    raise ValueError(mode)


def head_forward(z, P, W, mode='log', dtype=np.float32):
    """-> logits (B,h,w,C), distances (B,N,h,w), activations (M,N)   [model.py:264-283]"""
    d = l2_distances(z, P, dtype)
    B, N, h, w = d.shape
    dv = np.ascontiguousarray(d.transpose(0, 2, 3, 1)).reshape(-1, N)
    act = distance_2_similarity(dv, mode)
    logits = (act @ W.astype(dtype).T).astype(dtype)
    return logits.reshape(B, h, w, -1), d, act


def class_identity(N, C):
    """prototype j belongs to class j // (N/C)   [model.py:66-71]"""
    ident = np.zeros((N, C), dtype=np.float64)
    per = N // C
    for c in range(C):
        ident[c * per:(c + 1) * per, c] = 1
    return ident


def _logsumexp(a, axis):
    m = a.max(axis=axis, keepdims=True)
    return (m + np.log(np.exp(a - m).sum(axis=axis, keepdims=True))).squeeze(axis)


def kld_groups(target_img, C):
    """The (image, label) groups module.py:169-178 loops over: labels taken BEFORE the void shift."""
    for b in range(target_img.shape[0]):
        for cls in np.unique(target_img[b]):
            if cls < 0 or cls >= C:
                continue
            yield b, int(cls), np.nonzero(target_img[b] == cls)[0]


def step_loss(z, P, W, target, ignore_void=True, weights=(1.0, 0.25, 1e-4), ident=None,
              dtype=np.float64, want_grads=True, norm=None):
    """fwd + CE + KLD + L1 (+ analytic backward).  target (B,h,w) int, already resized.

    Returns dict(loss, cross_entropy, kld_loss, l1, n_correct, n_patches, [dz, dP, dW]).
    module.py:143-222; the backward is what loss.backward() (module.py:236) yields.
    norm=(n_patches, n_kld_terms) replaces the two mean normalisers by given (whole-batch) counts: a
    batch shard evaluated that way contributes its additive share of the whole-batch loss / gradients
    (used to check the data-parallel split; not a reference code path).
    """
    w_ce, w_kld, w_l1 = weights
    z = z.astype(dtype)
    P = P.astype(dtype)
    W = W.astype(dtype)
    B, K, h, w = z.shape
    N, C = P.shape[0], W.shape[0]
    if ident is None:
        ident = class_identity(N, C)
    hw = h * w
    M = B * hw
    zf = z.reshape(B, K, hw).transpose(0, 2, 1).reshape(M, K)           # (M,K)
    pre = (zf * zf).sum(1)[:, None] - 2 * zf @ P.T + (P * P).sum(1)[None, :]
    d = np.maximum(pre, 0)                                              # (M,N)
    s = np.log((d + 1) / (d + EPS))
    logits = s @ W.T                                                    # (M,C)
    tflat = target.reshape(-1).astype(np.int64)
    target_img = target.reshape(B, hw).astype(np.int64)
    if ignore_void:
        keep = np.nonzero(tflat != 0)[0]
        tt = tflat[keep] - 1
    else:
        keep = np.arange(M)
        tt = tflat
    lg = logits[keep]
    lse = _logsumexp(lg, 1)
    n_ce = len(keep) if norm is None else norm[0]
    ce = float((lse - lg[np.arange(len(keep)), tt]).sum() / n_ce)
    n_correct = int((lg.argmax(1) == tt).sum())

    # KLD over distances (module.py:167-208)
    terms = []   # (b, protos, pixel idx, i, j, kld)
    d_img = d.reshape(B, hw, N)
    for b, cls, pix in kld_groups(target_img, C):
        protos = np.nonzero(ident[:, cls])[0]
        if len(protos) < 2 or len(pix) < 2:
            continue
        a = d_img[b][np.ix_(pix, protos)]                               # (n, J) distances
        la = a - _logsumexp(a, 0)[None, :]                              # log_softmax over pixels
        q = np.exp(la)
        for i in range(len(protos)):
            for j in range(i + 1, len(protos)):
                kld1 = (q[:, j] * (la[:, j] - la[:, i])).sum()          # kl_div(log_p1, log_p2, log_target)
                kld2 = (q[:, i] * (la[:, i] - la[:, j])).sum()
                terms.append((b, protos, pix, i, j, (kld1 + kld2) / 2.0, q, a))
    if terms:
        kvals = np.array([t[5] for t in terms])
        n_terms = len(terms) if norm is None else norm[1]
        kld_loss = float(np.exp(-kvals).sum() / n_terms)
    else:
        kld_loss = 0.0
    l1_mask = 1 - ident.T
    l1 = float(np.abs(W * l1_mask).sum())
    loss = w_ce * ce + w_kld * kld_loss + w_l1 * l1
    out = dict(loss=loss, cross_entropy=ce, kld_loss=kld_loss, l1=l1, n_correct=n_correct,
               n_patches=int(len(keep)), n_kld_terms=len(terms), logits=logits.reshape(B, h, w, C),
               dist=d.reshape(B, hw, N).transpose(0, 2, 1).reshape(B, N, h, w))
    if not want_grads:
        return out

    # ---- backward ----
    dlogits = np.zeros((M, C), dtype=dtype)
    sm = np.exp(lg - lse[:, None])
    sm[np.arange(len(keep)), tt] -= 1
    dlogits[keep] = sm * (w_ce / n_ce)
    dW = dlogits.T @ s + w_l1 * np.sign(W) * l1_mask
    Gd = (dlogits @ W) * (1 / (d + 1) - 1 / (d + EPS))
    if terms:
        T = n_terms
        Gk = np.zeros((B, hw, N), dtype=dtype)
        for (b, protos, pix, i, j, kld, q, a) in terms:
            c = w_kld * (-np.exp(-kld)) / T
            delta = a[:, j] - a[:, i]
            Ej = (q[:, j] * delta).sum()
            Ei = (q[:, i] * delta).sum()
            gj = 0.5 * (q[:, j] * (delta - Ej) + q[:, j] - q[:, i])
            gi = 0.5 * (-q[:, i] * (delta - Ei) - q[:, j] + q[:, i])
            Gk[b, pix, protos[j]] += c * gj
            Gk[b, pix, protos[i]] += c * gi
        Gd = Gd + Gk.reshape(M, N)
    Gd = Gd * (pre > 0)                                                 # relu backward
    rs = Gd.sum(1)
    dzf = 2 * (rs[:, None] * zf - Gd @ P)
    dP = 2 * (Gd.sum(0)[:, None] * P - Gd.T @ zf)
    out.update(dz=dzf.reshape(B, hw, K).transpose(0, 2, 1).reshape(B, K, h, w), dP=dP, dW=dW,
               dlogits=dlogits, Gd=Gd)
    return out


# ----------------------------------------------------------------------------- push
def patch_luts(H, Wimg, h, w):
    """pixel -> patch index with the reference's float expression int(pixel / (H / h))  [push.py:201-202,217,219]"""
    ph = H / h
    pw = Wimg / w
    rows = np.array([int(i / ph) for i in range(H)], dtype=np.int32)
    cols = np.array([int(j / pw) for j in range(Wimg)], dtype=np.int32)
    return rows, cols


def patch_class_mask(labels, h, w):
    """labels (H,W) ints, 0 = void -> (h,w) uint32 bitmask: bit c set iff >=1 pixel of label c+1
    falls in the patch footprint   [push.py:214-223]"""
    H, Wimg = labels.shape
    rows, cols = patch_luts(H, Wimg, h, w)
    mask = np.zeros((h, w), dtype=np.uint32)
    lab = labels.astype(np.int64)
    bits = np.where(lab > 0, np.left_shift(np.uint32(1), np.clip(lab - 1, 0, 31).astype(np.uint32)),
                    np.uint32(0)).astype(np.uint32)
    np.bitwise_or.at(mask, (rows[:, None].repeat(Wimg, 1), cols[None, :].repeat(H, 0)), bits)
    return mask


def rf_box(patch_i, patch_j, H, Wimg, h, w):
    """[push.py:288-292]"""
    ph = H / h
    pw = Wimg / w
    return (int(patch_i * ph), int(patch_i * ph + ph) + 1, int(patch_j * pw), int(patch_j * pw + pw) + 1)


def push_search_image(z_img, P, mask, proto_class, dist=None):
    """One image: z (K,h,w), mask (h,w) uint32 -> per prototype (min distance, flat argmin or -1).
    Ties resolve to the LOWEST flat patch index (the reference's order is set-hash order,
    push.py:230,259-262 — SURVEY §7.4-6; exact in-image ties are logged by the tests)."""
    K, h, w = z_img.shape
    if dist is None:
        dist = l2_distances(z_img[None], P)[0]
    d = dist.reshape(P.shape[0], h * w)
    mflat = mask.reshape(-1)
    N = P.shape[0]
    dmin = np.full(N, np.inf, dtype=np.float64)
    amin = np.full(N, -1, dtype=np.int64)
    for j in range(N):
        cand = np.nonzero((mflat >> np.uint32(proto_class[j])) & np.uint32(1))[0]
        if len(cand) == 0:
            continue
        k = np.argmin(d[j, cand])
        dmin[j] = d[j, cand[k]]
        amin[j] = cand[k]
    return dmin, amin


def push_search(z_imgs, P, labels, C):
    """Whole dataset: strict '<' keeps the earliest image   [push.py:55-61,101-133,265-280]"""
    n_img, K, h, w = z_imgs.shape
    N = P.shape[0]
    proto_class = (np.arange(N) // (N // C)).astype(np.int64)
    gmin = np.full(N, np.inf)
    gpatch = np.zeros((N, K))
    gidx = np.full(N, -1, dtype=np.int64)          # img * h*w + patch
    rf = np.full((N, 6), -1, dtype=np.int64)
    H, Wimg = labels.shape[1:]
    for i in range(n_img):
        mask = patch_class_mask(labels[i], h, w)
        dmin, amin = push_search_image(z_imgs[i], P, mask, proto_class)
        for j in range(N):
            if amin[j] >= 0 and dmin[j] < gmin[j]:
                gmin[j] = dmin[j]
                gidx[j] = i * h * w + amin[j]
                pi, pj = divmod(int(amin[j]), w)
                gpatch[j] = z_imgs[i][:, pi, pj]
                rf[j] = (i,) + rf_box(pi, pj, H, Wimg, h, w) + (proto_class[j],)
    return gmin, gpatch, gidx, rf


def push_finalize(gpatch):
    """np.unique(axis=0) de-dup   [push.py:147-149] -> (sorted unique_index, duplicate_idx)"""
    _, unique_index = np.unique(gpatch, axis=0, return_index=True)
    dup = [i for i in range(gpatch.shape[0]) if i not in unique_index]
    return np.sort(unique_index), np.array(dup, dtype=np.int64)


# ---------------------------------------------------------------- evaluation epilogue
def bilinear_tables(in_size, out_size):
    """torch upsample_bilinear2d, align_corners=False (area_pixel_compute_source_index), float32:
    src = max(scale * (dst + 0.5) - 0.5, 0), scale = in / out  ->  (i0, i1, lambda1)"""
    scale = np.float32(in_size) / np.float32(out_size)
    dst = np.arange(out_size, dtype=np.float32)
    src = np.maximum(scale * (dst + np.float32(0.5)) - np.float32(0.5), np.float32(0)).astype(np.float32)
    i0 = np.minimum(src.astype(np.int64), in_size - 1)
    i1 = np.minimum(i0 + 1, in_size - 1)
    return i0, i1, (src - i0.astype(np.float32)).astype(np.float32)


def bilinear_upsample(planes, size):
    """planes (P,h,w) float32 -> (P,H,W): F.interpolate(mode='bilinear', align_corners=False)
    (eval_valid.py:171-172): h0 * (w0 * v00 + w1 * v01) + h1 * (w0 * v10 + w1 * v11)"""
    planes = np.asarray(planes, dtype=np.float32)
    _, h, w = planes.shape
    y0, y1, ly = bilinear_tables(h, size[0])
    x0, x1, lx = bilinear_tables(w, size[1])
    h1 = ly[None, :, None]
    h0 = (np.float32(1) - ly)[None, :, None]
    w1 = lx[None, None, :]
    w0 = (np.float32(1) - lx)[None, None, :]
    top = w0 * planes[:, y0][:, :, x0] + w1 * planes[:, y0][:, :, x1]
    bot = w0 * planes[:, y1][:, :, x0] + w1 * planes[:, y1][:, :, x1]
    return (h0 * top + h1 * bot).astype(np.float32)


def eval_image(logits_hwc, dist_nhw, ann, proto_class, C):
    """one image of eval_valid.py:165-199: upsample, argmax / argmin, counters.
    logits (h,w,C), dist (N,h,w) | None, ann (H,W) uint8 (0 = void) -> dict"""
    size = ann.shape
    up = bilinear_upsample(np.transpose(logits_hwc, (2, 0, 1)), size)
    pred = np.argmax(up, axis=0)                                       # eval_valid.py:176
    out = dict(pred=pred, up_logits=up)
    out["correct_pixels"] = int(np.sum(((pred + 1) == ann) & (ann != 0)))   # :178
    out["total_pixels"] = int(np.sum(ann != 0))                            # :180
    I, U = np.zeros(C, np.int64), np.zeros(C, np.int64)
    for c in range(C):                                                  # :182-189
        pr, gt = pred == c, ann == c + 1
        I[c] = np.sum(pr & gt)
        U[c] = np.sum((pr | gt) & (ann != 0))
    out["CLS_I"], out["CLS_U"] = I, U
    if dist_nhw is not None:
        upd = bilinear_upsample(dist_nhw, size)
        nearest = np.argmin(upd, axis=0)                                # :174
        out["nearest"], out["up_dist"] = nearest, upd
        counts = np.zeros(len(proto_class), np.int64)
        ok = np.asarray(proto_class)[nearest] == pred                   # :192-197
        np.add.at(counts, nearest[ok], 1)
        out["proto_counts"] = counts
    return out


# ---------------------------------------------------------------- k nearest patches / pruning
def pil_nearest_lut(in_size, out_size):
    """source index of every output pixel of PIL's Image.resize(..., NEAREST) (dataset.py:23-31 resize_label):
    Pillow's ImagingScaleAffine accumulates the coordinate in double, xo = a/2; xo += a, and truncates"""
    a = in_size / out_size
    xo = a * 0.5
    lut = np.empty(out_size, np.int64)
    for x in range(out_size):
        lut[x] = min(int(xo), in_size - 1)
        xo += a
    return lut


def find_k_nearest(z_imgs, P, labels, proto_class, k, dist=None):
    """find_nearest.py:92-226,331 (full_save label rule): z_imgs (n,K,h,w), labels (n,H,W) with 0 = void.
    -> labels_all_prototype (N,k) class ids (-1 = void) of the k nearest patches (ascending distance),
       distances (N,k), image ids (N,k), patch indices (N,k) (row-major in the h x w map)"""
    n, K, h, w = z_imgs.shape
    N = P.shape[0]
    H, W = labels.shape[1:]
    rl, cl = pil_nearest_lut(H, h), pil_nearest_lut(W, w)
    ph, pw = H / h, W / w                                       # find_nearest.py:122-123
    heaps = [[] for _ in range(N)]                              # entries (distance, image, patch, label) kept sorted
    for i in range(n):
        d = l2_distances(z_imgs[i:i + 1], P)[0] if dist is None else dist[i]     # (N,h,w)
        y = labels[i].astype(np.int64) - 1                      # :118-119
        void = (y[rl][:, cl] == -1)                             # :128-132 resize_label + void test
        d = d + 10e6 * void[None]
        for j in range(N):
            flat = int(np.argmin(d[j]))                         # :146
            dj = float(d[j].reshape(-1)[flat])
            pi, pj = divmod(flat, w)
            r0, r1, c0, c1 = int(pi * ph), int((pi + 1) * ph), int(pj * pw), int((pj + 1) * pw)   # :163-168
            lab = y[r0:r1, c0:c1]
            if lab.size == 0:                                   # :178-179
                continue
            tc = int(proto_class[j])
            if np.any(lab == tc):                               # :199-205
                label = tc
            else:
                vals, cnts = np.unique(lab, return_counts=True)
                label = int(vals[np.argmax(cnts)])
            hp = heaps[j]
            if len(hp) < k:                                     # :217-223 (heap of the k smallest distances;
                hp.append((dj, i, flat, label))                 #  a tie never displaces an earlier image)
            else:
                worst = max(range(k), key=lambda t: (hp[t][0], -t))
                if dj < hp[worst][0]:
                    hp[worst] = (dj, i, flat, label)
    out_l = np.full((N, k), -2, np.int64)
    out_d = np.full((N, k), np.inf)
    out_i = np.full((N, k), -1, np.int64)
    out_p = np.full((N, k), -1, np.int64)
    for j in range(N):
        for t, (dj, i, flat, label) in enumerate(sorted(heaps[j], key=lambda e: (e[0], e[1]))):
            out_l[j, t], out_d[j, t], out_i[j, t], out_p[j, t] = label, dj, i, flat
    return out_l, out_d, out_i, out_p


def prototypes_to_prune(nearest_labels, proto_class, prune_threshold):
    """prune.py:31-42: prototypes whose k nearest patches hold fewer than prune_threshold of their own class"""
    return [j for j in range(len(proto_class))
            if int(np.sum(np.asarray(nearest_labels[j]) == int(proto_class[j]))) < prune_threshold]


def min_pool_forward(z, P, W, mode='log', dtype=np.float32):
    """legacy classification branch   [model.py:284-299]
    -> logits (B,C), min_distances (B,N), argmin (B,N) flat patch index (first minimum, as max_pool2d)"""
    d = l2_distances(z, P, dtype)
    B, N, h, w = d.shape
    df = d.reshape(B, N, h * w)
    amin = df.argmin(axis=2)
    mind = np.take_along_axis(df, amin[:, :, None], axis=2)[:, :, 0]
    act = distance_2_similarity(mind, mode)
    logits = (act @ W.astype(dtype).T).astype(dtype)
    return logits, mind, amin


def min_pool_step(z, P, W, label, coefs=(1.0, 0.8, -0.08, 1e-4), eps=EPS):
    """one class-specific training batch of train_and_test.py:35-101 (use_l1_mask=True, default coefficients
    `cross_entropy + 0.8 cluster - 0.08 separation + 1e-4 l1`, :92) in float64, with the autograd backward
    (:101) written out.  label (B,) ints.  -> dict(loss terms, logits, min_distances, dz, dP, dW)"""
    z64, P64, W64 = z.astype(np.float64), P.astype(np.float64), W.astype(np.float64)
    B, K, h, w = z.shape
    N, C = P.shape[0], W.shape[0]
    zf = z64.reshape(B, K, h * w)
    pre = (zf * zf).sum(1)[:, None, :] - 2 * np.einsum('nk,bkm->bnm', P64, zf) + (P64 * P64).sum(1)[None, :, None]
    df = np.maximum(pre, 0)
    amin = df.argmin(axis=2)                                                 # (B,N)
    mind = np.take_along_axis(df, amin[:, :, None], axis=2)[:, :, 0]
    act = np.log((mind + 1) / (mind + eps))
    logits = act @ W64.T
    lse = _logsumexp(logits, axis=1)
    logp = logits - lse[:, None]
    ce = -logp[np.arange(B), label].mean()                                   # :38
    ident = class_identity(N, C)
    correct = ident[:, label].T                                              # (B,N)  :46
    wrong = 1 - correct
    max_dist = float(K)                                                      # :41-43, 1x1 prototypes
    inv_c = (max_dist - mind) * correct
    inv_w = (max_dist - mind) * wrong
    nc, nw = inv_c.argmax(axis=1), inv_w.argmax(axis=1)
    cluster = (max_dist - inv_c[np.arange(B), nc]).mean()                    # :47-48
    separation = (max_dist - inv_w[np.arange(B), nw]).mean()                 # :51-54
    l1_mask = 1 - ident.T                                                    # :62
    l1 = np.abs(W64 * l1_mask).sum()
    w_ce, w_cl, w_sep, w_l1 = coefs
    loss = w_ce * ce + w_cl * cluster + w_sep * separation + w_l1 * l1
    # backward
    dlogits = np.exp(logp)
    dlogits[np.arange(B), label] -= 1
    dlogits *= w_ce / B
    dW = dlogits.T @ act + w_l1 * np.sign(W64) * l1_mask
    dmind = (dlogits @ W64) * (1 / (mind + 1) - 1 / (mind + eps))
    dmind[np.arange(B), nc] += w_cl / B * correct[np.arange(B), nc]          # d(max_dist - (max_dist - m)) = dm
    dmind[np.arange(B), nw] += w_sep / B * wrong[np.arange(B), nw]
    gate = np.take_along_axis(pre, amin[:, :, None], axis=2)[:, :, 0] > 0    # relu, model.py:218
    g = np.where(gate, dmind, 0.0)
    dz = np.zeros_like(zf)
    dP = np.zeros_like(P64)
    for b in range(B):
        for n in range(N):
            diff = zf[b, :, amin[b, n]] - P64[n]
            dz[b, :, amin[b, n]] += 2 * g[b, n] * diff
            dP[n] -= 2 * g[b, n] * diff
    return dict(loss=loss, cross_entropy=ce, cluster=cluster, separation=separation, l1=l1, logits=logits,
                min_distances=mind, argmin=amin, dmind=dmind, dz=dz.reshape(z.shape), dP=dP, dW=dW)
