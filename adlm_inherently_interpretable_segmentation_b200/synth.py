"""Seeded synthetic inputs for the prototype head / push path.

numpy PCG64 only, so the same seed gives the same tensors in the build container and on the
GPU box.  Shapes follow the reference: z = add_on_layers output = Sigmoid (model.py:124-128),
prototypes = torch.rand init (model.py:54) or "post-push" copies of feature pixels
(push.py:142-145), last layer = +1 own class / -0.5 others (model.py:359-371).
"""
import numpy as np

BASE_SEED = 20220227  # the reference's train.random_seed (cityscapes_kld_imnet.gin:14)


def last_layer_pattern(N, C, incorrect=-0.5):
    per = N // C
    ident = np.zeros((N, C), dtype=np.float32)
    for c in range(C):
        ident[c * per:(c + 1) * per, c] = 1.0
    return (ident.T + incorrect * (1.0 - ident.T)).astype(np.float32)


def features(B, K, h, w, seed):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((B, K, h, w))
    return (1.0 / (1.0 + np.exp(-x))).astype(np.float32)


def head_inputs(B, K, h, w, N, C, regime="A", w_noise=0.0, seed=BASE_SEED):
    """z (B,K,h,w), P (N,K), W (C,N); regime A = fresh init, B = half the prototypes sit on pixels."""
    z = features(B, K, h, w, seed)
    rng = np.random.default_rng(seed + 1000003)
    P = rng.random((N, K)).astype(np.float32)
    if regime == "B":
        zf = z.transpose(0, 2, 3, 1).reshape(-1, K)
        idx = rng.integers(0, zf.shape[0], size=N)
        take = rng.random(N) < 0.5
        P[take] = zf[idx[take]]
    W = last_layer_pattern(N, C)
    if w_noise:
        W = (W + w_noise * rng.standard_normal(W.shape)).astype(np.float32)
    return z, P, W


def block_labels(B, H, W, C, block=32, void_frac=0.05, seed=BASE_SEED):
    """(B,H,W) uint8 labels in 0..C, 0 = void (push.py:222, module.py:156-160), block-constant."""
    rng = np.random.default_rng(seed + 7)
    hb, wb = -(-H // block), -(-W // block)
    lab = rng.integers(1, C + 1, size=(B, hb, wb))
    lab[rng.random((B, hb, wb)) < void_frac] = 0
    full = np.repeat(np.repeat(lab, block, axis=1), block, axis=2)[:, :H, :W]
    return np.ascontiguousarray(full).astype(np.uint8)
