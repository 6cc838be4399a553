#!/usr/bin/env python
"""Golden vectors at the FULL sizes BASELINE.json's configs quote, from the UNMODIFIED reference (/root/reference).

TEST INFRASTRUCTURE ONLY (build container; the reference does not exist on the GPU box).  Inputs regenerate from the
seeds stored in `meta` (package synth.py, numpy PCG64), so the fixtures hold outputs only — and only as much of them as
stays small: full dP / dW / scalars, strided samples plus per-plane float64 sums of the big tensors.

  step_cfg2      configs[1]: the reference's own PatchClassificationModule._step (module.py:119-261) on batch 16 of
                 64x65x65 features, 190 prototypes x 64-d, 19 classes, labels 513x513 resized by the reference's
                 resize_label (PIL NEAREST)
  head_fwd_cfg4  configs[3]: model.PPNet.forward_from_conv_features (model.py:259-283) on 64 slices of 64x65x65,
                 30 prototypes (3 classes x 10), per-pixel argmax as eval_valid.py:176 takes it
  head_fwd_cfg5  configs[4] on a 129x257 crop: 1000 prototypes x 256-d, 100 classes

Usage:  python oracle/make_golden_full.py
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

import make_golden as mg  # noqa: E402  (sets up the shims, imports the reference)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402

synth = mg.synth
ref_module = mg.ref_module
PX_STRIDE = 29          # sample every 29th pixel of the flattened h*w map


def gen_step_full(name, B, K, h, w, N, C, H, Wimg, seed, block):
    from collections import defaultdict
    z, P, W = synth.head_inputs(B, K, h, w, N, C, regime="A", w_noise=0.1, seed=seed)
    labels = synth.block_labels(B, H, Wimg, C, block=block, seed=seed + 1)
    net = mg.make_ref_ppnet(N, K, C, W=W, P=P)
    zt = torch.from_numpy(z).requires_grad_(True)
    net.conv_features = lambda x: zt
    mod = ref_module.PatchClassificationModule.__new__(ref_module.PatchClassificationModule)
    nn.Module.__init__(mod)
    mod.ppnet = net
    mod.iter_steps = 0
    mod.iter_size = 1
    mod.start_step = None
    mod.metrics = {k: ref_module.reset_metrics() for k in ['train', 'val']}
    mod.ignore_void_class = True
    mod.loss_weight_crs_ent, mod.loss_weight_kld, mod.loss_weight_l1 = 1.0, 0.25, 1e-4
    mod.lr_scheduler = None
    mod.batch_metrics = defaultdict(list)
    captured = {}
    object.__setattr__(mod, 'optimizers', lambda: mg._FakeOpt())
    object.__setattr__(mod, 'trainer', mg._FakeTrainer())
    object.__setattr__(mod, 'log', lambda *a, **k: None)

    def manual_backward(loss):
        captured['loss'] = float(loss.detach())
        loss.backward()
    object.__setattr__(mod, 'manual_backward', manual_backward)
    type(mod).device = property(lambda self: torch.device('cpu'))
    mod._step('train', (torch.zeros(B, 3, 8, 8), torch.from_numpy(labels.astype(np.int64))))
    m = mod.metrics['train']
    target = np.stack([ref_module.resize_label(l.astype(np.float32), size=(w, h)).numpy() for l in labels])
    dz = zt.grad.numpy().reshape(B, K, h * w)
    out = dict(target=target.astype(np.uint8), loss=np.float64(captured['loss']),
               cross_entropy=np.float64(m['cross_entropy']), kld_loss=np.float64(m['kld_loss']),
               n_correct=np.int64(int(m['n_correct'])), n_patches=np.int64(int(m['n_patches'])),
               dz_sample=dz[:, :, ::PX_STRIDE].copy(), dz_plane_sums=dz.astype(np.float64).sum(axis=2),
               dz_absmax=np.float64(np.abs(dz).max()),
               dP=net.prototype_vectors.grad.numpy().reshape(N, K), dW=net.last_layer.weight.grad.numpy(),
               weights=np.array([1.0, 0.25, 1e-4]), ignore_void=np.int64(1),
               meta=np.array([B, K, h, w, N, C, seed, PX_STRIDE], dtype=np.int64))
    np.savez_compressed(os.path.join(mg.OUT, name + ".npz"), **out)
    print("wrote", name, "loss", captured['loss'], "ce", m['cross_entropy'], "kld", m['kld_loss'],
          "n_correct", int(m['n_correct']), "/", int(m['n_patches']))


def gen_fwd_full(name, B, K, h, w, N, C, regime, seed, batch=8):
    z, P, W = synth.head_inputs(B, K, h, w, N, C, regime=regime, w_noise=0.1, seed=seed)
    net = mg.make_ref_ppnet(N, K, C, W=W, P=P)
    logits, dmin, dsum = [], [], []
    with torch.no_grad():
        for i in range(0, B, batch):
            lg, dist = net.forward_from_conv_features(torch.from_numpy(z[i:i + batch]))
            logits.append(lg.numpy())
            d = dist.numpy().reshape(lg.shape[0], N, h * w)
            dmin.append(d.min(axis=2))
            dsum.append(d.astype(np.float64).sum(axis=2))
    logits = np.concatenate(logits).reshape(B, h * w, C)
    srt = np.sort(logits, axis=2)
    out = dict(argmax=logits.argmax(axis=2).astype(np.uint8), top2_gap=(srt[:, :, -1] - srt[:, :, -2]).astype(np.float32),
               logits_sample=logits[:, ::PX_STRIDE].copy(), logits_absmax=np.float64(np.abs(logits).max()),
               dist_min=np.concatenate(dmin).astype(np.float32), dist_plane_sums=np.concatenate(dsum),
               meta=np.array([B, K, h, w, N, C, seed, PX_STRIDE, 1 if regime == "B" else 0], dtype=np.int64))
    np.savez_compressed(os.path.join(mg.OUT, name + ".npz"), **out)
    print("wrote", name, {k: v.shape for k, v in out.items()})


def main():
    S = synth.BASE_SEED
    torch.manual_seed(0)
    gen_step_full("step_cfg2", 16, 64, 65, 65, 190, 19, 513, 513, S + 200, block=32)
    gen_fwd_full("head_fwd_cfg4", 64, 64, 65, 65, 30, 3, "B", S + 210)
    gen_fwd_full("head_fwd_cfg5", 1, 256, 129, 257, 1000, 100, "A", S + 220, batch=1)


if __name__ == "__main__":
    main()
