#!/usr/bin/env python
"""Golden-vector generator: runs the UNMODIFIED reference (/root/reference) on seeded
synthetic inputs and writes small fixtures to tests/golden/*.npz.

TEST INFRASTRUCTURE ONLY.  Runs in the build container (the reference does not exist on the
GPU box); the fixtures it writes are committed and are what pins the oracle (oracle/head_oracle.py)
and the CUDA path.  The reference has no tests or golden vectors of its own,
so parity is pinned by executing the reference's own code here:

  * head forward      -> model.PPNet.forward_from_conv_features        (model.py:259-283)
  * fwd + loss + bwd  -> segmentation.module.PatchClassificationModule._step (module.py:119-261)
  * push search       -> segmentation.push.update_prototypes_on_image  (push.py:162-350)
                         + the 5 closing lines of push_prototypes        (push.py:142-152)

Usage:  python oracle/make_golden.py            (writes tests/golden/)
Inputs are produced by the package's synth.py (numpy PCG64, stable across machines) and are stored in
the fixture next to the outputs, so the GPU box never needs the reference or this script.
"""
import os
import sys
import json

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("ADLM_REFERENCE", "/root/reference")

for k in ("DATA_PATH", "TRAIN_DIR", "TEST_DIR", "TRAIN_PUSH_DIR"):
    os.environ.setdefault(k, "/tmp/adlm_ref_data")
os.environ.setdefault("LOG_DIR", "/tmp/adlm_ref_log")
sys.path.insert(0, os.path.join(HERE, "ref_shims"))
sys.path.insert(0, REF)
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402
from PIL import Image  # noqa: E402

import model as ref_model  # noqa: E402  (reference model.py)
import segmentation.module as ref_module  # noqa: E402
import segmentation.push as ref_push  # noqa: E402

from adlm_inherently_interpretable_segmentation_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
torch.set_num_threads(max(1, os.cpu_count() or 1))


class _StubBase(nn.Module):
    def __init__(self, k):
        super().__init__()
        self.c0 = nn.Conv2d(3, k, 1)
        self.c1 = nn.Conv2d(k, k, 1)

    def freeze_bn(self):
        pass


class MSC(nn.Module):
    """repr starts with 'MSC' and holds >=2 Conv2d in .base, as model.py:88-93 requires."""

    def __init__(self, k):
        super().__init__()
        self.base = _StubBase(k)

    def forward(self, x):
        raise RuntimeError("stub backbone: conv_features is patched")


def make_ref_ppnet(N, K, C, W=None, P=None):
    net = ref_model.PPNet(features=MSC(K), img_size=64, prototype_shape=(N, K, 1, 1),
                          proto_layer_rf_info=None, num_classes=C, init_weights=True,
                          prototype_activation_function='log', add_on_layers_type='deeplab_simple',
                          patch_classification=True)
    with torch.no_grad():
        if P is not None:
            net.prototype_vectors.copy_(torch.from_numpy(P).reshape(N, K, 1, 1))
        if W is not None:
            net.last_layer.weight.copy_(torch.from_numpy(W))
    return net


def gen_head_fwd(name, B, K, h, w, N, C, regime, w_noise, seed, store_dist=True):
    z, P, W = synth.head_inputs(B, K, h, w, N, C, regime=regime, w_noise=w_noise, seed=seed)
    net = make_ref_ppnet(N, K, C, W=W, P=P)
    with torch.no_grad():
        logits, dist = net.forward_from_conv_features(torch.from_numpy(z))
        _, act = net.forward_from_conv_features(torch.from_numpy(z), return_activations=True)
    out = dict(z=z, P=P, W=W, logits=logits.numpy(), meta=np.array(
        [B, K, h, w, N, C, seed], dtype=np.int64))
    if store_dist:
        out["dist"] = dist.numpy()
        out["act"] = act.numpy()
    else:
        # full-size case: inputs regenerate from the seed; keep logits + a strided distance sample
        del out["z"]
        out["dist_sample"] = dist.numpy().reshape(B, N, h * w)[:, ::7, ::13].copy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print("wrote", name, {k: v.shape for k, v in out.items()})


class _FakeOpt:
    param_groups = [{'lr': 0.0}]

    def zero_grad(self):
        pass

    def step(self):
        pass


class _FakeTrainer:
    global_step = 0


def gen_step(name, B, K, h, w, N, C, H, Wimg, regime, seed, ignore_void=True,
             w_ce=1.0, w_kld=0.25, w_l1=1e-4):
    z, P, W = synth.head_inputs(B, K, h, w, N, C, regime=regime, w_noise=0.1, seed=seed)
    # (B,H,W) labels: 0..C with 0 = void when the module ignores void, else plain classes 0..C-1
    labels = synth.block_labels(B, H, Wimg, C if ignore_void else C - 1, block=8, seed=seed + 1)
    net = make_ref_ppnet(N, K, C, W=W, P=P)
    zt = torch.from_numpy(z).requires_grad_(True)
    net.conv_features = lambda x: zt          # backbone is out of scope: feed synthetic z
    mod = ref_module.PatchClassificationModule.__new__(ref_module.PatchClassificationModule)
    nn.Module.__init__(mod)
    mod.ppnet = net
    mod.iter_steps = 0
    mod.iter_size = 1
    mod.start_step = None
    mod.metrics = {k: ref_module.reset_metrics() for k in ['train', 'val']}
    mod.ignore_void_class = ignore_void
    mod.loss_weight_crs_ent, mod.loss_weight_kld, mod.loss_weight_l1 = w_ce, w_kld, w_l1
    mod.lr_scheduler = None
    from collections import defaultdict
    mod.batch_metrics = defaultdict(list)
    captured = {}
    object.__setattr__(mod, 'optimizers', lambda: _FakeOpt())
    object.__setattr__(mod, 'trainer', _FakeTrainer())
    object.__setattr__(mod, 'log', lambda *a, **k: None)

    def manual_backward(loss):
        captured['loss'] = float(loss.detach())
        loss.backward()
    object.__setattr__(mod, 'manual_backward', manual_backward)
    type(mod).device = property(lambda self: torch.device('cpu'))
    image = torch.zeros(B, 3, 8, 8)
    mod._step('train', (image, torch.from_numpy(labels.astype(np.int64))))
    m = mod.metrics['train']
    # the target the reference derived (module.py:143-146): PIL NEAREST resize to (w,h)
    target = np.stack([ref_module.resize_label(l.astype(np.float32), size=(w, h)).numpy() for l in labels])
    out = dict(z=z, P=P, W=W, labels=labels.astype(np.uint8), target=target.astype(np.int32),
               loss=np.float64(captured['loss']), cross_entropy=np.float64(m['cross_entropy']),
               kld_loss=np.float64(m['kld_loss']), n_correct=np.int64(int(m['n_correct'])),
               n_patches=np.int64(int(m['n_patches'])),
               dz=zt.grad.numpy(), dP=net.prototype_vectors.grad.numpy().reshape(N, K),
               dW=net.last_layer.weight.grad.numpy(),
               weights=np.array([w_ce, w_kld, w_l1]), ignore_void=np.int64(ignore_void),
               meta=np.array([B, K, h, w, N, C, seed], dtype=np.int64))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print("wrote", name, "loss", captured['loss'], "ce", m['cross_entropy'], "kld", m['kld_loss'],
          "n_correct", int(m['n_correct']), "/", int(m['n_patches']))


class _FakeDataset:
    convert_targets = None


def gen_push(name, n_img, K, h, w, N, C, H, Wimg, seed, absent_class=None):
    rng_seed = seed
    zs, P, W = synth.head_inputs(n_img, K, h, w, N, C, regime="A", w_noise=0.0, seed=rng_seed)
    labels = synth.block_labels(n_img, H, Wimg, C, block=8, seed=seed + 1)
    if absent_class is not None:  # a class that never appears -> its prototypes keep the zero vector
        labels[labels == absent_class + 1] = 0
    net = make_ref_ppnet(N, K, C, W=W, P=P)
    net.eval()
    torch.Tensor.cuda = lambda self, *a, **k: self           # no GPU here (push.py:187)
    gmin = np.full(N, np.inf)
    gpatch = np.zeros([N, K, 1, 1])
    rf = np.full([N, 6], -1)
    bb = np.full([N, 6], -1)
    for i in range(n_img):
        zi = torch.from_numpy(zs[i:i + 1])
        net.conv_features = lambda x, zi=zi: zi
        img = Image.fromarray(np.zeros((H, Wimg, 3), dtype=np.uint8))
        with torch.no_grad():
            ref_push.update_prototypes_on_image(_FakeDataset(), img, i, net, gmin, gpatch, rf, bb,
                                                cls2name={}, img_y=labels[i].astype(np.int64),
                                                num_classes=C)
    # push.py:142-152
    upd = np.reshape(gpatch, (N, K, 1, 1))
    _, unique_index = np.unique(upd, axis=0, return_index=True)
    dup = [i for i in range(N) if i not in unique_index]
    out = dict(z=zs, P=P, labels=labels.astype(np.uint8), gmin=gmin, gpatch=gpatch.reshape(N, K),
               rf_boxes=rf.astype(np.int64), bound_boxes=bb.astype(np.int64),
               unique_index=np.sort(unique_index).astype(np.int64), duplicate_idx=np.array(dup, dtype=np.int64),
               meta=np.array([n_img, K, h, w, N, C, H, Wimg, seed], dtype=np.int64))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print("wrote", name, "finite mins", int(np.isfinite(gmin).sum()), "dups", len(dup))


def main():
    os.makedirs(OUT, exist_ok=True)
    S = synth.BASE_SEED
    # head forward: small cases with every tensor stored
    gen_head_fwd("head_fwd_small_A", 2, 64, 13, 17, 190, 19, "A", 0.0, S)
    gen_head_fwd("head_fwd_small_B", 2, 64, 13, 17, 190, 19, "B", 0.1, S + 10)
    gen_head_fwd("head_fwd_mds", 3, 64, 9, 11, 30, 3, "B", 0.1, S + 20)          # mds_new.gin:6-7
    gen_head_fwd("head_fwd_k32", 1, 32, 5, 7, 24, 4, "A", 0.1, S + 30)           # odd sizes
    # BASELINE cfg1 at full size: 2x64x65x65, 190x64, C=19 (inputs regenerate from the seed)
    gen_head_fwd("head_fwd_cfg1", 2, 64, 65, 65, 190, 19, "A", 0.0, S + 40, store_dist=False)
    # fwd + CE + KLD + L1 + bwd through the reference's own _step
    gen_step("step_small_A", 2, 64, 13, 17, 190, 19, 104, 136, "A", S + 50)
    gen_step("step_small_B", 2, 64, 13, 17, 190, 19, 100, 133, "B", S + 60)
    gen_step("step_novoid", 2, 64, 9, 11, 30, 3, 72, 88, "A", S + 70, ignore_void=False)
    # push search
    gen_push("push_small", 3, 64, 9, 13, 190, 19, 64, 96, S + 80, absent_class=7)
    gen_push("push_mds", 2, 64, 7, 9, 30, 3, 50, 70, S + 90)
    with open(os.path.join(OUT, "PROVENANCE.json"), "w") as f:
        json.dump({"generator": "oracle/make_golden.py", "reference": REF,
                   "torch": torch.__version__, "numpy": np.__version__,
                   "note": "outputs of the unmodified reference on CPU fp32"}, f, indent=1)


if __name__ == "__main__":
    main()
