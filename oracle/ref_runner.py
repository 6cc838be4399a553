"""Runs the reference's OWN training step on host cores, from the copy under oracle/_ref/ (oracle/build_ref.py).

TEST / BENCH INFRASTRUCTURE ONLY — imported by bench.py's cpu_baseline leg and `bench.py --impl reference`, never by the
product path.  Nothing here restates reference arithmetic: `step()` calls the unmodified
`segmentation.module.PatchClassificationModule._step('train', batch)` (module.py:119-261: forward through
`model.PPNet`, CE on non-void patches, the KLD loops, masked L1, `manual_backward`) on synthetic features, with the
backbone replaced by the given feature tensor (the backbone is out of scope, BASELINE.json north_star) and
Lightning / gin / matplotlib replaced by the import shims in oracle/ref_shims/.
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available():
    return os.path.exists(os.path.join(REF_DIR, "MANIFEST.json"))


_mods = {}


def _import_reference():
    if _mods:
        return _mods
    if not available():
        raise RuntimeError("oracle/_ref is not built (python oracle/build_ref.py in the build container)")
    for k in ("DATA_PATH", "TRAIN_DIR", "TEST_DIR", "TRAIN_PUSH_DIR"):
        os.environ.setdefault(k, "/tmp/adlm_ref_data")
    os.environ.setdefault("LOG_DIR", "/tmp/adlm_ref_log")
    os.makedirs(os.environ["LOG_DIR"], exist_ok=True)
    sys.path.insert(0, os.path.join(HERE, "ref_shims"))
    sys.path.insert(0, REF_DIR)
    import model as ref_model                       # oracle/_ref/model.py
    import segmentation.module as ref_module        # oracle/_ref/segmentation/module.py
    _mods.update(model=ref_model, module=ref_module)
    return _mods


def make_step(z, P, W, labels, weights=(1.0, 0.25, 1e-4), ignore_void=True):
    """z (B,K,h,w) float32 numpy features, P (N,K), W (C,N), labels (B,H,W) integer image-resolution labels
    (0 = void).  -> (step, info): step() runs one reference `_step` (fresh gradients each call) and returns the loss."""
    import numpy as np
    import torch
    import torch.nn as nn
    from collections import defaultdict
    m = _import_reference()
    ref_model, ref_module = m["model"], m["module"]
    B, K, h, w = z.shape
    N, C = P.shape[0], W.shape[0]

    class _StubBase(nn.Module):
        def __init__(self, k):
            super().__init__()
            self.c0 = nn.Conv2d(3, k, 1)
            self.c1 = nn.Conv2d(k, k, 1)

        def freeze_bn(self):                        # module.py:127 calls it on the backbone every training step
            pass

    class MSC(nn.Module):                           # repr starts with 'MSC', >= 2 Conv2d in .base (model.py:88-93)
        def __init__(self, k):
            super().__init__()
            self.base = _StubBase(k)

    net = ref_model.PPNet(features=MSC(K), img_size=64, prototype_shape=(N, K, 1, 1), proto_layer_rf_info=None,
                          num_classes=C, init_weights=True, prototype_activation_function='log',
                          add_on_layers_type='deeplab_simple', patch_classification=True)
    with torch.no_grad():
        net.prototype_vectors.copy_(torch.from_numpy(P).reshape(N, K, 1, 1))
        net.last_layer.weight.copy_(torch.from_numpy(W))
    zt = torch.from_numpy(z).requires_grad_(True)
    net.conv_features = lambda x: zt
    mod = ref_module.PatchClassificationModule.__new__(ref_module.PatchClassificationModule)
    nn.Module.__init__(mod)
    mod.ppnet = net
    mod.iter_steps = 0
    mod.iter_size = 1
    mod.start_step = None
    mod.metrics = {k: ref_module.reset_metrics() for k in ['train', 'val']}
    mod.ignore_void_class = ignore_void
    mod.loss_weight_crs_ent, mod.loss_weight_kld, mod.loss_weight_l1 = weights
    mod.lr_scheduler = None
    mod.batch_metrics = defaultdict(list)
    captured = {}

    class _FakeOpt:
        param_groups = [{'lr': 0.0}]

        def zero_grad(self):
            pass

        def step(self):
            pass

    class _FakeTrainer:
        global_step = 0

    def manual_backward(loss):
        captured['loss'] = float(loss.detach())
        loss.backward()
    object.__setattr__(mod, 'optimizers', lambda: _FakeOpt())
    object.__setattr__(mod, 'trainer', _FakeTrainer())
    object.__setattr__(mod, 'log', lambda *a, **k: None)
    object.__setattr__(mod, 'manual_backward', manual_backward)
    type(mod).device = property(lambda self: torch.device('cpu'))
    image = torch.zeros(B, 3, 8, 8)
    lab = torch.from_numpy(np.ascontiguousarray(labels).astype(np.int64))

    def step():
        zt.grad = None
        net.prototype_vectors.grad = None
        net.last_layer.weight.grad = None
        mod._step('train', (image, lab))
        return captured['loss']

    info = dict(pixels=B * h * w, grads=lambda: (zt.grad, net.prototype_vectors.grad, net.last_layer.weight.grad))
    return step, info
