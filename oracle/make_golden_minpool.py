#!/usr/bin/env python
"""Golden vectors for the legacy min-pool classification branch: runs the UNMODIFIED reference
model.PPNet with patch_classification=False (model.py:284-299) and the UNMODIFIED train_and_test._train_or_test
(train_and_test.py:7-125: cross entropy + cluster / separation costs + masked L1, loss.backward()) for one batch on
seeded inputs, with the backbone replaced by stored feature maps, and writes tests/golden/minpool_small_{A,B}.npz:
inputs, logits, min_distances, the logged loss terms and the gradients w.r.t. the features, the prototypes and the
last layer.  TEST INFRASTRUCTURE ONLY; runs in the build container.

Usage: python oracle/make_golden_minpool.py
"""
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
sys.argv = [sys.argv[0]]
import importlib.util
spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "make_golden.py"))
mg = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mg)                      # sets up shims / sys.path, imports the reference model

import numpy as np  # noqa: E402
import torch  # noqa: E402

import train_and_test as ref_tnt  # noqa: E402  (reference train_and_test.py)

synth = mg.synth
ref_model = mg.ref_model


class _Opt:
    def zero_grad(self):
        pass

    def step(self):
        pass


def gen(name, regime, seed):
    B, K, h, w, N, C = 6, 64, 11, 13, 40, 4
    z, P, W = synth.head_inputs(B, K, h, w, N, C, regime=regime, w_noise=0.1, seed=seed)
    label = np.random.default_rng(seed + 5).integers(0, C, size=B).astype(np.int64)
    net = ref_model.PPNet(features=mg.MSC(K), img_size=64, prototype_shape=(N, K, 1, 1),
                          proto_layer_rf_info=None, num_classes=C, init_weights=True,
                          prototype_activation_function='log', add_on_layers_type='deeplab_simple',
                          patch_classification=False)
    with torch.no_grad():
        net.prototype_vectors.copy_(torch.from_numpy(P).reshape(N, K, 1, 1))
        net.last_layer.weight.copy_(torch.from_numpy(W))
    zt = torch.from_numpy(z).requires_grad_(True)
    net.conv_features = lambda x: zt
    with torch.no_grad():
        logits, mind = net(torch.zeros(B, 3, 8, 8))
        _, _, dist = net.forward_from_conv_features(zt, return_distances=True)
    argmin = dist.reshape(B, N, h * w).argmin(dim=2)

    torch.Tensor.cuda = lambda self, *a, **kw: self
    lines = []
    par_call = type("Par", (), {"module": net, "__call__": lambda self, x: net(x)})()
    acc = ref_tnt._train_or_test(par_call, [(torch.zeros(B, 3, 8, 8), torch.from_numpy(label))], optimizer=_Opt(),
                                 class_specific=True, use_l1_mask=True, coefs=None, log=lines.append)
    terms = {}
    for ln in lines:
        key, _, val = ln.strip().partition(":")
        terms[key.strip()] = float(val.strip().rstrip("%"))
    path = os.path.join(mg.OUT, name + ".npz")
    np.savez_compressed(
        path, z=z, P=P, W=W, label=label, logits=logits.numpy(), min_distances=mind.numpy(),
        argmin=argmin.numpy().astype(np.int64),
        cross_entropy=np.float64(terms["cross ent"]), cluster=np.float64(terms["cluster"]),
        separation=np.float64(terms["separation"]), l1=np.float64(terms["l1"]), accuracy=np.float64(acc),
        dz=zt.grad.numpy(), dP=net.prototype_vectors.grad.numpy().reshape(N, K), dW=net.last_layer.weight.grad.numpy(),
        meta=np.array([B, K, h, w, N, C, seed], dtype=np.int64))
    print("wrote", path, terms, "acc", acc, "nonzero dz", int((zt.grad != 0).sum()))


def main():
    # regime A: fresh prototypes (well conditioned); regime B: half the prototypes sit on pixels, so many minima are
    # exactly 0 where the reference's own fp32 cancellation noise dominates log((d+1)/(d+eps))
    gen("minpool_small_A", "A", synth.BASE_SEED + 800)
    gen("minpool_small_B", "B", synth.BASE_SEED + 810)


if __name__ == "__main__":
    main()
