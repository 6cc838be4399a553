#!/usr/bin/env python
"""Golden vectors for k-nearest patches / pruning: runs the UNMODIFIED reference
find_nearest.find_k_nearest_patches_to_prototypes (find_nearest.py:66-342, full_save=True as prune.py:20-28 calls it)
on a synthetic on-disk dataset with the reference PPNet (features replaced by stored maps) and writes
tests/golden/nearest_small.npz.  TEST INFRASTRUCTURE ONLY; runs in the build container.

Usage: python oracle/make_golden_nearest.py
"""
import os
import sys
import tempfile
import types

HERE = os.path.dirname(os.path.abspath(__file__))
sys.argv = [sys.argv[0]]
import importlib.util
spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "make_golden.py"))
mg = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mg)                      # sets up shims / sys.path, imports the reference model

import numpy as np  # noqa: E402
import torch  # noqa: E402
from PIL import Image  # noqa: E402

import find_nearest as ref_fn  # noqa: E402  (reference find_nearest.py)

synth = mg.synth


def main():
    n_img, K, h, w, N, C, H, Wimg, k, seed = 7, 32, 9, 13, 20, 4, 70, 100, 3, synth.BASE_SEED + 700
    zs, P, W = synth.head_inputs(n_img, K, h, w, N, C, regime="A", w_noise=0.0, seed=seed)
    labels = synth.block_labels(n_img, H, Wimg, C, block=9, seed=seed + 1)          # 0 = void
    labels[1, :, :40] = 0                                                          # a large void area
    net = mg.make_ref_ppnet(N, K, C, W=W, P=P)
    net.eval()
    torch.Tensor.cuda = lambda self, *a, **kw: self
    it = iter(range(n_img))
    net.conv_features = lambda x: torch.from_numpy(zs[next(it)][None])
    root = tempfile.mkdtemp(prefix="adlm_nearest_")
    os.makedirs(os.path.join(root, "img")); os.makedirs(os.path.join(root, "ann")); os.makedirs(os.path.join(root, "out"))
    ids = [f"s{i:02d}" for i in range(n_img)]
    for i, sid in enumerate(ids):
        Image.fromarray(np.zeros((H, Wimg, 3), np.uint8)).save(os.path.join(root, "img", sid + ".png"))
        np.save(os.path.join(root, "ann", sid + ".npy"), labels[i].astype(np.int64))
    ds = types.SimpleNamespace(img_ids=ids, image_margin_size=0, annotations_dir=os.path.join(root, "ann"),
                               get_img_path=lambda sid: os.path.join(root, "img", sid + ".png"),
                               convert_targets=lambda t: t, __len__=lambda: n_img)

    class _DS:
        img_ids = ids
        image_margin_size = 0
        annotations_dir = os.path.join(root, "ann")

        def get_img_path(self, sid):
            return os.path.join(root, "img", sid + ".png")

        def convert_targets(self, t):
            return t

        def __len__(self):
            return n_img

    par = types.SimpleNamespace(module=net, eval=lambda: None)
    with torch.no_grad():
        out = ref_fn.find_k_nearest_patches_to_prototypes(_DS(), par, k=k, full_save=True,
                                                          root_dir_for_saving_images=os.path.join(root, "out"),
                                                          log=lambda *a: None)
    out = np.asarray(out)
    path = os.path.join(mg.OUT, "nearest_small.npz")
    np.savez_compressed(path, z=zs, P=P, labels=labels.astype(np.uint8), nearest_labels=out.astype(np.int64),
                        meta=np.array([n_img, K, h, w, N, C, H, Wimg, k, seed], dtype=np.int64))
    print("wrote", path, out.shape, "label histogram", np.unique(out, return_counts=True))


if __name__ == "__main__":
    main()
