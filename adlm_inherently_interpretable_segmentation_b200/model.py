"""Drop-in for the reference's `model.PPNet` whose prototype head runs on libadlm_head.so.

Same constructor, attributes, methods and return conventions as the reference class (model.py:39-386):
`forward`, `forward_from_conv_features`, `forward_with_features`, `prototype_distances`, `push_forward`,
`conv_features`, `_l2_convolution`, `distance_2_similarity`, `run_last_layer`, `prune_prototypes`,
`set_last_layer_incorrect_connection`; logits come back as (B,h,w,C), distances as (B,N,h,w), both
autograd-connected to the features, the prototypes and the last layer.  What changes is underneath:
one fused kernel instead of two 1x1 convolutions + ~10 elementwise passes + a transpose copy + an
SGEMM, and a backward that recomputes distances on chip instead of keeping five M x N activations.

`PrototypeHeadMixin` carries only the compute methods, so the reference's own `model.py` can keep its
class (and its pickled checkpoints their `model.PPNet` path) and inherit the fast head — see
INTEGRATION.md.  `PPNet` below is the standalone class with the reference's constructor.
"""
from typing import Optional

import torch
import torch.nn as nn

from . import head as _head

try:  # the reference is gin-configured (model.py:39); gin is optional here
    import gin
    _configurable = gin.configurable(allowlist=['bottleneck_stride', 'patch_classification'])
except Exception:  # pragma: no cover - gin not installed
    def _configurable(cls):
        return cls


class PrototypeHeadMixin:
    """Compute methods of the prototype head.  Expects the attributes the reference class owns:
    prototype_vectors (N,K,1,1), last_layer (Linear N->C, no bias), epsilon,
    prototype_activation_function, patch_classification, features, add_on_layers."""

    #: arithmetic of the z.p contraction: 'fp32' (CUDA cores), 'tf32' or 'tf32x3' (tcgen05)
    head_precision = None
    #: when False, forward() skips writing the (B,N,h,w) distance map and returns None in its place
    materialize_distances = True

    # -- helpers -------------------------------------------------------------------------------
    def _head_cfg(self):
        act = self.prototype_activation_function
        if not isinstance(act, str):
            raise NotImplementedError(
                "a callable prototype_activation_function cannot run inside the fused head; "
                "use 'log' or 'linear' (model.py:231-237)")
        return float(self.epsilon), act, self.head_precision

    def _run_head(self, conv_features, want_dist, want_sim):
        eps, act, prec = self._head_cfg()
        return _head.HeadFunction.apply(conv_features, self.prototype_vectors, self.last_layer.weight,
                                        eps, act, prec, want_dist, want_sim)

    # -- reference API -------------------------------------------------------------------------
    def run_last_layer(self, prototype_activations):
        return self.last_layer(prototype_activations)

    def conv_features(self, x):
        """backbone + add-on layers (model.py:164-175); stock torch, out of scope of the fused head"""
        x = self.features(x)
        if isinstance(x, list):   # multi-scale training
            return [self.add_on_layers(x_scaled) for x_scaled in x]
        return self.add_on_layers(x)

    def _l2_convolution(self, x):
        """(B,K,h,w) -> squared L2 distances (B,N,h,w)   (model.py:203-221)"""
        _, dist, _ = self._run_head(x, True, False)
        return dist

    def prototype_distances(self, x):
        return self._l2_convolution(self.conv_features(x))

    def distance_2_similarity(self, distances):
        """kept for callers that post-process distances themselves (model.py:231-237)"""
        act = self.prototype_activation_function
        if act == 'log':
            return torch.log((distances + 1) / (distances + self.epsilon))
        if act == 'linear':
            return -distances
        return act(distances)

    def forward(self, x, **kwargs):
        conv_features = self.conv_features(x)
        if isinstance(conv_features, list):
            return [self.forward_from_conv_features(c, **kwargs) for c in conv_features]
        return self.forward_from_conv_features(conv_features, **kwargs)

    def forward_with_features(self, x, **kwargs):
        conv_features = self.conv_features(x)
        if isinstance(conv_features, list):
            results = [self.forward_from_conv_features(c, **kwargs) for c in conv_features]
            return [(r[0], r[1], c) for r, c in zip(results, conv_features)]
        logits, distances = self.forward_from_conv_features(conv_features, **kwargs)
        return logits, distances, conv_features

    def forward_from_conv_features(self, conv_features, return_activations=False, return_distances=False):
        if isinstance(conv_features, list):
            return [self.forward_from_conv_features(c) for c in conv_features]
        if not getattr(self, 'patch_classification', False):
            return self._forward_min_pool(conv_features, return_distances)
        if return_activations:
            logits, _, sim = self._run_head(conv_features, False, True)
            return logits, sim
        want_dist = bool(self.materialize_distances)
        logits, dist, _ = self._run_head(conv_features, want_dist, False)
        return logits, (dist if want_dist else None)

    def _forward_min_pool(self, conv_features, return_distances):
        """legacy ProtoPNet classification branch (model.py:284-299): global min-pool of the distance map,
        then similarity + last layer on (B,N).  The minima come from the per-image scan kernel
        (adlm_min_pool_fwd: the (B,N,h,w) map is never written) and the backward scatters onto the arg-min
        patches; the (B,N)-sized tail is negligible and stays in torch.  Only `return_distances=True`
        materialises the map, because the caller asks for it."""
        if return_distances:
            distances = self._l2_convolution(conv_features)
            min_distances = distances.flatten(2).min(dim=2).values
        else:
            distances = None
            prec = self.head_precision if self.head_precision in ("fp32", "tf32x3") else None
            min_distances, _ = _head.MinPoolFunction.apply(conv_features, self.prototype_vectors, prec)
        prototype_activations = self.distance_2_similarity(min_distances)
        logits = self.run_last_layer(prototype_activations)
        if return_distances:
            return logits, min_distances, distances
        return logits, min_distances

    def push_forward(self, x):
        """features and distances for push / find_nearest (model.py:301-309)"""
        conv_output = self.conv_features(x)
        if isinstance(conv_output, list):
            return [(c, self._l2_convolution(c)) for c in conv_output]
        return conv_output, self._l2_convolution(conv_output)

    def predict(self, conv_features):
        """logits (B,h,w,C) + per-pixel argmax (B,h,w) in one launch, no distance map (eval path)"""
        eps, act, prec = self._head_cfg()
        out = _head.head_forward(conv_features, self.prototype_vectors, self.last_layer.weight, eps, act,
                                 prec, want_dist=False, want_sim=False, want_argmax=True)
        return out["logits"], out["argmax"]

    def proto_class(self):
        """(N,) int32 class of each prototype, on the prototypes' device.  Cached: the one-hot validation, the arg-max
        and the host -> device copy cost more than a training step; the cache is keyed on the identity tensor (object,
        version, shape) and the device, so prune_prototypes / a reassigned identity / .to(device) invalidate it."""
        ident = self.prototype_class_identity
        dev = self.prototype_vectors.device
        key = (id(ident), ident._version, tuple(ident.shape), str(dev))
        cached = self.__dict__.get('_proto_class_cache')
        if cached is None or cached[0] != key:
            pc = _head.proto_class_from_identity(ident).to(dev)
            per_class = int(torch.bincount(pc.to(torch.int64).cpu(), minlength=1).max()) if pc.numel() else 0
            cached = (key, pc, per_class)
            self.__dict__['_proto_class_cache'] = cached
        return cached[1]

    def max_prototypes_per_class(self):
        self.proto_class()
        return self.__dict__['_proto_class_cache'][2]

    def fused_step(self, conv_features, target, ignore_void_class=True, loss_weights=(1.0, 0.25, 1e-4),
                   grad_scale=1.0):
        """module.py:143-222 on given features in one C-ABI call: returns (loss, scalars) with autograd
        edges to conv_features, prototype_vectors and last_layer.weight (see head.FusedStepFunction)."""
        eps, act, prec = self._head_cfg()
        if act != 'log':
            raise NotImplementedError("fused_step implements the 'log' activation (all segmentation configs)")
        cfg = dict(ignore_void=ignore_void_class, weights=tuple(loss_weights), grad_scale=grad_scale,
                   eps=eps, precision=prec)
        return _head.FusedStepFunction.apply(conv_features, self.prototype_vectors, self.last_layer.weight,
                                             target, self.proto_class(), cfg)

    def add_on_is_sigmoid(self):
        """True when add_on_layers is exactly Sequential(Sigmoid()) — the 'deeplab_simple' add-on of every
        segmentation config (model.py:124-128)"""
        mods = list(self.add_on_layers.children()) if isinstance(self.add_on_layers, nn.Sequential) else None
        return mods is not None and len(mods) == 1 and isinstance(mods[0], nn.Sigmoid)

    def fused_step_from_backbone(self, backbone_features, target, ignore_void_class=True,
                                 loss_weights=(1.0, 0.25, 1e-4), grad_scale=1.0):
        """fused_step on the output of self.features(x), i.e. BEFORE the add-on layers (model.py:164-175).
        With the Sigmoid-only add-on and head_precision 'tf32' the sigmoid and its backward run inside the head
        kernels (adlm_step_fwd_bwd_pre); otherwise the add-on layers run as stock torch first."""
        eps, act, prec = self._head_cfg()
        K, N, C = self.prototype_shape[1], self.num_prototypes, self.num_classes
        if (self.add_on_is_sigmoid() and prec == 'tf32' and act == 'log' and K <= 64 and N <= 192 and C <= 32
                and not isinstance(backbone_features, list)):
            cfg = dict(ignore_void=ignore_void_class, weights=tuple(loss_weights), grad_scale=grad_scale,
                       eps=eps, precision=prec, pre_sigmoid=True)
            return _head.FusedStepFunction.apply(backbone_features, self.prototype_vectors, self.last_layer.weight,
                                                 target, self.proto_class(), cfg)
        return self.fused_step(self.add_on_layers(backbone_features), target, ignore_void_class, loss_weights,
                               grad_scale)


def _first_add_on_in_channels(features):
    """channel count entering the add-on layers, by backbone family (model.py:81-95)"""
    name = str(features).upper()
    convs = lambda m: [i for i in m.modules() if isinstance(i, nn.Conv2d)]
    if name.startswith('VGG') or name.startswith('RES'):
        return convs(features)[-1].out_channels
    if name.startswith('DENSE'):
        return [i for i in features.modules() if isinstance(i, nn.BatchNorm2d)][-1].num_features
    if name.startswith('DEEPLAB'):
        return convs(features)[-2].out_channels
    if name.startswith('MSC'):
        return convs(features.base)[-2].out_channels
    raise Exception(f'{name[:10]} base_architecture NOT implemented')


def _make_add_on_layers(kind, in_channels, proto_depth, bottleneck_stride):
    """model.py:97-136: 'bottleneck' halves channels down to the prototype depth, 'bottleneck_pool' puts a
    strided 3x3 conv in front, 'deeplab_simple' is a bare Sigmoid, anything else two 1x1 convs."""
    if kind == 'deeplab_simple':
        return nn.Sequential(nn.Sigmoid())
    if kind.startswith('bottleneck'):
        layers = []
        if kind == 'bottleneck_pool':
            layers += [nn.Conv2d(in_channels, in_channels, kernel_size=3, padding=1, stride=bottleneck_stride),
                       nn.ReLU()]
        cur = in_channels
        while cur > proto_depth or len(layers) == 0:
            out = max(proto_depth, cur // 2)
            layers += [nn.Conv2d(cur, out, kernel_size=1), nn.ReLU(), nn.Conv2d(out, out, kernel_size=1)]
            if out > proto_depth:
                layers.append(nn.ReLU())
            else:
                assert out == proto_depth
                layers.append(nn.Sigmoid())
            cur = cur // 2
        return nn.Sequential(*layers)
    return nn.Sequential(nn.Conv2d(in_channels, proto_depth, kernel_size=1), nn.ReLU(),
                         nn.Conv2d(proto_depth, proto_depth, kernel_size=1), nn.Sigmoid())


@_configurable
class PPNet(PrototypeHeadMixin, nn.Module):
    def __init__(self, features, img_size, prototype_shape,
                 proto_layer_rf_info, num_classes, init_weights=True,
                 prototype_activation_function='log',
                 add_on_layers_type='bottleneck',
                 bottleneck_stride: Optional[int] = None,
                 patch_classification: bool = False):
        super().__init__()
        self.img_size = img_size
        self.epsilon = 1e-4
        self.bottleneck_stride = bottleneck_stride
        self.patch_classification = patch_classification
        self.prototype_vectors = nn.Parameter(torch.rand(prototype_shape), requires_grad=True)
        self.prototype_activation_function = prototype_activation_function

        n_protos = self.num_prototypes
        assert n_protos % num_classes == 0
        per_class = n_protos // num_classes
        # one-hot class identity, the same number of prototypes for every class (model.py:66-73)
        self.prototype_class_identity = torch.zeros(n_protos, num_classes)
        self.prototype_class_identity[torch.arange(n_protos), torch.arange(n_protos) // per_class] = 1
        self.num_prototypes_per_class = per_class
        self.proto_layer_rf_info = proto_layer_rf_info

        self.features = features   # the name matters for loading pretrained weights
        self.add_on_layers = _make_add_on_layers(add_on_layers_type, _first_add_on_in_channels(features),
                                                 self.prototype_shape[1], bottleneck_stride)
        # unused by the fused kernels, kept because checkpoints and prune_prototypes expect it (model.py:140)
        self.ones = nn.Parameter(torch.ones(self.prototype_shape), requires_grad=False)
        self.last_layer = nn.Linear(n_protos, num_classes, bias=False)
        if init_weights:
            self._initialize_weights()

    @property
    def prototype_shape(self):
        return self.prototype_vectors.shape

    @property
    def num_prototypes(self):
        return self.prototype_vectors.shape[0]

    @property
    def num_classes(self):
        return self.prototype_class_identity.shape[1]

    def prune_prototypes(self, prototypes_to_prune):
        """drop prototypes by index (model.py:311-336); N changes, nothing is cached on the device"""
        keep = sorted(set(range(self.num_prototypes)) - set(int(i) for i in prototypes_to_prune))
        self.prototype_vectors = nn.Parameter(self.prototype_vectors.data[keep, ...], requires_grad=True)
        self.last_layer.in_features = self.num_prototypes
        self.last_layer.out_features = self.num_classes
        self.last_layer.weight.data = self.last_layer.weight.data[:, keep]
        self.ones = nn.Parameter(self.ones.data[keep, ...], requires_grad=False)
        self.prototype_class_identity = self.prototype_class_identity[keep, :]

    def __repr__(self):
        return ('PPNet(\n\tfeatures: {},\n\timg_size: {},\n\tprototype_shape: {},\n'
                '\tproto_layer_rf_info: {},\n\tnum_classes: {},\n\tepsilon: {}\n)').format(
            self.features, self.img_size, self.prototype_shape, self.proto_layer_rf_info,
            self.num_classes, self.epsilon)

    def set_last_layer_incorrect_connection(self, incorrect_strength):
        """+1 towards the prototype's own class, `incorrect_strength` elsewhere (model.py:359-371)"""
        own = torch.t(self.prototype_class_identity).to(self.last_layer.weight.device)
        self.last_layer.weight.data.copy_(own + incorrect_strength * (1 - own))

    def _initialize_weights(self):
        for m in self.add_on_layers.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode='fan_out', nonlinearity='relu')
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)
        self.set_last_layer_incorrect_connection(incorrect_strength=-0.5)
