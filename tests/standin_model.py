"""Stand-ins for the reference's two PPNetMultiScale classes, carrying the SAME attribute and method
names (segmentation/model/model_multiscale.py:71-245, model_multiscale_group.py:82-308) with the
oracle as their un-patched body.  They exist because /root/reference does not travel to the GPU box;
`install_on_classes` patches them exactly as `install()` patches the real classes."""
import torch
import torch.nn as nn

from oracle import head_oracle as ho


class Stage1Standin(nn.Module):
    def __init__(self, P, K, S=4, D=64, act="log"):
        super().__init__()
        self.epsilon = 1e-4
        self.patch_classification = True
        self.num_scales = S
        self.scale_head = None
        self.prototype_activation_function = act
        self.prototype_vectors = nn.Parameter(torch.rand(P, D, 1, 1))
        self.ones = nn.Parameter(torch.ones(P, D, 1, 1), requires_grad=False)
        self.prototype_class_identity = ho.make_identity(P, K, S)
        self.scale_num_prototypes = {s: r for s, r in enumerate(ho.uniform_ranges(P, S))}
        self.last_layer = nn.Linear(P, K, bias=False)
        self.last_layer.weight.data.copy_(ho.make_last_layer(self.prototype_class_identity))
        self.features = nn.Identity()
        self.add_on_layers = nn.Sequential(nn.Sigmoid())

    @property
    def num_prototypes(self):
        return self.prototype_vectors.shape[0]

    @property
    def num_classes(self):
        return self.prototype_class_identity.shape[1]

    def _ranges(self):
        return [self.scale_num_prototypes[s] for s in range(self.num_scales)]

    def conv_features(self, x):
        x = self.features(x)
        if isinstance(x, list):
            return [self.add_on_layers(v) for v in x]
        return self.add_on_layers(x)

    def _scale_l2_convolution(self, x):
        return ho.scale_l2_convolution(x, self.prototype_vectors, self._ranges())

    def prototype_distances(self, x):
        return self._scale_l2_convolution(self.conv_features(x))

    def distance_2_similarity(self, d):
        return ho.distance_2_similarity(d, self.prototype_activation_function, self.epsilon)

    def run_last_layer(self, a):
        return self.last_layer(a)

    def forward(self, x, **kw):
        c = self.conv_features(x)
        if isinstance(c, list):
            return [self.forward_from_conv_features(v, **kw) for v in c]
        return self.forward_from_conv_features(c, **kw)

    def forward_from_conv_features(self, conv_features, return_activations=False, return_distances=False):
        if isinstance(conv_features, list):
            return [self.forward_from_conv_features(c) for c in conv_features]
        d = self._scale_l2_convolution(conv_features)
        B, P, H, W = d.shape
        a = self.distance_2_similarity(d.permute(0, 2, 3, 1).contiguous().reshape(-1, P))
        logits = self.run_last_layer(a).reshape(B, H, W, -1)
        if return_activations and not return_distances:
            return logits, a
        if return_activations and return_distances:
            return logits, d, a
        return logits, d

    def push_forward(self, x):
        c = self.conv_features(x)
        return c, self._scale_l2_convolution(c)

    def prune_prototypes(self, prune):
        """model_multiscale.py:400-432."""
        keep = sorted(set(range(self.num_prototypes)) - set(prune))
        for s in range(self.num_scales):
            lo, hi = self.scale_num_prototypes[s]
            n = len(set(range(lo, hi)) - set(prune))
            start = 0 if s == 0 else self.scale_num_prototypes[s - 1][1]
            self.scale_num_prototypes[s] = (start, start + n)
        self.prototype_vectors = nn.Parameter(self.prototype_vectors.data[keep, ...])
        self.last_layer.in_features = self.num_prototypes
        self.last_layer.weight.data = self.last_layer.weight.data[:, keep]
        self.ones = nn.Parameter(self.ones.data[keep, ...], requires_grad=False)
        self.prototype_class_identity = self.prototype_class_identity[keep, :]


class Stage2Standin(Stage1Standin):
    def __init__(self, P, K, S=4, D=64, num_groups=3):
        super().__init__(P, K, S, D)
        self.num_groups = num_groups
        self._initialize_groups()

    def _initialize_groups(self):
        sizes = [int(self.prototype_class_identity[:, c].sum()) for c in range(self.num_classes)]
        sizes = [n for n in sizes if n > 0]
        self.group_projection = nn.ModuleList([nn.Linear(n, self.num_groups, bias=False) for n in sizes])
        g = torch.Generator().manual_seed(3)
        for gp in self.group_projection:
            w = torch.rand(gp.weight.shape, generator=g)
            gp.weight.data = w / w.sum(dim=1, keepdim=True)  # rows on the simplex
        self.last_layer_group = nn.Linear(len(sizes) * self.num_groups, self.num_classes, bias=False)
        self.last_layer_group.weight.data.uniform_(-0.5, 1.0, generator=g)

    def compute_group(self, a):
        return ho.compute_group(a, self.prototype_class_identity, [gp.weight for gp in self.group_projection])

    def run_last_layer(self, a):
        return self.last_layer_group(torch.cat(self.compute_group(a), dim=-1))


class SingleScaleStandin(nn.Module):
    """Attribute and method names of the single-scale ProtoSeg `PPNet` (segmentation/model/model.py:74-354): no
    num_scales / scale_num_prototypes, `_l2_convolution` instead of `_scale_l2_convolution`, and the two-way return
    convention of its forward_from_conv_features (:327-330)."""

    def __init__(self, P, K, D=64, act="log", patch_classification=True):
        super().__init__()
        self.epsilon = 1e-4
        self.patch_classification = patch_classification
        self.prototype_activation_function = act
        self.prototype_vectors = nn.Parameter(torch.rand(P, D, 1, 1))
        self.ones = nn.Parameter(torch.ones(P, D, 1, 1), requires_grad=False)
        self.prototype_class_identity = ho.make_identity(P, K, 1)
        self.last_layer = nn.Linear(P, K, bias=False)
        self.last_layer.weight.data.copy_(ho.make_last_layer(self.prototype_class_identity))
        self.features = nn.Identity()
        self.add_on_layers = nn.Sequential(nn.Sigmoid())

    @property
    def num_prototypes(self):
        return self.prototype_vectors.shape[0]

    @property
    def num_classes(self):
        return self.prototype_class_identity.shape[1]

    def conv_features(self, x):
        x = self.features(x)
        if isinstance(x, list):
            return [self.add_on_layers(v) for v in x]
        return self.add_on_layers(x)

    def _l2_convolution(self, x):
        return ho.l2_convolution(x, self.prototype_vectors)

    def prototype_distances(self, x):
        return self._l2_convolution(self.conv_features(x))

    def distance_2_similarity(self, d):
        return ho.distance_2_similarity(d, self.prototype_activation_function, self.epsilon)

    def run_last_layer(self, a):
        return self.last_layer(a)

    def forward(self, x, **kw):
        c = self.conv_features(x)
        if isinstance(c, list):
            return [self.forward_from_conv_features(v, **kw) for v in c]
        return self.forward_from_conv_features(c, **kw)

    def forward_from_conv_features(self, conv_features, return_activations=False, return_distances=False):
        if isinstance(conv_features, list):
            return [self.forward_from_conv_features(c) for c in conv_features]
        d = self._l2_convolution(conv_features)
        B, P, H, W = d.shape
        if self.patch_classification:
            a = self.distance_2_similarity(d.permute(0, 2, 3, 1).contiguous().reshape(-1, P))
            logits = self.run_last_layer(a).reshape(B, H, W, -1)
            return (logits, a) if return_activations else (logits, d)
        mins = d.reshape(B, P, -1).min(dim=2).values          # model.py:331-343, global min pooling
        logits = self.run_last_layer(self.distance_2_similarity(mins))
        return (logits, mins, d) if return_distances else (logits, mins)

    def push_forward(self, x):
        c = self.conv_features(x)
        return c, self._l2_convolution(c)
