"""Flat parameter storage: one contiguous fp32 buffer with the 30 state_dict tensors as views
(state_dict order of models/model_res_sigmoid.py:159-181), so checkpoints of the reference
round-trip with load_state_dict(strict=True) while the kernels see a single pointer."""
import math
from collections import OrderedDict

import torch

from . import capi

# (name, rows, cols) in state_dict order
_MODULES = (("encoder.0", 128, 256), ("encoder.1", 128, 128), ("encoder.2", 128, 128), ("encoder.3", 128, 128),
            ("encoder1.0", 128, 256), ("encoder1.1", 128, 128), ("encoder1.2", 128, 128),
            ("generator.0", 256, 256), ("generator.1", 256, 256), ("generator.2", 256, 256),
            ("generator.3", 128, 256), ("generator.4", 1, 128),
            ("generator1.0", 256, 256), ("generator1.1", 256, 256), ("generator1.2", 256, 256))
TENSOR_NAMES = tuple(n + s for n, _, _ in _MODULES for s in (".weight", ".bias"))
# tensors the forward pass never touches: autograd leaves their .grad None, AdamW skips them
UNUSED = ("encoder1.0.weight", "encoder1.0.bias")


def tensor_shapes():
    out = OrderedDict()
    for n, r, c in _MODULES:
        out[n + ".weight"] = (r, c)
        out[n + ".bias"] = (r,)
    return out


def tensor_offsets():
    off, out = 0, OrderedDict()
    for k, shp in tensor_shapes().items():
        out[k] = off
        off += int(math.prod(shp))
    assert off == capi.PARAM_COUNT
    return out


class FlatParams:
    """Flat fp32 parameter buffer + named views, gradient buffer, AdamW state and update mask."""

    def __init__(self, device):
        self.device = torch.device(device)
        self.flat = torch.zeros(capi.PARAM_COUNT, dtype=torch.float32, device=self.device)
        # gradient ++ the step's two loss scalars in one buffer: the kernels write both, the data-parallel
        # all-reduce sends it as one message without packing copies (dist.GradAllReduce.reduce_flat)
        self.gradbuf = torch.zeros(capi.PARAM_COUNT + 2, dtype=torch.float32, device=self.device)
        self.grad = self.gradbuf[:capi.PARAM_COUNT]
        self.loss_slot = self.gradbuf[capi.PARAM_COUNT:]
        self.shapes = tensor_shapes()
        self.offsets = tensor_offsets()
        self.views = OrderedDict()
        self.grad_views = OrderedDict()
        for k, shp in self.shapes.items():
            o, n = self.offsets[k], int(math.prod(shp))
            self.views[k] = self.flat[o:o + n].view(shp)
            self.grad_views[k] = self.grad[o:o + n].view(shp)
        mask = torch.ones(capi.PARAM_COUNT, dtype=torch.uint8)
        for k in UNUSED:
            o, n = self.offsets[k], int(math.prod(self.shapes[k]))
            mask[o:o + n] = 0
        self.mask = mask.to(self.device)

    def scratch_grad(self):
        """Scratch gradient of the autograd path (nn._LossFn): forward writes it, backward adds it to `grad`."""
        if getattr(self, "_scratch", None) is None:
            self._scratch = torch.zeros(capi.PARAM_COUNT, dtype=torch.float32, device=self.device)
        return self._scratch

    def init_weights(self, generator=None):
        """U(+-2/sqrt(fan_in)) for weights AND biases (models/model_res_sigmoid.py:183-189)."""
        for n, r, c in _MODULES:
            bound = 2.0 / math.sqrt(c)
            for k in (n + ".weight", n + ".bias"):
                v = self.views[k]
                v.copy_(((torch.rand(v.shape, generator=generator) * 2 - 1) * bound).to(self.device))

    def state_dict(self):
        return OrderedDict((k, v.detach().clone()) for k, v in self.views.items())

    def load_state_dict(self, sd, strict=True):
        missing = [k for k in self.views if k not in sd]
        unexpected = [k for k in sd if k not in self.views]
        if strict and (missing or unexpected):
            raise RuntimeError("Error(s) in loading state_dict: missing %s unexpected %s" % (missing, unexpected))
        for k, v in self.views.items():
            if k in sd:
                t = torch.as_tensor(sd[k])
                if tuple(t.shape) != tuple(v.shape):
                    raise RuntimeError("size mismatch for %s: %s vs %s" % (k, tuple(t.shape), tuple(v.shape)))
                v.copy_(t.to(device=self.device, dtype=torch.float32))
