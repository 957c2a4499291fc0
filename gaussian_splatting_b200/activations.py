"""Fused activations (second half): the reference's model keeps log-scales, unnormalised
quaternions and opacity logits and activates them with torch kernels on every step
(scene/gaussian_model.py:32-55 `setup_functions`, :130-157 `get_scaling / get_rotation / get_opacity`, called from
gaussian_renderer/__init__.py:77,89-90), followed by their three backward kernels in autograd.  Here the
per-Gaussian kernels apply exp / normalize / sigmoid in registers and chain their derivatives before the stores
(GsrParams.raw_inputs, include/gsrast.h), so the rasterizer takes the raw parameters and returns raw gradients.

Two ways in, neither edits a reference file:

* `GaussianRasterizer.forward(..., raw_inputs=RAW_SCALES | RAW_ROTATIONS | RAW_OPACITIES)` with the raw tensors;
* `install(GaussianModel)` (same mechanism as `densify.install`): rebinds the three properties so that they return a
  `Deferred` proxy.  `render()` hands the proxies to the rasterizer, which unwraps them to the raw leaves; any other
  consumer (`densify_and_split`, `reset_opacity`, TensorBoard histograms ...) sees the activated tensor, computed
  on first use by the reference's own activation function.
"""
import torch

from ._lib import RAW_OPACITIES, RAW_ROTATIONS, RAW_SCALES

__all__ = ["Deferred", "install", "uninstall", "unwrap", "RAW_SCALES", "RAW_ROTATIONS", "RAW_OPACITIES"]


class Deferred:
    """An activated parameter that has not been computed yet: `raw` is the model's leaf, `fn` its activation, `bit`
    the GSR_RAW_* flag under which the rasterizer's kernels apply the same activation themselves."""

    __slots__ = ("raw", "fn", "bit", "_value")

    def __init__(self, raw, fn, bit):
        self.raw, self.fn, self.bit, self._value = raw, fn, bit, None

    def value(self):
        if self._value is None:
            self._value = self.fn(self.raw)
        return self._value

    # --- behave like the activated tensor everywhere else
    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        def conv(a):
            if isinstance(a, Deferred):
                return a.value()
            if isinstance(a, (list, tuple)):
                return type(a)(conv(x) for x in a)
            return a
        return func(*conv(args), **{k: conv(v) for k, v in (kwargs or {}).items()})

    def __getattr__(self, name):
        return getattr(self.value(), name)

    def __getitem__(self, idx):
        return self.value()[idx]

    def __len__(self):
        return len(self.value())

    def __repr__(self):
        return f"Deferred({self.fn.__name__ if hasattr(self.fn, '__name__') else self.fn}, raw={tuple(self.raw.shape)})"


def _binary(name):
    def op(self, other):
        return getattr(self.value(), name)(other.value() if isinstance(other, Deferred) else other)
    op.__name__ = name
    return op


for _n in ("__lt__", "__le__", "__gt__", "__ge__", "__eq__", "__ne__", "__add__", "__radd__", "__sub__", "__rsub__",
           "__mul__", "__rmul__", "__truediv__", "__rtruediv__", "__pow__", "__matmul__"):
    setattr(Deferred, _n, _binary(_n))
Deferred.__neg__ = lambda self: -self.value()
Deferred.__hash__ = object.__hash__


def unwrap(opacities, scales, rotations, raw_inputs=0):
    """Replace Deferred proxies by their raw leaves and set their bits (GaussianRasterizer.forward)."""
    raw_inputs = int(raw_inputs or 0)
    out = []
    for t in (opacities, scales, rotations):
        if isinstance(t, Deferred):
            raw_inputs |= t.bit
            t = t.raw
        out.append(t)
    return out[0], out[1], out[2], raw_inputs


_SAVED = {}


def _fusable(model):
    """The kernels implement exactly the reference's default activations (gaussian_model.py:44-55)."""
    return (getattr(model, "scaling_activation", None) is torch.exp
            and getattr(model, "opacity_activation", None) is torch.sigmoid
            and getattr(model, "rotation_activation", None) is torch.nn.functional.normalize)


def install(model_cls):
    """Rebind get_scaling / get_rotation / get_opacity of the reference's GaussianModel class (scene/gaussian_model.py:
    130-157) to return Deferred proxies.  A model whose activation functions were replaced keeps the reference's
    behaviour.  Returns the class."""
    if model_cls in _SAVED:
        return model_cls
    _SAVED[model_cls] = {n: model_cls.__dict__.get(n) for n in ("get_scaling", "get_rotation", "get_opacity")}

    def get_scaling(self):
        return Deferred(self._scaling, self.scaling_activation, RAW_SCALES) if _fusable(self) \
            else self.scaling_activation(self._scaling)

    def get_rotation(self):
        return Deferred(self._rotation, self.rotation_activation, RAW_ROTATIONS) if _fusable(self) \
            else self.rotation_activation(self._rotation)

    def get_opacity(self):
        return Deferred(self._opacity, self.opacity_activation, RAW_OPACITIES) if _fusable(self) \
            else self.opacity_activation(self._opacity)

    model_cls.get_scaling = property(get_scaling)
    model_cls.get_rotation = property(get_rotation)
    model_cls.get_opacity = property(get_opacity)
    return model_cls


def uninstall(model_cls):
    for n, v in _SAVED.pop(model_cls, {}).items():
        if v is not None:
            setattr(model_cls, n, v)
