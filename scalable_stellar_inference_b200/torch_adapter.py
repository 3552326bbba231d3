"""``torch.autograd.Function`` adapters over the C ABI: the CUDA potential and the emulator
operators as differentiable torch functions.  This is the adapter that can be exercised on this image (the jax.ffi one,
``jax_ffi.py``, needs jaxlib); it has the same shape: ONE kernel yields value and gradient, ``backward`` only scales.

    U = potential_function(model)(z_flat_cuda)          # float64 scalar tensor; U.backward() fills z.grad
    y = emulator_function(model)(x_cuda)                # [n,5] -> [n,4]; autograd uses csi_net_vjp for dL/dx
"""
from __future__ import annotations

import torch

from . import _lib


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream or None


def _handle_of(owner):
    """(ctypes library, csi_handle*) of a model, an ``Emulator("cuda")`` / ``BolometricCorrections`` or a DeviceOps."""
    if hasattr(owner, "_h"):
        return owner._lib, owner._h
    ops = getattr(owner, "_ops", owner)
    return ops.lib, ops.h


class _Potential(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, model):
        if not (z.is_cuda and z.dtype == torch.float32 and z.is_contiguous() and z.numel() == model.flat_size):
            raise ValueError(f"z must be a contiguous float32 CUDA tensor of {model.flat_size} elements")
        if z.device.index != model.device:
            raise ValueError("z lives on another device than the model")
        out = torch.empty(8, dtype=torch.float64, device=z.device)
        grad = torch.empty_like(z)
        with torch.cuda.device(z.device):
            model.potential_device(z.data_ptr(), out.data_ptr(), grad.data_ptr(), _stream(z.device))
        nh = len(model.hyper_sites)
        if nh:   # the hyper slots of grad hold fp32 local partials; the fp64 sums are in out[1:]
            grad[:nh] = out[1:1 + nh].to(grad.dtype)
        ctx.save_for_backward(grad)
        return out[0].clone()

    @staticmethod
    def backward(ctx, ct):
        (grad,) = ctx.saved_tensors
        return (grad * ct.to(grad.dtype)), None


def potential_function(model):
    """z (float32 CUDA vector in ``csi_layout`` order) -> U(z) as a float64 scalar tensor with autograd support.
    For a sharded model U and the hyper-gradient are the LOCAL sums unless the in-kernel peer exchange is on."""
    return lambda z: _Potential.apply(z, model)


def potential_fn(model):
    """dict site -> unconstrained CUDA tensor, as numpyro's ``potential_fn`` takes it."""
    f = potential_function(model)
    n = model.n_local

    def fn(z: dict):
        parts = [z[s].reshape(1) for s in model.hyper_sites] + [z[s].expand(n) for s in model.star_sites]
        return f(torch.cat(parts).to(torch.float32).contiguous())
    return fn


class _Net(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, model, which, n_out):
        x = x.contiguous()
        if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 2):
            raise ValueError("x must be a float32 CUDA matrix [n, n_in]")
        y = torch.empty(x.shape[0], n_out, dtype=torch.float32, device=x.device)
        lib, h = _handle_of(model)
        with torch.cuda.device(x.device):
            _lib.check(lib.csi_net_forward(h, which, x.data_ptr(), y.data_ptr(), x.shape[0], _stream(x.device)))
        ctx.save_for_backward(x)
        ctx.model, ctx.which = model, which
        return y

    @staticmethod
    def backward(ctx, gy):
        (x,) = ctx.saved_tensors
        gy = gy.contiguous().to(torch.float32)
        gx = torch.empty_like(x)
        lib, h = _handle_of(ctx.model)
        with torch.cuda.device(x.device):
            _lib.check(lib.csi_net_vjp(h, ctx.which, x.data_ptr(), gy.data_ptr(), gx.data_ptr(), x.shape[0],
                                       _stream(x.device)))
        return gx, None, None, None


def emulator_function(owner):
    """``Emulator.model`` (src/celestify/_jax/emulator.py:35-50): x[n,5] -> y[n,4], differentiable in x.  ``owner`` is
    any object holding a libcsi handle (``Emulator("cuda")`` or a model)."""
    return lambda x: _Net.apply(x, owner, 0, 4)


def bc_emulator_function(owner):
    """``BolometricCorrections.model`` (src/celestify/_jax/bolometric_corrections.py:17-25): x[n,4] -> y[n,3]."""
    return lambda x: _Net.apply(x, owner, 1, 3)
