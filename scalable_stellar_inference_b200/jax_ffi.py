"""JAX side of the drop-in: ``jax.ffi`` custom calls over libcsi + ``jax.custom_vjp`` rules, so that numpyro's NUTS
drives the CUDA path unchanged (BASELINE north_star; the reference calls ``NUTS(model, ...)`` in
``scripts/model_star.py:134-136`` and numpyro differentiates ``potential_fn`` with ``jax.value_and_grad``).

Needs ``jax`` with the CUDA plugin and ``libcsi_xla_ffi.so`` (``make -C scalable_stellar_inference_b200/csrc xla_ffi``,
built only where jaxlib's headers are installed).  Neither exists on the image this repository was developed on
("jax.ffi: unavailable on this box"), so importing this module raises ImportError there; the C ABI it wraps is what
the GPU tests exercise through ctypes and through the torch adapter (``torch_adapter.py``).

    from scalable_stellar_inference_b200 import jax_ffi, HierarchicalStarModel
    model = HierarchicalStarModel(n, ["log_Teff", "log_g"]).condition(obs, diag, M_H_loc=..., M_H_scale=...)
    potential_fn = jax_ffi.potential_fn(model)            # dict of unconstrained sites -> scalar, differentiable
    kernel = numpyro.infer.NUTS(potential_fn=potential_fn, target_accept_prob=0.8, ...)
    emulator = jax_ffi.emulator(model)                    # x[n,5] -> y[n,4] with the input-VJP rule (Emulator backend)
"""
from __future__ import annotations

import ctypes
import os

import jax  # noqa: F401  (ImportError here = "jax.ffi: unavailable on this box")
import jax.numpy as jnp

from . import _lib

_FFI_LIB = os.path.join(_lib.HERE, "libcsi_xla_ffi.so")
_TARGETS = {"csi_potential": "CsiPotential", "csi_net_forward": "CsiNetForward", "csi_net_vjp": "CsiNetVjp",
            "csi_grid_forward": "CsiGridForward", "csi_grid_vjp": "CsiGridVjp"}
_registered = False


def register():
    """Register the XLA-FFI targets once per process."""
    global _registered
    if _registered:
        return
    if not os.path.exists(_FFI_LIB):
        raise _lib.CsiLibraryError(f"{_FFI_LIB} is missing: build it with `make -C {os.path.join(_lib.HERE, 'csrc')} xla_ffi`")
    _lib.load()                                   # libcsi.so first: the shim links against it
    lib = ctypes.CDLL(_FFI_LIB)
    for target, symbol in _TARGETS.items():
        jax.ffi.register_ffi_target(target, jax.ffi.pycapsule(getattr(lib, symbol)), platform="CUDA")
    _registered = True


def _handle(model) -> int:
    return int(model._h.value)


def potential_flat(model):
    """f32[n_hyper + n_sites * n_local] (csi_layout order) -> scalar U, with dU/dz as the custom VJP.  One kernel
    yields value and gradient: ``fwd`` keeps the gradient as the residual, ``bwd`` scales it by the cotangent."""
    register()
    h = _handle(model)

    def _call(z):
        return jax.ffi.ffi_call("csi_potential", (jax.ShapeDtypeStruct((6,), jnp.float64),
                                                  jax.ShapeDtypeStruct(z.shape, jnp.float32)))(z, handle=h)

    @jax.custom_vjp
    def potential(z):
        return _call(z)[0][0]

    def fwd(z):
        out6, g = _call(z)
        nh = len(model.hyper_sites)
        if nh:                                     # hyper slots of grad hold fp32 LOCAL partials: take the fp64 sums
            g = g.at[:nh].set(out6[1:1 + nh].astype(g.dtype))
        return out6[0], g

    def bwd(g, ct):
        return ((g * ct).astype(g.dtype),)

    potential.defvjp(fwd, bwd)
    return potential


def potential_fn(model):
    """The ``potential_fn`` numpyro's ``NUTS(potential_fn=...)`` takes: dict site -> unconstrained array."""
    flat = potential_flat(model)
    n = model.n_local

    def fn(z: dict):
        parts = [jnp.reshape(z[s], (1,)) for s in model.hyper_sites]
        parts += [jnp.broadcast_to(z[s], (n,)) for s in model.star_sites]
        return flat(jnp.concatenate(parts).astype(jnp.float32))
    return fn


def _net(model, which, n_in, n_out):
    register()
    h = _handle(model)

    @jax.custom_vjp
    def net(x):
        return jax.ffi.ffi_call("csi_net_forward", jax.ShapeDtypeStruct((x.shape[0], n_out), x.dtype))(x, handle=h, which=which)

    def fwd(x):
        return net(x), x

    def bwd(x, gy):   # input-VJP only: the weights are constants of the model
        return (jax.ffi.ffi_call("csi_net_vjp", jax.ShapeDtypeStruct(x.shape, x.dtype))(x, gy, handle=h, which=which),)

    net.defvjp(fwd, bwd)
    return net


def emulator(model):
    """``Emulator.model`` (src/celestify/_jax/emulator.py:35-50) as a JAX function with its input-VJP rule."""
    return _net(model, 0, 5, 4)


def bc_emulator(model):
    """``BolometricCorrections.model`` (src/celestify/_jax/bolometric_corrections.py:17-25)."""
    return _net(model, 1, 4, 3)
