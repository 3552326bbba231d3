"""Dense ReLU networks of the reference: stellar emulator and BC emulator (oracle; tests only).

Follows ``src/celestify/_jax/emulator.py:44-50`` (input z-score, 6x relu(x@W+b), linear head,
output affine) and ``src/celestify/_jax/bolometric_corrections.py:17-25`` (min-max input scaling,
same body).  Weights come from the ``.npz`` conversions of the reference's HDF5 files (dataset
names kept, ``src/celestify/emulator.py:9-19``, ``src/celestify/bolometric_corrections.py:15-23``).
The input-VJP is the hand-written transpose of the same graph (what JAX autodiff produces).
"""
from __future__ import annotations

import os
import numpy as np

DATA_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                        "scalable_stellar_inference_b200", "data")


class DenseNet:
    """y = out_loc + out_scale * MLP((x - in_loc) / in_scale); ReLU hidden layers."""

    def __init__(self, weights, bias, in_loc, in_scale, out_loc, out_scale, dtype=np.float64):
        self.dtype = np.dtype(dtype)
        self.weights = [np.asarray(w, dtype) for w in weights]
        self.bias = [np.asarray(b, dtype) for b in bias]
        self.in_loc = np.asarray(in_loc, dtype)
        self.in_scale = np.asarray(in_scale, dtype)
        self.out_loc = np.asarray(out_loc, dtype)
        self.out_scale = np.asarray(out_scale, dtype)

    def astype(self, dtype):
        return DenseNet(self.weights, self.bias, self.in_loc, self.in_scale,
                        self.out_loc, self.out_scale, dtype)

    def forward(self, x, cache=False):
        x = np.atleast_2d(np.asarray(x, self.dtype))
        h = (x - self.in_loc) / self.in_scale
        masks = []
        for w, b in zip(self.weights[:-1], self.bias[:-1]):
            h = h @ w + b
            m = h > 0
            h = np.where(m, h, 0).astype(self.dtype)
            masks.append(m)
        y = h @ self.weights[-1] + self.bias[-1]
        out = self.out_loc + self.out_scale * y
        return (out, masks) if cache else out

    __call__ = forward

    def vjp(self, masks, g_out):
        """Cotangent of the inputs given the cotangent of the outputs (shape [N, n_out])."""
        g = np.asarray(g_out, self.dtype) * self.out_scale
        g = g @ self.weights[-1].T
        for w, m in zip(self.weights[-2::-1], masks[::-1]):
            g = np.where(m, g, 0).astype(self.dtype)
            g = g @ w.T
        return g / self.in_scale

    def flops_per_row(self):
        return sum(2 * w.shape[0] * w.shape[1] for w in self.weights)


def load_emulator(dtype=np.float64) -> DenseNet:
    """src/celestify/emulator.py:8-21: dense_7..13, normalization mean/variance**0.5, rescaling."""
    z = np.load(os.path.join(DATA_DIR, "emulator_weights.npz"))
    keys = [f"dense_{i}" for i in range(7, 14)]
    # the reference takes variance**0.5 in the file's fp32 before any widening
    in_scale = z["normalization/variance:0"] ** 0.5
    return DenseNet([z[f"{k}/{k}/kernel:0"] for k in keys], [z[f"{k}/{k}/bias:0"] for k in keys],
                    z["normalization/mean:0"], in_scale,
                    z["rescaling/offset"], z["rescaling/scale"], dtype)


def load_bc_net(dtype=np.float64) -> DenseNet:
    """src/celestify/bolometric_corrections.py:15-26: dense, dense_1..6, min-max scaling, loc/scale."""
    z = np.load(os.path.join(DATA_DIR, "bc_weights.npz"))
    keys = ["dense"] + [f"dense_{i}" for i in range(1, 7)]
    minval, maxval = z["input_scaling/minval"], z["input_scaling/maxval"]
    return DenseNet([z[f"{k}/{k}/kernel:0"] for k in keys], [z[f"{k}/{k}/bias:0"] for k in keys],
                    minval, maxval - minval,
                    z["output_rescaling/loc"], z["output_rescaling/scale"], dtype)
