"""One-off conversion of the reference's weight files into the package's data directory.

Reads ``/root/reference/src/celestify/data/{emulator_weights,bc_weights}.h5`` with the
in-repo HDF5 reader (h5py is not installed) and writes ``.npz`` archives that keep every
array's original dtype and the reference's dataset names (``src/celestify/emulator.py:9-19``,
``src/celestify/bolometric_corrections.py:15-23``).  ``emulator_error.json`` is copied verbatim
(``src/celestify/numpyro_models.py:107-115`` reads it).  Data only — no reference source is copied.

    python tools/convert_weights.py [/root/reference]
"""
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from scalable_stellar_inference_b200.h5lite import H5File  # noqa: E402


def main(ref="/root/reference"):
    src = os.path.join(ref, "src", "celestify", "data")
    dst = os.path.join(os.path.dirname(HERE), "scalable_stellar_inference_b200", "data")
    os.makedirs(dst, exist_ok=True)
    for name in ("emulator_weights", "bc_weights"):
        f = H5File(os.path.join(src, name + ".h5"))
        arrays = {k: f[k] for k in f.keys()}
        np.savez(os.path.join(dst, name + ".npz"), **arrays)
        print(name, {k: (v.shape, str(v.dtype)) for k, v in arrays.items()})
    shutil.copyfile(os.path.join(src, "emulator_error.json"), os.path.join(dst, "emulator_error.json"))


if __name__ == "__main__":
    main(*sys.argv[1:])
