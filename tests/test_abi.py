"""CPU-side checks of the boundary: the library loads, exports every symbol include/csi.h declares, the
ctypes mirror of the structs has the C layout, and the product never imports the oracle.  No GPU needed."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "csi.h")
PKG = os.path.join(ROOT, "scalable_stellar_inference_b200")


@pytest.fixture(scope="module")
def lib():
    from scalable_stellar_inference_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    return _lib


def _declared():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(csi_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    names = _declared()
    assert len(names) >= 20
    out = subprocess.run(["nm", "-D", "--defined-only", lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (csi_[a-z0-9_]+)", out))
    missing = [n for n in names if n not in exported]
    assert not missing, missing
    assert set(lib.EXPORTS) == set(names), set(lib.EXPORTS) ^ set(names)
    handle = lib.load()
    assert handle.csi_version() == 200


def test_struct_layout_matches_header(lib, tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "csi.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(csi_config),sizeof(csi_net),sizeof(csi_grid),sizeof(csi_tn),offsetof(csi_config,emulator),'
                   'offsetof(csi_config,device));printf("%zu %zu\\n",sizeof(csi_nuts_stats),offsetof(csi_nuts_stats,depth));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(t) for t in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(lib.Config), C.sizeof(lib.Net), C.sizeof(lib.Grid), C.sizeof(lib.Tn), lib.Config.emulator.offset,
            lib.Config.device.offset, C.sizeof(lib.NutsStats), lib.NutsStats.depth.offset]
    assert got == want


def test_no_device_is_an_error_not_a_fallback(lib):
    """Without a GPU, creation must fail loudly (CSI_ENODEVICE) -- never compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import scalable_stellar_inference_b200 as ssi
    with pytest.raises(lib.CsiError) as e:
        ssi.HierarchicalStarModel(10, ["log_Teff", "log_g"])
    assert e.value.code == lib.ENODEVICE
    with pytest.raises(lib.CsiError):
        ssi.Emulator("cuda")


def test_product_never_imports_oracle():
    bad = []
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M) or "oracle/" in txt and f.endswith(".py") and "import" in txt and re.search(r"sys\.path.*oracle", txt):
                    bad.append(f)
    assert not bad, bad
    # bench.py may only touch the oracle in its cpu_baseline / --impl reference legs
    b = open(os.path.join(ROOT, "bench.py")).read() if os.path.exists(os.path.join(ROOT, "bench.py")) else ""
    for m in re.finditer(r"^(\s*)(from oracle|import oracle)", b, flags=re.M):
        assert len(m.group(1)) > 0, "bench.py imports the oracle at module level"


def test_python_surface_mirrors_reference():
    import inspect
    import scalable_stellar_inference_b200 as ssi
    sig = inspect.signature(ssi.SingleStarModel.__init__)
    assert list(sig.parameters)[:5] == ["self", "observables", "const", "kind", "prior"]
    sig = inspect.signature(ssi.MultiStarModel.__init__)
    assert list(sig.parameters)[:5] == ["self", "num_stars", "observables", "const", "kind"]
    sig = inspect.signature(ssi.HierarchicalStarModel.__init__)
    assert list(sig.parameters)[:5] == ["self", "num_stars", "observables", "const", "kind"]
    assert ssi.SingleStarModel.outputs == ["log_age", "log_Teff", "log_R", "log_Dnu", "log_L", "log_g", "log_numax"]
    for f in (ssi.Emulator, ssi.Star, ssi.BolometricCorrections):
        assert "backend" in inspect.signature(f).parameters


def test_xla_ffi_shim_compiles_against_an_api_stub(tmp_path):
    """csrc/csi_xla_ffi.cc (the jax.ffi handlers) cannot be built here -- jaxlib's headers are absent -- but it must at
    least be valid C++ against the C ABI: compiled with a stand-in for xla/ffi/api/ffi.h that models the names the shim
    uses and checks every handler's argument list against its binding (tests/ffi_stub).  The Makefile rule `xla_ffi`
    builds the real library wherever `import jax` works."""
    src = os.path.join(PKG, "csrc", "csi_xla_ffi.cc")
    cuda_inc = "/usr/local/cuda/include"
    if not os.path.exists(os.path.join(cuda_inc, "cuda_runtime.h")):
        pytest.skip("no CUDA headers")
    out = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(ROOT, "tests", "ffi_stub"), "-I", cuda_inc,
                          src], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-3000:]
    mk = subprocess.run(["make", "-C", os.path.join(PKG, "csrc"), "xla_ffi"], capture_output=True, text=True)
    assert mk.returncode == 0 and ("unavailable on this box" in mk.stdout or os.path.exists(os.path.join(PKG, "libcsi_xla_ffi.so")))
