"""The C-ABI library loads and exports every symbol include/gzcuda.h declares; without a
CUDA device it refuses to build a batch (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "gzcuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gzcuda_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_are_exported(gzlib):
    from gazebo_b200 import batch
    out = subprocess.check_output(["nm", "-D", "--defined-only", batch.LIB_PATH]).decode()
    exported = set(line.split()[-1] for line in out.splitlines() if line.strip())
    decl = declared_symbols()
    assert len(decl) >= 20
    missing = [s for s in decl if s not in exported]
    assert not missing, missing
    assert sorted(batch.EXPORTS) == decl    # the Python binding covers the whole header


def test_adapter_header_symbols_are_exported():
    """include/gzcuda_ode.h (the reference-facing step interface over live ODE objects) against libgzcuda_ode.so."""
    path = os.path.join(ROOT, "gazebo_b200", "libgzcuda_ode.so")
    if not os.path.exists(path):
        pytest.skip("libgzcuda_ode.so needs the reference's headers to build (make -C host)")
    text = open(os.path.join(ROOT, "include", "gzcuda_ode.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    decl = sorted(set(re.findall(r"\b(gzcuda_[a-z_0-9]+)\s*\(", text)) - {"gzcuda_surface_fn"})
    out = subprocess.check_output(["nm", "-D", "--defined-only", path]).decode()
    exported = set(line.split()[-1] for line in out.splitlines() if line.strip())
    assert len(decl) >= 8 and not [s for s in decl if s not in exported]
    undefined = subprocess.check_output(["nm", "-D", "--undefined-only", path]).decode()
    # nothing of the oracle is linked in: the only ODE symbol it needs comes from the host process
    assert "gzref" not in undefined and "dGeomMoved" in undefined


def test_no_torch_or_cxx_types_in_signatures():
    text = open(os.path.join(ROOT, "include", "gzcuda.h")).read()
    assert "std::" not in text and "torch" not in text and "at::" not in text


def test_library_loads_and_reports_version(gzlib):
    assert gzlib.gzcuda_version() >= 100
    assert gzlib.gzcuda_device_count() >= 0


def test_create_without_gpu_fails_loudly(gzlib):
    if gzlib.gzcuda_device_count() > 0:
        pytest.skip("a CUDA device is present")
    from gazebo_b200.batch import Batch, GzCudaError
    from tests import scenes
    sc, p, _ = scenes.box_drop()
    with pytest.raises(GzCudaError) as e:
        Batch(sc, p)
    assert "no CPU fallback" in str(e.value)


def test_null_arguments_return_error_codes(gzlib):
    gzlib.gzcuda_batch_step.restype = C.c_int
    assert gzlib.gzcuda_batch_step(None, 1) < 0
    assert gzlib.gzcuda_batch_sync(None) < 0
    assert b"batch" in gzlib.gzcuda_last_error()
    assert gzlib.gzcuda_world_step(None, C.c_double(0.001)) == 0   # dWorldQuickStep convention: 0 = failure


def test_product_never_imports_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/."""
    bad = re.compile(r"from\s+oracle|import\s+oracle|#include\s+\".*oracle|libgzref|liboderef|refharness|dlopen")
    for pkg in (os.path.join(ROOT, "gazebo_b200"), os.path.join(ROOT, "host"), os.path.join(ROOT, "include")):
        for dirpath, _, files in os.walk(pkg):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                    src = open(os.path.join(dirpath, f)).read()
                    assert not bad.search(src), f
