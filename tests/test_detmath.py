"""include/gz_detmath.h on the host: accuracy against mpmath and IEEE special cases.
(Device-vs-host bit equality is covered by the GPU parity tests: any mismatch in atan2 or
sin would show up as a non-zero trajectory difference on the actuated robots.)"""
import ctypes as C
import math
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def dm(tmp_path_factory):
    d = tmp_path_factory.mktemp("dm")
    src = d / "t.c"
    src.write_text('#include "%s/include/gz_detmath.h"\n'
                   "double c_atan2(double y,double x){return gzdet_atan2(y,x);}\n"
                   "double c_sin(double x){return gzdet_sin(x);}\n"
                   "double c_cos(double x){return gzdet_cos(x);}\n" % ROOT)
    so = d / "libdm.so"
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-o", str(so), str(src)])
    L = C.CDLL(str(so))
    for f in (L.c_atan2, L.c_sin, L.c_cos):
        f.restype = C.c_double
    L.c_atan2.argtypes = [C.c_double, C.c_double]
    L.c_sin.argtypes = [C.c_double]
    L.c_cos.argtypes = [C.c_double]
    return L


def _ulps(got, exact):
    import mpmath as mp
    e = float(exact)
    u = math.ulp(e) if e != 0 else 5e-324
    return abs(float((mp.mpf(got) - exact) / u))


def test_atan2_accuracy(dm):
    mp = pytest.importorskip("mpmath")
    mp.mp.dps = 40
    rng = np.random.default_rng(1)
    worst = 0.0
    for _ in range(5000):
        y = rng.normal() * 10 ** rng.uniform(-6, 3)
        x = rng.normal() * 10 ** rng.uniform(-6, 3)
        worst = max(worst, _ulps(dm.c_atan2(y, x), mp.atan2(mp.mpf(y), mp.mpf(x))))
    assert worst < 1.6


def test_sin_cos_accuracy(dm):
    mp = pytest.importorskip("mpmath")
    mp.mp.dps = 40
    rng = np.random.default_rng(2)
    worst = 0.0
    for i in range(5000):
        x = rng.uniform(-40, 40) if i % 2 else rng.uniform(-1e4, 1e4)
        worst = max(worst, _ulps(dm.c_sin(x), mp.sin(mp.mpf(x))), _ulps(dm.c_cos(x), mp.cos(mp.mpf(x))))
    assert worst < 1.0


def test_special_values(dm):
    assert dm.c_atan2(0.0, 1.0) == 0.0
    assert dm.c_atan2(0.0, -1.0) == math.pi
    assert dm.c_atan2(-0.0, -1.0) == -math.pi
    assert dm.c_atan2(1.0, 0.0) == math.pi / 2
    assert dm.c_atan2(-1.0, 0.0) == -math.pi / 2
    assert dm.c_atan2(1.0, 1.0) == math.pi / 4
    assert dm.c_atan2(math.inf, math.inf) == math.pi / 4
    assert math.isnan(dm.c_atan2(math.nan, 1.0))
    assert dm.c_sin(0.0) == 0.0 and dm.c_cos(0.0) == 1.0
    assert math.isnan(dm.c_sin(math.nan))
