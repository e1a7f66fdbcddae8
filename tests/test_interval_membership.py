"""The 3-sigma membership test as an exact float interval (raytracer-rpf_b200/csrc/rpf_kernels.cuh: three_sigma_interval).

The reference rejects a candidate when |f - mu| >= 3 sigma in DOUBLE arithmetic for some feature (ops.h:100-107 allLessThan,
rpf.cpp:579-586), f a float sample value. membership_kernel instead rejects iff (f < lo || f > hi) with a float interval found
once per pixel and feature. This test compiles THE SOURCE TEXT of that device function for the host (the CUDA intrinsics it uses
replaced by their IEEE meaning) and checks, by brute force around the interval ends and on special values, that the two tests
agree for every float -- no GPU needed; the GPU parity tests then hold the kernel to exact neighbourhood counts."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "raytracer-rpf_b200", "csrc", "rpf_kernels.cuh")

SHIM = r"""
#include <cstring>
#include <cstdint>
#define __device__
#define __forceinline__ inline
static inline float __int_as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
static inline double __dsub_rn(double a, double b) { return a - b; }   // IEEE round-to-nearest subtraction (-ffp-contract=off)
%s
extern "C" void interval(double mu, double sd3, float *lo, float *hi) { three_sigma_interval(mu, sd3, *lo, *hi); }
"""


@pytest.fixture(scope="module")
def interval(tmp_path_factory):
    text = open(SRC).read()
    m = re.search(r"(__device__ __forceinline__ float ordered_float.*?\n}\n)(?=\n__global__ void pixel_stats_kernel)", text, re.S)
    assert m, "three_sigma_interval not found in rpf_kernels.cuh"
    d = tmp_path_factory.mktemp("interval")
    (d / "shim.cpp").write_text(SHIM % m.group(1))
    so = d / "libinterval.so"
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-o", str(so), str(d / "shim.cpp")], check=True)
    lib = C.CDLL(str(so))
    lib.interval.argtypes = [C.c_double, C.c_double, C.POINTER(C.c_float), C.POINTER(C.c_float)]

    def f(mu, sd3):
        lo, hi = C.c_float(), C.c_float()
        lib.interval(mu, sd3, C.byref(lo), C.byref(hi))
        return np.float32(lo.value), np.float32(hi.value)
    return f


def reference_rejects(f, mu, sd3):
    """ops.h:100-107 on one feature: reject when |f - mu| >= 3 sigma, all in double (NaN compares false: accepted)."""
    with np.errstate(invalid="ignore", over="ignore"):
        return np.abs(f.astype(np.float64) - np.float64(mu)) >= np.float64(sd3)


def interval_rejects(f, lo, hi):
    with np.errstate(invalid="ignore"):
        return (f < lo) | (f > hi)


def neighbours(x, n=6):
    """x and the n floats on either side of it (finite x)."""
    out = [np.float32(x)]
    a = b = np.float32(x)
    with np.errstate(over="ignore"):  # stepping past FLT_MAX gives inf: a value worth testing too
        for _ in range(n):
            a = np.nextafter(a, np.float32(-np.inf))
            b = np.nextafter(b, np.float32(np.inf))
            out += [a, b]
    return out


SPECIAL = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, np.finfo(np.float32).max, -np.finfo(np.float32).max,
                    np.finfo(np.float32).tiny, 1e-45, -1e-45, 1.0, -1.0], np.float32)


def check(interval, mu, sd3, extra=()):
    lo, hi = interval(mu, sd3)
    pts = list(SPECIAL) + list(extra)
    for e in (lo, hi):
        if np.isfinite(e):
            pts += neighbours(e)
    with np.errstate(over="ignore", invalid="ignore"):
        for c in (mu - sd3, mu + sd3, mu):
            c32 = np.float32(c)
            if np.isfinite(c32):
                pts += neighbours(c32)
    f = np.array(pts, np.float32)
    got, want = interval_rejects(f, lo, hi), reference_rejects(f, mu, sd3)
    assert np.array_equal(got, want), (mu, sd3, lo, hi, f[got != want][:5])


def test_interval_equals_reference_test_on_random_statistics(interval):
    rng = np.random.default_rng(5)
    for _ in range(400):
        scale = 10.0 ** rng.uniform(-30, 30)
        mu = float(rng.normal() * scale)
        sd3 = float(abs(rng.normal()) * scale * 10.0 ** rng.uniform(-12, 3))
        extra = (mu + sd3 * rng.uniform(-2, 2, 32)).astype(np.float32)
        check(interval, mu, sd3, extra)


def test_interval_means_of_floats_and_tiny_sigmas(interval):
    """mu as the reference computes it (a double mean of floats), sigma far below the float spacing at mu: the interval may
    hold a single float or none."""
    rng = np.random.default_rng(6)
    for _ in range(200):
        v = rng.normal(size=8).astype(np.float32) * np.float32(10.0 ** rng.uniform(-3, 3))
        mu = float(v.astype(np.float64).sum() / 8)
        for sd3 in (0.0, 1e-300, 1e-12 * abs(mu), 1e-8 * abs(mu), float(3 * v.astype(np.float64).std())):
            check(interval, mu, sd3, v)


def test_interval_special_statistics(interval):
    inf, nan, fmax = float("inf"), float("nan"), float(np.finfo(np.float32).max)
    lo, hi = interval(1.0, nan)            # NaN sigma: everything passes (rpf.cpp:579 on a NaN stdDev), infinities included
    assert lo == -np.inf and hi == np.inf
    lo, hi = interval(1.0, 0.0)            # sigma 0: nothing passes but a NaN feature
    assert lo == np.inf and hi == -np.inf
    lo, hi = interval(0.0, inf)            # infinite sigma: every finite float passes, the infinities do not
    assert lo == -np.float32(fmax) and hi == np.float32(fmax)
    for mu, sd3 in ((0.0, 1.0), (-0.0, 1e-40), (1e38, 3e38), (-3e38, 1e38), (5.0, 1e-3), (0.0, 5e-324), (1e-45, 1e-45),
                    (123.456, 0.0), (1.0, nan), (0.0, inf), (fmax, 1.0), (-fmax, fmax)):
        check(interval, mu, sd3)
