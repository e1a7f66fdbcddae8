"""Film reconstruction after the filter (rpf.cpp:776-804 -> core/film.h:121-161, film.cpp:69-77,
117-130,169-205). Fixtures tests/golden/film_case*.npz come from the reference's own lines compiled in place
(oracle/_ref/libref_film.so, tests/make_golden.py film). Float arithmetic: the bar is bit-identical."""
import glob
import os

import numpy as np
import pytest

import oracle_bind as ob

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "film_case*.npz")))


def _case(path):
    d = np.load(path)
    W, H, P, fk, x0, y0, x1, y1 = [int(v) for v in d["meta"]]
    rx, ry, alpha, scale, mx = [float(v) for v in d["fmeta"]]
    rw = d["ray_weight"] if d["ray_weight"].size else None
    return d, (W, H, P), fk, (rx, ry, alpha), (x0, y0, x1, y1), scale, mx, rw


def test_fixtures_exist():
    assert len(GOLD) >= 4


@pytest.mark.parametrize("path", GOLD, ids=os.path.basename)
def test_oracle_film_matches_reference_fixture(path):
    d, (W, H, P), fk, (rx, ry, alpha), bounds, scale, mx, rw = _case(path)
    table = ob.oracle_film_table(fk, rx, ry, alpha)
    assert np.array_equal(table, d["table"])
    rgb = ob.oracle_film(W, H, P, d["pfilm"], d["colors"], rw, table, rx, ry, bounds, scale, mx)
    assert np.array_equal(rgb, d["rgb"])


@pytest.mark.parametrize("path", GOLD, ids=os.path.basename)
def test_product_filter_table_matches_reference_fixture(path):
    """rpf_film_filter_table is host code of the product library (no GPU needed)."""
    import rpf_b200
    d, _, fk, (rx, ry, alpha), *_ = _case(path)
    assert np.array_equal(rpf_b200.film_filter_table(fk, rx, ry, alpha), d["table"])


def test_live_reference_film_when_built():
    """Where oracle/_ref/libref_film.so exists (build container), a fresh random case against the restatement."""
    if ob.ref_film() is None:
        pytest.skip("oracle/_ref/libref_film.so not built here")
    rng = np.random.default_rng(3)
    W, H, P = 13, 11, 5
    S = W * H * P
    px = np.repeat(np.arange(W * H) % W, P) + rng.uniform(size=S)
    py = np.repeat(np.arange(W * H) // W, P) + rng.uniform(size=S)
    pfilm = np.stack([px, py]).astype(np.float32)
    col = rng.uniform(0, 3, size=(3, S))
    rw = rng.uniform(size=S).astype(np.float32)
    for fk, rx, ry, alpha in [(0, 0.5, 0.5, 0.0), (1, 2.0, 2.0, 2.0), (1, 3.0, 1.0, 0.5)]:
        ref_rgb, ref_table = ob.ref_film_reconstruct(W, H, P, pfilm, col, rw, fk, rx, ry, alpha, (1, 0, W, H - 2), 0.7, 1.5)
        table = ob.oracle_film_table(fk, rx, ry, alpha)
        assert np.array_equal(table, ref_table)
        assert np.array_equal(ob.oracle_film(W, H, P, pfilm, col, rw, table, rx, ry, (1, 0, W, H - 2), 0.7, 1.5), ref_rgb)


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLD, ids=os.path.basename)
def test_gpu_film_matches_reference_fixture(path):
    import rpf_b200
    d, (W, H, P), fk, (rx, ry, alpha), bounds, scale, mx, rw = _case(path)
    planes = np.zeros((19, W * H * P), np.float32)
    planes[0:2] = d["pfilm"]
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    f.upload_colors_f64(d["colors"])
    rgb = f.reconstruct_film(rpf_b200.film_filter_table(fk, rx, ry, alpha), rx, ry, pixel_bounds=bounds, scale=scale,
                             max_sample_luminance=mx, ray_weight=rw)
    f.close()
    assert np.array_equal(rgb, d["rgb"])


@pytest.mark.gpu
def test_gpu_film_after_filter_matches_oracle():
    """Filter, then reconstruct with sample.pbrt's gaussian filter: the image the reference would write."""
    import rpf_b200
    W, H, P = 48, 36, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=12, kind=0)
    o = ob.oracle_filter(planes, W, H, P, [7], seed=0.5)
    table = ob.oracle_film_table(1, 2.0, 2.0, 2.0)
    want = ob.oracle_film(W, H, P, planes[0:2], o["colors"], None, table, 2.0, 2.0, (0, 0, W, H), 1.0, float("inf"))
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    f.apply([7], 0.5)
    got = f.reconstruct_film(rpf_b200.film_filter_table(1, 2.0, 2.0, 2.0), 2.0, 2.0)
    f.close()
    # colours differ from the oracle's at the 1e-12 level before the cast to Float: allow one float ulp of the result
    assert np.abs(got - want).max() <= 2e-6 * max(1.0, np.abs(want).max())


@pytest.mark.gpu
def test_host_cpp_filter_and_film():
    """Product host C++ (ApplyRPFFilterAndFilmB200 + MakeFilmDescB200) over a reference SamplingFilm with ray weights."""
    import ctypes as C
    import rpf_b200
    path = os.path.join(ob.ORACLE_DIR, "_ref", "libhost_harness.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libhost_harness.so was not built (needs /root/reference at build time)")
    lib = C.CDLL(path)
    W, H, P, boxes, seed = 26, 20, 8, [7], 0.5
    planes = rpf_b200.synth_planes(W, H, P, seed=33, kind=0)
    rw = np.random.default_rng(1).uniform(0.5, 1.0, W * H * P).astype(np.float32)
    table = ob.oracle_film_table(1, 2.0, 2.0, 2.0)
    rgb = np.empty((H, W, 3), np.float32)
    bx = np.asarray(boxes, np.int32)
    err = C.create_string_buffer(512)
    lib.host_apply_rpf_film_b200.argtypes = [C.c_int] * 3 + [C.c_void_p] * 3 + [C.c_int, C.c_double, C.c_float, C.c_float, C.c_void_p,
                                                                              C.c_float, C.c_float, C.c_void_p, C.c_char_p, C.c_int]
    rc = lib.host_apply_rpf_film_b200(W, H, P, planes.ctypes.data, rw.ctypes.data, bx.ctypes.data, len(bx), seed, 2.0, 2.0,
                                      table.ctypes.data, 1.0, float("inf"), rgb.ctypes.data, err, 512)
    assert rc == 0, err.value
    o = ob.oracle_filter(planes, W, H, P, boxes, seed=seed)
    want = ob.oracle_film(W, H, P, planes[0:2], o["colors"], rw, table, 2.0, 2.0, (0, 0, W, H), 1.0, float("inf"))
    assert np.abs(rgb - want).max() <= 2e-6 * max(1.0, np.abs(want).max())
