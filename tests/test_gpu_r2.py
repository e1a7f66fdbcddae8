"""GPU parity, round 2: the fixtures that pin the oracle where the big kernels live (tests/make_golden_r2.py), non-finite
input, crops of the RENDERED feature buffer of scenes/sample.pbrt, the component-level MI probe, row ranges in any order,
and row bands inside one process (rpf_config.n_gpus). Everything goes through the C ABI; the oracle is only the checker."""
import glob
import os

import numpy as np
import pytest

import oracle_bind as ob
from test_gpu_parity import TOL_TIGHT, _compare, _gpu

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _close(gc, oc, tol=TOL_TIGHT):
    assert np.array_equal(np.isnan(gc), np.isnan(oc)), "NaN patterns differ"
    m = np.isfinite(oc)
    assert np.array_equal(gc[~m & ~np.isnan(oc)], oc[~m & ~np.isnan(oc)]), "infinities differ"
    assert np.abs(gc[m] - oc[m]).max(initial=0.0) <= tol * max(1.0, np.abs(oc[m]).max(initial=0.0))


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "big_case*.npz"))), ids=os.path.basename)
def test_big_fixtures_gpu_vs_reference_output(path):
    """The fixture's colours are the REFERENCE's own (REF_DEFINED build); the GPU is held to them directly, and to the
    oracle for N / alpha / beta / W_r_c."""
    g = np.load(path)
    W, H, P, _ = (int(v) for v in g["meta"])
    boxes, seed = [int(b) for b in g["boxes"]], float(g["seed"])
    o = ob.oracle_filter(g["planes"], W, H, P, boxes, seed=seed, debug=True, threads=8)
    r = _gpu(g["planes"], W, H, P, boxes, seed)
    _compare(r, o, W, H, P)
    _close(r["colors"], g["colors"])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "nonfinite_case*.npz"))), ids=os.path.basename)
@pytest.mark.parametrize("general", [False, True], ids=["default", "general"])
def test_nonfinite_input(path, general):
    """Inf / NaN colours and NaN features: same NaN pattern, same NaN count, same finite values as the reference build."""
    g = np.load(path)
    W, H, P, box = (int(v) for v in g["meta"])
    r = _gpu(g["planes"], W, H, P, [box], float(g["seed"]), general=general)
    _close(r["colors"], g["colors"])
    assert r["stats"]["nan_channels"] == int(np.isnan(g["colors"]).sum())


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "real_crop*.npz"))), ids=os.path.basename)
@pytest.mark.parametrize("general", [False, True], ids=["default", "general"])
def test_rendered_feature_buffer(path, general):
    """Real content (scenes/sample.pbrt through the replacement integrator's dump hook): sigma = 0 features, black
    background -- where P2 / P3 and the exact-zero MI recompute matter."""
    g = np.load(path)
    W, H, P = (int(v) for v in g["meta"][:3])
    boxes, seed = [int(b) for b in g["boxes"]], float(g["seed"])
    o = ob.oracle_filter(g["planes"], W, H, P, boxes, seed=seed, debug=True, threads=8)
    r = _gpu(g["planes"], W, H, P, boxes, seed, general=general)
    _compare(r, o, W, H, P)
    _close(r["colors"], g["colors"])
    assert r["stats"]["nan_channels"] == int(np.isnan(g["colors"]).sum())


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "probe_case*.npz"))), ids=os.path.basename)
@pytest.mark.parametrize("route", ["default", "general", "large"])
def test_mi_probe_counts_exact(path, route):
    """histogram counts exact, MI known answers -- on every dependency path (fused / general / large)."""
    import rpf_b200
    g = np.load(path)
    W, H, P, box, px, py, kind, seed = (int(v) for v in g["meta"])
    planes = rpf_b200.synth_planes(W, H, P, seed=seed, kind=kind)
    o = ob.oracle_pixel_probe(planes, W, H, P, box, px, py, seed=0.5)
    f = rpf_b200.RPFFilter(W, H, P, force_general=route == "general", force_large=route == "large")
    f.upload(planes)
    f.set_debug_pixel(px, py)
    f.apply([box], 0.5)
    r = f.download_debug_mi()
    f.close()
    assert r["N"] == o["N"] and r["bins"] == o["bins"]
    assert np.array_equal(r["marginals"].astype(np.int64), o["marginals"].astype(np.int64))
    # the GPU sums n ln n in fixed point instead of p log(p / (px py)) cell by cell: equal to mi.cpp within 1e-12
    assert np.abs(r["mi"] - g["ref_mi"]).max() <= 1e-12, np.abs(r["mi"] - g["ref_mi"]).max()
    assert np.array_equal(r["mi"] == 0.0, g["ref_mi"] == 0.0)


@pytest.mark.parametrize("boxes", [[7], [17]])
def test_row_ranges_in_any_order(boxes):
    """rows [0,a), [b,H), then [a,b) -- the per-pixel statistics must be valid for every row whatever the order
    (they once were cached as one merged interval)."""
    import rpf_b200
    W, H, P = 28, 30, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=21, kind=0)
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    f.apply(boxes, 0.5)
    whole = f.download_colors_f64()
    f.close()
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    for y0, y1 in ((0, 9), (22, H), (9, 22)):
        f.filter_rows(boxes[0], y0, y1, 0.5)
    f.swap_colors()
    parts = f.download_colors_f64()
    f.close()
    assert np.array_equal(whole, parts)


@pytest.mark.parametrize("box,ranges", [(17, (0, 8, 24, 32)), (17, (8, 16, 40, 48)), (7, (0, 8, 40, 48)), (35, (0, 16, 32, 48))])
def test_two_row_ranges_in_one_set_of_launches(box, ranges, monkeypatch):
    """The band schedule filters the rows next to both band edges as ONE chunk with a gap (rpf_filter_rows2), then the
    interior: together bit-identical to the whole frame. A tiny scratch budget puts chunk boundaries and the gap in
    different places."""
    import rpf_b200
    W, H, P = 40, 48, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=31, kind=0)
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    f.apply([box], 0.5)
    whole = f.download_colors_f64()
    f.close()
    y0, y1, y2, y3 = ranges
    for budget in (None, "1"):
        if budget:
            monkeypatch.setenv("RPF_SCRATCH_MB", budget)
        f = rpf_b200.RPFFilter(W, H, P)
        f.upload(planes)
        f.filter_rows2(box, y0, y1, y2, y3, 0.5)
        if y0 > 0:
            f.filter_rows(box, 0, y0, 0.5)
        f.filter_rows(box, y1, y2, 0.5)
        if y3 < H:
            f.filter_rows(box, y3, H, 0.5)
        f.swap_colors()
        parts = f.download_colors_f64()
        f.close()
        assert np.array_equal(whole, parts), (box, ranges, budget)


def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("n", [2, 4])
def test_row_bands_inside_one_process(n):
    """rpf_config.n_gpus: one band per device, peer copies of the colour halos, edge rows first. Bit-identical to one GPU."""
    import rpf_b200
    if _n_gpus() < n:
        pytest.skip(f"needs {n} GPUs")
    W, H, P = 36, 20 * n, 8
    boxes = [9, 7, 5]
    planes = rpf_b200.synth_planes(W, H, P, seed=99, kind=0)
    one = rpf_b200.RPFFilter(W, H, P, debug=True)
    one.upload(planes)
    one.apply(boxes, 0.5)
    ref = one.download_colors_f64()
    refdbg = one.download_debug()
    one.close()
    f = rpf_b200.RPFFilter(W, H, P, debug=True, devices=list(range(n)), max_box=9)
    f.upload(planes)
    f.apply(boxes, 0.5)
    got = f.download_colors_f64()
    dbg = f.download_debug()
    st = f.stats()
    f.close()
    assert np.array_equal(got, ref)
    assert np.array_equal(dbg["N"], refdbg["N"])
    assert st["halo_bytes"] > 0 and st["pixels_filtered"] == W * H * len(boxes)
    o = ob.oracle_filter(planes, W, H, P, boxes, seed=0.5, threads=8)
    _close(got, o["colors"])


def test_single_device_band_schedule_equals_plain_filter():
    """rpf_band_attach with world 1 and the edge-first row order of the band schedule on ONE GPU (the schedule itself, no
    exchange): bit-identical to rpf_filter."""
    import rpf_b200
    W, H, P = 30, 26, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=5, kind=0)
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    f.apply([17, 7], 0.5)
    a = f.download_colors_f64()
    f.close()
    f = rpf_b200.RPFFilter(W, H, P)
    f.attach_band(H, 0, 0, H, 0, 1)
    f.upload(planes)
    f.apply([17, 7], 0.5)
    b = f.download_colors_f64()
    f.close()
    assert np.array_equal(a, b)
