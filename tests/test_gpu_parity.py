"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): per-pixel max abs error <= 1e-3 and image relMSE <= 1e-5 in linear RGB on the
per-pixel mean; this suite also holds the FP64 path to much tighter bars: neighbourhood size N exact,
alpha/beta/W_r_c within 1e-9, per-sample colours within 1e-9 relative."""
import numpy as np
import pytest

import oracle_bind as ob

pytestmark = pytest.mark.gpu

TOL_MAXABS = 1e-3   # north_star
TOL_RELMSE = 1e-5   # north_star
TOL_TIGHT = 1e-9    # what the FP64 kernels actually deliver


def _gpu(planes, W, H, P, boxes, seed, flags=7, colors=None, general=False, large=False):
    import rpf_b200
    f = rpf_b200.RPFFilter(W, H, P, flags=flags, debug=True, force_general=general, force_large=large)
    f.upload(planes)
    if colors is not None:
        f.upload_colors_f64(colors)
    f.apply(boxes, seed)
    out = dict(colors=f.download_colors_f64(), **f.download_debug())
    out["stats"] = f.stats()
    out["image"] = f.reconstruct_box()
    f.close()
    return out


def _compare(g, o, W, H, P):
    assert np.array_equal(g["N"], o["N"]), "neighbourhood sizes differ"
    for k in ("alpha", "beta", "wrc"):
        a, b = g[k], o[k]
        assert np.array_equal(np.isnan(a), np.isnan(b)), k
        assert np.nanmax(np.abs(a - b), initial=0.0) <= TOL_TIGHT, (k, np.nanmax(np.abs(a - b)))
    gc, oc = g["colors"], o["colors"]
    assert np.array_equal(np.isnan(gc), np.isnan(oc))
    m = ~np.isnan(oc)
    err = np.abs(gc[m] - oc[m])
    assert err.max(initial=0.0) <= TOL_TIGHT * max(1.0, np.abs(oc[m]).max(initial=0.0)), err.max()
    # the north_star metric: per-pixel mean image
    gi = np.nan_to_num(gc.reshape(3, H * W, P).mean(axis=2))
    oi = np.nan_to_num(oc.reshape(3, H * W, P).mean(axis=2))
    assert np.abs(gi - oi).max() <= TOL_MAXABS
    relmse = np.mean((gi - oi) ** 2 / (oi ** 2 + 1e-2))
    assert relmse <= TOL_RELMSE
    assert g["stats"]["nan_channels"] == o["nan_count"]


PATHS = pytest.mark.parametrize("general", [False, True], ids=["fused", "general"])


@pytest.fixture(autouse=True)
def _fused_kernel_up_to_its_limit(request, monkeypatch):
    """By default neighbourhoods above 512 candidates go to the general path (rpf_fused_max_candidates: it is faster there);
    the cases labelled 'fused' keep the fused kernel covered up to its hard limit of 1,279 candidates."""
    if "fused" in request.node.name:
        monkeypatch.setenv("RPF_FUSED_MAX", "1279")


def test_default_crossover_between_the_fused_and_the_general_path(monkeypatch):
    """box 7 x 8 spp (392 candidates) runs fused, box 9 x 8 spp (648) general; RPF_FUSED_MAX moves the line; same colours."""
    import rpf_b200
    monkeypatch.delenv("RPF_FUSED_MAX", raising=False)
    assert rpf_b200.lib().rpf_fused_max_candidates() == 512
    W, H, P = 24, 20, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=3, kind=0)

    def run(box):
        f = rpf_b200.RPFFilter(W, H, P)
        f.set_profiling(True)
        f.upload(planes)
        f.apply([box], 0.3)
        f.synchronize()
        kt = f.kernel_times()
        col = f.download_colors_f64()
        f.close()
        return kt, col

    kt7, _ = run(7)
    kt9, c9 = run(9)
    assert kt7["rpf_pixel_kernel"][1] == 1 and kt7["dependency_kernel"][1] == 0
    assert kt9["rpf_pixel_kernel"][1] == 0 and kt9["dependency_kernel"][1] >= 1
    monkeypatch.setenv("RPF_FUSED_MAX", "1279")
    kt9f, c9f = run(9)
    assert kt9f["rpf_pixel_kernel"][1] == 1 and kt9f["dependency_kernel"][1] == 0
    m = ~np.isnan(c9)
    assert np.array_equal(np.isnan(c9), np.isnan(c9f)) and np.abs(c9[m] - c9f[m]).max() <= TOL_TIGHT * max(1.0, np.abs(c9[m]).max())
    monkeypatch.setenv("RPF_FUSED_MAX", "0")
    kt7g, _ = run(7)
    assert kt7g["rpf_pixel_kernel"][1] == 0 and kt7g["dependency_kernel"][1] >= 1


@PATHS
@pytest.mark.parametrize("seed", [0.002, 0.2, 2.0])
@pytest.mark.parametrize("kind", [0, 1])
def test_box7_matches_oracle(seed, kind, general):
    import rpf_b200
    W, H, P = 40, 28, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=1234, kind=kind)
    o = ob.oracle_filter(planes, W, H, P, [7], seed=seed, debug=True)
    g = _gpu(planes, W, H, P, [7], seed, general=general)
    _compare(g, o, W, H, P)
    if kind == 0 and seed > 0.1:
        assert np.abs(g["colors"] - planes[2:5]).max() > 1e-2  # the filter is not the identity here


@PATHS
@pytest.mark.parametrize("boxes,P", [([3], 8), ([5], 4), ([9, 7, 5], 8), ([7], 16), ([7, 7], 3), ([11], 1)])
def test_box_and_spp_variants(boxes, P, general):
    import rpf_b200
    W, H = 24, 20
    planes = rpf_b200.synth_planes(W, H, P, seed=7, kind=0)
    o = ob.oracle_filter(planes, W, H, P, boxes, seed=0.5, debug=True)
    g = _gpu(planes, W, H, P, boxes, 0.5, general=general)
    _compare(g, o, W, H, P)


# Large neighbourhoods (general path). What each case exercises in the dependency kernel's plan (rpf_api.cu plan_general):
#   box 17 x 8 spp   N <= 2312   bins in shared memory, 4 histograms at a time
#   box 35 x 8 spp   ncap 9800   bins in the L2-resident scratch, 4 histograms
#   box 57 x 8 spp   ncap 25992  2 histograms at a time
#   box 41 x 32 spp  ncap 53792  1 histogram at a time
#   box 13 x 64 spp  ncap 10816  64 own samples per pixel (8 blend passes)
@pytest.mark.parametrize("W,H,P,boxes,seed,kind", [
    (40, 36, 8, [17], 0.5, 0),
    (36, 30, 8, [17, 7], 0.02, 1),
    (24, 20, 8, [35], 1.0, 0),
    (24, 20, 8, [57, 35], 0.5, 0),
    (20, 16, 32, [41], 0.5, 0),
    (14, 12, 64, [13], 0.2, 0),
    (30, 26, 16, [21, 9], 2.0, 0),
])
def test_large_neighbourhoods(W, H, P, boxes, seed, kind):
    import rpf_b200
    planes = rpf_b200.synth_planes(W, H, P, seed=31, kind=kind)
    o = ob.oracle_filter(planes, W, H, P, boxes, seed=seed, debug=True)
    g = _gpu(planes, W, H, P, boxes, seed)
    _compare(g, o, W, H, P)


def test_box55_full_neighbourhood():
    """north_star's largest box at 8 spp on a frame big enough that interior pixels see all 24,200 candidates."""
    import rpf_b200
    W, H, P = 60, 58, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=55, kind=0)
    o = ob.oracle_filter(planes, W, H, P, [55], seed=0.5, debug=True)
    g = _gpu(planes, W, H, P, [55], 0.5)
    assert g["N"].max() > 5000
    _compare(g, o, W, H, P)


# The large-neighbourhood dependency kernel (rpf_large.cuh: u32 candidate numbers, u16 bins, u32 cells, slabbed tables).
@pytest.mark.parametrize("W,H,P,boxes,seed,kind", [
    (24, 20, 8, [7], 0.5, 0),          # forced: four small tables per pass
    (30, 24, 8, [17, 9], 0.2, 1),      # forced, degenerate content (flat channels, exact-zero MI recomputation)
    (20, 16, 32, [41], 0.5, 0),        # forced, same case as the general path above
    (12, 10, 64, [35], 0.5, 0),        # 78,400 candidates: taken without forcing (BASELINE.json configs[4], box 35)
    (8, 6, 64, [55], 1.0, 0),          # 193,600 candidates (configs[4], box 55)
])
def test_large_kernel_matches_oracle(W, H, P, boxes, seed, kind):
    import rpf_b200
    planes = rpf_b200.synth_planes(W, H, P, seed=77, kind=kind)
    o = ob.oracle_filter(planes, W, H, P, boxes, seed=seed, debug=True)
    g = _gpu(planes, W, H, P, boxes, seed, large=True)
    _compare(g, o, W, H, P)


@pytest.mark.parametrize("W,H,P,boxes,seed", [
    (12, 10, 64, [35], 0.5),      # 78,400 candidates, N <= 7,680: dependency_kernel with the 32-bit candidate list in HBM
    (10, 8, 64, [55, 35], 1.0),   # 193,600 / 78,400 candidates
    (16, 12, 32, [47], 0.3),      # 70,688 candidates
])
def test_many_candidates_few_accepted(W, H, P, boxes, seed):
    """box*box*spp beyond 65,535 while every pixel accepts fewer samples: the regular dependency kernel serves them."""
    import rpf_b200
    planes = rpf_b200.synth_planes(W, H, P, seed=91, kind=0)
    o = ob.oracle_filter(planes, W, H, P, boxes, seed=seed, debug=True)
    g = _gpu(planes, W, H, P, boxes, seed)
    _compare(g, o, W, H, P)


@pytest.mark.parametrize("rows", [(18, 22), (0, 2)], ids=["all-large", "mixed"])
def test_large_kernel_beyond_255_bins(rows):
    """N > 65,536 accepted samples: more than 255 bins per axis (u16 bins), and a joint table (nb*nb u32 cells) that
    no longer fits in shared memory at once, so it is built in row slabs. Smooth features make every candidate pass
    the 3-sigma test; 40x40 px x 64 spp inside a 55 box gives N up to 102,400 -> 320 bins, 409 KB per table."""
    import rpf_b200
    W, H, P = 40, 40, 64
    planes = rpf_b200.synth_planes(W, H, P, seed=9, kind=0)
    planes[7:19] = (planes[7:19] * 0.0 + np.random.default_rng(5).normal(size=(12, W * H * P))).astype(np.float32)
    # rows 18-21: every pixel accepts more than 65,535 samples (dependency_large_kernel); rows 0-1: corner pixels stay
    # below (dependency_kernel), the middle of the row above, so both kernels share one pass
    o = ob.oracle_filter(planes, W, H, P, [55], seed=0.5, debug=True, rows=rows)
    import rpf_b200 as rb
    f = rb.RPFFilter(W, H, P, debug=True)
    f.upload(planes)
    f.filter_rows(55, rows[0], rows[1], 0.5)
    f.swap_colors()
    got = f.download_colors_f64()
    dbg = f.download_debug()
    f.close()
    sel = np.zeros(W * H, bool)
    sel[rows[0] * W:rows[1] * W] = True
    assert np.array_equal(dbg["N"][sel], o["N"][sel])
    assert dbg["N"][sel].max() > 65536
    if rows[0] == 0:
        assert dbg["N"][sel].min() <= 65535
    for k in ("alpha", "beta", "wrc"):
        assert np.abs(dbg[k][sel] - o[k][sel]).max() <= TOL_TIGHT, k
    ssel = np.repeat(sel, P)
    assert np.abs(got[:, ssel] - o["colors"][:, ssel]).max() <= TOL_TIGHT


def test_unsupported_neighbourhood_is_an_error():
    import rpf_b200
    f = rpf_b200.RPFFilter(4, 4, 64)
    f.upload(rpf_b200.synth_planes(4, 4, 64, seed=1, kind=0))
    with pytest.raises(rpf_b200.RPFError, match="exceeds this build"):
        f.apply([129], 0.5)   # 129*129*64 = 1,065,024 candidates > 2^20
    f.close()


def test_verbatim_flags_match_reference_semantics():
    """Without P1-P3 the reference leaves beta[3..11] undefined (NaN here) and 0/0 stays NaN."""
    import rpf_b200
    W, H, P = 20, 16, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=3, kind=0)
    o = ob.oracle_filter(planes, W, H, P, [7], seed=0.002, flags=0, debug=True)
    g = _gpu(planes, W, H, P, [7], 0.002, flags=rpf_b200.FLAGS_VERBATIM)
    assert np.array_equal(g["N"], o["N"])
    assert np.abs(g["alpha"] - o["alpha"]).max() <= TOL_TIGHT
    assert np.abs(g["beta"][:, :3] - o["beta"][:, :3]).max() <= TOL_TIGHT
    assert np.isnan(g["beta"][:, 3:]).all()


@PATHS
def test_tiny_and_ragged_frames(general):
    import rpf_b200
    for (W, H, P) in [(1, 1, 8), (2, 9, 8), (7, 3, 5), (1, 1, 1), (3, 2, 200)]:
        planes = rpf_b200.synth_planes(W, H, P, seed=11, kind=0)
        o = ob.oracle_filter(planes, W, H, P, [7], seed=1.0, debug=True)
        g = _gpu(planes, W, H, P, [7], 1.0, general=general)
        _compare(g, o, W, H, P)


@PATHS
def test_row_band_equals_full_frame(general):
    """filter_rows over two bands (what two ranks would do) equals one full-frame pass."""
    import rpf_b200
    W, H, P = 32, 24, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=5, kind=0)
    f = rpf_b200.RPFFilter(W, H, P, force_general=general)
    f.upload(planes)
    f.apply([7], 0.5)
    full = f.download_colors_f64()
    f.upload(planes)
    f.filter_rows(7, 0, 10, 0.5)
    f.filter_rows(7, 10, H, 0.5)
    f.swap_colors()
    bands = f.download_colors_f64()
    f.close()
    assert np.array_equal(full, bands)


def test_host_cpp_through_reference_samplingfilm():
    """The product's host C++ (raytracer-rpf_b200/host/rpf_b200_host.cpp) packing a real reference SamplingFilm, as Render would."""
    import ctypes as C
    import os
    import rpf_b200
    path = os.path.join(ob.ORACLE_DIR, "_ref", "libhost_harness.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libhost_harness.so was not built (needs /root/reference at build time)")
    lib = C.CDLL(path)
    W, H, P, boxes, seed = 28, 22, 8, [9, 7], 0.5
    planes = rpf_b200.synth_planes(W, H, P, seed=21, kind=0)
    col = planes[2:5].astype(np.float64).copy()
    bx = np.asarray(boxes, np.int32)
    mean, nans, err = C.c_double(), C.c_longlong(), C.create_string_buffer(512)
    rc = lib.host_apply_rpf_b200(W, H, P, planes.ctypes.data_as(C.c_void_p), col.ctypes.data_as(C.c_void_p),
                                 bx.ctypes.data_as(C.c_void_p), len(bx), C.c_double(seed), 0, C.byref(mean), C.byref(nans), err, 512)
    assert rc == 0, err.value
    o = ob.oracle_filter(planes, W, H, P, boxes, seed=seed)
    assert np.abs(col - o["colors"]).max() <= TOL_TIGHT
    assert nans.value == o["nan_count"] and mean.value > P


def test_row_chunks_and_staging_blocks_do_not_change_the_result(monkeypatch):
    """The general path works through the frame in chunks of rows sized by a scratch budget (RPF_SCRATCH_MB), and the
    membership kernel stages a column of neighbour pixels in blocks of rows sized by RPF_MEMB_SMEM_KB. Tiny budgets
    force many chunks and blocks; the colours must be bit-identical to the default configuration."""
    import rpf_b200
    W, H, P = 40, 33, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=17, kind=0)

    def run():
        f = rpf_b200.RPFFilter(W, H, P)
        f.upload(planes)
        f.apply([17, 9], 0.5)
        out = f.download_colors_f64()
        launches = f.stats()["kernel_launches"]
        f.close()
        return out, launches

    ref, n_ref = run()
    monkeypatch.setenv("RPF_SCRATCH_MB", "1")     # 1 MiB of mask + records: a few rows per chunk
    monkeypatch.setenv("RPF_MEMB_SMEM_KB", "4")   # a handful of neighbour rows per staging block
    got, n_got = run()
    assert n_got > n_ref                          # really chunked
    assert np.array_equal(ref, got)


def test_full_size_frame_spot_checked_against_oracle():
    """BASELINE.json configs[1] at full size (1280x720, 8 spp, boxes 55/35/17/7). The oracle cannot filter the whole
    frame in test time, but the filter is local: a pixel after the four passes depends on samples within 27+17+8+3 = 55
    pixels. So the oracle filters a 134x134 crop generated on its own (the synthetic generator is counter based) and its
    central 24x24 pixels must equal the GPU's full-frame result. Also: two GPU runs are bit-identical."""
    import rpf_b200
    W, H, P, boxes, seed = 1280, 720, 8, [55, 35, 17, 7], 0.5
    halo = sum((b - 1) // 2 for b in boxes)
    win, x0, y0 = 24, 611, 333
    planes = rpf_b200.synth_planes(W, H, P, seed=1234, kind=0)
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    f.apply(boxes, seed)
    got = f.download_colors_f64()
    f.upload(planes)
    f.apply(boxes, seed)
    again = f.download_colors_f64()
    st = f.stats()
    f.close()
    assert np.array_equal(got, again), "two runs differ: the atomics-ordered sums are supposed to be order independent"
    assert st["nan_channels"] == 0
    cw = win + 2 * halo
    crop = rpf_b200.synth_planes(cw, cw, P, seed=1234, kind=0, x_off=x0 - halo, y_off=y0 - halo)
    full = planes.reshape(19, H, W, P)[:, y0 - halo:y0 + win + halo, x0 - halo:x0 + win + halo]
    assert np.array_equal(crop.reshape(19, cw, cw, P), full)  # the crop is the same data
    o = ob.oracle_filter(crop, cw, cw, P, boxes, seed=seed)["colors"].reshape(3, cw, cw, P)[:, halo:halo + win, halo:halo + win]
    g = got.reshape(3, H, W, P)[:, y0:y0 + win, x0:x0 + win]
    assert np.abs(g - o).max() <= TOL_TIGHT * max(1.0, np.abs(o).max())
    assert np.abs(g - planes[2:5].reshape(3, H, W, P)[:, y0:y0 + win, x0:x0 + win]).max() > 1e-3


# Neighbour subsampling (rpf_set_subsampling; not in the reference): the oracle and the kernels call the same
# counter-based RNG (include/rpf_subsample.h), so they must agree on N exactly and on everything else as tightly as before.
@pytest.mark.parametrize("W,H,P,boxes,keep,general", [
    (40, 28, 8, [7], 0.5, False),        # fused kernel
    (40, 28, 8, [7], 0.5, True),         # membership kernel
    (36, 30, 8, [17, 7], 0.25, False),   # general pass then fused pass: a different subset per box size
    (12, 10, 64, [35], 0.1, False),      # more than 65,535 candidates, 64 samples per pixel (two staging chunks of 32)
    (24, 20, 8, [9], 1.0, False),        # keep = 1: off
])
def test_subsampled_neighbourhoods_match_oracle(W, H, P, boxes, keep, general):
    import rpf_b200
    planes = rpf_b200.synth_planes(W, H, P, seed=51, kind=0)
    o = ob.oracle_filter(planes, W, H, P, boxes, seed=0.5, debug=True, subsample=(keep, 99))
    f = rpf_b200.RPFFilter(W, H, P, debug=True, force_general=general)
    f.set_subsampling(keep, 99)
    f.upload(planes)
    f.apply(boxes, 0.5)
    g = dict(colors=f.download_colors_f64(), **f.download_debug())
    g["stats"] = f.stats()
    f.close()
    _compare(g, o, W, H, P)
    full = ob.oracle_filter(planes, W, H, P, boxes, seed=0.5, debug=True)
    if keep < 1.0:
        assert g["N"].sum() < 0.8 * full["N"].sum()
    else:
        assert np.array_equal(g["N"], full["N"])


def test_subsampling_is_the_same_in_a_row_band():
    """A band with global_row0 set draws the stream of the full frame (global coordinates)."""
    import rpf_b200
    W, H, P, y0, y1, halo = 32, 30, 8, 10, 20, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=5, kind=0)
    f = rpf_b200.RPFFilter(W, H, P)
    f.set_subsampling(0.4, 7)
    f.upload(planes)
    f.apply([17], 0.5)
    full = f.download_colors_f64().reshape(3, H, W * P)[:, y0:y1]
    f.close()
    b0, b1 = y0 - halo, y1 + halo
    band_planes = np.ascontiguousarray(planes.reshape(19, H, W * P)[:, b0:b1]).reshape(19, -1)
    f = rpf_b200.RPFFilter(W, b1 - b0, P)
    f.set_subsampling(0.4, 7, global_row0=b0)
    f.upload(band_planes)
    f.filter_rows(17, y0 - b0, y1 - b0, 0.5)
    f.swap_colors()
    band = f.download_colors_f64().reshape(3, b1 - b0, W * P)[:, y0 - b0:y1 - b0]
    f.close()
    assert np.array_equal(full, band)
