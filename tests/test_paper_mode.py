"""Paper mode: epsilon in the W denominators, the low-variance guard and the world-position factor of
the 3-sigma test, one sigma per iteration, spike removal. None of it is in the reference, so there is no reference output
to match as shipped. The in-filter options (epsilon, guard, factor, sigma per pass) are pinned against the reference's
own rpf.cpp:302-733 compiled with two one-line hooks (oracle/ref_paper.sed -> oracle/_ref/libref_rpf_paper.so; fixtures
tests/golden/paper_case*.npz), bit for bit; spike removal, which has no counterpart at all, against a numpy twin; plus
properties (all-zero options = the reference arithmetic, limits, monotonicity, chaining). The GPU tests hold the kernels
to the oracle as tightly as everywhere else."""
import numpy as np
import pytest

import oracle_bind as ob


def _frame(W=20, H=16, P=8, seed=11, kind=0):
    import rpf_b200
    return rpf_b200.synth_planes(W, H, P, seed=seed, kind=kind)


# ---------------------------------------------------------------- oracle (CPU) ----------------------------------------
def test_all_zero_options_are_the_reference_arithmetic():
    W, H, P = 20, 16, 8
    planes = _frame(W, H, P)
    a = ob.oracle_filter(planes, W, H, P, [7, 5], seed=0.3, debug=True)
    b = ob.oracle_filter(planes, W, H, P, [7, 5], seed=0.3, debug=True,
                         paper=dict(w_epsilon=0.0, min_sigma=0.0, world_pos_sigmas=0.0, hdr_clamp_sigmas=0.0))
    for k in ("colors", "N", "alpha", "beta", "wrc"):
        assert np.array_equal(a[k], b[k], equal_nan=True), k


def test_epsilon_limits():
    """A huge epsilon drives every W ratio to 0: alpha = 1, beta = 0, W_r_c = 0. A tiny one changes almost nothing."""
    W, H, P = 16, 12, 8
    planes = _frame(W, H, P)
    big = ob.oracle_filter(planes, W, H, P, [7], seed=0.3, debug=True, paper=dict(w_epsilon=1e30))
    assert np.allclose(big["alpha"], 1.0, atol=1e-12) and np.allclose(big["beta"], 0.0, atol=1e-12)
    assert np.allclose(big["wrc"], 0.0, atol=1e-12)
    base = ob.oracle_filter(planes, W, H, P, [7], seed=0.3, debug=True)
    tiny = ob.oracle_filter(planes, W, H, P, [7], seed=0.3, debug=True, paper=dict(w_epsilon=1e-12))
    assert np.array_equal(base["N"], tiny["N"])
    assert np.nanmax(np.abs(base["alpha"] - tiny["alpha"])) < 1e-6
    # epsilon removes every 0/0 even without the P2 guard
    verb = ob.oracle_filter(planes, W, H, P, [7], seed=0.3, flags=ob.P1 | ob.P3, debug=True,
                            paper=dict(w_epsilon=1e-7))
    assert not np.isnan(verb["alpha"]).any() and not np.isnan(verb["beta"]).any()


def test_guard_and_world_position_factor_only_ever_add_neighbours():
    W, H, P = 20, 16, 8
    planes = _frame(W, H, P, kind=1)
    n0 = ob.oracle_filter(planes, W, H, P, [9], seed=0.3, debug=True)["N"]
    n1 = ob.oracle_filter(planes, W, H, P, [9], seed=0.3, debug=True, paper=dict(min_sigma=0.1))["N"]
    n2 = ob.oracle_filter(planes, W, H, P, [9], seed=0.3, debug=True, paper=dict(min_sigma=0.1, world_pos_sigmas=30.0))["N"]
    n3 = ob.oracle_filter(planes, W, H, P, [9], seed=0.3, debug=True, paper=dict(min_sigma=1e30))["N"]
    assert (n1 >= n0).all() and (n2 >= n1).all() and n1.sum() > n0.sum() and n2.sum() >= n1.sum()
    # a guard above every distance accepts the whole box (clipped at the frame border)
    xs, ys = np.meshgrid(np.arange(W), np.arange(H))
    cnt = lambda c, n: np.minimum(c + 4, n - 1) - np.maximum(c - 4, 0) + 1
    assert np.array_equal(n3.reshape(H, W), cnt(xs, W) * cnt(ys, H) * P)


def test_sigma_schedule_equals_chained_single_passes():
    W, H, P = 16, 12, 8
    planes = _frame(W, H, P)
    sched = ob.oracle_filter(planes, W, H, P, [9, 7, 5], paper=dict(sigma_seeds=[2.0, 1.0, 0.5]))["colors"]
    col = None
    for box, sg in zip([9, 7, 5], [2.0, 1.0, 0.5]):
        col = ob.oracle_filter(planes, W, H, P, [box], seed=sg, colors=col)["colors"]
    assert np.array_equal(sched, col, equal_nan=True) and np.isfinite(sched).mean() > 0.9
    same = ob.oracle_filter(planes, W, H, P, [9, 7, 5], seed=2.0)["colors"]
    assert not np.array_equal(sched, same, equal_nan=True)


def _clamp_numpy(col, W, H, P, k):
    c = col.reshape(3, H * W, P).copy()
    mu = np.empty((3, H * W))
    lim = np.empty((3, H * W))
    for ch in range(3):
        s = np.zeros(H * W)
        q = np.zeros(H * W)
        for i in range(P):  # sequential, like the oracle
            s = s + c[ch, :, i]
            q = q + c[ch, :, i] * c[ch, :, i]
        mu[ch] = s / P
        var = q / P - mu[ch] * mu[ch]
        lim[ch] = k * np.sqrt(np.where(var > 0, var, 0.0))
    spike = (np.abs(c - mu[:, :, None]) > lim[:, :, None]).any(axis=0)
    for ch in range(3):
        c[ch] = np.where(spike, mu[ch][:, None], c[ch])
    return c.reshape(3, -1), spike


def test_hdr_clamp_restatement():
    W, H, P = 12, 9, 8
    rng = np.random.default_rng(3)
    col = rng.random((3, W * H * P))
    col[:, ::37] *= 50.0  # fireflies
    want, spike = _clamp_numpy(col, W, H, P, 1.5)
    got = col.copy()
    ob.oracle().rpf_oracle_hdr_clamp(ob._p(got), W, H, P, 0, H, 1.5)
    assert spike.any() and not spike.all()
    assert np.array_equal(got, want)
    # rows outside [y0,y1) are untouched
    part = col.copy()
    ob.oracle().rpf_oracle_hdr_clamp(ob._p(part), W, H, P, 3, 5, 1.5)
    v = lambda a: a.reshape(3, H, W * P)
    assert np.array_equal(v(part)[:, 3:5], v(want)[:, 3:5])
    assert np.array_equal(v(part)[:, :3], v(col)[:, :3]) and np.array_equal(v(part)[:, 5:], v(col)[:, 5:])
    # the filter applies it after the last pass
    planes = _frame(W, H, P)
    a = ob.oracle_filter(planes, W, H, P, [5], seed=0.3)["colors"]
    b = ob.oracle_filter(planes, W, H, P, [5], seed=0.3, paper=dict(hdr_clamp_sigmas=1.0))["colors"]
    assert np.array_equal(b, _clamp_numpy(a, W, H, P, 1.0)[0])


# ---- pinned against the reference's own code carrying the hooks (oracle/_ref/libref_rpf_paper.so, ref_paper.sed) --------
import glob
import os

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _oracle_for_fixture(g):
    W, H, P, _ = (int(v) for v in g["meta"])
    eps, ms, pos = (float(v) for v in g["opts"])
    sub = (float(g["keep"]), int(g["rng_seed"])) if "keep" in g.files else None
    return ob.oracle_filter(g["planes"], W, H, P, [int(b) for b in g["boxes"]], subsample=sub,
                            paper=dict(w_epsilon=eps, min_sigma=ms, world_pos_sigmas=pos, sigma_seeds=list(g["sigmas"])))


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "paper_*case*.npz"))), ids=os.path.basename)
def test_oracle_reproduces_the_patched_reference_bit_for_bit(path):
    """Fixtures come from rpf.cpp:302-733 itself with REF_DEFINED + the two paper hooks (tests/make_golden.py paper)."""
    g = np.load(path)
    o = _oracle_for_fixture(g)
    assert np.array_equal(o["colors"], g["colors"], equal_nan=True)
    assert (o["nan_count"] > 0) == bool(g["nan_exit"])
    assert np.isfinite(g["colors"]).mean() > 0.99 and not np.array_equal(g["colors"], g["planes"][2:5].astype(np.float64))


def test_fixture_set_is_complete():
    assert len(glob.glob(os.path.join(GOLD, "paper_case*.npz"))) == 4
    assert len(glob.glob(os.path.join(GOLD, "paper_sub_case*.npz"))) == 2


def test_live_patched_reference_on_fresh_cases():
    """Where oracle/_ref exists (this container, and the GPU box through the snapshot): fresh random cases, and the
    hooks are no-ops at 0 (the paper build then IS the REF_DEFINED build)."""
    import make_golden as mg
    if not os.path.exists(os.path.join(ob.ORACLE_DIR, "_ref", "libref_rpf_paper.so")):
        pytest.skip("oracle/_ref/libref_rpf_paper.so not built")
    lib = mg.paper_lib()
    rng = np.random.default_rng(17)
    for i in range(3):
        W, H, P = int(rng.integers(10, 18)), int(rng.integers(8, 14)), int(rng.choice([2, 4, 8]))
        boxes = [int(b) for b in rng.choice([5, 7, 9], size=2)]
        sigmas = [float(v) for v in rng.uniform(0.1, 1.0, size=2)]
        eps, ms, pos = float(10.0 ** rng.uniform(-9, -2)), float(rng.uniform(0.01, 0.3)), float(rng.choice([0.0, 10.0, 30.0]))
        planes = _frame(W, H, P, seed=500 + i, kind=i % 2)
        keep, rs = (1.0, 0) if i == 0 else (float(rng.uniform(0.2, 0.8)), int(rng.integers(1, 1 << 40)))
        want, _ = mg.paper_reference(lib, planes, W, H, P, boxes, sigmas, eps, ms, pos, keep, rs)
        got = ob.oracle_filter(planes, W, H, P, boxes, subsample=None if keep >= 1.0 else (keep, rs),
                               paper=dict(w_epsilon=eps, min_sigma=ms, world_pos_sigmas=pos, sigma_seeds=sigmas))["colors"]
        assert np.array_equal(got, want, equal_nan=True), (W, H, P, boxes, sigmas, eps, ms, pos, keep, rs)
    planes = _frame(14, 10, 8, seed=9)
    off, _ = mg.paper_reference(lib, planes, 14, 10, 8, [7], [0.4], 0.0, 0.0, 0.0)
    dlib = mg.defined_lib()
    col = planes[2:5].astype(np.float64).copy()
    dlib.ref_apply_rpf_filter(14, 10, 8, planes.ctypes.data, col.ctypes.data, 7, 8, 0.4)
    assert np.array_equal(off, col, equal_nan=True)


# ---------------------------------------------------------------- kernels (GPU) ---------------------------------------
PAPER = dict(w_epsilon=1e-7, min_sigma=0.1, world_pos_sigmas=30.0)


def _gpu_paper(planes, W, H, P, boxes, sigmas, paper, clamp=0.0, **kw):
    import rpf_b200
    f = rpf_b200.RPFFilter(W, H, P, debug=True, **kw)
    f.set_paper_mode(**paper)
    f.upload(planes)
    f.apply_sigmas(boxes, sigmas)
    if clamp > 0:
        f.hdr_clamp(clamp)
    g = dict(colors=f.download_colors_f64(), **f.download_debug())
    g["stats"] = f.stats()
    f.close()
    return g


@pytest.mark.gpu
@pytest.mark.parametrize("W,H,P,boxes,sigmas,kw", [
    (40, 28, 8, [7], [0.3], {}),                                        # fused kernel
    (40, 28, 8, [7], [0.3], dict(force_general=True)),                  # general path
    (36, 30, 8, [17, 9, 7], [0.5, 0.25, 0.125], {}),                    # sigma per iteration over both paths
    (20, 16, 8, [9], [0.3], dict(force_large=True)),                    # dependency_large_kernel
    (14, 12, 32, [11], [0.3], {}),                                      # blend_wide_kernel
], ids=["fused", "general", "schedule", "large", "wide"])
def test_paper_mode_matches_oracle(W, H, P, boxes, sigmas, kw):
    from test_gpu_parity import _compare
    planes = _frame(W, H, P, seed=61, kind=1)
    o = ob.oracle_filter(planes, W, H, P, boxes, debug=True, paper=dict(PAPER, sigma_seeds=sigmas))
    g = _gpu_paper(planes, W, H, P, boxes, sigmas, PAPER, **kw)
    _compare(g, o, W, H, P)
    base = ob.oracle_filter(planes, W, H, P, boxes[-1:], seed=sigmas[-1], debug=True)
    assert g["N"].sum() > base["N"].sum()  # the guard did widen the neighbourhoods


@pytest.mark.gpu
def test_hdr_clamp_is_bit_identical_and_local_to_rows():
    import rpf_b200
    W, H, P = 24, 18, 8
    rng = np.random.default_rng(8)
    planes = _frame(W, H, P)
    col = rng.random((3, W * H * P))
    col[:, ::41] *= 80.0
    want = col.copy()
    ob.oracle().rpf_oracle_hdr_clamp(ob._p(want), W, H, P, 4, 11, 1.0)
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    f.upload_colors_f64(col)
    f.hdr_clamp(1.0, 4, 11)
    got = f.download_colors_f64()
    assert np.array_equal(got, want) and not np.array_equal(got, col)
    with pytest.raises(rpf_b200.RPFError):
        f.hdr_clamp(0.0)
    f.close()


@pytest.mark.gpu
def test_paper_mode_off_again_is_the_reference_arithmetic():
    """rpf_set_paper_mode rebuilds the per-pixel thresholds: switching the options off restores the default result."""
    import rpf_b200
    W, H, P = 24, 18, 8
    planes = _frame(W, H, P)
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    f.apply([7], 0.3)
    ref = f.download_colors_f64()
    f.set_paper_mode(**PAPER)
    f.upload(planes)
    f.apply([7], 0.3)
    on = f.download_colors_f64()
    f.set_paper_mode()
    f.upload(planes)
    f.apply([7], 0.3)
    off = f.download_colors_f64()
    f.close()
    assert np.array_equal(ref, off) and not np.array_equal(ref, on)


@pytest.mark.gpu
def test_row_band_driver_takes_the_schedule_and_the_clamp():
    """bands.run_passes with one sigma per pass and spike removal on the owned rows = rpf_filter_sigmas + rpf_hdr_clamp."""
    import bands
    import rpf_b200
    W, H, P, boxes, sigmas = 28, 22, 8, [9, 7], [0.5, 0.25]
    planes = _frame(W, H, P, seed=5)
    o = ob.oracle_filter(planes, W, H, P, boxes, paper=dict(PAPER, sigma_seeds=sigmas, hdr_clamp_sigmas=1.0))["colors"]
    f = rpf_b200.RPFFilter(W, H, P)
    f.set_paper_mode(**PAPER)
    f.upload(planes)
    bands.run_passes(f, bands.make_band(0, 1, H, boxes), boxes, sigmas, hdr_clamp_sigmas=1.0)
    g = f.download_colors_f64()
    f.close()
    m = ~np.isnan(o)
    assert np.array_equal(np.isnan(g), np.isnan(o))
    assert np.abs(g[m] - o[m]).max() <= 1e-9 * max(1.0, np.abs(o[m]).max())
