"""CPU: the oracle against the round-2 fixtures (tests/make_golden_r2.py): large boxes and many samples per pixel,
non-finite input, crops of the RENDERED feature buffer of scenes/sample.pbrt. The fixtures hold the output of the
reference's own rpf.cpp:302-733 + mi.cpp (REF_DEFINED build); the oracle must reproduce them bit for bit. Where
oracle/_ref exists (the build container) the same comparison also runs live on fresh random cases."""
import glob
import os

import numpy as np
import pytest

import oracle_bind as ob

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def same_bits(a, b):
    """Bit-identical doubles, NaN == NaN (payload and sign ignored: the CPU's default NaN vs a propagated one)."""
    a, b = np.asarray(a), np.asarray(b)
    na, nb = np.isnan(a), np.isnan(b)
    return np.array_equal(na, nb) and np.array_equal(a[~na], b[~nb])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "big_case*.npz"))), ids=os.path.basename)
def test_big_boxes_bit_exact(path):
    g = np.load(path)
    W, H, P, _ = (int(v) for v in g["meta"])
    out = ob.oracle_filter(g["planes"], W, H, P, [int(b) for b in g["boxes"]], seed=float(g["seed"]), threads=8)
    assert same_bits(out["colors"], g["colors"])
    assert (out["nan_count"] > 0) == bool(g["nan_exit"])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "nonfinite_case*.npz"))), ids=os.path.basename)
def test_nonfinite_input_bit_exact(path):
    g = np.load(path)
    W, H, P, box = (int(v) for v in g["meta"])
    out = ob.oracle_filter(g["planes"], W, H, P, [box], seed=float(g["seed"]), threads=8)
    assert same_bits(out["colors"], g["colors"]), str(g["what"])
    assert out["nan_count"] == int(np.isnan(g["colors"]).sum())


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "real_crop*.npz"))), ids=os.path.basename)
def test_rendered_buffer_crops_bit_exact(path):
    g = np.load(path)
    W, H, P = (int(v) for v in g["meta"][:3])
    out = ob.oracle_filter(g["planes"], W, H, P, [int(b) for b in g["boxes"]], seed=float(g["seed"]), threads=8, debug=True)
    assert same_bits(out["colors"], g["colors"])
    assert out["nan_count"] == int(np.isnan(g["colors"]).sum())
    # what real content looks like: flat surfaces have sigma 0 in the normal features and reject every
    # neighbour, so most pixels keep only their own samples
    assert out["N"].min() == P


def test_fixture_sets_are_present():
    for pat, n in (("big_case*.npz", 6), ("nonfinite_case*.npz", 4), ("real_crop*.npz", 3)):
        assert len(glob.glob(os.path.join(GOLD, pat))) == n


@pytest.mark.skipif(not os.path.exists(os.path.join(ob.ORACLE_DIR, "_ref", "libref_rpf_defined.so")), reason="oracle/_ref not built here")
@pytest.mark.parametrize("W,H,P,box,seed,kind", [(22, 19, 8, 17, 0.5, 0), (20, 16, 8, 35, 2.0, 0), (12, 9, 16, 9, 0.002, 1), (18, 14, 8, 17, 0.2, 1)])
def test_live_reference_large_boxes(W, H, P, box, seed, kind):
    import make_golden as mg
    planes = mg.synth(W, H, P, 9100 + box + P, kind)
    lib = mg.defined_lib()
    col = planes[2:5].astype(np.float64).copy()
    lib.ref_apply_rpf_filter(W, H, P, planes.ctypes.data, col.ctypes.data, box, 8, seed)
    out = ob.oracle_filter(planes, W, H, P, [box], seed=seed, threads=8)
    assert same_bits(out["colors"], col)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "probe_case*.npz"))), ids=os.path.basename)
def test_pixel_probe_mi_equals_the_references_mi(path):
    """MI known answers. The fixture holds mi.cpp's own values for the 96 pairs of one pixel."""
    import rpf_b200
    g = np.load(path)
    W, H, P, box, px, py, kind, seed = (int(v) for v in g["meta"])
    planes = rpf_b200.synth_planes(W, H, P, seed=seed, kind=kind)
    r = ob.oracle_pixel_probe(planes, W, H, P, box, px, py, seed=0.5)
    assert np.array_equal(r["z"], g["z"])
    assert np.array_equal(r["mi"], g["ref_mi"])
    assert (r["marginals"].sum(axis=1) == r["N"]).all()
