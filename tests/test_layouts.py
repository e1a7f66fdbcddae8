"""Run-time channel table: a sample record is position (2) | colour (3) | n_random random parameters |
n_features features; the reference's SampleData is 2 / 12 (sd.h:40-46,62-94). north_star's wider buffer (time and light samples
as random parameters, a texture colour as features) runs through the same formulas with the loop bounds lifted.

There is no reference output for a wider record, so it is pinned the way the paper mode is: the oracle's generalisation is
tied to the reference-pinned default layout by an exact invariance (random parameters that carry no information change
nothing, bit for bit), and the GPU is held to that oracle."""
import numpy as np
import pytest

import oracle_bind as ob

LAYOUTS = [(5, 12), (5, 15)]


def test_layout_table_of_this_build():
    import rpf_b200
    L = rpf_b200.lib()
    assert L.rpf_layout_supported(0, 0) == 1 and L.rpf_layout_supported(2, 12) == 1
    for nr, nf in LAYOUTS:
        assert L.rpf_layout_supported(nr, nf) == 1
    assert L.rpf_layout_supported(3, 13) == 0
    assert L.rpf_num_channels(None) == 19 and L.rpf_num_pairs(None) == 96   # rpf.cpp:416-442: 12*4 + 3*16


def test_synthetic_wider_record_contains_the_reference_record():
    import rpf_b200
    W, H, P = 12, 9, 4
    base = rpf_b200.synth_planes(W, H, P, seed=3, kind=0)
    wide = rpf_b200.synth_planes(W, H, P, seed=3, kind=0, n_random=5, n_features=15)
    assert wide.shape == (25, W * H * P)
    assert np.array_equal(wide[:7], base[:7]) and np.array_equal(wide[10:22], base[7:19])
    assert wide[7:10].min() >= 0 and wide[7:10].max() < 1 and wide[22:25].std() > 0.05


@pytest.mark.parametrize("box,kind", [(7, 0), (9, 1)])
def test_oracle_uninformative_random_parameters_change_nothing(box, kind):
    """Three more random parameters that are constant: their mutual information with everything is exactly 0, they do not
    enter the 3-sigma test or the weights -- the generalised loops must give the reference layout's result bit for bit."""
    import rpf_b200
    W, H, P = 20, 16, 8
    base = rpf_b200.synth_planes(W, H, P, seed=11, kind=kind)
    wide = np.concatenate([base[:7], np.full((3, W * H * P), 0.5, np.float32), base[7:]])
    a = ob.oracle_filter(base, W, H, P, [box], seed=0.5, debug=True)
    b = ob.oracle_filter(wide, W, H, P, [box], seed=0.5, debug=True, layout=(5, 12))
    assert np.array_equal(a["N"], b["N"])
    for k in ("colors", "alpha", "beta", "wrc"):
        assert np.array_equal(np.isnan(a[k]), np.isnan(b[k])) and np.array_equal(np.nan_to_num(a[k]), np.nan_to_num(b[k])), k


def test_oracle_wider_record_matters_when_it_carries_information():
    import rpf_b200
    W, H, P = 20, 16, 8
    base = rpf_b200.synth_planes(W, H, P, seed=11, kind=0)
    wide = rpf_b200.synth_planes(W, H, P, seed=11, kind=0, n_random=5, n_features=15)
    a = ob.oracle_filter(base, W, H, P, [7], seed=0.5, debug=True)
    b = ob.oracle_filter(wide, W, H, P, [7], seed=0.5, debug=True, layout=(5, 15))
    assert b["beta"].shape == (W * H, 15)
    assert b["N"].mean() < a["N"].mean()            # three more features to pass the 3-sigma test on
    assert np.abs(a["colors"] - b["colors"]).max() > 1e-3


def _gpu(planes, W, H, P, boxes, seed, layout):
    import rpf_b200
    f = rpf_b200.RPFFilter(W, H, P, debug=True, n_random=layout[0], n_features=layout[1])
    f.upload(planes)
    f.apply(boxes, seed)
    out = dict(colors=f.download_colors_f64(), **f.download_debug())
    out["stats"] = f.stats()
    f.close()
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("layout", LAYOUTS)
@pytest.mark.parametrize("boxes,seed,kind", [([7], 0.5, 0), ([17, 7], 0.5, 0), ([9], 2.0, 1), ([7], 0.002, 0)])
def test_gpu_wider_layouts_match_the_oracle(layout, boxes, seed, kind):
    import rpf_b200
    from test_gpu_parity import TOL_TIGHT
    W, H, P = 30, 24, 8
    planes = rpf_b200.synth_planes(W, H, P, seed=77, kind=kind, n_random=layout[0], n_features=layout[1])
    o = ob.oracle_filter(planes, W, H, P, boxes, seed=seed, debug=True, layout=layout, threads=8)
    g = _gpu(planes, W, H, P, boxes, seed, layout)
    assert np.array_equal(g["N"], o["N"])
    for k in ("alpha", "beta", "wrc"):
        assert np.array_equal(np.isnan(g[k]), np.isnan(o[k])), k
        assert np.nanmax(np.abs(g[k] - o[k]), initial=0.0) <= TOL_TIGHT, (k, np.nanmax(np.abs(g[k] - o[k])))
    gc, oc = g["colors"], o["colors"]
    assert np.array_equal(np.isnan(gc), np.isnan(oc))
    m = ~np.isnan(oc)
    assert np.abs(gc[m] - oc[m]).max(initial=0.0) <= TOL_TIGHT * max(1.0, np.abs(oc[m]).max(initial=0.0))
    assert g["stats"]["nan_channels"] == o["nan_count"]


@pytest.mark.gpu
def test_gpu_uninformative_random_parameters_change_nothing():
    """The invariance of the oracle test above, on the device: layout 5 / 12 with three constant random parameters is
    bit-identical to the reference layout (through the general path for both)."""
    import rpf_b200
    W, H, P = 26, 20, 8
    base = rpf_b200.synth_planes(W, H, P, seed=5, kind=0)
    wide = np.ascontiguousarray(np.concatenate([base[:7], np.full((3, W * H * P), 0.5, np.float32), base[7:]]))
    f = rpf_b200.RPFFilter(W, H, P, debug=True, force_general=True)
    f.upload(base)
    f.apply([9, 7], 0.5)
    a = dict(colors=f.download_colors_f64(), **f.download_debug())
    f.close()
    b = _gpu(wide, W, H, P, [9, 7], 0.5, (5, 12))
    assert np.array_equal(a["N"], b["N"]) and np.array_equal(a["colors"], b["colors"]) and np.array_equal(a["beta"], b["beta"])


@pytest.mark.gpu
def test_gpu_mi_probe_of_a_wider_layout():
    import rpf_b200
    W, H, P, box = 24, 20, 8, 9
    layout = (5, 15)
    planes = rpf_b200.synth_planes(W, H, P, seed=8, kind=0, n_random=5, n_features=15)
    o = ob.oracle_pixel_probe(planes, W, H, P, box, 11, 9, seed=0.5, layout=layout)
    f = rpf_b200.RPFFilter(W, H, P, n_random=5, n_features=15)
    assert f.nch == 25 and f.n_pairs == 15 * 7 + 3 * 22
    f.upload(planes)
    f.set_debug_pixel(11, 9)
    f.apply([box], 0.5)
    r = f.download_debug_mi()
    f.close()
    assert r["N"] == o["N"] and r["bins"] == o["bins"]
    assert np.array_equal(r["marginals"].astype(np.int64), o["marginals"].astype(np.int64))
    assert np.abs(r["mi"] - o["mi"]).max() <= 1e-12


@pytest.mark.gpu
def test_gpu_wider_layouts_refuse_what_only_the_reference_layout_has():
    import rpf_b200
    planes = rpf_b200.synth_planes(8, 6, 16, seed=1, kind=0, n_random=5, n_features=12)
    f = rpf_b200.RPFFilter(8, 6, 16, n_random=5, n_features=12)
    f.upload(planes)
    with pytest.raises(rpf_b200.RPFError, match="reference channel layout"):
        f.apply([7], 0.5)          # 16 spp: the wide blend kernel
    f.close()
    with pytest.raises(rpf_b200.RPFError, match="no kernels"):
        rpf_b200.RPFFilter(8, 6, 8, n_random=4, n_features=12)
