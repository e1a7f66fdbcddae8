"""CPU: the C-ABI library loads and exports every symbol include/rpf_b200.h declares; host-side logic (synthetic
generator, band planning, error behaviour without a GPU). No device compute here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lib_or_skip():
    import rpf_b200
    if not os.path.exists(rpf_b200.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return rpf_b200


def test_header_symbols_are_exported():
    rpf_b200 = _lib_or_skip()
    hdr = open(os.path.join(ROOT, "include", "rpf_b200.h")).read()
    declared = set(re.findall(r"\b(rpf_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"rpf_ctx"}
    assert declared == set(rpf_b200.ABI_SYMBOLS), declared ^ set(rpf_b200.ABI_SYMBOLS)
    L = C.CDLL(rpf_b200.LIB_PATH)
    for s in declared:
        assert hasattr(L, s), s


def test_no_cpu_fallback_without_device():
    rpf_b200 = _lib_or_skip()
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(rpf_b200.RPFError, match="no CUDA device|no CPU fallback"):
        rpf_b200.RPFFilter(8, 8, 8)


def test_synth_is_deterministic_and_band_consistent():
    rpf_b200 = _lib_or_skip()
    W, H, P = 12, 10, 4
    a = rpf_b200.synth_planes(W, H, P, seed=9, kind=0)
    b = rpf_b200.synth_planes(W, H, P, seed=9, kind=0)
    assert np.array_equal(a, b) and a.dtype == np.float32
    band = rpf_b200.synth_planes(W, 4, P, seed=9, kind=0, y_off=3)   # rows 3..6 generated on their own
    assert np.array_equal(band.reshape(19, 4, W, P), a.reshape(19, H, W, P)[:, 3:7])
    xs = a[0].reshape(H, W, P)
    assert np.all(np.floor(xs) == np.arange(W)[None, :, None])        # pFilm.x = pixel + [0,1)
    d = rpf_b200.synth_planes(W, H, P, seed=9, kind=1)
    assert (d[7:19] == 0).mean() > 0.3                                # degenerate kind has flat / zero features
    assert rpf_b200.lib().rpf_synth_planes(0, 1, 1, 1, 0, 0, 0, None) < 0


def test_max_neighbourhood_covers_reference_box():
    rpf_b200 = _lib_or_skip()
    assert rpf_b200.lib().rpf_max_neighbourhood() >= 7 * 7 * 16      # box 7 (rpf.cpp:767) up to 16 spp


def test_band_planning():
    import bands
    boxes = [55, 35, 17, 7]
    H = 1080
    bs = [bands.make_band(r, 8, H, boxes) for r in range(8)]
    assert bs[0].own0 == 0 and bs[-1].own1 == H
    assert all(bs[i].own1 == bs[i + 1].own0 for i in range(7))
    assert bs[3].buf0 == bs[3].own0 - 27 and bs[0].buf0 == 0 and bs[7].buf1 == H
    # sends of one rank match the receives of its neighbour, row for row (in global coordinates)
    for nb in (35, 17, 7):
        for r in range(7):
            lo, hi = bs[r], bs[r + 1]
            s = [(p, sr, rr) for p, sr, rr in bands.halo_plan(lo, nb) if p == r + 1][0]
            t = [(p, sr, rr) for p, sr, rr in bands.halo_plan(hi, nb) if p == r][0]
            g = lambda band, rows: (rows[0] + band.buf0, rows[1] + band.buf0)
            assert g(lo, s[1]) == g(hi, t[2]) and g(hi, t[1]) == g(lo, s[2])
            assert s[1][1] - s[1][0] == bands.b_delta(nb)
    assert bands.halo_plan(bands.make_band(0, 1, H, boxes), 7) == []


def test_host_cpp_reports_missing_device():
    """The C++ host shim (built against the reference's SamplingFilm) surfaces the C ABI's error instead of falling back."""
    import torch
    path = os.path.join(ROOT, "oracle", "_ref", "libhost_harness.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libhost_harness.so not built")
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    _lib_or_skip()
    lib = C.CDLL(path)
    W, H, P = 4, 3, 2
    planes = np.zeros((19, W * H * P), np.float32)
    col = np.zeros((3, W * H * P), np.float64)
    bx = np.asarray([3], np.int32)
    err = C.create_string_buffer(512)
    rc = lib.host_apply_rpf_b200(W, H, P, planes.ctypes.data_as(C.c_void_p), col.ctypes.data_as(C.c_void_p),
                                 bx.ctypes.data_as(C.c_void_p), 1, C.c_double(0.5), 0, None, None, err, 512)
    assert rc == -1 and b"rpf_create" in err.value


def test_feature_buffer_file_round_trip(tmp_path):
    """lossless dump / load of the per-sample feature buffer (host code, no GPU)."""
    import numpy as np
    import rpf_b200
    W, H, P = 9, 7, 3
    planes = rpf_b200.synth_planes(W, H, P, seed=5, kind=1)
    rw = np.linspace(0.1, 1.0, W * H * P).astype(np.float32)
    path = str(tmp_path / "frame.rpfb")
    rpf_b200.write_buffer(path, planes, W, H, P, ray_weight=rw, sample_origin=(-2, 3))
    got, grw, info = rpf_b200.read_buffer(path)
    assert np.array_equal(got, planes) and np.array_equal(grw, rw)
    assert info == dict(width=W, height=H, spp=P, sample_x0=-2, sample_y0=3, has_ray_weight=1, n_random=2, n_features=12)
    # a wider record carries its channel layout in the header
    wide = rpf_b200.synth_planes(W, H, P, seed=5, kind=0, n_random=5, n_features=15)
    rpf_b200.write_buffer(path, wide, W, H, P, n_random=5, n_features=15)
    got, _, info = rpf_b200.read_buffer(path)
    assert got.shape == (25, W * H * P) and np.array_equal(got, wide) and (info["n_random"], info["n_features"]) == (5, 15)
    # no ray weights stored -> ones
    rpf_b200.write_buffer(path, planes, W, H, P)
    _, grw, info = rpf_b200.read_buffer(path)
    assert info["has_ray_weight"] == 0 and np.all(grw == 1.0)
    # truncated file and foreign file are errors, not garbage
    raw = open(path, "rb").read()
    open(path, "wb").write(raw[:len(raw) // 2])
    import pytest
    with pytest.raises(rpf_b200.RPFError, match="truncated"):
        rpf_b200.read_buffer(path)
    open(path, "wb").write(b"EXR!" + raw[4:])
    with pytest.raises(rpf_b200.RPFError, match="not a feature-buffer"):
        rpf_b200.read_buffer(path)


def test_host_cpp_dump_hook_round_trip(tmp_path):
    """DumpSamplingFilmB200 over a real reference SamplingFilm -> file -> read_buffer gives the samples back."""
    import ctypes as C
    import os
    import numpy as np
    import pytest
    import oracle_bind as ob
    import rpf_b200
    path = os.path.join(ob.ORACLE_DIR, "_ref", "libhost_harness.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libhost_harness.so was not built (needs /root/reference at build time)")
    lib = C.CDLL(path)
    if not hasattr(lib, "host_dump_sampling_film"):
        pytest.skip("prebuilt harness predates the dump hook")
    W, H, P = 6, 5, 4
    planes = rpf_b200.synth_planes(W, H, P, seed=8, kind=0)
    rw = np.linspace(0.5, 1.0, W * H * P).astype(np.float32)
    out = str(tmp_path / "dump.rpfb")
    lib.host_dump_sampling_film.argtypes = [C.c_int] * 3 + [C.c_void_p] * 2 + [C.c_char_p]
    assert lib.host_dump_sampling_film(W, H, P, planes.ctypes.data, rw.ctypes.data, out.encode()) == 0
    got, grw, info = rpf_b200.read_buffer(out)
    assert np.array_equal(got, planes) and np.array_equal(grw, rw) and info["spp"] == P


def test_fused_limit_and_its_override(monkeypatch):
    """rpf_fused_max_candidates: 512 by default, RPF_FUSED_MAX moves it within [0, 1279] (no compute, no GPU)."""
    rpf_b200 = _lib_or_skip()
    f = rpf_b200.lib().rpf_fused_max_candidates
    monkeypatch.delenv("RPF_FUSED_MAX", raising=False)
    assert f() == 512
    assert 7 * 7 * 8 <= f() < 9 * 9 * 8      # the shipped box 7 at 8 spp stays fused, box 9 does not
    for env, want in (("0", 0), ("1279", 1279), ("100000", 1279), ("-5", 0), ("700", 700)):
        monkeypatch.setenv("RPF_FUSED_MAX", env)
        assert f() == want


def test_new_entry_points_reject_a_null_context():
    """Paper-mode entry points follow the error convention (negative status, nothing thrown) without a context."""
    import ctypes as C
    rpf_b200 = _lib_or_skip()
    L = rpf_b200.lib()
    o = rpf_b200.PaperOpts(1e-7, 0.1, 30.0)
    bx = (C.c_int * 1)(7)
    sg = (C.c_double * 1)(0.3)
    assert L.rpf_set_paper_mode(None, C.byref(o)) < 0
    assert L.rpf_filter_sigmas(None, bx, sg, 1) < 0
    assert L.rpf_hdr_clamp(None, C.c_double(1.0), 0, 0) < 0


def test_band_planning_small_frames_and_many_ranks():
    """Bands tile the frame exactly for every world size, also when there are barely more rows than ranks; the buffered
    rows cover the largest box's halo, clipped at the frame."""
    import bands
    for H, world, boxes in ((720, 8, [55, 35, 17, 7]), (9, 8, [7]), (64, 4, [17, 7]), (10, 3, [55])):
        bs = [bands.make_band(r, world, H, boxes) for r in range(world)]
        assert bs[0].own0 == 0 and bs[-1].own1 == H
        assert all(a.own1 == b.own0 for a, b in zip(bs, bs[1:]))
        assert all(b.own1 > b.own0 for b in bs)
        d = bands.b_delta(max(boxes))
        for b in bs:
            assert b.buf0 == max(0, b.own0 - d) and b.buf1 == min(H, b.own1 + d)
            assert b.local_own == (b.own0 - b.buf0, b.own1 - b.buf0) and b.local_height == b.buf1 - b.buf0


def test_bands_thinner_than_the_halo_are_refused():
    """Halo exchange is between adjacent ranks only; a plan that would need colours from two bands away fails loudly."""
    import bands
    import pytest
    with pytest.raises(ValueError, match="thinner"):
        bands.make_band(0, 8, 64, [55, 35, 17, 7])     # 8 rows per band, the box-35 pass reaches 17 rows
    bands.make_band(0, 8, 64, [55])                    # one pass: nothing is exchanged
    bands.make_band(0, 2, 64, [55, 35, 17, 7])         # 32 rows per band


def test_headers_are_plain_c_and_link_against_the_library(tmp_path):
    """The boundary is a C ABI: include/*.h compile as pedantic C99, and a C program links against librpf_b200.so and calls
    an entry point that needs no device (the cgo / JNI / ctypes binding of INTEGRATION.md sees exactly this)."""
    import shutil
    import subprocess
    rpf_b200 = _lib_or_skip()
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        import pytest
        pytest.skip("no C compiler")
    src = tmp_path / "abi.c"
    src.write_text('#include "rpf_b200.h"\n#include "rpf_subsample.h"\n#include <stdio.h>\n'
                   'int main(void) {\n'
                   '    rpf_paper_opts o = {0, 0, 0};\n'
                   '    rpf_ctx *ctx = 0;\n'
                   '    rpf_config cfg = {0};\n'
                   '    float table[256];\n'
                   '    (void)o; (void)ctx; (void)cfg;\n'
                   '    if (rpf_film_filter_table(0, 0.5f, 0.5f, 0.f, table) != RPF_OK) return 2;\n'
                   '    if (rpf_subsample_keep(rpf_subsample_centre_key(1, 7, 3, 4), RPF_SUBSAMPLE_ALL, 1, 2, 0) != 1) return 3;\n'
                   '    printf("%d %d\\n", rpf_fused_max_candidates(), rpf_max_neighbourhood());\n'
                   '    return 0;\n}\n')
    exe = tmp_path / "abi"
    libdir = os.path.dirname(rpf_b200.LIB_PATH)
    inc = os.path.join(ROOT, "include")
    r = subprocess.run([cc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", f"-I{inc}", str(src), "-o", str(exe),
                        f"-L{libdir}", "-lrpf_b200", f"-Wl,-rpath,{libdir}"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    env = dict(os.environ)
    env.pop("RPF_FUSED_MAX", None)
    r = subprocess.run([str(exe)], capture_output=True, text=True, env=env)
    assert r.returncode == 0, (r.returncode, r.stderr)
    assert r.stdout.split() == ["512", str(1 << 20)]


def test_band_plan_from_measured_work():
    """rpf_band_plan_weighted: bands of equal MEASURED cost. A uniform cost gives equal heights; a band that measured slower
    gets fewer rows; the constraint (every band owns the rows later passes reach) still holds."""
    import numpy as np
    rpf_b200 = _lib_or_skip()
    H, boxes = 1080, [55, 35, 17, 7]
    assert rpf_b200.band_plan(H, 8, boxes, np.ones(H)) == [0, 135, 270, 405, 540, 675, 810, 945, 1080]
    w = np.ones(H)
    w[:135] = 1.1                                   # the first band measured 10 % slower per row
    plan = rpf_b200.band_plan(H, 8, boxes, w)
    assert plan[0] == 0 and plan[-1] == H and plan[1] < 135 and all(b - a >= 17 for a, b in zip(plan, plan[1:]))
    cost = [w[a:b].sum() for a, b in zip(plan, plan[1:])]
    assert max(cost) - min(cost) <= 2.2             # within two rows' worth of cost
    # degenerate weights cannot break the adjacency constraint
    w2 = np.zeros(H)
    w2[500:520] = 1.0
    plan = rpf_b200.band_plan(H, 8, boxes, w2)
    assert plan[0] == 0 and plan[-1] == H and all(b - a >= 17 for a, b in zip(plan, plan[1:]))
    # the geometric plan: the two edge bands a little taller at most
    g = rpf_b200.band_plan(H, 8, boxes)
    assert g[0] == 0 and g[-1] == H and all(130 <= b - a <= 140 for a, b in zip(g, g[1:]))
