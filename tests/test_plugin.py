"""The plugin surface: `Integrator "rpf"` of an UNMODIFIED reference pbrt resolved to the replacement integrator
(raytracer-rpf_b200/host/rpf_integrator.cpp linked ahead of libpbrt.a -> host/_build/pbrt_rpf_b200, built by
host/Makefile in the build container; the binary travels to the GPU box with the snapshot).

CPU: the scene-file parameters, the dump hook and the Film tail on a scene of our own (tests/scenes/spheres.pbrt.in).
GPU: the whole chain -- pbrt renders, the B200 filters, pbrt's Film writes the image -- against the oracle + the Film
oracle on the dumped feature buffer."""
import os
import subprocess

import numpy as np
import pytest

import oracle_bind as ob

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PBRT = os.path.join(ROOT, "raytracer-rpf_b200", "host", "_build", "pbrt_rpf_b200")
SCENE = os.path.join(ROOT, "tests", "scenes", "spheres.pbrt.in")

needs_binary = pytest.mark.skipif(not os.path.exists(PBRT), reason="host/_build/pbrt_rpf_b200 not built (needs the reference tree)")


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = (int(v) for v in f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), dtype="<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return data[::-1].astype(np.float32)  # PFM rows run bottom to top


def render(tmp_path, integrator, W=64, H=48, spp=8, extra_args=()):
    out = tmp_path / "out.pfm"
    text = open(SCENE).read().replace("@INTEGRATOR@", integrator).replace("@W@", str(W)).replace("@H@", str(H))
    text = text.replace("@SPP@", str(spp)).replace("@OUT@", str(out))
    scene = tmp_path / "scene.pbrt"
    scene.write_text(text)
    r = subprocess.run([PBRT, "--nthreads", "4", *extra_args, str(scene)], capture_output=True, text=True, timeout=300)
    return r, out


def film_of(planes, colors, W, H, P):
    """pbrt's default box filter of radius 0.5 through the Film oracle (bit-identical to the reference's FilmTile path)."""
    table = ob.oracle_film_table(0, 0.5, 0.5, 0.0)
    return ob.oracle_film(W, H, P, planes[0:2], colors, None, table, 0.5, 0.5, (0, 0, W, H), 1.0, float("inf"))


@needs_binary
def test_plugin_resolves_and_dumps_without_a_gpu(tmp_path):
    """Integrator "rpf" reaches the replacement CreateRPFIntegrator; every parameter is looked up (no "unused" warning from
    api.cpp:1713); the dump hook writes the feature buffer; with rpfdumponly the unfiltered samples go through pbrt's Film."""
    import rpf_b200
    dump = tmp_path / "frame.rpfb"
    integ = (f'Integrator "rpf" "integer maxdepth" [4] "integer boxsizes" [17 7] "float sigmaseed" [0.5] "integer ngpus" [1] '
             f'"string rpfmode" "defined" "string rpfdump" "{dump}" "bool rpfdumponly" "true"')
    r, out = render(tmp_path, integ)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "unused" not in (r.stdout + r.stderr).lower()
    planes, rw, info = rpf_b200.read_buffer(str(dump))
    assert (info["width"], info["height"], info["spp"]) == (64, 48, 8) and (rw == 1).all()
    A = planes.reshape(19, 48, 64, 8)
    assert np.isfinite(planes).all()
    # pFilm lies in its pixel, the lens sample in the unit square (thin lens: a real random parameter)
    assert (np.floor(A[0]) == np.arange(64)[None, :, None]).all() and (np.floor(A[1]) == np.arange(48)[:, None, None]).all()
    assert A[5].min() >= 0 and A[5].max() < 1 and A[5].std() > 0.2
    # first-hit normals are unit vectors wherever something was hit
    n0 = np.sqrt((A[7:10] ** 2).sum(axis=0))
    assert np.all((np.abs(n0 - 1) < 1e-4) | (n0 == 0))
    img = read_pfm(out)
    ref = film_of(planes, planes[2:5].astype(np.float64), 64, 48, 8)
    assert np.array_equal(img, ref), np.abs(img - ref).max()


@needs_binary
def test_plugin_without_device_fails_loudly(tmp_path):
    """No CPU fallback: without a usable B200 the filter stage reports the C ABI's error and the process exits non-zero."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r, out = render(tmp_path, 'Integrator "rpf"', W=32, H=24, spp=4)
    assert r.returncode != 0
    assert "RPF B200 filter failed" in r.stdout + r.stderr
    assert not out.exists()


@needs_binary
@pytest.mark.gpu
@pytest.mark.parametrize("boxes,seed,ngpus", [("7", 0.5, 1), ("17 7", 0.5, 1), ("7", 0.002, 1), ("9 7", 0.5, 2)])
def test_pbrt_renders_b200_filters_film_writes(tmp_path, boxes, seed, ngpus):
    import rpf_b200
    import torch
    if torch.cuda.device_count() < ngpus:
        pytest.skip(f"needs {ngpus} GPUs")
    dump = tmp_path / "frame.rpfb"
    W, H, P = 64, 48 if ngpus == 1 else 64, 8
    integ = (f'Integrator "rpf" "integer boxsizes" [{boxes}] "float sigmaseed" [{seed}] "integer ngpus" [{ngpus}] "string rpfdump" "{dump}"')
    r, out = render(tmp_path, integ, W=W, H=H, spp=P)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert f"RPF on {ngpus} B200 GPU(s)" in r.stdout
    planes, _, info = rpf_b200.read_buffer(str(dump))
    bx = [int(b) for b in boxes.split()]
    o = ob.oracle_filter(planes, W, H, P, bx, seed=seed, threads=8)
    img = read_pfm(out)
    ref = film_of(planes, o["colors"], W, H, P)
    # the GPU colours are within 1e-9 of the oracle's; through the Film's float accumulation that is an ulp or two
    tol = 1e-5 * max(1.0, float(np.abs(ref[np.isfinite(ref)]).max()))
    m = np.isfinite(ref)
    assert np.array_equal(np.isfinite(img), m)
    assert np.abs(img[m] - ref[m]).max() <= tol
    assert np.abs(ref[m] - film_of(planes, planes[2:5].astype(np.float64), W, H, P)[m]).max() > 0 or seed < 0.01


@needs_binary
def test_wide_record_producer_keeps_the_reference_record(tmp_path):
    """SURVEY.md 8(f) N3, producer side: "string rpfrecord" "wide" adds time, the first light sample and the first-hit
    texture colour (25 channels) without touching the sampler's request order -- the 19 reference channels of the wide
    frame are the reference frame bit for bit."""
    import rpf_b200
    d19, d25 = tmp_path / "ref.rpfb", tmp_path / "wide.rpfb"
    r, _ = render(tmp_path, f'Integrator "rpf" "string rpfdump" "{d19}" "bool rpfdumponly" "true"')
    assert r.returncode == 0, r.stderr[-2000:]
    r, _ = render(tmp_path, f'Integrator "rpf" "string rpfrecord" "wide" "string rpfdump" "{d25}" "bool rpfdumponly" "true"')
    assert r.returncode == 0, r.stderr[-2000:]
    assert "unused" not in (r.stdout + r.stderr).lower()
    a, _, ia = rpf_b200.read_buffer(str(d19))
    b, _, ib = rpf_b200.read_buffer(str(d25))
    assert (ia["n_random"], ia["n_features"]) == (2, 12) and (ib["n_random"], ib["n_features"]) == (5, 15) and b.shape[0] == 25
    assert np.array_equal(b[:7], a[:7]) and np.array_equal(b[10:22], a[7:19])
    t, lu, lv, tex = b[7], b[8], b[9], b[22:25]
    assert 0 <= t.min() and t.max() < 1 and t.std() > 0.2                      # the sample's time
    hit = np.abs(a[7:10]).sum(axis=0) > 0                                      # a first hit exists (n0 != 0)
    assert lu[hit].std() > 0.2 and lv[hit].std() > 0.2 and lu.max() < 1        # the light sample is a real random parameter
    assert (tex >= 0).all() and tex[:, hit].max() > 0.3 and (tex[:, ~hit] == 0).all()   # reflectance of what was hit


@needs_binary
@pytest.mark.gpu
def test_wide_record_through_the_b200(tmp_path):
    import rpf_b200
    dump = tmp_path / "wide.rpfb"
    W, H, P = 64, 48, 8
    r, out = render(tmp_path, f'Integrator "rpf" "string rpfrecord" "wide" "integer boxsizes" [9 7] "float sigmaseed" [0.5] "string rpfdump" "{dump}"')
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    planes, _, info = rpf_b200.read_buffer(str(dump))
    assert planes.shape[0] == 25
    o = ob.oracle_filter(planes, W, H, P, [9, 7], seed=0.5, threads=8, layout=(5, 15))
    img = read_pfm(out)
    ref = film_of(planes, o["colors"], W, H, P)
    m = np.isfinite(ref)
    assert np.array_equal(np.isfinite(img), m)
    assert np.abs(img[m] - ref[m]).max() <= 1e-5 * max(1.0, float(np.abs(ref[m]).max()))
