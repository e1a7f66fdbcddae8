"""Round-2 fixtures, generated like tests/make_golden.py from the REFERENCE'S OWN sources compiled in place
(oracle/_ref/libref_rpf_defined.so = rpf.cpp:302-733 + mi.cpp with the REF_DEFINED patch list). Build container only.

    python tests/make_golden_r2.py big                 large boxes / many samples: where the general, large and wide-blend
                                                       GPU kernels live (the whole-filter pin stopped at box 9)
    python tests/make_golden_r2.py nonfinite           Inf / NaN colours and a NaN feature (P3 on non-finite input)
    python tests/make_golden_r2.py probe               one pixel's normalised neighbourhood + the 96 values of the reference's
                                                       own MutualInformation (mi.cpp) on it
    python tests/make_golden_r2.py real <dump.rpfb>    crops of a RENDERED feature buffer: scenes/sample.pbrt through the
                                                       replacement integrator's dump hook (host/rpf_integrator.cpp,
                                                       "string rpfdump"), identical to the reference producer's buffer
                                                       (oracle/_ref/pbrt_ref_dump)
"""
import os
import sys

import numpy as np

import make_golden as mg

GOLD = mg.GOLD

# (W, H, P, boxes, sigma seed, synth kind)
BIG_CASES = [
    (30, 26, 8, [17], 0.5, 0),        # general path: membership / ordered statistics / dependency / blend
    (24, 20, 8, [35], 0.5, 0),        # box 35: bins in the L2 scratch
    (14, 12, 16, [9], 0.5, 0),        # 16 spp: blend_wide_kernel
    (10, 8, 32, [9], 1.0, 1),         # 32 spp, degenerate input
    (26, 22, 8, [17, 7], 0.5, 0),     # two passes, general then fused
    (22, 18, 8, [17], 0.002, 1),      # the reference's literal seed on degenerate input
]


def run_defined(lib, planes, W, H, P, boxes, seed, colors=None):
    col = (planes[2:5].astype(np.float64) if colors is None else colors).copy()
    flag = 0
    for b in boxes:
        flag |= lib.ref_apply_rpf_filter(W, H, P, planes.ctypes.data, col.ctypes.data, b, 8, seed)
    return col, flag


def big_main():
    lib = mg.defined_lib()
    for i, (W, H, P, boxes, seed, kind) in enumerate(BIG_CASES):
        planes = mg.synth(W, H, P, 5000 + i, kind)
        col, flag = run_defined(lib, planes, W, H, P, boxes, seed)
        np.savez_compressed(os.path.join(GOLD, f"big_case{i}.npz"), planes=planes, colors=col, meta=np.array([W, H, P, kind]),
                            boxes=np.array(boxes), seed=seed, nan_exit=flag)
        print(f"big_case{i}: {W}x{H}x{P} boxes {boxes} nan_exit {flag}")


# (W, H, P, box, seed, what)
NONFINITE_CASES = [
    (16, 12, 8, 7, 0.5, "inf_colour"),
    (16, 12, 8, 7, 0.5, "nan_colour"),
    (16, 12, 8, 7, 0.5, "nan_feature"),
    (20, 16, 8, 17, 0.5, "inf_and_nan_colour"),   # general path
]


def poison(planes, W, H, P, what, case):
    """Fireflies: a few samples get a non-finite value (renderers produce them; the pinned build must agree on them)."""
    S = W * H * P
    rng = np.random.default_rng(700 + case)
    idx = rng.choice(S, size=3, replace=False)
    planes = planes.copy()
    if what in ("inf_colour", "inf_and_nan_colour"):
        planes[2, idx[0]] = np.inf
        planes[4, idx[1]] = np.inf
    if what in ("nan_colour", "inf_and_nan_colour"):
        planes[3, idx[2]] = np.nan
    if what == "nan_feature":
        planes[8, idx[0]] = np.nan
        planes[16, idx[1]] = np.nan
    return planes


def nonfinite_main():
    lib = mg.defined_lib()
    for i, (W, H, P, box, seed, what) in enumerate(NONFINITE_CASES):
        planes = poison(mg.synth(W, H, P, 6000 + i, 0), W, H, P, what, i)
        col, flag = run_defined(lib, planes, W, H, P, [box], seed)
        np.savez_compressed(os.path.join(GOLD, f"nonfinite_case{i}.npz"), planes=planes, colors=col, meta=np.array([W, H, P, box]),
                            seed=seed, nan_exit=flag, what=what)
        print(f"nonfinite_case{i} ({what}): NaN colours in the reference output {int(np.isnan(col).sum())}, nan_exit {flag}")


# crops of scenes/sample.pbrt (700x700, halton 8 spp): (x0, y0, w, h, boxes, seed)
REAL_CROPS = [
    (530, 530, 64, 64, [7], 0.002),      # killeroo silhouette against the back wall and the black light-free corner
    (300, 280, 48, 48, [17, 7], 0.002),  # body of the killeroo: curved normals, the general path first
    (120, 600, 40, 32, [7], 2.0),        # floor: flat normals (sigma 0 rejects every neighbour), wide sigma
]


def real_main(dump):
    sys.path.insert(0, os.path.join(os.path.dirname(mg.HERE), "raytracer-rpf_b200"))
    import rpf_b200
    pl, rw, info = rpf_b200.read_buffer(dump)
    W, H, P = info["width"], info["height"], info["spp"]
    A = pl.reshape(19, H, W, P)
    lib = mg.defined_lib()
    for i, (x0, y0, w, h, boxes, seed) in enumerate(REAL_CROPS):
        planes = np.ascontiguousarray(A[:, y0:y0 + h, x0:x0 + w, :]).reshape(19, w * h * P)
        col, flag = run_defined(lib, planes, w, h, P, boxes, seed)
        np.savez_compressed(os.path.join(GOLD, f"real_crop{i}.npz"), planes=planes, colors=col, meta=np.array([w, h, P, x0, y0]),
                            boxes=np.array(boxes), seed=seed, nan_exit=flag, scene="scenes/sample.pbrt 700x700 halton 8 spp")
        print(f"real_crop{i}: {w}x{h}x{P} at ({x0},{y0}) boxes {boxes}: NaN colours {int(np.isnan(col).sum())} of {col.size}, nan_exit {flag}")


# one pixel, component level: the normalised neighbourhood the oracle hands to MutualInformation and what the REFERENCE'S
# OWN mi.cpp returns for the 96 pairs of rpf.cpp:416-442 on it. (W, H, P, box, pixel x, pixel y, synth kind)
PROBE_CASES = [(24, 20, 8, 7, 11, 9, 0), (24, 20, 8, 17, 12, 10, 0), (16, 12, 8, 9, 3, 2, 1)]
PROBE_PAIRS = ([(7 + i, b) for i in range(12) for b in (5, 6, 0, 1)] +
               [(2 + i, b) for i in range(3) for b in (5, 6, 0, 1) + tuple(range(7, 19))])


def probe_main():
    import oracle_bind as ob
    for i, (W, H, P, box, px, py, kind) in enumerate(PROBE_CASES):
        planes = mg.synth(W, H, P, 7000 + i, kind)
        r = ob.oracle_pixel_probe(planes, W, H, P, box, px, py, seed=0.5)
        ref_mi = np.array([ob.ref_mi(r["z"][:, a], r["z"][:, b]) for a, b in PROBE_PAIRS])
        np.savez_compressed(os.path.join(GOLD, f"probe_case{i}.npz"), meta=np.array([W, H, P, box, px, py, kind, 7000 + i]),
                            z=r["z"], ref_mi=ref_mi)
        print(f"probe_case{i}: N {r['N']} bins {r['bins']} oracle == mi.cpp: {np.array_equal(ref_mi, r['mi'])}")


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else ""
    if what == "big":
        big_main()
    elif what == "nonfinite":
        nonfinite_main()
    elif what == "real":
        real_main(sys.argv[2])
    elif what == "probe":
        probe_main()
    else:
        print(__doc__)
