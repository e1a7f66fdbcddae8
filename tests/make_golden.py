"""Generates tests/golden/*.npz from the REFERENCE'S OWN sources compiled in place (oracle/_ref/*.so, built by
oracle/Makefile from /root/reference). Run in the build container only; the fixtures are committed so that the
oracle can be pinned where /root/reference does not exist.

    python tests/make_golden.py
"""
import ctypes as C
import os

import numpy as np

import oracle_bind as ob

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


def defined_lib():
    lib = C.CDLL(os.path.join(ob.ORACLE_DIR, "_ref", "libref_rpf_defined.so"))
    lib.ref_apply_rpf_filter.argtypes = [C.c_int] * 3 + [C.c_void_p] * 2 + [C.c_int] * 2 + [C.c_double]
    lib.ref_cf_weights.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


def synth(W, H, P, seed, kind):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "raytracer-rpf_b200"))
    import rpf_b200
    return rpf_b200.synth_planes(W, H, P, seed=seed, kind=kind)


def main():
    os.makedirs(GOLD, exist_ok=True)
    rng = np.random.default_rng(20261104)
    ref = ob.ref()
    assert ref is not None and ref.ref_has_filter(), "build oracle/_ref first (make -C oracle)"
    # 1. MutualInformation known answers (mi.cpp verbatim)
    xs, ys, ns, mis = [], [], [], []
    for n in [1, 2, 3, 8, 9, 64, 392, 2312]:
        for mode in range(4):
            x = rng.normal(size=n)
            y = [0.7 * x + rng.normal(size=n), rng.normal(size=n), np.full(n, 0.25), np.round(x)][mode]
            if mode == 3:
                x = np.round(rng.normal(size=n) * 2) / 2
            xs.append(x); ys.append(y); ns.append(n); mis.append(ob.ref_mi(x, y))
    np.savez_compressed(os.path.join(GOLD, "mi_known_answers.npz"), x=np.concatenate(xs), y=np.concatenate(ys),
                        n=np.array(ns), mi=np.array(mis))
    # 2. mean / sigma (ops.h verbatim), including cancellation cases
    vs, out = [], []
    for n in [1, 8, 8, 8, 392]:
        v = rng.normal(size=n).astype(np.float32).astype(np.float64)
        vs.append(v)
    vs.append(np.full(8, np.float32(0.1), dtype=np.float64))
    vs.append(np.full(7, np.float32(1234.567), dtype=np.float64))
    vs.append(np.array([1e8, 1e8 + 1, 1e8, 1e8 + 1], dtype=np.float64))
    for v in vs:
        out.append(ob.mean_std(ref.ref_mean_std, v))
    np.savez_compressed(os.path.join(GOLD, "mean_std.npz"), v=np.concatenate(vs), n=np.array([len(v) for v in vs]),
                        mean_sd=np.array(out))
    # 3. ComputeCFWeights: verbatim (alpha, W_r_c, beta[0..2] defined) and REF_DEFINED build
    dlib = defined_lib()
    zs, res_v, res_d = [], [], []
    for n, tweak in [(8, 0), (64, 1), (392, 2), (392, 3)]:
        z = rng.normal(size=(n, 19))
        if tweak >= 1:
            z[:, 7] += z[:, 5]; z[:, 2] += z[:, 0] + z[:, 8]
        if tweak >= 2:
            z[:, 9] = 0.5                      # constant feature -> 0/0 in the verbatim code
        if tweak >= 3:
            z[:, 13] = np.sign(z[:, 13]); z[:, 3] = 0.0
        zs.append(z)
        a, b, w = ob.ref_cf(z)
        res_v.append(np.concatenate([a, b[:3], [w]]))
        al, be, ww = np.zeros(3), np.zeros(12), C.c_double()
        dlib.ref_cf_weights(z.ctypes.data, n, al.ctypes.data, be.ctypes.data, C.byref(ww))
        res_d.append(np.concatenate([al, be, [ww.value]]))
    np.savez_compressed(os.path.join(GOLD, "cf_weights.npz"), z=np.concatenate([z.ravel() for z in zs]),
                        n=np.array([len(z) for z in zs]), verbatim=np.array(res_v), defined=np.array(res_d))
    # 4. whole filter, REF_DEFINED build of rpf.cpp:497-733, several seeds / kinds / boxes
    cases = [(20, 16, 8, 7, 0.002, 0), (20, 16, 8, 7, 0.5, 0), (20, 16, 8, 7, 2.0, 1), (16, 12, 4, 5, 0.5, 0),
             (12, 10, 8, 9, 1.0, 1)]
    for i, (W, H, P, box, seed, kind) in enumerate(cases):
        planes = synth(W, H, P, 1000 + i, kind)
        col = planes[2:5].astype(np.float64).copy()
        flag = dlib.ref_apply_rpf_filter(W, H, P, planes.ctypes.data, col.ctypes.data, box, 8, seed)
        np.savez_compressed(os.path.join(GOLD, f"filter_case{i}.npz"), planes=planes, colors=col,
                            meta=np.array([W, H, P, box, kind]), seed=seed, nan_exit=flag)
    # 5. verbatim filter in the identity regime (rpf.cpp as shipped, seed literal 0.002)
    W, H, P = 20, 16, 8
    planes = synth(W, H, P, 77, 0)
    col, flag = ob.ref_filter(planes, W, H, P, 7, threads=8)
    np.savez_compressed(os.path.join(GOLD, "filter_verbatim.npz"), planes=planes, colors=col, meta=np.array([W, H, P, 7, 0]),
                        nan_exit=flag)
    print("golden fixtures written to", GOLD)


# Paper mode: the reference's own rpf.cpp:302-733 with the REF_DEFINED patches plus the two hooks of
# oracle/ref_paper.sed (oracle/_ref/libref_rpf_paper.so), chained over the box sizes with one sigma per pass.
# (W, H, P, boxes, sigmas, epsilon, min sigma, world-position factor, synth kind)
PAPER_CASES = [
    (20, 16, 8, [7], [0.3], 1e-7, 0.0, 0.0, 0),                    # epsilon alone
    (20, 16, 8, [7], [0.3], 0.0, 0.1, 0.0, 1),                     # low-variance guard alone
    (18, 14, 8, [9, 7, 5], [0.5, 0.25, 0.125], 1e-7, 0.1, 30.0, 1),  # everything, sigma per pass
    (14, 12, 4, [5], [1.0], 1e-3, 0.05, 30.0, 0),
]
# the same plus random neighbour subsampling: (..., keep probability, rng seed)
PAPER_SUB_CASES = [
    (20, 16, 8, [7], [0.3], 0.0, 0.0, 0.0, 0, 0.5, 99),             # subsampling alone
    (18, 14, 8, [9, 7], [0.5, 0.25], 1e-7, 0.1, 30.0, 1, 0.25, 7),  # with everything else
]


def paper_lib():
    lib = C.CDLL(os.path.join(ob.ORACLE_DIR, "_ref", "libref_rpf_paper.so"))
    lib.ref_apply_rpf_filter.argtypes = [C.c_int] * 3 + [C.c_void_p] * 2 + [C.c_int] * 2 + [C.c_double]
    lib.ref_set_paper.argtypes = [C.c_double] * 3
    lib.ref_set_paper.restype = None
    lib.ref_set_subsampling.argtypes = [C.c_uint, C.c_ulonglong]
    lib.ref_set_subsampling.restype = None
    return lib


def paper_reference(lib, planes, W, H, P, boxes, sigmas, eps, min_sigma, pos, keep=1.0, rng_seed=0):
    """Colours after the passes and whether the reference would have exited on a NaN in any of them."""
    col = planes[2:5].astype(np.float64).copy()
    lib.ref_set_paper(eps, min_sigma, pos)
    lib.ref_set_subsampling(int(keep * (1 << 24)) if keep < 1.0 else 0, rng_seed)
    flag = 0
    for box, sg in zip(boxes, sigmas):
        flag |= lib.ref_apply_rpf_filter(W, H, P, planes.ctypes.data, col.ctypes.data, box, 8, sg)
    lib.ref_set_paper(0.0, 0.0, 0.0)
    lib.ref_set_subsampling(0, 0)
    return col, flag


def paper_main():
    lib = paper_lib()
    for i, (W, H, P, boxes, sigmas, eps, ms, pos, kind) in enumerate(PAPER_CASES):
        planes = synth(W, H, P, 3000 + i, kind)
        col, flag = paper_reference(lib, planes, W, H, P, boxes, sigmas, eps, ms, pos)
        np.savez_compressed(os.path.join(GOLD, f"paper_case{i}.npz"), planes=planes, colors=col, meta=np.array([W, H, P, kind]),
                            boxes=np.array(boxes), sigmas=np.array(sigmas), opts=np.array([eps, ms, pos]), nan_exit=flag)
    for i, (W, H, P, boxes, sigmas, eps, ms, pos, kind, keep, rs) in enumerate(PAPER_SUB_CASES):
        planes = synth(W, H, P, 3100 + i, kind)
        col, flag = paper_reference(lib, planes, W, H, P, boxes, sigmas, eps, ms, pos, keep, rs)
        np.savez_compressed(os.path.join(GOLD, f"paper_sub_case{i}.npz"), planes=planes, colors=col, meta=np.array([W, H, P, kind]),
                            boxes=np.array(boxes), sigmas=np.array(sigmas), opts=np.array([eps, ms, pos]), nan_exit=flag,
                            keep=keep, rng_seed=rs)
    print("paper-mode fixtures written to", GOLD)


# Film reconstruction cases, from oracle/_ref/libref_film.so = the reference's own film.h / film.cpp /
# filters lines. (W, H, P, filter kind, rx, ry, alpha, pixel bounds, scale, max luminance, ray weights?, synth kind)
FILM_CASES = [
    (24, 18, 8, 0, 0.5, 0.5, 0.0, (0, 0, 24, 18), 1.0, float("inf"), False, 0),    # pbrt default box filter
    (24, 18, 8, 1, 2.0, 2.0, 2.0, (0, 0, 24, 18), 1.0, float("inf"), False, 0),    # sample.pbrt's gaussian
    (20, 14, 4, 1, 1.5, 2.5, 1.0, (2, 1, 18, 12), 0.5, 0.8, True, 1),              # crop, scale, luminance clamp, ray weights
    (9, 7, 3, 0, 1.0, 1.0, 0.0, (0, 0, 9, 7), 2.0, float("inf"), True, 0),         # wide box filter
]


def film_inputs(i, W, H, P, kind, with_rw):
    planes = synth(W, H, P, 4000 + i, kind)
    rng = np.random.default_rng(900 + i)
    col = planes[2:5].astype(np.float64) * rng.uniform(0.5, 1.5, size=(3, W * H * P))  # doubles that are not floats
    rw = rng.uniform(0.25, 1.0, size=W * H * P).astype(np.float32) if with_rw else None
    return planes[0:2].copy(), col, rw


def film_main():
    os.makedirs(GOLD, exist_ok=True)
    assert ob.ref_film() is not None, "build oracle/_ref first (make -C oracle)"
    for i, (W, H, P, fk, rx, ry, alpha, bounds, scale, mx, with_rw, kind) in enumerate(FILM_CASES):
        pfilm, col, rw = film_inputs(i, W, H, P, kind, with_rw)
        rgb, table = ob.ref_film_reconstruct(W, H, P, pfilm, col, rw, fk, rx, ry, alpha, bounds, scale, mx)
        np.savez_compressed(os.path.join(GOLD, f"film_case{i}.npz"), pfilm=pfilm, colors=col,
                            ray_weight=rw if rw is not None else np.zeros(0, np.float32), rgb=rgb, table=table,
                            meta=np.array([W, H, P, fk, *bounds]), fmeta=np.array([rx, ry, alpha, scale, mx], np.float32))
    print("film fixtures written to", GOLD)


if __name__ == "__main__":
    import sys
    if len(sys.argv) > 1 and sys.argv[1] == "paper":
        paper_main()
    elif len(sys.argv) > 1 and sys.argv[1] == "film":
        film_main()
    else:
        main()
        film_main()
