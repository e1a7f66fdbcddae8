"""ctypes bindings to the CPU oracle (oracle/liboracle_rpf.so) and, when present, to the reference's own
sources compiled in place (oracle/_ref/libref_rpf.so). Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
P1, P2, P3, DEFINED = 1, 2, 4, 7
NCH = 19


class OracleArgs(C.Structure):
    _fields_ = [
        ("width", C.c_int), ("height", C.c_int), ("spp", C.c_int),
        ("planes", C.c_void_p), ("colors", C.c_void_p), ("box_sizes", C.c_void_p),
        ("n_iter", C.c_int), ("sigma_fc_seed", C.c_double), ("flags", C.c_int), ("n_threads", C.c_int),
        ("y0", C.c_int), ("y1", C.c_int), ("x0", C.c_int), ("x1", C.c_int),
        ("nbh_size", C.c_void_p), ("alpha", C.c_void_p), ("beta", C.c_void_p), ("wrc", C.c_void_p),
        ("nan_count", C.c_longlong),
        ("subsample_keep_q24", C.c_uint), ("subsample_seed", C.c_ulonglong), ("global_row0", C.c_int),
        ("w_epsilon", C.c_double), ("min_sigma", C.c_double), ("world_pos_sigmas", C.c_double),
        ("sigma_seeds", C.c_void_p), ("hdr_clamp_sigmas", C.c_double),
        ("n_random", C.c_int), ("n_features", C.c_int),
    ]


_oracle = None
_ref = None


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True, stdout=subprocess.DEVNULL)


def oracle():
    global _oracle
    if _oracle is None:
        path = os.path.join(ORACLE_DIR, "liboracle_rpf.so")
        if not os.path.exists(path):
            build_oracle()
        lib = C.CDLL(path)
        lib.rpf_oracle_filter.argtypes = [C.POINTER(OracleArgs)]
        lib.rpf_oracle_filter.restype = C.c_int
        lib.rpf_oracle_mi.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        lib.rpf_oracle_mi.restype = C.c_double
        lib.rpf_oracle_mean_std.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        lib.rpf_oracle_cf_weights.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.rpf_oracle_hdr_clamp.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double]
        lib.rpf_oracle_hdr_clamp.restype = None
        _oracle = lib
    return _oracle


def ref():
    """The reference's own code (None when oracle/_ref was not built, e.g. never on a box without it)."""
    global _ref
    if _ref is None:
        path = os.path.join(ORACLE_DIR, "_ref", "libref_rpf.so")
        if not os.path.exists(path):
            return None
        lib = C.CDLL(path)
        lib.ref_mi.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        lib.ref_mi.restype = C.c_double
        lib.ref_mean_std.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        lib.ref_within_3sigma.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.ref_within_3sigma.restype = C.c_int
        lib.ref_has_filter.restype = C.c_int
        if lib.ref_has_filter():
            lib.ref_cf_weights.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
            lib.ref_apply_rpf_filter.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double]
            lib.ref_apply_rpf_filter.restype = C.c_int
        _ref = lib
    return _ref


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def oracle_filter(planes, W, H, P, boxes, seed=0.002, flags=DEFINED, colors=None, threads=0,
                  rows=None, cols=None, debug=False, subsample=None, paper=None, layout=None):
    # layout = (n_random, n_features) of a wider record or None = the reference's (2, 12)
    # paper = dict(w_epsilon=, min_sigma=, world_pos_sigmas=, sigma_seeds=[per iteration], hdr_clamp_sigmas=) or None
    # subsample = (keep probability, rng seed, global row of local row 0) or None
    """planes: float32 [19, W*H*P]. Returns dict(colors=[3,S] float64, N, alpha, beta, wrc, nan_count)."""
    planes = np.ascontiguousarray(planes, dtype=np.float32)
    S = W * H * P
    nr, nf = layout or (2, 12)
    assert planes.shape == (5 + nr + nf, S)
    col = np.ascontiguousarray(planes[2:5].astype(np.float64) if colors is None else colors, dtype=np.float64).copy()
    bx = np.asarray(boxes, dtype=np.int32)
    a = OracleArgs()
    a.width, a.height, a.spp = W, H, P
    a.planes, a.colors, a.box_sizes = _p(planes), _p(col), _p(bx)
    a.n_iter, a.sigma_fc_seed, a.flags, a.n_threads = len(bx), seed, flags, threads
    a.n_random, a.n_features = nr, nf
    a.y0, a.y1 = rows if rows else (0, 0)
    a.x0, a.x1 = cols if cols else (0, 0)
    if subsample is not None:
        a.subsample_keep_q24 = int(subsample[0] * (1 << 24))
        a.subsample_seed = subsample[1]
        a.global_row0 = subsample[2] if len(subsample) > 2 else 0
    if paper:
        a.w_epsilon = paper.get("w_epsilon", 0.0)
        a.min_sigma = paper.get("min_sigma", 0.0)
        a.world_pos_sigmas = paper.get("world_pos_sigmas", 0.0)
        a.hdr_clamp_sigmas = paper.get("hdr_clamp_sigmas", 0.0)
        if paper.get("sigma_seeds") is not None:
            sg = np.asarray(paper["sigma_seeds"], np.float64)
            assert len(sg) == len(bx)
            a.sigma_seeds = _p(sg)
    out = {}
    if debug:
        out["N"] = np.zeros(W * H, np.int32)
        out["alpha"] = np.zeros((W * H, 3))
        out["beta"] = np.zeros((W * H, nf))
        out["wrc"] = np.zeros(W * H)
        a.nbh_size, a.alpha, a.beta, a.wrc = _p(out["N"]), _p(out["alpha"]), _p(out["beta"]), _p(out["wrc"])
    rc = oracle().rpf_oracle_filter(C.byref(a))
    assert rc == 0
    out["colors"] = col
    out["nan_count"] = a.nan_count
    return out


def oracle_mi(x, y):
    x = np.ascontiguousarray(x, np.float64)
    y = np.ascontiguousarray(y, np.float64)
    return oracle().rpf_oracle_mi(_p(x), _p(y), len(x))


def ref_mi(x, y):
    x = np.ascontiguousarray(x, np.float64)
    y = np.ascontiguousarray(y, np.float64)
    return ref().ref_mi(_p(x), _p(y), len(x))


def mean_std(lib_fn, v):
    v = np.ascontiguousarray(v, np.float64)
    m, s = C.c_double(), C.c_double()
    lib_fn(_p(v), len(v), C.byref(m), C.byref(s))
    return m.value, s.value


def oracle_cf(z, flags=DEFINED):
    z = np.ascontiguousarray(z, np.float64)
    al, be, w = np.zeros(3), np.zeros(12), C.c_double()
    oracle().rpf_oracle_cf_weights(_p(z), z.shape[0], flags, _p(al), _p(be), C.byref(w))
    return al, be, w.value


def ref_cf(z):
    z = np.ascontiguousarray(z, np.float64)
    al, be, w = np.zeros(3), np.zeros(12), C.c_double()
    ref().ref_cf_weights(_p(z), z.shape[0], _p(al), _p(be), C.byref(w))
    return al, be, w.value


def ref_filter(planes, W, H, P, box, colors=None, threads=1):
    planes = np.ascontiguousarray(planes, dtype=np.float32)
    col = np.ascontiguousarray(planes[2:5].astype(np.float64) if colors is None else colors, dtype=np.float64).copy()
    flag = ref().ref_apply_rpf_filter(W, H, P, _p(planes), _p(col), box, threads, 0.002)
    return col, flag


# ---- Film reconstruction -----------------------------------------------------------------------
def ref_film():
    """oracle/_ref/libref_film.so: the reference's FilmTile::AddSample / MergeFilmTile / WriteImage lines, or None."""
    path = os.path.join(ORACLE_DIR, "_ref", "libref_film.so")
    if not os.path.exists(path):
        return None
    lib = C.CDLL(path)
    lib.ref_film_reconstruct.argtypes = ([C.c_int] * 5 + [C.c_void_p] * 3 + [C.c_int] + [C.c_float] * 3 + [C.c_int] * 4 +
                                         [C.c_float] * 2 + [C.c_void_p] * 2)
    return lib


def ref_film_reconstruct(W, H, P, pfilm, colors, ray_weight, kind, rx, ry, alpha, bounds, scale, max_lum):
    pfilm = np.ascontiguousarray(pfilm, np.float32)
    colors = np.ascontiguousarray(colors, np.float64)
    rw = None if ray_weight is None else np.ascontiguousarray(ray_weight, np.float32)
    x0, y0, x1, y1 = bounds
    rgb = np.empty((y1 - y0, x1 - x0, 3), np.float32)
    table = np.empty(256, np.float32)
    ref_film().ref_film_reconstruct(W, H, P, 0, 0, _p(pfilm), _p(colors), None if rw is None else _p(rw), kind, rx, ry, alpha,
                                    x0, y0, x1, y1, scale, max_lum, _p(rgb), _p(table))
    return rgb, table


def oracle_film_table(kind, rx, ry, alpha):
    lib = oracle()
    lib.rpf_oracle_film_table.argtypes = [C.c_int] + [C.c_float] * 3 + [C.c_void_p]
    t = np.empty(256, np.float32)
    lib.rpf_oracle_film_table(kind, rx, ry, alpha, _p(t))
    return t


def oracle_film(W, H, P, pfilm, colors, ray_weight, table, rx, ry, bounds, scale, max_lum):
    lib = oracle()
    lib.rpf_oracle_film.argtypes = ([C.c_int] * 3 + [C.c_void_p] * 4 + [C.c_float] * 2 + [C.c_int] * 4 + [C.c_float] * 2 + [C.c_void_p])
    pfilm = np.ascontiguousarray(pfilm, np.float32)
    colors = np.ascontiguousarray(colors, np.float64)
    rw = None if ray_weight is None else np.ascontiguousarray(ray_weight, np.float32)
    table = np.ascontiguousarray(table, np.float32)
    x0, y0, x1, y1 = bounds
    rgb = np.empty((y1 - y0, x1 - x0, 3), np.float32)
    lib.rpf_oracle_film(W, H, P, _p(pfilm), _p(colors), None if rw is None else _p(rw), _p(table), rx, ry, x0, y0, x1, y1,
                        scale, max_lum, _p(rgb))
    return rgb


def oracle_pixel_probe(planes, W, H, P, box, px, py, seed=0.002, flags=DEFINED, colors=None, bins_cap=1024, layout=None):
    """One pixel of one pass: dict(mi [96], N, bins, marginals [19, bins], z [N, 19])."""
    planes = np.ascontiguousarray(planes, dtype=np.float32)
    bx = np.asarray([box], dtype=np.int32)
    a = OracleArgs()
    a.width, a.height, a.spp = W, H, P
    col = None if colors is None else np.ascontiguousarray(colors, np.float64)
    a.planes, a.colors, a.box_sizes = _p(planes), (None if col is None else _p(col)), _p(bx)
    a.n_iter, a.sigma_fc_seed, a.flags = 1, seed, flags
    nr, nf = layout or (2, 12)
    a.n_random, a.n_features = nr, nf
    nch, npairs = 5 + nr + nf, nf * (nr + 2) + 3 * (nr + 2 + nf)
    lib = oracle()
    lib.rpf_oracle_pixel_probe.argtypes = [C.POINTER(OracleArgs), C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                           C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    mi, n, nb = np.zeros(npairs), C.c_int(), C.c_int()
    marg = np.zeros((nch, bins_cap), np.int32)
    ncap = box * box * P
    z = np.zeros((ncap, nch))
    rc = lib.rpf_oracle_pixel_probe(C.byref(a), px, py, _p(mi), C.byref(n), C.byref(nb), _p(marg), bins_cap, _p(z), ncap)
    assert rc == 0, rc
    return dict(mi=mi, N=n.value, bins=nb.value, marginals=marg[:, :nb.value].copy(), z=z[:n.value].copy())
