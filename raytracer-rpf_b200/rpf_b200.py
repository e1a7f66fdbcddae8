"""ctypes binding of include/rpf_b200.h (lib/librpf_b200.so) for tests and bench.py.

The product is the C ABI; this file only marshals numpy arrays / device pointers into it. There is no
fallback of any kind: if the CUDA library is missing or no sm_100 device is present, calls raise RPFError.
Names mirror the reference's seam: RPFFilter.apply() is the loop `for box_size in box_sizes: ApplyRPFFilter(...)`
of reference src/custom/rpf.cpp:767-775.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RPF_B200_LIB", os.path.join(HERE, "lib", "librpf_b200.so"))
NCH = 19
FLAGS_DEFINED = 7
FLAGS_VERBATIM = 0x100
FLAG_FORCE_GENERAL = 0x200
FLAG_FORCE_LARGE = 0x400

# every symbol include/rpf_b200.h declares
ABI_SYMBOLS = [
    "rpf_create", "rpf_destroy", "rpf_last_error", "rpf_upload", "rpf_upload_device", "rpf_upload_colors_f64",
    "rpf_filter", "rpf_filter_rows", "rpf_swap_colors", "rpf_color_device_ptr", "rpf_stream", "rpf_synchronize",
    "rpf_download_colors", "rpf_download_colors_f64", "rpf_download_debug", "rpf_reconstruct_box",
    "rpf_get_stats", "rpf_reset_stats", "rpf_set_profiling", "rpf_get_kernel_times", "rpf_max_neighbourhood",
    "rpf_synth_planes", "rpf_reconstruct_film", "rpf_film_filter_table",
    "rpf_buffer_write", "rpf_buffer_info_read", "rpf_buffer_read", "rpf_set_subsampling",
    "rpf_set_paper_mode", "rpf_filter_sigmas", "rpf_hdr_clamp", "rpf_fused_max_candidates",
    "rpf_band_plan", "rpf_band_unique_id", "rpf_band_attach", "rpf_host_alloc", "rpf_host_free",
    "rpf_set_debug_pixel", "rpf_download_debug_mi", "rpf_reset_colors",
    "rpf_num_channels", "rpf_num_pairs", "rpf_layout_supported", "rpf_synth_planes_layout", "rpf_filter_rows2",
    "rpf_band_plan_weighted", "rpf_filter_pass_noalloc",
]


class RPFError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("spp", C.c_int), ("device", C.c_int),
                ("flags", C.c_int), ("debug", C.c_int), ("n_gpus", C.c_int), ("device_ids", C.c_int * 8), ("n_random", C.c_int),
                ("n_features", C.c_int), ("max_box", C.c_int)]


class Stats(C.Structure):
    _fields_ = [("pixels_filtered", C.c_longlong), ("nbh_sum", C.c_longlong), ("nbh_min", C.c_int),
                ("nbh_max", C.c_int), ("nan_channels", C.c_longlong), ("kernel_launches", C.c_longlong),
                ("last_filter_ms", C.c_double), ("halo_bytes", C.c_longlong)]


class Band(C.Structure):
    """rpf_band: this context is one row band of a larger frame, held by one process of `world`."""
    _fields_ = [("global_height", C.c_int), ("buf0", C.c_int), ("own0", C.c_int), ("own1", C.c_int), ("rank", C.c_int),
                ("world", C.c_int)]


class PaperOpts(C.Structure):
    """rpf_paper_opts: Sen & Darabi's additions, none of them in the reference; all 0 = reference arithmetic."""
    _fields_ = [("w_epsilon", C.c_double), ("min_sigma", C.c_double), ("world_pos_sigmas", C.c_double)]


class FilmDesc(C.Structure):
    """rpf_film_desc: what pbrt's Film holds (core/film.h:58-105)."""
    _fields_ = [("sample_x0", C.c_int), ("sample_y0", C.c_int), ("pixel_x0", C.c_int), ("pixel_y0", C.c_int),
                ("pixel_x1", C.c_int), ("pixel_y1", C.c_int), ("radius_x", C.c_float), ("radius_y", C.c_float),
                ("filter_table", C.c_float * 256), ("scale", C.c_float), ("max_sample_luminance", C.c_float)]


def film_filter_table(kind, rx, ry, alpha=2.0):
    """Film::filterTable (core/film.cpp:69-77) for pbrt's box (kind 0) or gaussian (kind 1) filter."""
    t = np.empty(256, np.float32)
    rc = lib().rpf_film_filter_table(kind, C.c_float(rx), C.c_float(ry), C.c_float(alpha), _p(t))
    if rc != 0:
        raise RPFError(f"rpf_film_filter_table failed ({rc})")
    return t


class BufferInfo(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("spp", C.c_int), ("sample_x0", C.c_int), ("sample_y0", C.c_int),
                ("has_ray_weight", C.c_int), ("n_random", C.c_int), ("n_features", C.c_int)]


def write_buffer(path, planes, width, height, spp, ray_weight=None, sample_origin=(0, 0), n_random=2, n_features=12):
    """Feature-buffer file (csrc/rpf_io.cpp): 5 + n_random + n_features float32 planes + optional ray weights."""
    nch = 5 + n_random + n_features
    assert planes.dtype == np.float32 and planes.shape == (nch, width * height * spp) and planes.flags.c_contiguous
    info = BufferInfo(width, height, spp, sample_origin[0], sample_origin[1], 0, n_random, n_features)
    ptrs = (C.c_void_p * nch)(*[planes[c].ctypes.data for c in range(nch)])
    rw = None if ray_weight is None else np.ascontiguousarray(ray_weight, np.float32)
    rc = hostio().rpf_buffer_write(path.encode(), C.byref(info), ptrs, None if rw is None else _p(rw))
    if rc != 0:
        raise RPFError(f"rpf_buffer_write({path}) failed ({rc})")


def read_buffer(path):
    """-> (planes [19, S] float32, ray_weight [S] float32, info dict)"""
    info = BufferInfo()
    rc = hostio().rpf_buffer_info_read(path.encode(), C.byref(info))
    if rc != 0:
        raise RPFError(f"{path}: not a feature-buffer file ({rc})")
    S = info.width * info.height * info.spp
    planes, rw = np.empty((5 + info.n_random + info.n_features, S), np.float32), np.empty(S, np.float32)
    rc = hostio().rpf_buffer_read(path.encode(), _p(planes), _p(rw))
    if rc != 0:
        raise RPFError(f"{path}: truncated or unreadable ({rc})")
    return planes, rw, {k: getattr(info, k) for k, _ in BufferInfo._fields_}


KERNEL_NAMES = ["pixel_stats_kernel", "rpf_pixel_kernel", "membership_kernel", "ordered_stats_kernel", "dependency_kernel",
                "blend_kernel"]


class KernelTimes(C.Structure):
    _fields_ = [("ms", C.c_double * 6), ("launches", C.c_longlong * 6)]


_lib = None
_hostio = None


def hostio():
    """lib/librpf_hostio.so: the two host-only helpers of the ABI (rpf_synth_planes, rpf_buffer_*), without CUDA. A process
    that only generates or reads feature buffers (bench.py's CPU reference arm) never maps the GPU library."""
    global _hostio
    if _hostio is None:
        path = os.path.join(os.path.dirname(LIB_PATH), "librpf_hostio.so")
        if not os.path.exists(path):
            raise RPFError(f"{path} is not built (run `python -c 'import __graft_entry__ as g; g.build()'`)")
        L = C.CDLL(path)
        L.rpf_synth_planes.argtypes = [C.c_int, C.c_int, C.c_int, C.c_ulonglong, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.rpf_synth_planes_layout.argtypes = [C.c_int, C.c_int, C.c_int, C.c_ulonglong, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        _hostio = L
    return _hostio


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RPFError(f"{LIB_PATH} is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                           "there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        vp, i, d = C.c_void_p, C.c_int, C.c_double
        L.rpf_create.argtypes = [C.POINTER(vp), C.POINTER(Config)]
        L.rpf_destroy.argtypes = [vp]
        L.rpf_destroy.restype = None
        L.rpf_last_error.argtypes = [vp]
        L.rpf_last_error.restype = C.c_char_p
        L.rpf_upload.argtypes = [vp, C.POINTER(vp)]
        L.rpf_upload_device.argtypes = [vp, C.POINTER(vp)]
        L.rpf_upload_colors_f64.argtypes = [vp, vp]
        L.rpf_filter.argtypes = [vp, vp, i, d]
        L.rpf_filter_rows.argtypes = [vp, i, d, i, i]
        L.rpf_filter_rows2.argtypes = [vp, i, d, i, i, i, i]
        L.rpf_set_paper_mode.argtypes = [vp, vp]
        L.rpf_filter_sigmas.argtypes = [vp, vp, vp, i]
        L.rpf_hdr_clamp.argtypes = [vp, d, i, i]
        L.rpf_swap_colors.argtypes = [vp]
        L.rpf_color_device_ptr.argtypes = [vp, i, C.POINTER(vp)]
        L.rpf_stream.argtypes = [vp, C.POINTER(vp)]
        L.rpf_synchronize.argtypes = [vp]
        L.rpf_download_colors.argtypes = [vp, vp, vp, vp]
        L.rpf_download_colors_f64.argtypes = [vp, vp]
        L.rpf_download_debug.argtypes = [vp, vp, vp, vp, vp]
        L.rpf_reconstruct_box.argtypes = [vp, vp]
        L.rpf_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.rpf_reset_stats.argtypes = [vp]
        L.rpf_set_profiling.argtypes = [vp, i]
        L.rpf_get_kernel_times.argtypes = [vp, C.POINTER(KernelTimes)]
        L.rpf_max_neighbourhood.argtypes = []
        L.rpf_synth_planes.argtypes = [i, i, i, C.c_ulonglong, i, i, i, vp]
        L.rpf_band_plan.argtypes = [i, i, vp, i, vp]
        L.rpf_band_plan_weighted.argtypes = [i, i, vp, i, vp, vp]
        L.rpf_band_unique_id.argtypes = [vp]
        L.rpf_band_attach.argtypes = [vp, C.POINTER(Band), vp]
        L.rpf_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
        L.rpf_host_free.argtypes = [vp]
        L.rpf_host_free.restype = None
        L.rpf_reset_colors.argtypes = [vp]
        L.rpf_num_channels.argtypes = [vp]
        L.rpf_num_pairs.argtypes = [vp]
        L.rpf_layout_supported.argtypes = [i, i]
        L.rpf_set_debug_pixel.argtypes = [vp, i, i]
        L.rpf_download_debug_mi.argtypes = [vp, vp, C.POINTER(i), C.POINTER(i), vp, i]
        L.rpf_filter_pass_noalloc.argtypes = [vp, vp, vp, i, i, i, i, d, i, vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def band_plan(height, n_bands, box_sizes, row_weight=None):
    """rpf_band_plan(_weighted): first owned row of each of n_bands equal-work row bands, [n_bands + 1] ints (host only).
    row_weight: measured cost per row ([height] doubles) instead of the geometric model."""
    bx = np.asarray(box_sizes, np.int32)
    out = np.zeros(n_bands + 1, np.int32)
    if row_weight is None:
        rc = lib().rpf_band_plan(height, n_bands, _p(bx), len(bx), _p(out))
    else:
        w = np.ascontiguousarray(row_weight, np.float64)
        assert w.shape == (height,)
        rc = lib().rpf_band_plan_weighted(height, n_bands, _p(bx), len(bx), _p(w), _p(out))
    if rc != 0:
        raise RPFError(f"rpf_band_plan({height} rows, {n_bands} bands, boxes {list(box_sizes)}) failed ({rc})")
    return [int(v) for v in out]


NOALLOC_MAX_PIXELS = 65536  # RPF_NOALLOC_MAX_PIXELS (include/rpf_b200.h)


def filter_pass_noalloc(d_planes, d_colors_in, d_colors_out, width, height, spp, box, seed, flags=0, stream=0):
    """rpf_filter_pass_noalloc: one pass over DEVICE buffers (raw addresses) with the library's static scratch -- no
    allocation anywhere in the call. Only enqueues on `stream` (a cudaStream_t value, 0 = default stream)."""
    rc = lib().rpf_filter_pass_noalloc(d_planes, d_colors_in, d_colors_out, width, height, spp, box, float(seed), flags, stream or None)
    if rc != 0:
        raise RPFError(f"rpf_filter_pass_noalloc failed ({rc}): {lib().rpf_last_error(None).decode()}")


def nccl_unique_id():
    buf = (C.c_char * 128)()
    rc = lib().rpf_band_unique_id(buf)
    if rc != 0:
        raise RPFError(f"rpf_band_unique_id failed ({rc}): {lib().rpf_last_error(None).decode()}")
    return bytes(buf)


def synth_planes(width, height, spp, seed=1234, kind=0, x_off=0, y_off=0, out=None, n_random=2, n_features=12):
    """Synthetic feature buffer of SURVEY.md 8(d): float32 [5 + n_random + n_features, W*H*spp] (host side, deterministic);
    19 planes for the reference's record."""
    nch = 5 + n_random + n_features
    planes = out if out is not None else np.empty((nch, width * height * spp), np.float32)
    assert planes.dtype == np.float32 and planes.flags.c_contiguous and planes.size == nch * width * height * spp
    rc = hostio().rpf_synth_planes_layout(width, height, spp, seed, kind, x_off, y_off, n_random, n_features, _p(planes))
    if rc != 0:
        raise RPFError(f"rpf_synth_planes failed ({rc})")
    return planes


class RPFFilter:
    """One device context = one SamplingFilm (reference src/custom/sample_film.h:28-45) resident in HBM."""

    def __init__(self, width, height, spp, device=0, flags=FLAGS_DEFINED, debug=False, force_general=False, force_large=False,
                 devices=None, max_box=0, n_random=0, n_features=0):
        """devices: list of CUDA ordinals -> one process, one row band per device (rpf_config.n_gpus / device_ids)."""
        if force_general:
            flags |= FLAG_FORCE_GENERAL
        if force_large:
            flags |= FLAG_FORCE_LARGE
        self.W, self.H, self.P = width, height, spp
        self.S = width * height * spp
        self._h = C.c_void_p()
        devs = list(devices) if devices else []
        cfg = Config(width, height, spp, device, flags, 1 if debug else 0, len(devs) if len(devs) > 1 else 0,
                     (C.c_int * 8)(*(devs + [0] * (8 - len(devs)))), n_random, n_features, max_box)
        rc = lib().rpf_create(C.byref(self._h), C.byref(cfg))
        if rc != 0:
            raise RPFError(f"rpf_create failed ({rc}): {lib().rpf_last_error(None).decode()}")
        self.nch, self.n_pairs = lib().rpf_num_channels(self._h), lib().rpf_num_pairs(self._h)
        self.nf = (n_features or 12)

    def _check(self, rc):
        if rc != 0:
            raise RPFError(f"rpf call failed ({rc}): {lib().rpf_last_error(self._h).decode()}")

    def close(self):
        if self._h:
            lib().rpf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- data movement --------------------------------------------------------------------------
    def upload(self, planes):
        """planes: float32 [nch, S] host array (pinned or pageable); nch = 19 for the reference's record."""
        assert planes.dtype == np.float32 and planes.shape == (self.nch, self.S) and planes.flags.c_contiguous
        ptrs = (C.c_void_p * self.nch)(*[planes[c].ctypes.data for c in range(self.nch)])
        self._check(lib().rpf_upload(self._h, ptrs))

    def upload_device(self, dev_ptr):
        """dev_ptr: integer device address of a contiguous float32 [19, S] buffer."""
        ptrs = (C.c_void_p * self.nch)(*[dev_ptr + 4 * self.S * c for c in range(self.nch)])
        self._check(lib().rpf_upload_device(self._h, ptrs))

    def reset_colors(self):
        """Working colours <- the uploaded planes[2..4] (device side)."""
        self._check(lib().rpf_reset_colors(self._h))

    def upload_colors_f64(self, rgb):
        rgb = np.ascontiguousarray(rgb, np.float64)
        assert rgb.shape == (3, self.S)
        self._check(lib().rpf_upload_colors_f64(self._h, _p(rgb)))

    def download_colors(self, out=None):
        out = out if out is not None else np.empty((3, self.S), np.float32)
        self._check(lib().rpf_download_colors(self._h, _p(out[0]), _p(out[1]), _p(out[2])))
        return out

    def download_colors_f64(self):
        out = np.empty((3, self.S), np.float64)
        self._check(lib().rpf_download_colors_f64(self._h, _p(out)))
        return out

    def download_debug(self):
        n = self.W * self.H
        N, al, be, w = np.zeros(n, np.int32), np.zeros((n, 3)), np.zeros((n, self.nf)), np.zeros(n)
        self._check(lib().rpf_download_debug(self._h, _p(N), _p(al), _p(be), _p(w)))
        return dict(N=N, alpha=al, beta=be, wrc=w)

    def set_debug_pixel(self, x, y):
        self._check(lib().rpf_set_debug_pixel(self._h, x, y))

    def download_debug_mi(self, bins_cap=1024):
        """-> dict(mi [96], N, bins, marginals [19, bins]) of the pixel chosen with set_debug_pixel."""
        mi, n, nb = np.zeros(self.n_pairs), C.c_int(), C.c_int()
        marg = np.zeros((self.nch, bins_cap), np.uint32)
        self._check(lib().rpf_download_debug_mi(self._h, _p(mi), C.byref(n), C.byref(nb), _p(marg), bins_cap))
        return dict(mi=mi, N=n.value, bins=nb.value, marginals=marg[:, :nb.value].copy())

    def reconstruct_box(self):
        out = np.empty((self.H, self.W, 3), np.float32)
        self._check(lib().rpf_reconstruct_box(self._h, _p(out)))
        return out

    def reconstruct_film(self, table, rx, ry, pixel_bounds=None, sample_origin=(0, 0), scale=1.0,
                         max_sample_luminance=float("inf"), ray_weight=None):
        """Film reconstruction with any filter (rpf.cpp:776-804 -> core/film.h:121-161, film.cpp:117-130,169-205).
        pixel_bounds = (x0, y0, x1, y1) croppedPixelBounds, default the context's extent. Returns [h, w, 3] float32."""
        x0, y0, x1, y1 = pixel_bounds or (sample_origin[0], sample_origin[1], sample_origin[0] + self.W, sample_origin[1] + self.H)
        d = FilmDesc(sample_origin[0], sample_origin[1], x0, y0, x1, y1, rx, ry, (C.c_float * 256)(*np.asarray(table, np.float32)),
                     scale, max_sample_luminance)
        out = np.empty((y1 - y0, x1 - x0, 3), np.float32)
        rw = None if ray_weight is None else np.ascontiguousarray(ray_weight, np.float32)
        self._check(lib().rpf_reconstruct_film(self._h, C.byref(d), None if rw is None else _p(rw), _p(out)))
        return out

    # -- the filter --------------------------------------------------------------------------------
    def apply(self, box_sizes=(7,), sigma_fc_seed=0.002):
        bx = np.asarray(box_sizes, np.int32)
        self._check(lib().rpf_filter(self._h, _p(bx), len(bx), sigma_fc_seed))

    def apply_sigmas(self, box_sizes, sigma_seeds):
        """One sigma_fc_seed per iteration (paper mode; the reference uses 0.002 throughout, rpf.cpp:533)."""
        bx = np.asarray(box_sizes, np.int32)
        sg = np.asarray(sigma_seeds, np.float64)
        assert len(bx) == len(sg)
        self._check(lib().rpf_filter_sigmas(self._h, _p(bx), _p(sg), len(bx)))

    def set_paper_mode(self, w_epsilon=0.0, min_sigma=0.0, world_pos_sigmas=0.0):
        """Paper-mode options (include/rpf_b200.h rpf_paper_opts); no arguments restores the reference's arithmetic."""
        o = PaperOpts(w_epsilon, min_sigma, world_pos_sigmas)
        self._check(lib().rpf_set_paper_mode(self._h, C.byref(o)))

    def hdr_clamp(self, k_sigmas=1.0, y0=0, y1=0):
        """Spike removal on the current colours of rows [y0,y1) (y1 <= 0: all rows)."""
        self._check(lib().rpf_hdr_clamp(self._h, C.c_double(k_sigmas), y0, y1))

    def filter_rows(self, box, y0, y1, sigma_fc_seed=0.002):
        self._check(lib().rpf_filter_rows(self._h, box, sigma_fc_seed, y0, y1))

    def filter_rows2(self, box, y0, y1, y2, y3, sigma_fc_seed=0.002):
        """Rows [y0,y1) and [y2,y3) in one set of launches (the two edge slabs of a band)."""
        self._check(lib().rpf_filter_rows2(self._h, box, sigma_fc_seed, y0, y1, y2, y3))

    def set_subsampling(self, keep_probability, rng_seed=0, global_row0=0):
        """Random neighbour subsampling (include/rpf_subsample.h); keep_probability >= 1 switches it off."""
        self._check(lib().rpf_set_subsampling(self._h, C.c_double(keep_probability), C.c_ulonglong(rng_seed), global_row0))

    def attach_band(self, global_height, buf0, own0, own1, rank, world, nccl_unique_id=None):
        """One band per process (rpf_band_attach): apply() then runs the band schedule with NCCL halo exchange."""
        b = Band(global_height, buf0, own0, own1, rank, world)
        uid = None if nccl_unique_id is None else (C.c_char * 128).from_buffer_copy(bytes(nccl_unique_id))
        self._check(lib().rpf_band_attach(self._h, C.byref(b), uid))

    def swap_colors(self):
        self._check(lib().rpf_swap_colors(self._h))

    def color_ptr(self, which=0):
        p = C.c_void_p()
        self._check(lib().rpf_color_device_ptr(self._h, which, C.byref(p)))
        return p.value

    def stream(self):
        p = C.c_void_p()
        self._check(lib().rpf_stream(self._h, C.byref(p)))
        return p.value or 0

    def synchronize(self):
        self._check(lib().rpf_synchronize(self._h))

    def stats(self):
        s = Stats()
        self._check(lib().rpf_get_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in Stats._fields_}

    def set_profiling(self, on=True):
        self._check(lib().rpf_set_profiling(self._h, 1 if on else 0))

    def kernel_times(self):
        """{kernel name: (total ms, launches)} since the last reset_stats; needs set_profiling(True)."""
        kt = KernelTimes()
        self._check(lib().rpf_get_kernel_times(self._h, C.byref(kt)))
        return {KERNEL_NAMES[k]: (kt.ms[k], kt.launches[k]) for k in range(len(KERNEL_NAMES))}

    def reset_stats(self):
        self._check(lib().rpf_reset_stats(self._h))
