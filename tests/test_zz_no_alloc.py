"""rpf_filter_pass_noalloc (include/rpf_b200.h): the entry point that allocates nothing -- scratch in file-scope static
__device__ arrays of the library. No counterpart in the reference, so the
check is against the existing entry points (rpf_create / rpf_upload / rpf_filter) on the same inputs: bit-identical colours,
on calls in a row that share the static scratch. Collected last on purpose.

The test allocates the caller's buffers with torch; the entry point itself must not allocate: free device memory is the
same before and after the later calls."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _existing(planes, W, H, P, box, seed):
    import rpf_b200
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    f.apply([box], seed)
    c = f.download_colors_f64()
    f.close()
    return c


def _noalloc(planes, W, H, P, box, seed, stream=None):
    import rpf_b200
    import torch
    d_planes = torch.from_numpy(planes).cuda()
    d_in = d_planes[2:5].to(torch.float64).contiguous()
    d_out = torch.full_like(d_in, float("nan"))
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    s = stream or torch.cuda.current_stream()
    rpf_b200.filter_pass_noalloc(d_planes.data_ptr(), d_in.data_ptr(), d_out.data_ptr(), W, H, P, box, seed, stream=s.cuda_stream)
    s.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return d_out.cpu().numpy(), free0, free1


def test_noalloc_pass_equals_rpf_filter_on_calls_in_a_row():
    import rpf_b200
    import torch
    # different sizes, boxes and content in a row: every call finds the previous call's statistics and counters in the scratch
    cases = [(40, 36, 8, 7, 0.5, 11, 0), (24, 20, 8, 5, 0.002, 12, 0), (40, 36, 8, 7, 0.5, 11, 0), (33, 17, 4, 9, 2.0, 13, 1)]
    frees = []
    for W, H, P, box, seed, rng, kind in cases:
        planes = rpf_b200.synth_planes(W, H, P, seed=rng, kind=kind)
        got, free0, free1 = _noalloc(planes, W, H, P, box, seed)
        frees.append((free0, free1))
        ref = _existing(planes, W, H, P, box, seed)
        # same kernels, same inputs: the tolerance the GPU-vs-GPU tests use is exact equality (NaN patterns included)
        assert np.array_equal(got, ref, equal_nan=True), (W, H, P, box, np.nanmax(np.abs(got - ref)))
        assert np.isfinite(got).all() or kind == 1
    # the first call on a device loads the module and does the one-time set-up; the later ones must not take device memory
    for free0, free1 in frees[1:]:
        assert free1 >= free0, (free0, free1)
    # a non-default stream
    W, H, P, box, seed, rng, kind = cases[0]
    planes = rpf_b200.synth_planes(W, H, P, seed=rng, kind=kind)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    got, _, _ = _noalloc(planes, W, H, P, box, seed, stream=side)
    assert np.array_equal(got, _existing(planes, W, H, P, box, seed), equal_nan=True)


def test_noalloc_pass_rejects_what_the_static_scratch_cannot_hold():
    import rpf_b200
    import torch
    buf = torch.zeros(1024, device="cuda")
    a = buf.data_ptr()
    with pytest.raises(rpf_b200.RPFError, match="RPF_NOALLOC_MAX_PIXELS"):   # 257 x 256 pixels > 65,536
        rpf_b200.filter_pass_noalloc(a, a, a + 8, 257, 256, 8, 7, 0.5)
    with pytest.raises(rpf_b200.RPFError, match="candidates exceed"):        # box 17 x 8 spp = 2,312 candidates: general path only
        rpf_b200.filter_pass_noalloc(a, a, a + 8, 16, 16, 8, 17, 0.5)
    with pytest.raises(rpf_b200.RPFError, match="must differ"):
        rpf_b200.filter_pass_noalloc(a, a, a, 16, 16, 8, 7, 0.5)
    with pytest.raises(rpf_b200.RPFError, match="not memory of the current device"):
        host = np.zeros(16, np.float32)
        rpf_b200.filter_pass_noalloc(host.ctypes.data, a, a + 8, 1, 1, 1, 1, 0.5)
    assert rpf_b200.NOALLOC_MAX_PIXELS == 65536
