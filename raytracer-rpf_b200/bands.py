"""Row-band sharding of the RPF filter across ranks.

Pixels are independent within one ApplyRPFFilter pass (Jacobi: reference src/custom/rpf.cpp:527,732), so the image is
split into contiguous row bands, one per rank. Features, positions and random parameters never change
(rpf.cpp:715 writes colours only): each rank holds its band plus `halo` rows of all 19 channels once. Between passes
only the 3 colour planes of the rows within b_delta(next box) of a band edge move to the neighbouring rank
(point-to-point send/recv; there is no collective). This module is backend agnostic: torch.distributed with NCCL on
GPUs, gloo in the CPU tests.

Since round 2 the schedule itself lives behind the C ABI (csrc/rpf_bands.inl): `attach` hands a context its band and an
NCCL communicator (rpf_band_attach) and rpf_filter then filters the edge rows first, ships them with ncclSend / ncclRecv
on a copy stream and filters the interior meanwhile, with no host synchronisation. What is left here is launcher
plumbing for tests and bench.py: the band layout (rpf_band_plan, equal work) and, for the gloo CPU twin of the exchange,
`halo_plan` / `exchange_halos`.
"""
from dataclasses import dataclass


def b_delta(box):
    return (box - 1) // 2  # rpf.cpp:561


@dataclass
class Band:
    rank: int
    world: int
    height: int      # global image height
    own0: int        # first owned global row
    own1: int        # one past the last owned global row
    buf0: int        # first global row held in the local buffer (own0 - halo, clipped)
    buf1: int        # one past the last global row held

    @property
    def local_height(self):
        return self.buf1 - self.buf0

    @property
    def local_own(self):
        """Owned rows in local buffer coordinates."""
        return self.own0 - self.buf0, self.own1 - self.buf0


def make_band(rank, world, height, box_sizes, row_weight=None):
    """Equal-WORK bands (rpf_band_plan of the C ABI: rows near the top and bottom edge see truncated boxes and weigh
    less); halo = largest b_delta over all passes (feature rows are needed at that depth)."""
    import rpf_b200
    halo = max(b_delta(b) for b in box_sizes)
    # the exchange after a pass talks to the two adjacent ranks only: every band must own at least the rows the next
    # pass reaches across its edge, or a pixel would need colours from a rank two bands away
    reach = max([b_delta(b) for b in box_sizes[1:]], default=0)
    if world > 1 and height // world < reach:
        raise ValueError(f"{height} rows over {world} ranks leaves bands of {height // world} rows, thinner than the "
                         f"{reach}-row halo of the later passes: use fewer ranks")
    plan = rpf_b200.band_plan(height, world, box_sizes, row_weight)
    own0, own1 = plan[rank], plan[rank + 1]
    return Band(rank, world, height, own0, own1, max(0, own0 - halo), min(height, own1 + halo))


def halo_plan(band, next_box):
    """Transfers after a pass so that the NEXT pass (box next_box) sees up-to-date colours.

    Returns a list of (peer, send_rows, recv_rows) with row ranges in LOCAL buffer coordinates:
    send_rows are owned rows this rank ships to `peer`, recv_rows are halo rows it receives from `peer`."""
    d = b_delta(next_box)
    plan = []
    if d == 0:
        return plan
    lo0, lo1 = band.local_own
    if band.rank > 0 and band.own1 > band.own0:
        n_recv = min(d, band.own0 - band.buf0)
        n_send = min(d, band.own1 - band.own0)
        plan.append((band.rank - 1, (lo0, lo0 + n_send), (lo0 - n_recv, lo0)))
    if band.rank < band.world - 1 and band.own1 > band.own0:
        n_recv = min(d, band.buf1 - band.own1)
        n_send = min(d, band.own1 - band.own0)
        plan.append((band.rank + 1, (lo1 - n_send, lo1), (lo1, lo1 + n_recv)))
    return plan


def exchange_halos(dist, band, next_box, color_next, row_elems):
    """color_next: torch tensor [3, local_height*row_elems] (the NEXT colour buffer, owned rows just written).
    Sends/receives the halo rows for the next pass with batched point-to-point ops. row_elems = width*spp."""
    ops, keep = [], []
    for peer, (s0, s1), (r0, r1) in halo_plan(band, next_box):
        for k in range(3):
            if s1 > s0:
                t = color_next[k, s0 * row_elems:s1 * row_elems]
                ops.append(dist.P2POp(dist.isend, t, peer))
            if r1 > r0:
                t = color_next[k, r0 * row_elems:r1 * row_elems]
                ops.append(dist.P2POp(dist.irecv, t, peer))
        keep.append(peer)
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return len(ops)


def halo_bytes(band, next_box, row_elems, itemsize=8):
    n = 0
    for _, (s0, s1), _ in halo_plan(band, next_box):
        n += 3 * (s1 - s0) * row_elems * itemsize
    return n


class DeviceColors:
    """Zero-copy view of a context's colour planes ([3][S] doubles in HBM) for torch.distributed."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (3, n), "typestr": "<f8", "data": (ptr, False), "version": 2}


def attach(ctx, band, dist=None):
    """Give a context (holding rows [band.buf0, band.buf1)) its band; with world > 1 rank 0 creates the NCCL unique id and
    torch.distributed carries its 128 bytes to the other ranks (launcher plumbing; the data path never touches torch)."""
    import rpf_b200
    uid = None
    if band.world > 1:
        import torch
        box = [rpf_b200.nccl_unique_id() if band.rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]
    ctx.attach_band(band.height, band.buf0, band.own0, band.own1, band.rank, band.world, uid)


def run_passes_host_exchange(ctx, band, boxes, seed, dist=None, device=0, hdr_clamp_sigmas=0.0):
    """The loop of rpf.cpp:767-775 over one row band on one GPU: each pass filters the owned rows, then the freshly
    written colour rows within b_delta(next box) of a band edge go to the neighbouring ranks (NCCL send/recv over
    NVLink) before the buffers swap. ctx: rpf_b200.RPFFilter holding the band plus its halo rows.
    Paper mode: seed may be a sequence (one sigma per pass, rpf_filter_sigmas); hdr_clamp_sigmas > 0 runs the spike
    removal on the owned rows after the last pass (local to a pixel: no exchange)."""
    import torch
    o0, o1 = band.local_own
    row = ctx.W * ctx.P
    seeds = list(seed) if hasattr(seed, "__len__") else [seed] * len(boxes)
    if len(seeds) != len(boxes):
        raise ValueError("one sigma per box size")
    for i, b in enumerate(boxes):
        ctx.filter_rows(b, o0, o1, seeds[i])
        if dist is not None and band.world > 1 and i + 1 < len(boxes):
            ctx.synchronize()
            nxt = torch.as_tensor(DeviceColors(ctx.color_ptr(1), ctx.S), device=f"cuda:{device}")
            exchange_halos(dist, band, boxes[i + 1], nxt, row)
            torch.cuda.synchronize(device)
        ctx.swap_colors()
    if hdr_clamp_sigmas > 0 and len(boxes):
        ctx.hdr_clamp(hdr_clamp_sigmas, o0, o1)


def run_passes(ctx, band, boxes, seed, dist=None, device=0, hdr_clamp_sigmas=0.0):
    """All passes over an ATTACHED band context (see `attach`): one rpf_filter / rpf_filter_sigmas call, the exchange
    happens inside the library. run_passes_host_exchange is the round-1 driver (host-synchronised torch.distributed
    exchange), kept for A/B measurements."""
    if hasattr(seed, "__len__"):
        ctx.apply_sigmas(boxes, list(seed))
    else:
        ctx.apply(boxes, seed)
    if hdr_clamp_sigmas > 0 and len(boxes):
        o0, o1 = band.local_own
        ctx.hdr_clamp(hdr_clamp_sigmas, o0, o1)
