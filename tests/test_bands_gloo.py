"""CPU, world_size 2 over gloo: the row-band driver (bands.py: band layout + halo exchange between passes) with the
CPU oracle standing in for the device filter. The sharded result must equal the single-process full frame."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_bind as ob

W, H, P = 16, 22, 4
BOXES = [9, 5, 7]  # not monotone on purpose: the halo depth follows the NEXT box
SEED = 0.7


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, tmp):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "raytracer-rpf_b200"))
    import bands
    import rpf_b200
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    band = bands.make_band(rank, world, H, BOXES)
    lh = band.local_height
    planes = rpf_b200.synth_planes(W, lh, P, seed=21, kind=0, y_off=band.buf0)
    cur = planes[2:5].astype(np.float64).copy()
    o0, o1 = band.local_own
    for i, box in enumerate(BOXES):
        # one pass over the owned rows only; halo rows keep whatever they hold
        nxt = ob.oracle_filter(planes, W, lh, P, [box], seed=SEED, colors=cur, rows=(o0, o1), threads=2)["colors"]
        if i + 1 < len(BOXES):
            t = torch.from_numpy(nxt)
            bands.exchange_halos(dist, band, BOXES[i + 1], t, W * P)
        cur = nxt
    own = cur.reshape(3, lh, W * P)[:, o0:o1]
    np.save(os.path.join(tmp, f"band{rank}.npy"), own)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_bands_equal_full_frame(tmp_path):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "raytracer-rpf_b200"))
    import rpf_b200
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    planes = rpf_b200.synth_planes(W, H, P, seed=21, kind=0)
    full = ob.oracle_filter(planes, W, H, P, BOXES, seed=SEED)["colors"].reshape(3, H, W * P)
    got = np.concatenate([np.load(tmp_path / "band0.npy"), np.load(tmp_path / "band1.npy")], axis=1)
    assert got.shape == full.shape
    assert np.array_equal(got, full)
    assert np.abs(full - planes[2:5].reshape(3, H, W * P)).max() > 1e-3
