#!/usr/bin/env python
"""Row bands on real GPUs: run under torchrun with N ranks (one per GPU, NCCL); every rank filters its band with halo
exchange between passes inside the library (rpf_band_attach + rpf_filter: edge rows first, ncclSend / ncclRecv on the copy
stream, interior rows meanwhile), rank 0 gathers the owned rows and checks them against the CPU oracle's
full-frame result. Test infrastructure (uses oracle/): tests/test_multigpu.py launches it.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/band_parity.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracer-rpf_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bands  # noqa: E402
import rpf_b200  # noqa: E402

W, H, P = 48, 40, 8
BOXES = [17, 9, 7]   # general path, then two fused passes; halo depth follows the next box
SEED = 0.3


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    band = bands.make_band(rank, world, H, BOXES)
    planes = rpf_b200.synth_planes(W, band.local_height, P, seed=21, kind=0, y_off=band.buf0)
    ctx = rpf_b200.RPFFilter(W, band.local_height, P, device=lr)
    bands.attach(ctx, band, dist)  # rpf_band_attach: the exchange runs inside the library (ncclSend / ncclRecv)
    ctx.upload(planes)
    bands.run_passes(ctx, band, BOXES, SEED, dist, lr)
    halo_bytes = ctx.stats()["halo_bytes"]
    assert world == 1 or halo_bytes > 0, "no halo traffic recorded"
    o0, o1 = band.local_own
    own = ctx.download_colors_f64().reshape(3, band.local_height, W * P)[:, o0:o1]
    ctx.close()
    t = torch.from_numpy(np.ascontiguousarray(own)).cuda()
    parts = [torch.empty((3, bands.make_band(r, world, H, BOXES).own1 - bands.make_band(r, world, H, BOXES).own0, W * P),
                         dtype=torch.float64, device="cuda") for r in range(world)]
    if rank == 0:
        parts[0] = t
        for r in range(1, world):
            dist.recv(parts[r], src=r)
    else:
        dist.send(t, dst=0)
    ok = True
    if rank == 0:
        import oracle_bind as ob
        got = torch.cat(parts, dim=1).cpu().numpy()
        full_planes = rpf_b200.synth_planes(W, H, P, seed=21, kind=0)
        ref = ob.oracle_filter(full_planes, W, H, P, BOXES, seed=SEED)["colors"].reshape(3, H, W * P)
        err = float(np.abs(got - ref).max())
        # one GPU, same frame: the band result must be BIT-identical to it
        one = rpf_b200.RPFFilter(W, H, P, device=lr)
        one.upload(full_planes)
        one.apply(BOXES, SEED)
        same = bool(np.array_equal(one.download_colors_f64().reshape(3, H, W * P), got))
        one.close()
        print(f"band_parity world={world} bit-identical to one GPU: {same}")
        ok = same and err <= 1e-9 and float(np.abs(ref - full_planes[2:5].reshape(3, H, W * P)).max()) > 1e-3
        print(f"band_parity world={world} max|gpu-oracle|={err:.3e} {'OK' if ok else 'FAIL'}")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
