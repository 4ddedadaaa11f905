"""worker of tests/test_multigpu_host.py: one rank of a world_size-2 gloo job on CPU.

The product has no CPU renderer, so the ORACLE stands in for the device here (tests may use it);
what is under test is the host logic of volumetricrenderer_b200/multigpu.py: the cost-balanced
band plan, its broadcast, the collective gather, and the view sharding."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    from oracle import oracle
    from volumetricrenderer_b200 import multigpu, scene
    from util import oracle_params

    out_path = sys.argv[1]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    W, H = 160, 90
    rng = np.random.default_rng(5)
    vol = rng.integers(0, 256, size=(8, 8, 8, 4), dtype=np.uint8)
    ov, _k = oracle.make_volume(vol, oracle.VOL_RGBA8)
    u0, u1 = scene.default_scene(20.0, 10.0)
    p = oracle_params(oracle, u0, u1)

    # ---- one frame, row bands
    g = multigpu.BandGather(dist, rank, world, H, mode="collective")
    cost = None
    if rank == 0:
        _, _, steps = oracle.render(ov, p, W, H, want_steps=True)
        cost = np.maximum(steps, 0).sum(axis=1).astype(np.float32) + 0.05 * W
    bounds = g.plan(row_cost=cost)
    r0, r1 = g.rows
    band, st = oracle.render(ov, p, W, H, rows=(r0, r1))
    band8 = oracle.encode(band, oracle.OUT_RGBA8_UNORM)
    frame = g.gather(band8, W, 4, np.uint8, torch.from_numpy)
    steps_mine = torch.tensor([st["steps"]], dtype=torch.int64)
    all_steps = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(all_steps, steps_mine)

    # ---- view shards
    shards = multigpu.shard_views(5, world)
    v0, v1 = shards[rank]
    views = scene.orbit_views(5)
    mine = []
    for k in range(v0, v1):
        pk = oracle_params(oracle, views[k, :192].view(np.float32), views[k, 192:332].view(np.float32))
        img, _ = oracle.render(ov, pk, 64, 36)
        mine.append(oracle.encode(img, oracle.OUT_RGBA8_UNORM))
    gathered = [None] * world
    dist.all_gather_object(gathered, (v0, v1, mine))

    if rank == 0:
        full, _ = oracle.render(ov, p, W, H)
        full8 = oracle.encode(full, oracle.OUT_RGBA8_UNORM)
        per_view = []
        for k in range(5):
            pk = oracle_params(oracle, views[k, :192].view(np.float32), views[k, 192:332].view(np.float32))
            img, _ = oracle.render(ov, pk, 64, 36)
            per_view.append(oracle.encode(img, oracle.OUT_RGBA8_UNORM))
        got_views = {}
        for (a, b, imgs) in gathered:
            for k, im in zip(range(a, b), imgs):
                got_views[k] = im
        np.savez(out_path, bounds=np.array(bounds), frame_ok=np.array_equal(frame, full8),
                 band_steps=np.array([int(t.item()) for t in all_steps]),
                 views_ok=all(np.array_equal(got_views[k], per_view[k]) for k in range(5)),
                 n_views=len(got_views), shards=np.array(shards))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
