"""three ways of getting the row bands of one frame onto GPU 0, same march, same bands.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29517 tools/gather_compare.py [--workload c3]

  peer   the march kernel of every rank stores its pixels straight into GPU 0's frame (CUDA IPC mapping, NVLink)
  copy   every rank marches into its own frame, then one 2-D device-to-device copy of the band into the mapping
  nccl   every rank marches into its own frame, then a batched NCCL send/recv of the bands (torch.distributed)
Wall clock per frame between barriers (all ranks done, every band on GPU 0), median of the timed frames.
"""
import argparse
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import volumetricrenderer_b200 as vr  # noqa: E402
from volumetricrenderer_b200 import capi, multigpu, scene  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--frames", type=int, default=30)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, H, desc, kw = bench.WORKLOADS[args.workload]
    r = vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, device=local)
    bench.setup_volume(r, args.workload, "texture")
    ubo0, ubo1 = scene.default_scene()
    r.set_params(ubo0, ubo1, **kw)
    bg = multigpu.BandGather(dist, rank, world, H, mode="peer")
    row_cost = r.row_cost()
    bg.plan(row_cost=row_cost)
    mapped = bg.connect_peer(r)

    def barrier():
        r.wait()
        torch.cuda.synchronize()
        dist.barrier(device_ids=[local])

    for _ in range(4):                                     # same feedback re-cut as bench.py
        r.render(rows=bg.rows, to_host=False); r.wait()
        bg.rebalance(row_cost, r.stats()["gpu_ms"])
    rows, bounds = bg.rows, bg.bounds
    frame_ptr, pitch, _ = r.frame_info()
    # torch view of this rank's libvolrt frame for the NCCL variant (no copy: same device memory)
    nbytes = pitch * H

    class _Ext:                                            # __cuda_array_interface__ wrapper over the frame
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (frame_ptr, False), "version": 2}
    frame_t = torch.as_tensor(_Ext(), device="cuda").view(H, pitch)

    def peer():
        if rank == 0:
            r.render(rows=rows, to_host=False)
        else:
            r.render_device(mapped, rows=rows, flags=capi.RENDER_DIRECT)

    def copy():
        if rank == 0:
            r.render(rows=rows, to_host=False)
        else:
            r.render_device(mapped, rows=rows, flags=0)    # march into the own frame + cudaMemcpy2DAsync D2D into GPU 0's

    def nccl():
        r.render(rows=rows, to_host=False)
        r.wait()                                           # the march runs on libvolrt's stream, NCCL on torch's
        ops = []
        if rank == 0:
            for src in range(1, world):
                if bounds[src + 1] > bounds[src]:
                    ops.append(dist.P2POp(dist.irecv, frame_t[bounds[src]:bounds[src + 1]], src))
        elif rows[1] > rows[0]:
            ops.append(dist.P2POp(dist.isend, frame_t[rows[0]:rows[1]], 0))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    out = {}
    ref = None
    for name, fn in (("peer", peer), ("copy", copy), ("nccl", nccl)):
        for _ in range(5):
            fn(); barrier()
        ts = []
        for _ in range(args.frames):
            r.flush_l2(); barrier()
            t0 = time.perf_counter()
            fn()
            barrier()
            ts.append(time.perf_counter() - t0)
        out[name] = float(np.median(ts)) * 1e3
        if rank == 0:                                      # every variant must leave the same frame on GPU 0
            img = r.render(flags=capi.RENDER_NO_MARCH)[0].copy()
            if ref is None:
                ref = img
            assert np.array_equal(ref, img), name
        # clear GPU 0's frame between variants so that a variant that moved nothing cannot pass
        if rank == 0:
            frame_t.zero_()
        barrier()
    t = torch.tensor([out["peer"], out["copy"], out["nccl"]], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print("gather_compare %s %d GPUs bands %s: peer stores %.3f ms | march + D2D band copy %.3f ms | march + NCCL send/recv %.3f ms "
              "(wall per frame incl. one barrier, median of %d)" % (args.workload, world, bounds, t[0], t[1], t[2], args.frames), flush=True)
    bg.disconnect_peer(r)
    barrier()
    r.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
