"""Host-side logic of the multi-GPU path: one process per GPU, every ray independent.

The reference renders on a single GPU (VulkanDevice.cpp:42-49) and has no parallel code; the only
parallelism libvolrt adds is screen-space / view data parallelism:
  * one view  -> contiguous ROW BANDS of nearly equal estimated cost (coverage is ~21 % of the
                 pixels and confined to the middle rows, so equal-height bands would idle the
                 outer ranks); the cut is `vr_split_rows` over a per-row cost vector;
  * many views -> contiguous VIEW SHARDS (views are independent frames).
The volume is replicated (8 MiB .. 2 GiB against 180 GB of HBM).  The single exchange per frame is the
gather of the disjoint bands on rank 0:
  * "peer"  — every rank's march kernel stores its pixels straight into rank 0's frame through an
              NVLink peer mapping (CUDA IPC; vr_frame_export / vr_frame_import + VR_RENDER_DIRECT):
              the gather is fused into the march, no copy kernel and no collective launch;
  * "collective" — torch.distributed gather of the band tensors (NCCL on GPUs, gloo in the CPU tests).
`torch.distributed` is plumbing only (rendezvous, barrier, handle broadcast).
"""
import ctypes as C

import numpy as np

from . import capi


def split_rows(cost, parts):
    """bounds[parts+1] of contiguous row bands of nearly equal summed cost (C: vr_split_rows)."""
    cost = np.ascontiguousarray(cost, np.float32)
    b = (C.c_uint32 * (parts + 1))()
    st = capi.load().vr_split_rows(cost.ctypes.data_as(C.POINTER(C.c_float)), cost.size, parts, b)
    if st != capi.VR_OK:
        raise ValueError("vr_split_rows: status %d" % st)
    return [int(x) for x in b]


def shard_views(view_count, world):
    """-> [(begin, end)] per rank: contiguous, sizes differ by at most one, earlier ranks get the extra."""
    base, extra = divmod(view_count, world)
    out, b = [], 0
    for r in range(world):
        e = b + base + (1 if r < extra else 0)
        out.append((b, e))
        b = e
    return out


def rescale_and_split(row_cost, bounds, times):
    """row_cost: model cost per row; bounds: current cut; times: measured time per band -> new bounds"""
    cost = np.array(row_cost, np.float64, copy=True)
    for r, t in enumerate(times):
        b0, b1 = bounds[r], bounds[r + 1]
        s = cost[b0:b1].sum()
        if b1 > b0 and s > 0 and t > 0:
            cost[b0:b1] *= t / s
    return split_rows(cost.astype(np.float32), len(times))


def gather_ticket(frame_index, world, slots=2):
    """Counters of the device-side gather (include/volrt.h: vr_gather_*) for the frame_index-th frame sent through it:
    -> (slot, releases, arrivals): the frame goes into `slot`; a rank may write the slot once GPU 0 has released it
    `releases` times (= earlier frames that used it); GPU 0 copies the slot out once `arrivals` bands have landed in it
    over its whole life (every rank signals once per frame, empty bands included)."""
    if frame_index < 0 or world < 1 or slots < 1:
        raise ValueError("gather_ticket(%r, %r, %r)" % (frame_index, world, slots))
    slot, uses = frame_index % slots, frame_index // slots
    return slot, uses, world * (uses + 1)


class BandGather:
    """Row-band partition of ONE frame over `world` ranks and its gather on rank 0.

    render_band(rows, dst) is the device hook: it must render rows [r0, r1) and, when `dst` is not
    None ("peer" mode), store them straight into that mapped frame.  In "collective" mode it returns
    the band as an array shaped (r1-r0, W, C).
    """

    def __init__(self, dist, rank, world, height, mode="collective"):
        if mode not in ("peer", "collective"):
            raise ValueError(mode)
        self.dist, self.rank, self.world, self.height, self.mode = dist, rank, world, height, mode
        self.bounds = [0, height] if world == 1 else None
        self.last_times = self.last_bounds = None
        self.mapped = None

    # ---- partition (rank 0 decides, everyone learns)
    def plan(self, row_cost=None, bounds=None):
        """rank 0 passes a per-row cost vector (or ready bounds); all ranks return the same bounds."""
        box = [None]
        if self.rank == 0:
            box[0] = list(bounds) if bounds is not None else split_rows(row_cost, self.world)
        if self.world > 1:
            self.dist.broadcast_object_list(box, src=0)
        self.bounds = box[0]
        if len(self.bounds) != self.world + 1 or self.bounds[0] != 0 or self.bounds[-1] != self.height:
            raise ValueError("bad band bounds %r" % (self.bounds,))
        return self.bounds

    def rebalance(self, row_cost, my_ms):
        """Feedback re-cut between frames: every rank reports the device time of its band; the model
        cost of each row is rescaled by its band's measured time per unit of model cost, and the rows are
        cut again.  (The analytic model counts primary steps only; how many of them carry a light-march
        depends on the density field, which the previous frame has just measured.)"""
        if self.world == 1:
            return self.bounds
        times = [None] * self.world
        self.dist.all_gather_object(times, float(my_ms))
        self.last_times, self.last_bounds = times, list(self.bounds)      # what was measured, and for which cut
        self.bounds = rescale_and_split(row_cost, self.bounds, times)
        return self.bounds

    @property
    def rows(self):
        return self.bounds[self.rank], self.bounds[self.rank + 1]

    # ---- peer mode: rank 0 exports its frame, the others map it
    def connect_peer(self, renderer):
        box = [renderer.frame_export() if self.rank == 0 else None]
        if self.world > 1:
            self.dist.broadcast_object_list(box, src=0)
        if self.rank != 0:
            self.mapped = renderer.frame_import(box[0])
        return self.mapped

    def disconnect_peer(self, renderer):
        if self.mapped is not None:
            renderer.frame_unmap(self.mapped)
            self.mapped = None

    # ---- collective mode
    def gather(self, band, width, channels, dtype, as_tensor, device=None):
        """band: this rank's rows (r1-r0, W, C).  Returns the full frame on rank 0, None elsewhere.
        Bands are padded to the tallest one because gather wants equal shapes."""
        import torch
        tallest = max(b1 - b0 for b0, b1 in zip(self.bounds, self.bounds[1:]))
        pad = np.zeros((tallest, width, channels), dtype)
        r0, r1 = self.rows
        pad[:r1 - r0] = band
        t = as_tensor(pad)
        if self.world == 1:
            return np.asarray(band)
        outs = [torch.empty_like(t) for _ in range(self.world)] if self.rank == 0 else None
        self.dist.gather(t, outs, dst=0)
        if self.rank != 0:
            return None
        frame = np.zeros((self.height, width, channels), dtype)
        for r, o in enumerate(outs):
            b0, b1 = self.bounds[r], self.bounds[r + 1]
            frame[b0:b1] = o.cpu().numpy()[:b1 - b0]
        return frame
