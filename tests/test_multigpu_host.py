"""Host logic of the multi-GPU path on CPU: band planning (C: vr_split_rows), view sharding, and a
world_size-2 gloo job that plans, renders (oracle as the stand-in device) and gathers one frame."""
import os
import subprocess
import sys

import numpy as np
import pytest

from volumetricrenderer_b200 import multigpu

HERE = os.path.dirname(os.path.abspath(__file__))


def test_split_rows_balances_cost():
    H = 720
    y = np.arange(H)
    cost = np.where((y >= 116) & (y <= 663), 1.0 + 40.0 * np.sin((y - 116) / 547.0 * np.pi), 0.05).astype(np.float32)
    for parts in (1, 2, 3, 4, 8):
        b = multigpu.split_rows(cost, parts)
        assert b[0] == 0 and b[-1] == H and len(b) == parts + 1
        assert all(b1 >= b0 for b0, b1 in zip(b, b[1:]))
        sums = [cost[b0:b1].sum() for b0, b1 in zip(b, b[1:])]
        assert max(sums) <= cost.sum() / parts + cost.max() + 1e-3      # off by at most one row's cost
    # equal-height bands would be badly unbalanced for this profile — the reason the cut is by cost
    eq = [cost[i * H // 8:(i + 1) * H // 8].sum() for i in range(8)]
    cut = multigpu.split_rows(cost, 8)
    bal = [cost[b0:b1].sum() for b0, b1 in zip(cut, cut[1:])]
    assert max(eq) / (cost.sum() / 8) > 1.5 and max(bal) / (cost.sum() / 8) < 1.1


def test_split_rows_edge_cases():
    assert multigpu.split_rows(np.zeros(10, np.float32), 4) == [0, 2, 5, 7, 10]          # no cost: equal heights
    assert multigpu.split_rows(np.ones(3, np.float32), 8)[-1] == 3                       # more parts than rows
    b = multigpu.split_rows(np.array([0, 0, 9, 0, 0], np.float32), 2)
    assert b == [0, 3, 5]
    one = multigpu.split_rows(np.array([5.0], np.float32), 1)
    assert one == [0, 1]


def test_feedback_rebalance_converges():
    """the analytic cost misjudges the bands (it cannot see the density field); re-cutting from measured band
    times converges on the true balance within a few frames"""
    H = 400
    y = np.arange(H)
    model = np.where((y > 60) & (y < 340), 10.0, 0.1).astype(np.float32)            # what the set-up predicts
    true = model * (1.0 + 3.0 * (y > 200))                                          # lower half is 4x denser
    for parts in (2, 4, 8):
        b = multigpu.split_rows(model, parts)
        for _ in range(5):
            times = [float(true[b0:b1].sum()) for b0, b1 in zip(b, b[1:])]
            b = multigpu.rescale_and_split(model, b, times)
        times = np.array([true[b0:b1].sum() for b0, b1 in zip(b, b[1:])])
        assert times.max() / times.mean() < 1.12, (parts, b, times)


def test_shard_views():
    assert multigpu.shard_views(64, 8) == [(8 * i, 8 * i + 8) for i in range(8)]
    assert multigpu.shard_views(5, 2) == [(0, 3), (3, 5)]
    assert multigpu.shard_views(3, 4) == [(0, 1), (1, 2), (2, 3), (3, 3)]
    for n, w in [(0, 3), (7, 7), (100, 8)]:
        s = multigpu.shard_views(n, w)
        assert s[0][0] == 0 and s[-1][1] == n and all(a[1] == b[0] for a, b in zip(s, s[1:]))


def test_world_size_2_gloo_band_gather(tmp_path):
    out = str(tmp_path / "gloo_result.npz")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(HERE, "_gloo_worker.py"), out]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    z = np.load(out)
    b = z["bounds"].tolist()
    assert b[0] == 0 and b[-1] == 90 and 0 < b[1] < 90
    assert bool(z["frame_ok"])                       # gathered bands == the full frame, bit for bit
    s = z["band_steps"].astype(np.float64)
    assert abs(s[0] - s[1]) / s.sum() < 0.08         # the two bands carry nearly equal work
    assert bool(z["views_ok"]) and int(z["n_views"]) == 5


def test_gather_tickets_model_the_device_protocol():
    """host arithmetic of vr_gather_*: replay the counters for 3 ranks and 7 frames through 2 slots — every frame finds its
    slot released exactly as often as it was used before, and GPU 0's arrival target is reached by exactly this frame"""
    world, slots = 3, 2
    arrived, released = [0] * slots, [0] * slots
    for i in range(7):
        slot, releases, arrivals = multigpu.gather_ticket(i, world, slots)
        assert slot == i % slots and released[slot] == releases          # acquire passes without waiting for a later frame
        arrived[slot] += world                                            # every rank signals once
        assert arrived[slot] == arrivals                                  # collect fires for this frame, not earlier
        released[slot] += 1
    with pytest.raises(ValueError):
        multigpu.gather_ticket(-1, 2)
