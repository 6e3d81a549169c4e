"""World-size-2 test of the multi-GPU decomposition on CPU (gloo).

The CUDA path shards TARGET tiles over ranks (NbSimPlan: an equal share of the massive tiles and of
the massless tiles per rank), every rank computes its targets against ALL sources from the OLD
positions, and only the new positions of the massive particles are re-replicated before the next
step.  Here each rank plays one GPU with the oracle as its "kernel" and torch.distributed/gloo as
its "NVLink": if the plan covered a target twice or not at all, or the exchange missed a range, the
gathered result would differ from the single-rank oracle -- it must be bit-identical.
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _particle_ranges(ranges, rank, tile, total_len):
    a0, a1, b0, b1 = ranges[rank]
    out = []
    for t0, t1 in ((a0, a1), (b0, b1)):
        lo, hi = min(t0 * tile, total_len), min(t1 * tile, total_len)
        if hi > lo:
            out.append((lo, hi))
    return out


def _rank_main(rank, world, port, n, galaxies, steps, dt, result_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import nbody_b200 as nb
    from _oracle import Oracle

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    oracle = Oracle()
    w0, m = oracle.partition(nb.make_galaxies(n, galaxies))      # identical on every rank
    info, ranges = nb.plan(len(w0), m, world)
    tile = info["tile_targets"]
    mine = _particle_ranges(ranges, rank, tile, len(w0))

    # SetSimulationData, sharded: a rank reads only its own particles out of the array it is handed (the foreign
    # ones are poisoned here), then the source data (x, y, mass of the massive particles) is replicated rank to rank
    arr = w0.copy()   # this rank's replica: own targets complete, foreign massive positions current
    own = np.zeros(len(w0), dtype=bool)
    for lo, hi in mine:
        own[lo:hi] = True
    for k in ("px", "py", "vx", "vy", "ax", "ay", "mass", "radius"):
        arr[k][~own] = np.nan
    for q in range(world):
        q_lo, q_hi = min(ranges[q][0] * tile, m), min(ranges[q][1] * tile, m)
        buf = torch.from_numpy(np.stack([arr[k][q_lo:q_hi] for k in ("px", "py", "mass")], axis=1).copy())
        dist.broadcast(buf, src=q)
        for c, k in enumerate(("px", "py", "mass")):
            arr[k][q_lo:q_hi] = buf[:, c].numpy()
    assert not np.isnan(arr["px"][:m]).any() and not np.isnan(arr["mass"][:m]).any()
    for _ in range(steps):
        new = arr.copy()
        for lo, hi in mine:                                      # "kernel": own targets vs all OLD sources
            new[lo:hi] = oracle.step_range(arr, m, lo, hi, dt)[lo:hi]
        # "all-gather": new positions of the massive particles of every rank's massive range
        a_lo, a_hi = min(ranges[rank][0] * tile, m), min(ranges[rank][1] * tile, m)
        send = torch.from_numpy(np.stack([new["px"][a_lo:a_hi], new["py"][a_lo:a_hi]], axis=1).copy())
        for q in range(world):
            q_lo, q_hi = min(ranges[q][0] * tile, m), min(ranges[q][1] * tile, m)
            buf = send if q == rank else torch.empty((q_hi - q_lo, 2), dtype=torch.float32)
            dist.broadcast(buf, src=q)
            if q != rank:
                new["px"][q_lo:q_hi] = buf[:, 0].numpy()
                new["py"][q_lo:q_hi] = buf[:, 1].numpy()
        arr = new

    # GetSimulationData: gather the ranges each rank owns
    pieces = [None] * world
    dist.all_gather_object(pieces, [(lo, hi, arr[lo:hi].tobytes()) for lo, hi in mine])
    if rank == 0:
        full = np.zeros_like(w0)
        seen = np.zeros(len(w0), dtype=np.int32)
        for plist in pieces:
            for lo, hi, raw in plist:
                full[lo:hi] = np.frombuffer(raw, dtype=w0.dtype)
                seen[lo:hi] += 1
        ref = oracle.update(w0, m, dt, steps, 8)
        np.savez(result_path, full=full, ref=ref, seen=seen)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n,galaxies,world", [(3000, 2, 2), (5000, 3, 2), (2600, 2, 3)])
def test_sharded_step_with_position_exchange_equals_single_rank(tmp_path, n, galaxies, world):
    import torch.multiprocessing as mp
    path = str(tmp_path / "out.npz")
    mp.spawn(_rank_main, args=(world, _free_port(), n, galaxies, 3, 0.01, path), nprocs=world, join=True)
    out = np.load(path)
    assert np.all(out["seen"] == 1), "every target must be computed by exactly one rank"
    assert out["full"].tobytes() == out["ref"].tobytes()


def test_plan_covers_all_tiles_once_and_is_independent_of_shard_count():
    import nbody_b200 as nb
    for n, m in [(1, 1), (300, 0), (3000, 1504), (16384, 8086), (262144, 130916), (1048576, 523884),
                 (4194304, 2097152), (1000003, 17)]:
        base = None
        for shards in (1, 2, 3, 4, 8):
            info, ranges = nb.plan(n, m, shards)
            # summation order (tile size, slab length) must not depend on the number of GPUs
            key = (info["variant"], info["tile_targets"], info["slab_len"], info["slabs"])
            base = base or key
            assert key == base
            tiles = np.zeros(info["tiles"], dtype=np.int32)
            for a0, a1, b0, b1 in ranges:
                assert a0 <= a1 <= info["massive_tiles"] <= b0 <= b1 <= info["tiles"]
                tiles[a0:a1] += 1
                tiles[b0:b1] += 1
            assert np.all(tiles == 1)
            # balance: shares differ by at most one tile per class
            for lo, hi in ((0, 1), (2, 3)):
                sizes = [r[hi] - r[lo] for r in ranges]
                assert max(sizes) - min(sizes) <= 1
            assert info["tiles"] * info["tile_targets"] >= n
            assert info["slabs"] * info["slab_len"] >= m


def test_source_packs_of_the_shards_partition_the_source_array():
    """The sharded upload replicates, per shard, the packs [lo/4, ceil(min(hi, M)/4)) of its massive range:
    together they must cover every data pack exactly once, whatever the shard count."""
    import nbody_b200 as nb
    for n, m in [(3000, 1504), (16384, 8086), (100000, 49944), (1048576, 523884), (1000003, 17), (5000, 4999)]:
        for shards in (2, 3, 4, 8):
            info, ranges = nb.plan(n, m, shards)
            tile = info["tile_targets"]
            assert tile % 4 == 0
            packs = np.zeros((m + 3) // 4, dtype=np.int32)
            for a0, a1, _, _ in ranges:
                lo, hi = min(a0 * tile, m), min(a1 * tile, m)
                if hi > lo:
                    packs[lo // 4:(hi + 3) // 4] += 1
            assert np.all(packs == 1), (n, m, shards)


def test_plan_invariants_on_random_sizes():
    """Property test of the host-only planner (hypothesis): geometry covers the World, slab boundaries are whole
    packs, shares are balanced, and nothing depends on the shard count."""
    from hypothesis import given, settings, strategies as st
    import nbody_b200 as nb

    @settings(max_examples=150, deadline=None)
    @given(n=st.integers(1, 6_000_000), frac=st.floats(0.0, 1.0), shards=st.sampled_from([1, 2, 3, 4, 8]))
    def check(n, frac, shards):
        m = min(n, int(n * frac))
        info, ranges = nb.plan(n, m, shards)
        tile = info["tile_targets"]
        assert info["tiles"] * tile >= n > (info["tiles"] - 1) * tile
        assert info["massive_tiles"] * tile >= m
        assert info["slab_len"] % 16 == 0 and info["slabs"] >= 1
        assert info["slabs"] * info["slab_len"] >= m
        if m > 0:
            assert (info["slabs"] - 1) * info["slab_len"] < m          # no empty slab
            assert info["slab_len"] <= 16384                           # accuracy floor: chains of <= 8192 terms
        one, _ = nb.plan(n, m, 1)
        assert (one["variant"], one["tile_targets"], one["slab_len"], one["slabs"]) == \
            (info["variant"], tile, info["slab_len"], info["slabs"])
        covered = 0
        for a0, a1, b0, b1 in ranges:
            assert a0 <= a1 <= info["massive_tiles"] <= b0 <= b1 <= info["tiles"]
            covered += (a1 - a0) + (b1 - b0)
        assert covered == info["tiles"]

    check()
