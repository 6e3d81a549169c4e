"""CPU model of the CUDA kernel's summation order (no GPU needed).

The plan (NbSimPlan: tile size, slab length) fixes the order in which the step kernel adds the pair terms of a target:
inside a slab, even-numbered sources go to one FP32 accumulator and odd-numbered ones to another, in source order
(FMA); the slab partial is their sum; slab partials are added in slab order (step_kernel.cuh, force_kernel.cuh).
With slab groups (large Worlds, NbSimPlanInfo.group_slabs > 1) the slabs of a group are added in slab order first,
then the group sums in group order.
Here that order is replayed in NumPy FP32 (exact 1/sqrt instead of MUFU.RSQ, which changes single terms by ~1 ulp but
not the growth of the rounding error with the chain length) on the reference's golden fixtures, so that a planner
change that lengthens the chains shows up as lost accuracy on the build machine, before any GPU run.
"""
import numpy as np
import pytest

from _oracle import acc_of, mass_len_of, rel_acc_error

F32 = np.float32



def _order(info):
    """What fixes the summation order (everything but HOW chained steps of a tiny World on one GPU are launched)."""
    return {k: v for k, v in info.items() if not k.startswith("resident_")}

def _fma(a, b, c):
    # exact for FP32 inputs: the product is exact in FP64, one rounding to FP64 and one to FP32
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(F32)


def replay_step_kernel_acc(w0, m, slab_len, nslab, group_slabs=1):
    n = len(w0)
    x, y, rho = w0["px"], w0["py"], w0["radius"]
    gm = (w0["mass"][:m] * F32(10.0)).astype(F32)          # NB_G folded once per source (layout_kernels.cuh)
    total, group = None, None
    for s in range(nslab):
        lo, hi = s * slab_len, min((s + 1) * slab_len, m)
        acc = [[np.zeros(n, F32), np.zeros(n, F32)], [np.zeros(n, F32), np.zeros(n, F32)]]
        for j in range(lo, hi):
            par = j & 1
            dx = (x[j] - x).astype(F32)
            dy = (y[j] - y).astype(F32)
            r2 = _fma(dx, dx, _fma(dy, dy, rho))
            inv = (F32(1.0) / np.sqrt(r2)).astype(F32)
            f = ((inv * inv).astype(F32) * (gm[j] * inv).astype(F32)).astype(F32)
            acc[par][0] = _fma(f, dx, acc[par][0])
            acc[par][1] = _fma(f, dy, acc[par][1])
        part = np.stack([(acc[0][0] + acc[1][0]).astype(F32), (acc[0][1] + acc[1][1]).astype(F32)], axis=1)
        group = part if group is None else (group + part).astype(F32)
        if (s + 1) % group_slabs == 0 or s + 1 == nslab:        # the CTA's group sum goes to the partial buffer
            total = group if total is None else (total + group).astype(F32)
            group = None
    return total


@pytest.mark.parametrize("name", ["galaxy_1000_2.npz", "galaxy_4003_2.npz", "galaxy_6000_3.npz"])
def test_planned_summation_order_is_at_least_as_accurate_as_the_reference(nb, oracle, golden, name):
    g = golden(name)
    w0, ref1 = g["world0"], g["step1"]
    m = mass_len_of(w0)
    info, _ = nb.plan(len(w0), m, 1)
    assert info["slabs"] * info["slab_len"] >= m
    got = replay_step_kernel_acc(w0, m, info["slab_len"], info["slabs"])
    truth, _ = oracle.acc_f64(w0, m)
    cuda_ref = rel_acc_error(got, acc_of(ref1))
    ref_f64 = rel_acc_error(acc_of(ref1), truth)
    cuda_f64 = rel_acc_error(got, truth)
    print(f"\n[summation model] {name}: {info['slabs']} slabs x {info['slab_len']}: cuda-vs-ref p99 "
          f"{np.percentile(cuda_ref, 99):.2e}, ref-vs-f64 p99 {np.percentile(ref_f64, 99):.2e}, "
          f"cuda-vs-f64 p99 {np.percentile(cuda_f64, 99):.2e}")
    assert np.percentile(cuda_ref, 99) <= 1e-5                      # north_star tolerance
    assert np.percentile(cuda_f64, 99) <= max(np.percentile(ref_f64, 99), 2e-6) * 1.05
    assert np.median(cuda_f64) <= max(np.median(ref_f64), 1e-6) * 1.05


def test_longer_chains_lose_accuracy_in_the_model(nb, oracle, golden):
    """The reason for the planner's accuracy floor: one slab (chains of M/2 terms) is measurably worse than 16."""
    g = golden("galaxy_6000_3.npz")
    w0 = g["world0"]
    m = mass_len_of(w0)
    truth, _ = oracle.acc_f64(w0, m)
    one = rel_acc_error(replay_step_kernel_acc(w0, m, m + (-m) % 16, 1), truth)
    many = rel_acc_error(replay_step_kernel_acc(w0, m, 192, -(-m // 192)), truth)
    assert np.percentile(many, 99) < np.percentile(one, 99)


def test_slab_groups_bound_the_partial_buffer(nb, monkeypatch):
    """Advisor finding (round 1): one partial per (slab, target) is O(N x M) bytes.  Past a fixed budget the plan
    groups consecutive slabs per CTA; the buffer stays within the budget at every size, the chains do not get
    longer, and the grouping is a function of (N, M) only (same for every shard count)."""
    budget = 8 << 30
    for n in (1 << 20, 1 << 21, 1 << 22, 1 << 24, 1 << 26, 3 << 26):
        m = n // 2
        info, _ = nb.plan(n, m, 1)
        tiles_pad = info["tiles"] * info["tile_targets"]
        assert info["groups"] == -(-info["slabs"] // info["group_slabs"])
        assert info["groups"] * tiles_pad * 8 <= budget or info["groups"] == 1, (n, info)
        assert info["slab_len"] <= 16384 and info["slabs"] * info["slab_len"] >= m
        assert (info["group_slabs"] > 1) == (n > 1 << 22)      # N <= 4M: one CTA per slab, as measured
        if info["group_slabs"] > 1:
            assert info["slab_len"] % 512 == 0 or info["variant"] == "t128x2"     # whole source tiles
        for shards in (2, 8):
            other, _ = nb.plan(n, m, shards)
            assert _order(other) == _order(info)
    one_m, _ = nb.plan(1 << 20, 523884, 1)
    assert one_m["group_slabs"] == 1 and one_m["slabs"] == 37      # the measured N = 1M plan is untouched
    # forced at small N (what the GPU tests use)
    monkeypatch.setenv("NBODY_SMALL_MAX_N", "0")   # N = 20 000 would otherwise take the balanced kernel
    base, _ = nb.plan(20000, 9956, 1)
    per_partial = base["tiles"] * base["tile_targets"] * 8
    for want in (1, 3, 5):
        monkeypatch.setenv("NBODY_PART_BUDGET", str(want * per_partial + 8))
        info, _ = nb.plan(20000, 9956, 1)
        assert info["group_slabs"] > 1 and info["groups"] <= want, (want, info)
        assert info["slabs"] * info["slab_len"] >= 9956 and info["slab_len"] <= base["slab_len"]


@pytest.mark.parametrize("name", ["galaxy_4003_2.npz", "galaxy_6000_3.npz"])
def test_grouped_summation_order_in_the_model(nb, oracle, golden, name, monkeypatch):
    """The grouped order (slabs of a group first, then the groups) is as accurate as the ungrouped one."""
    g = golden(name)
    w0 = g["world0"]
    m = mass_len_of(w0)
    monkeypatch.setenv("NBODY_PART_BUDGET", str(len(w0) * 8 * 3))
    monkeypatch.setenv("NBODY_SMALL_MAX_N", "0")     # not the small-World kernel: it has no partial buffer to bound
    info, _ = nb.plan(len(w0), m, 1)
    assert info["group_slabs"] > 1 and info["groups"] <= 3
    got = replay_step_kernel_acc(w0, m, info["slab_len"], info["slabs"], info["group_slabs"])
    truth, _ = oracle.acc_f64(w0, m)
    ref_f64 = rel_acc_error(acc_of(g["step1"]), truth)
    cuda_f64 = rel_acc_error(got, truth)
    assert np.percentile(cuda_f64, 99) <= max(np.percentile(ref_f64, 99), 2e-6) * 1.05


def test_small_worlds_split_their_slabs_over_the_warps_of_one_cta(nb, monkeypatch):
    """Worlds of at most 4 x 32 x 148 targets take the small-World kernel (small_kernel.cuh): one CTA per tile of
    32 T targets, T = ceil(N / (32 x 148)), at most one CTA per SM; T = 3, 4 only when most SMs get a CTA.  Slabs
    of equal length, at most 512 sources, one (or k) per warp; the order of summation is still 'slabs in slab order'
    (group_slabs == 1) and does not depend on the shard count."""
    want = {1: "w32x1", 250: "w32x1", 1000: "w32x1", 4003: "w32x1", 4736: "w32x1", 4737: "w16x2", 6000: "w16x2",
            9472: "w16x2", 9473: "b16x2", 10000: "b16x2", 12000: "w16x3", 14208: "w16x3", 14209: "b16x2",
            16384: "w16x4", 18944: "w16x4", 18945: "b16x2", 24000: "b16x2", 24001: "t128x2"}
    for n, variant in want.items():
        m = max(1, n // 2)
        info, _ = nb.plan(n, m, 1)
        assert info["variant"] == variant, (n, info)
        if not variant.startswith("w"):
            continue
        t = int(variant[-1])
        warps = int(variant[1:-2])
        assert info["tile_targets"] == 32 * t and info["tiles"] <= 148 and info["group_slabs"] == 1
        assert info["slabs"] * info["slab_len"] >= m and (info["slabs"] - 1) * info["slab_len"] < m
        assert info["slab_len"] <= 512 and info["slab_len"] % 16 == 0
        assert info["slabs"] <= warps * -(-m // (warps * 512))
        assert _order(nb.plan(n, m, 2)[0]) == _order(info)
    monkeypatch.setenv("NBODY_SMALL_MAX_N", "0")
    assert nb.plan(1000, 485, 1)[0]["variant"] == "t128x2"
