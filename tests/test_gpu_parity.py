"""GPU parity tests: the CUDA path, called through the C ABI, against the oracle and the golden
fixtures produced by the reference.  Run on the B200 box with `pytest -m gpu`.

Tolerance (BASELINE.json north_star): per-particle acceleration relative error <= 1e-5 after one
step against the reference CPU update.  The reference's own FP32 rounding noise against an FP64
restatement already reaches 1e-5 (max) at N = 16K, so every comparison reports
three columns -- CUDA vs reference, reference vs FP64, CUDA vs FP64 -- and asserts
  (a) the strict bound on the 99th percentile and a 5e-5 cap on the worst particle, and
  (b) that the CUDA path is at least as close to FP64 as the reference is (p99 and max).
"""
import numpy as np
import pytest

from _oracle import PARTICLE, acc_of, mass_len_of, rel_acc_error

pytestmark = pytest.mark.gpu

TOL_ACC = 1e-5        # north_star tolerance, asserted on p99
TOL_ACC_WORST = 5e-5  # worst single particle (cancellation-dominated outliers)


def _stats(e):
    return dict(median=float(np.median(e)), p99=float(np.percentile(e, 99)), max=float(e.max()))


def _report(tag, cuda_ref, ref_f64, cuda_f64):
    print(f"\n[parity] {tag}")
    for name, e in (("cuda-vs-ref", cuda_ref), ("ref-vs-f64", ref_f64), ("cuda-vs-f64", cuda_f64)):
        s = _stats(e)
        print(f"   {name:12s} median {s['median']:.2e}  p99 {s['p99']:.2e}  max {s['max']:.2e}")


def _check_one_step(nb, oracle, w0, ref1, dt, tag, strict=True):
    """strict=True: the north_star bound itself (N <= ~20K, where the reference's own FP32 noise is
    below it).  strict=False (larger N): the reference's rounding noise against FP64 exceeds 1e-5
    (BASELINE.md section 2), so the bound on CUDA-vs-reference is the reference's own distance to FP64
    (triangle inequality) and the 1e-5 bound is asserted on CUDA-vs-FP64 instead."""
    m = mass_len_of(w0)
    with nb.World(w0) as w:          # w0 is already partitioned: CreateWorld keeps its order
        assert w.mass_len == m
        w.update_gpu(dt, 1)
        got = w.particles()
    truth, _ = oracle.acc_f64(w0, m)
    cuda_ref = rel_acc_error(acc_of(got), acc_of(ref1))
    ref_f64 = rel_acc_error(acc_of(ref1), truth)
    cuda_f64 = rel_acc_error(acc_of(got), truth)
    _report(tag, cuda_ref, ref_f64, cuda_f64)
    if strict:
        assert np.percentile(cuda_ref, 99) <= TOL_ACC
        assert cuda_ref.max() <= TOL_ACC_WORST
    else:
        assert np.percentile(cuda_f64, 99) <= TOL_ACC
        assert np.percentile(cuda_ref, 99) <= 1.1 * np.percentile(ref_f64, 99) + 2e-6
        assert np.median(cuda_ref) <= 1.1 * np.median(ref_f64) + 1e-6
    assert np.percentile(cuda_f64, 99) <= max(np.percentile(ref_f64, 99), 2e-6) * 1.05
    assert cuda_f64.max() <= max(ref_f64.max(), 1e-5) * 1.05
    # integrator: v = v0 + a dt, x = x0 + v dt with the reference's rounding, given OUR acc
    a = acc_of(got)
    for c, (vk, pk) in enumerate((("vx", "px"), ("vy", "py"))):
        v = (w0[vk] + a[:, c].astype(np.float32) * np.float32(dt)).astype(np.float32)
        x = (w0[pk] + v * np.float32(dt)).astype(np.float32)
        assert np.array_equal(v, got[vk]) and np.array_equal(x, got[pk])
    # mass and radius are never modified
    assert np.array_equal(got["mass"], w0["mass"]) and np.array_equal(got["radius"], w0["radius"])
    return got


@pytest.mark.parametrize("name", ["galaxy_1000_2.npz", "galaxy_4003_2.npz", "galaxy_6000_3.npz"])
def test_one_step_against_reference_fixture(nb, oracle, golden, name):
    g = golden(name)
    _check_one_step(nb, oracle, g["world0"], g["step1"], float(g["dt"]), name)


@pytest.mark.parametrize("n,gal", [(16384, 2), (20000, 2), (50000, 2)])
def test_one_step_against_oracle(nb, oracle, n, gal):
    ps = nb.make_galaxies(n, gal)
    w0, m = oracle.partition(ps)
    ref1 = oracle.update(w0, m, 0.01, 1, 8)   # bit-exact restatement of the reference AVX build
    # strict (p99 of CUDA-vs-reference <= 1e-5) where the reference's own p99 against FP64 is below the bound; at
    # N = 20 000 it is 1.05e-5, and a CUDA result close to FP64 (p99 8.8e-7) is therefore ~1.0e-5 from the reference
    got = _check_one_step(nb, oracle, w0, ref1, 0.01, f"galaxy {n}/{gal}", strict=n <= 16384)
    if n <= 20000:
        assert rel_acc_error(acc_of(got), acc_of(ref1)).max() <= TOL_ACC_WORST


# The tolerance exactly as north_star words it -- EVERY particle's acceleration within 1e-5 of the reference's after one
# step -- with the measured worst particle on B200 (round 2).  Where the reference's own FP32 rounding noise against FP64
# is already above the bound (third column) no implementation with another summation order can meet it: those
# cases are strict xfails, so the miss stays visible and a change that closes it is noticed.
LITERAL_CASES = [
    # n, galaxies, worst cuda-vs-ref, worst ref-vs-f64, meets the literal bound
    (1000, 2, 1.1e-06, 8.7e-07, True),
    (4003, 2, 3.0e-06, 2.8e-06, True),
    (6000, 3, 3.4e-05, 5.1e-05, False),
    (16384, 2, 7.2e-06, 1.0e-05, True),
    (20000, 2, 1.5e-05, 1.5e-05, False),
    (50000, 2, 6.7e-05, 8.0e-05, False),
]


@pytest.mark.parametrize("n,gal,worst,ref_noise,meets", LITERAL_CASES)
def test_literal_north_star_bound_on_every_particle(nb, oracle, request, n, gal, worst, ref_noise, meets):
    if not meets:
        request.applymarker(pytest.mark.xfail(strict=True, reason=(
            f"literal bound missed: worst particle {worst:.1e} vs the reference (the reference itself is {ref_noise:.1e} "
            "from FP64 on its worst particle)")))
    w0, m = oracle.partition(nb.make_galaxies(n, gal))
    ref1 = oracle.update(w0, m, 0.01, 1, 8)
    with nb.World(w0) as w:
        w.update_gpu(0.01, 1)
        got = w.particles()
    err = rel_acc_error(acc_of(got), acc_of(ref1))
    print(f"\n[literal] N={n}/{gal}: worst particle {err.max():.2e} (recorded {worst:.1e}), p99 {np.percentile(err, 99):.2e}")
    assert err.max() <= TOL_ACC


@pytest.mark.parametrize("want_groups,n", [(1, 20000), (3, 20000), (5, 50000), (2, 4003)])
def test_slab_groups_against_oracle(nb, oracle, monkeypatch, want_groups, n):
    """Large Worlds (N >= 2M) run the GROUPED kernel: a CTA walks several slabs and adds their sums itself, so that the
    partial buffer stays within a fixed budget (nslab x N x 8 B otherwise).  NBODY_PART_BUDGET forces the same plan
    shape at sizes the oracle finishes in seconds: same parity bar as the ungrouped plan, chained == single steps."""
    monkeypatch.setenv("NBODY_SMALL_MAX_N", "0")   # N = 4003 would otherwise take the small-World kernel (no partials)
    ps = nb.make_galaxies(n, 2)
    w0, m = oracle.partition(ps)
    base, _ = nb.plan(n, m, 1)
    monkeypatch.setenv("NBODY_PART_BUDGET", str(want_groups * base["tiles"] * base["tile_targets"] * 8 + 8))
    info, _ = nb.plan(n, m, 1)
    assert info["group_slabs"] > 1 and info["groups"] <= want_groups
    ref1 = oracle.update(w0, m, 0.01, 1, 8)
    with nb.World(w0) as w:
        w.update_gpu(0.01, 1)
        pinfo = w.pipeline_info()
    assert pinfo["group_slabs"] == info["group_slabs"] and pinfo["slabs"] == info["slabs"]
    _check_one_step(nb, oracle, w0, ref1, 0.01, f"galaxy {n}/2, {info['slabs']} slabs in {info['groups']} group(s)",
                    strict=n <= 20000)
    with nb.World(ps) as a, nb.World(ps) as b:
        a.update_gpu(0.01, 4)
        for _ in range(4):
            b.update_gpu(0.01, 1)
        assert a.particles().tobytes() == b.particles().tobytes()


@pytest.mark.parametrize("name", ["galaxy_1000_2.npz", "galaxy_4003_2.npz"])
def test_ten_steps_against_reference_fixture(nb, golden, name):
    g = golden(name)
    with nb.World(g["input"]) as w:
        w.update_gpu(float(g["dt"]), 10)
        got = w.particles()
    ref = g["step10"]
    scale = np.abs(ref["px"]).max()
    assert np.abs(got["px"] - ref["px"]).max() <= 1e-6 * scale
    assert np.abs(got["py"] - ref["py"]).max() <= 1e-6 * scale
    vs = np.hypot(ref["vx"], ref["vy"]).max()
    assert np.hypot(got["vx"] - ref["vx"], got["vy"] - ref["vy"]).max() <= 1e-5 * vs


def test_two_body_closed_form(nb):
    d, m0, m1, r0, r1 = 100.0, 5.0, 7.0, 2.0, 3.0
    ps = np.zeros(2, dtype=PARTICLE)
    ps["px"] = [0.0, d]
    ps["mass"] = [m0, m1]
    ps["radius"] = [r0, r1]
    with nb.World(ps) as w:
        w.update_gpu(0.5, 1)
        out = w.particles()
    a0 = 10.0 * m1 * d / (d * d + r0) ** 1.5
    a1 = -10.0 * m0 * d / (d * d + r1) ** 1.5
    assert out["ax"][0] == pytest.approx(a0, rel=1e-6)
    assert out["ax"][1] == pytest.approx(a1, rel=1e-6)
    assert out["ay"][0] == 0.0 and out["ay"][1] == 0.0
    assert out["vx"][0] == pytest.approx(a0 * 0.5, rel=1e-6)
    assert out["px"][0] == pytest.approx(a0 * 0.25, rel=1e-6)


def test_massless_targets_and_self_term(nb):
    ps = np.zeros(4, dtype=PARTICLE)
    ps["px"] = [50.0, 0.0, -30.0, 1e4]
    ps["py"] = [0.0, 0.0, 0.0, 1e4]
    ps["mass"] = [0.0, 100.0, 0.0, 0.0]
    ps["radius"] = [0.5, 1.0, 0.5, 0.5]
    with nb.World(ps) as w:
        assert w.mass_len == 1
        w.update_gpu(0.1, 1)
        out = w.particles()
    i = int(np.flatnonzero(out["mass"] > 0)[0])
    assert i == 0
    assert out["ax"][0] == 0.0 and out["ay"][0] == 0.0      # only source is itself: 0 * f
    assert out["px"][0] == 0.0
    for k in range(1, 4):                                   # massless particles are pulled
        assert np.hypot(out["ax"][k], out["ay"][k]) > 0


def test_no_massive_particles(nb):
    ps = np.zeros(300, dtype=PARTICLE)
    ps["px"] = np.arange(300)
    ps["vx"] = 1.0
    ps["radius"] = 0.5
    with nb.World(ps) as w:
        w.update_gpu(0.5, 2)
        out = w.particles()
    assert np.all(out["ax"] == 0) and np.all(out["ay"] == 0)
    assert np.array_equal(out["px"], ps["px"] + 1.0)


def test_chained_steps_equal_single_steps_bitwise(nb):
    ps = nb.make_galaxies(5000, 2)
    with nb.World(ps) as a, nb.World(ps) as b:
        a.update_gpu(0.01, 7)
        for _ in range(7):
            b.update_gpu(0.01, 1)
        assert a.particles().tobytes() == b.particles().tobytes()


@pytest.mark.parametrize("n,steps", [(250, 2), (250, 100), (1000, 3), (2000, 17), (4003, 10), (6000, 5), (10000, 4)])
def test_persistent_multi_step_kernel_equals_one_launch_per_step_bitwise(nb, oracle, monkeypatch, n, steps):
    """Opt-in path (NB_SIM_FLAG_PERSIST): n >= 2 chained steps of a small World as ONE cooperative launch with a grid
    barrier between steps (persist_kernel.cuh; measured slower than n launches, hence not the default).  Same plan,
    same summation order: not a bit may differ from n launches."""
    from nbody_b200.api import FLAG_PERSIST
    monkeypatch.setenv("NBODY_SMALL_MAX_N", "0")   # the persistent kernel is a build of the (tile, slab)-per-CTA kernel
    arr, m = oracle.partition(nb.make_galaxies(n, 2))
    with nb.SimPipeline(n, m, flags=FLAG_PERSIST) as a, nb.SimPipeline(n, m) as b:   # NBODY_SMALL_MAX_N=0: no resident path either
        a.set_data(arr)
        b.set_data(arr)
        a.update(steps, 0.01)
        b.update(steps, 0.01)
        ia, ib = a.info(), b.info()
        assert ia["persistent_launches"] == 1 and ia["kernels_launched"] < ib["kernels_launched"]
        assert ib["persistent_launches"] == 0
        assert a.get_data().tobytes() == b.get_data().tobytes()
        # a second call continues from the right buffer (odd and even step counts) and the barrier counter carries on
        a.update(3, 0.02)
        b.update(3, 0.02)
        a.update(1, 0.01)
        b.update(1, 0.01)
        a.update(2, 0.01)
        b.update(2, 0.01)
        assert a.get_data().tobytes() == b.get_data().tobytes()
        assert a.info()["persistent_launches"] == 3


def test_large_worlds_do_not_take_the_persistent_path(nb, oracle):
    from nbody_b200.api import FLAG_PERSIST
    n = 50000
    arr, m = oracle.partition(nb.make_galaxies(n, 2))
    with nb.SimPipeline(n, m, flags=FLAG_PERSIST) as a:
        a.set_data(arr)
        a.update(3, 0.01)
        assert a.info()["persistent_launches"] == 0


@pytest.mark.parametrize("n", [1, 31, 32, 33, 250, 1000, 2049, 4003, 6000, 12000])
@pytest.mark.parametrize("variant", ["w32x1", "w16x2", "w16x3", "w16x4", "w8x1"])
def test_small_world_kernel_equals_step_kernel_bitwise(nb, oracle, monkeypatch, n, variant):
    """small_kernel.cuh splits a tile's slabs over the warps of one CTA and adds the slab sums in shared memory;
    step_kernel.cuh gives each slab a CTA and adds them through global memory.  Same slab boundaries, same chains:
    the results must not differ by a bit -- one step and chained steps, with massless tails and ragged sizes."""
    arr, m = oracle.partition(nb.make_galaxies(n, 2) if n >= 250 else _cloud(n))
    monkeypatch.setenv("NBODY_VARIANT", variant)
    with nb.SimPipeline(n, m) as a:
        info = a.info()
        a.set_data(arr)
        a.update(1, 0.01)
        one_small = a.get_data()
        a.update(5, 0.02)
        six_small = a.get_data()
    assert info["threads"] == 32 * int(variant[1:-2]) and info["tile_targets"] == 32 * int(variant[-1])
    monkeypatch.setenv("NBODY_VARIANT", "t128x2")
    with nb.SimPipeline(n, m, slab_len=info["slab_len"]) as b:
        assert b.info()["slab_len"] == info["slab_len"] and b.info()["slabs"] == info["slabs"]
        b.set_data(arr)
        b.update(1, 0.01)
        one_big = b.get_data()
        b.update(5, 0.02)
        six_big = b.get_data()
    assert one_small.tobytes() == one_big.tobytes()
    assert six_small.tobytes() == six_big.tobytes()


@pytest.mark.parametrize("n", [1, 33, 250, 1000, 4003, 6000, 12000, 16384, 20000])
@pytest.mark.parametrize("variant,slab", [("b16x2", 0), ("b16x2", 128), ("b16x1", 0), ("b16x4", 256), ("b16x3", 512), ("b16x4", 0)])
def test_balanced_kernel_equals_step_kernel_bitwise(nb, oracle, monkeypatch, n, variant, slab):
    """balanced_kernel.cuh shares (target row, slab) units out evenly over one CTA per SM: rows whose slabs all fall
    into one CTA are added up in shared memory, rows that straddle a CTA boundary through the partial buffer and a
    ticket.  Which CTA computed which unit must not show: same slab boundaries, same chains as step_kernel.cuh, the
    same bits -- one step and chained steps, ragged sizes, massless tails, planned and forced slab lengths."""
    arr, m = oracle.partition(nb.make_galaxies(n, 2) if n >= 250 else _cloud(n))
    monkeypatch.setenv("NBODY_VARIANT", variant)
    with nb.SimPipeline(n, m, slab_len=slab) as a:
        info = a.info()
        a.set_data(arr)
        a.update(1, 0.01)
        one_bal = a.get_data()
        a.update(5, 0.02)
        six_bal = a.get_data()
    assert info["variant"] == variant and info["tile_targets"] == 32 * int(variant[-1])
    monkeypatch.setenv("NBODY_VARIANT", "t128x2")
    with nb.SimPipeline(n, m, slab_len=info["slab_len"]) as b:
        assert b.info()["slab_len"] == info["slab_len"] and b.info()["slabs"] == info["slabs"]
        b.set_data(arr)
        b.update(1, 0.01)
        one_big = b.get_data()
        b.update(5, 0.02)
        six_big = b.get_data()
    assert one_bal.tobytes() == one_big.tobytes()
    assert six_bal.tobytes() == six_big.tobytes()


def _cloud(n):
    rng = np.random.default_rng(1000 + n)
    ps = np.zeros(n, dtype=PARTICLE)
    ps["px"] = rng.normal(0, 1000, n)
    ps["py"] = rng.normal(0, 1000, n)
    ps["vx"] = rng.normal(0, 1, n)
    ps["mass"] = np.where(rng.random(n) < 0.6, rng.uniform(10, 1000, n), 0.0)
    ps["radius"] = rng.uniform(0.5, 9.5, n)
    return ps


@pytest.mark.parametrize("n", [250, 1000, 4003])
def test_cuda_graph_chunks_equal_stream_launches_bitwise(nb, oracle, n):
    """Chained steps of a small World are replayed as CUDA-graph chunks of 20 steps (step_async); the remainder
    and Worlds above 4 736 targets go launch by launch.  Same kernels, same arguments: same bits -- across odd
    step counts (the chunk must start from either source buffer) and a change of dt (the graph is re-captured)."""
    from nbody_b200.api import FLAG_NO_GRAPH, FLAG_NO_RESIDENT    # (tiny Worlds would take the resident kernel instead)
    arr, m = oracle.partition(nb.make_galaxies(n, 2))
    with nb.SimPipeline(n, m, flags=FLAG_NO_RESIDENT) as a, nb.SimPipeline(n, m, flags=FLAG_NO_GRAPH | FLAG_NO_RESIDENT) as b:
        a.set_data(arr)
        b.set_data(arr)
        for steps, dt in ((45, 0.01), (1, 0.01), (61, 0.01), (40, 0.02), (3, 0.02), (20, 0.01)):
            a.update(steps, dt)
            b.update(steps, dt)
            assert a.get_data().tobytes() == b.get_data().tobytes(), (steps, dt)
        ia, ib = a.info(), b.info()
    assert ia["graph_steps"] == 20 and ib["graph_steps"] == 0
    assert ia["graph_launches"] == 2 + 3 + 2 + 1 and ib["graph_launches"] == 0
    assert ia["kernels_launched"] == ib["kernels_launched"]        # replayed kernels are counted like launched ones


@pytest.mark.parametrize("n", [1, 31, 33, 250, 500, 640, 800, 1000, 1200, 1536])
def test_resident_cluster_kernel_equals_launch_per_step_bitwise(nb, oracle, monkeypatch, n):
    """Chained steps of a tiny World run as ONE launch of ONE thread-block cluster that keeps the whole state on chip
    (resident_kernel.cuh): positions travel between the CTAs through distributed shared memory, a cluster barrier
    separates the steps, global memory is read at the start and written at the end.  Same plan and summation order
    as the small-World kernel: the same bits -- odd and even step counts (either source buffer may be the last),
    a change of dt, single steps in between (they go launch by launch), massless tails, ragged sizes."""
    from nbody_b200.api import FLAG_NO_GRAPH, FLAG_NO_RESIDENT
    monkeypatch.setenv("NBODY_RESIDENT_MAX_N", "1536")   # planned up to N = 640; the kernel itself is checked beyond
    arr, m = oracle.partition(nb.make_galaxies(n, 2) if n >= 250 else _cloud(n))
    with nb.SimPipeline(n, m) as a, nb.SimPipeline(n, m, flags=FLAG_NO_GRAPH | FLAG_NO_RESIDENT) as b:
        a.set_data(arr)
        b.set_data(arr)
        for steps, dt in ((1, 0.01), (7, 0.01), (2, 0.02), (1, 0.02), (100, 0.01), (3, 0.01)):
            a.update(steps, dt)
            b.update(steps, dt)
            assert a.get_data().tobytes() == b.get_data().tobytes(), (steps, dt)
        ia, ib = a.info(), b.info()
    rows = -(-n // 32)
    per_cta = -(-rows // 16)
    assert ia["resident_ctas"] == -(-rows // per_cta) and ib["resident_ctas"] == 0
    assert ia["resident_launches"] == 4 and ib["resident_launches"] == 0
    assert ia["kernels_launched"] < ib["kernels_launched"]


@pytest.mark.parametrize("n,want", [(250, 1), (800, 1), (1000, 4), (1200, 3), (1600, 2), (2000, 2), (2368, 2), (2400, 1), (4003, 1)])
def test_cluster_split_small_kernel_equals_one_cta_per_tile_bitwise(nb, oracle, monkeypatch, n, want):
    """Small Worlds whose tiles would leave more than half of the SMs idle share each tile's slabs over a cluster of
    2 ... 4 CTAs (split_kernel.cuh): the slab sums meet in the first CTA's shared memory through DSMEM and are added
    in slab order there.  A launch detail, not part of the plan: same bits as one CTA per tile -- single steps,
    chained steps (CUDA-graph chunks on both sides), a change of dt."""
    from nbody_b200.api import FLAG_NO_RESIDENT
    arr, m = oracle.partition(nb.make_galaxies(n, 2) if n >= 250 else _cloud(n))
    with nb.SimPipeline(n, m, flags=FLAG_NO_RESIDENT) as a:
        monkeypatch.setenv("NBODY_NO_SPLIT", "1")
        with nb.SimPipeline(n, m, flags=FLAG_NO_RESIDENT) as b:
            monkeypatch.delenv("NBODY_NO_SPLIT")
            ia, ib = a.info(), b.info()
            assert ia["cluster_split"] == want and ib["cluster_split"] == 1
            assert (ia["variant"], ia["slab_len"], ia["slabs"]) == (ib["variant"], ib["slab_len"], ib["slabs"])
            a.set_data(arr)
            b.set_data(arr)
            for steps, dt in ((1, 0.01), (45, 0.01), (2, 0.02), (1, 0.02), (23, 0.01)):
                a.update(steps, dt)
                b.update(steps, dt)
                assert a.get_data().tobytes() == b.get_data().tobytes(), (steps, dt)


@pytest.mark.parametrize("massive", ["none", "all"])
def test_resident_and_split_paths_without_massless_or_without_massive_particles(nb, oracle, massive):
    """mass_len = 0 (only padding sources: every acceleration is exactly zero, particles drift) and mass_len =
    total_len, through the resident cluster (N = 300, chained) and the cluster split (N = 1 900, single and chained)."""
    from nbody_b200.api import FLAG_NO_GRAPH, FLAG_NO_RESIDENT
    for n in (300, 1900):
        ps = _cloud(n)
        ps["mass"] = 0.0 if massive == "none" else np.maximum(ps["mass"], 5.0)
        arr, m = oracle.partition(ps)
        assert m == (0 if massive == "none" else n)
        with nb.SimPipeline(n, m) as a:
            import os
            os.environ["NBODY_NO_SPLIT"] = "1"
            try:
                b = nb.SimPipeline(n, m, flags=FLAG_NO_GRAPH | FLAG_NO_RESIDENT)
            finally:
                del os.environ["NBODY_NO_SPLIT"]
            with b:
                a.set_data(arr)
                b.set_data(arr)
                for steps in (1, 6, 3):
                    a.update(steps, 0.01)
                    b.update(steps, 0.01)
                    assert a.get_data().tobytes() == b.get_data().tobytes(), (n, steps)
                if massive == "none":
                    assert not a.get_data()["ax"].any()


def test_worlds_above_the_limit_do_not_take_the_resident_path(nb, oracle):
    arr, m = oracle.partition(nb.make_galaxies(800, 2))
    with nb.SimPipeline(800, m) as a:
        a.set_data(arr)
        a.update(5, 0.01)
        assert a.info()["resident_ctas"] == 0 and a.info()["resident_launches"] == 0


def test_get_between_updates_does_not_disturb_state(nb):
    ps = nb.make_galaxies(3000, 2)
    with nb.World(ps) as a, nb.World(ps) as b:
        a.update_gpu(0.01, 4)
        for _ in range(4):
            b.update_gpu(0.01, 1)
            b.particles()
        assert a.particles().tobytes() == b.particles().tobytes()


def test_zero_steps_is_noop_and_dt_may_change(nb, golden):
    g = golden("galaxy_1000_2.npz")
    with nb.World(g["input"]) as w:
        w.update_gpu(0.01, 0)
        assert w.particles().tobytes() == g["world0"].tobytes()
        w.update_gpu(0.01, 1)
        w.update_gpu(0.02, 1)
        out = w.particles()
    assert np.isfinite(out["px"]).all()


def test_cpu_gpu_interleaving_keeps_state(nb, oracle, golden):
    """TAB in the reference GUI switches paths on one World (ref src/main.c:112-116)."""
    g = golden("galaxy_1000_2.npz")
    dt = float(g["dt"])
    with nb.World(g["input"]) as w:
        w.update_gpu(dt, 2)
        w.update_cpu(dt, 3)
        w.update_gpu(dt, 2)
        w.update_cpu(dt, 3)
        got = w.particles()
    ref = g["step10"]
    scale = np.abs(ref["px"]).max()
    assert np.abs(got["px"] - ref["px"]).max() <= 1e-6 * scale
    assert np.abs(got["py"] - ref["py"]).max() <= 1e-6 * scale


@pytest.mark.parametrize("n", [1, 2, 3, 5, 127, 128, 129, 255, 256, 257, 511, 513, 1023, 1024, 1025, 2049])
def test_ragged_sizes_against_oracle(nb, oracle, n):
    rng = np.random.default_rng(n)
    ps = np.zeros(n, dtype=PARTICLE)
    ps["px"] = rng.normal(0, 1000, n)
    ps["py"] = rng.normal(0, 1000, n)
    ps["vx"] = rng.normal(0, 1, n)
    ps["mass"] = np.where(rng.random(n) < 0.6, rng.uniform(10, 1000, n), 0.0)
    ps["radius"] = rng.uniform(0.5, 9.5, n)
    w0, m = oracle.partition(ps)
    ref = oracle.update(w0, m, 0.01, 1, 8)
    with nb.World(ps) as w:
        w.update_gpu(0.01, 1)
        got = w.particles()
    assert np.array_equal(got["mass"], w0["mass"])
    # Random clouds contain targets whose acceleration is a near-complete cancellation of large
    # terms; for those the relative error of ANY FP32 evaluation is unbounded.  So: the 1e-5 bound
    # on the 99th percentile, and for every particle the error normalised by sum_j |term_ij| (the
    # conditioning-free scale the oracle reports) within a few FP32 ulps.
    truth, nrm = oracle.acc_f64(w0, m)
    err = rel_acc_error(acc_of(got), acc_of(ref))
    if n >= 100:
        assert np.percentile(err, 99) <= 1e-5, np.percentile(err, 99)
    d = np.hypot(*(acc_of(got).astype(np.float64) - truth).T)
    d_ref = np.hypot(*(acc_of(ref).astype(np.float64) - truth).T)
    scale = np.where(nrm > 0, nrm, 1.0)
    assert (d / scale).max() <= 1e-6, (d / scale).max()
    assert (d / scale).max() <= max(3.0 * (d_ref / scale).max(), 5e-7)
    assert np.allclose(got["px"], ref["px"], rtol=1e-6, atol=1e-6)
