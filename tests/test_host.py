"""CPU-side tests of the product library (no GPU needed): the partition permutation, the galaxy
generator, UpdateWorld_CPU, and that libnbody.so exports every symbol the headers declare."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from _oracle import PARTICLE, RefLib, mass_len_of

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# The seven cases of the reference's only unit test (ref test/test_particle_sort.c:27-111):
# input -> (expected order, expected count of non-zeros).
SORT_CASES = [
    ([1, 2, 3, 4, 5], [1, 2, 3, 4, 5], 5),       # test_on_non_zeros
    ([0, 0, 0, 0, 0], [0, 0, 0, 0, 0], 0),       # test_on_zeros
    ([1, 2, 3, 0, 0], [1, 2, 3, 0, 0], 3),       # test_on_sorted
    ([0, 0, 1, 2, 3], [3, 2, 1, 0, 0], 3),       # test_on_reverse_sorted_odd
    ([0, 0, 0, 1, 2, 3], [3, 2, 1, 0, 0, 0], 3),  # test_on_reverse_sorted_even
    ([0, 1, 2, 0, 3], [3, 1, 2, 0, 0], 3),       # test_on_unsorted_odd
    ([0, 1, 2, 0, 3, 0], [3, 1, 2, 0, 0, 0], 3),  # test_on_unsorted_even
]


def _particles_from_masses(ms):
    ps = np.zeros(len(ms), dtype=PARTICLE)
    ps["mass"] = ms
    ps["px"] = np.arange(len(ms)) + 1000.0   # identity tag that survives the partition
    ps["radius"] = 1.0
    return ps


@pytest.mark.parametrize("inp,exp,count", SORT_CASES)
def test_oracle_partition_int_reference_cases(oracle, inp, exp, count):
    out, m = oracle.partition_int(inp)
    assert m == count and out.tolist() == exp


@pytest.mark.parametrize("inp,exp,count", SORT_CASES)
def test_createworld_partition_reference_cases(nb, inp, exp, count):
    """The same seven cases through the product's CreateWorld + GetWorldParticles."""
    with nb.World(_particles_from_masses(inp)) as w:
        got = w.particles()
        assert w.mass_len == count
    assert got["mass"].tolist() == [float(x) for x in exp]


def test_createworld_partition_random_matches_oracle(nb, oracle):
    rng = np.random.default_rng(7)
    for n in (1, 2, 3, 17, 256, 1001):
        ms = rng.integers(0, 3, size=n).astype(np.float32) * rng.random(n).astype(np.float32)
        ms[rng.random(n) < 0.1] = -1.0       # negative mass counts as massless (mass <= 0)
        ps = _particles_from_masses(ms)
        exp, m = oracle.partition(ps)
        with nb.World(ps) as w:
            got = w.particles()
            assert w.mass_len == m
        assert got.tobytes() == exp.tobytes()


def test_createworld_copies_and_size_query(nb):
    ps = _particles_from_masses([1, 0, 2])
    before = ps.copy()
    with nb.World(ps) as w:
        ps["px"] = -1.0                       # caller keeps ownership; World must not alias it
        got = w.particles()
    assert sorted(got["px"].tolist()) == sorted(before["px"].tolist())
    # size pointer may be NULL (ref world.c:93)
    L = nb.lib()
    w = nb.CreateWorld(before)
    assert L.GetWorldParticles(w, None)
    nb.DestroyWorld(w)
    nb.DestroyWorld(None)                     # NULL is ignored (ref world.c:68)


@pytest.mark.parametrize("name", ["galaxy_1000_2.npz", "galaxy_4003_2.npz", "galaxy_6000_3.npz"])
def test_make_galaxies_bit_identical_to_reference_fixture(nb, golden, name):
    g = golden(name)
    n = len(g["input"])
    gal = int(name.split("_")[2].split(".")[0])
    mine = nb.make_galaxies(n, gal, seed=int(g["seed"]))
    assert mine.tobytes() == g["input"].tobytes()


@pytest.mark.parametrize("n,gal,seed", [(200, 2, 1), (777, 3, 99), (20000, 2, 11037), (5000, 5, 3)])
def test_make_galaxies_bit_identical_to_compiled_reference(nb, n, gal, seed):
    if not RefLib.available("avx"):
        pytest.skip("oracle/_ref not built")
    r = RefLib("avx")
    assert nb.make_galaxies(n, gal, seed=seed).tobytes() == r.make_galaxies(n, gal, seed=seed).tobytes()


def test_make_galaxies_value_ranges(nb):
    ps = nb.make_galaxies(16384, 2)
    massless = ps["mass"] == 0
    assert 0.4 < massless.mean() < 0.6                     # about half are tracer particles
    assert np.all(ps["radius"][massless] == 0.5)
    assert np.all(ps["radius"] > 0)                        # softening never zero
    cores = ps["radius"] >= 200
    assert cores.sum() == 2


@pytest.mark.parametrize("name", ["galaxy_1000_2.npz", "galaxy_4003_2.npz"])
def test_update_world_cpu_matches_reference_fixture(nb, golden, name):
    """UpdateWorld_CPU of the product rounds like the reference's AVX build."""
    g = golden(name)
    with nb.World(g["input"]) as w:
        w.update_cpu(float(g["dt"]), 1)
        one = w.particles()
        w.update_cpu(float(g["dt"]), 9)
        ten = w.particles()
    assert one.tobytes() == g["step1"].tobytes()
    assert ten.tobytes() == g["step10"].tobytes()


def test_update_world_cpu_zero_steps_is_noop(nb, golden):
    g = golden("galaxy_1000_2.npz")
    with nb.World(g["input"]) as w:
        w.update_cpu(0.01, 0)
        assert w.particles().tobytes() == g["world0"].tobytes()


def _declared_functions(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+)?[A-Za-z_][A-Za-z0-9_ ]*?[\s\*]+([A-Z][A-Za-z0-9_]+)\s*\([^;{]*\)\s*;", text,
                       flags=re.M)
    return [n for n in names if not n.startswith("V2")]


def test_library_exports_every_declared_symbol(nb):
    declared = _declared_functions("nbody.h") + _declared_functions("galaxy.h") + _declared_functions("nbody_sim.h")
    for must in ("CreateWorld", "UpdateWorld_GPU", "MakeGalaxies", "CreateSimPipeline", "PerformSimUpdate",
                 "NbSimCreate", "NbSimGetUniqueId"):
        assert must in declared, (must, declared)
    L = C.CDLL(nb.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/ but not exported by libnbody.so"


def test_library_is_sm100a_only(nb):
    out = subprocess.run(["cuobjdump", "-lelf", nb.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_\d+a?", out.stdout))
    assert archs == {"sm_100a"}, archs


def test_pair_loop_schedule_is_the_measured_one(nb):
    """The order ptxas gives the pair loop is worth a few per cent, so the opcode sequence of
    the built kernels must be the one the committed measurements were taken with (tools/sass_fingerprint.py)."""
    import importlib.util
    import json
    spec = importlib.util.spec_from_file_location(
        "sass_fingerprint", os.path.join(os.path.dirname(__file__), "..", "tools", "sass_fingerprint.py"))
    fpmod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fpmod)
    if fpmod.cuobjdump() is None:
        pytest.skip("cuobjdump unavailable")
    want = json.load(open(fpmod.FINGERPRINT))
    got = fpmod.fingerprint(nb.LIB_PATH)
    meta = want.get("_meta", {})
    for name, w in want.items():
        if name.startswith("_"):
            continue
        g = got[name]
        # per 32 pairs (two packs of 4 sources x 4 targets): 9 packed FP32 instructions and 2 MUFU per 2 pairs
        assert g["loop_f32x2"] == w["loop_f32x2"] and g["loop_mufu"] == w["loop_mufu"], (name, g)
        # This assertion is MEANT to fail on any nvcc/ptxas or kernel-source change that reorders the loop: the
        # committed timings belong to the recorded schedule.  It is a prompt to re-measure, not a correctness failure.
        assert g["loop_sha1"] == w["loop_sha1"], (
            f"{name}: ptxas scheduled the pair loop differently from the schedule the committed numbers were "
            f"measured with (recorded with CUDA toolkit '{meta.get('toolkit', 'unknown')}', this build uses "
            f"'{fpmod.toolkit_version()}').  Procedure: {meta.get('procedure', 're-measure on a B200, then python tools/sass_fingerprint.py --update')}")


def test_device_count_without_gpu_is_zero_or_more(nb):
    assert nb.device_count() >= 0


def test_empty_world_is_legal_on_the_host_side(nb):
    """size == 0: the reference's CPU side handles it (ref sim_cpu.c:109-111); here every call is a no-op."""
    w = nb.CreateWorld(np.zeros(0, dtype=PARTICLE), 0)
    assert len(nb.GetWorldParticles(w)) == 0
    nb.UpdateWorld_CPU(w, 0.01, 3)
    nb.UpdateWorld_GPU(w, 0.01, 3)        # no particles: returns before any device is touched
    assert len(nb.GetWorldParticles(w)) == 0
    nb.DestroyWorld(w)


def test_two_worlds_coexist(nb, golden):
    """bench.c creates a CPU and a GPU World from the same array (ref src/bench.c:55-56)."""
    g = golden("galaxy_1000_2.npz")
    with nb.World(g["input"]) as a, nb.World(g["input"]) as b:
        a.update_cpu(float(g["dt"]), 1)
        assert a.particles().tobytes() == g["step1"].tobytes()
        assert b.particles().tobytes() == g["world0"].tobytes()


def test_c_program_against_the_world_api(nb, tmp_path):
    """A plain C user of include/nbody.h: the reference unit test's seven cases (tests/c/test_world_partition.c)."""
    exe = str(tmp_path / "test_world_partition")
    lib_dir = os.path.dirname(nb.LIB_PATH)
    subprocess.run(["/usr/bin/gcc", "-std=gnu11", "-O1", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "c", "test_world_partition.c"), "-o", exe,
                    "-L", lib_dir, "-lnbody", f"-Wl,-rpath,{lib_dir}", "-lm"], check=True)
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.count(" ok") == 7


def test_empty_pipeline_is_inert_through_the_reference_seam(nb):
    """The reference's own world.c creates the pipeline eagerly in CreateWorld (ref world.c:52), also for an empty
    World that its CPU path accepts (ref sim_cpu.c:109-111): CreateSimPipeline(total_len == 0) must return a
    pipeline on which Set/Get/Perform are no-ops -- without touching any device, so this runs on the build box."""
    import ctypes as C
    L = nb.lib()
    sim = L.CreateSimPipeline(nb.api.WorldData(0, 0, 0.0))
    assert sim
    L.SetSimulationData(sim, None)
    L.PerformSimUpdate(sim, 3, 0.01)
    L.NbSimStepAsync(sim, 2, 0.01)
    L.NbSimSynchronize(sim)
    L.GetSimulationData(sim, None)
    assert L.NbSimLastStepMs(sim) == 0.0
    info = nb.api.SimInfo()
    L.NbSimGetInfo(sim, C.byref(info))
    assert info.total_len == 0 and info.kernels_launched == 0
    L.DestroySimPipeline(sim)


def test_tiny_worlds_are_planned_onto_one_resident_cluster(nb, monkeypatch):
    """Host-only half of the resident-kernel decision (NbSimPlan): Worlds up to 640 particles on one GPU run chained
    steps as one launch of one cluster of at most 16 CTAs; every target row has a CTA, every target of a CTA a thread,
    every CTA enough warps for its (row, slab) units or the full 32."""
    for n in list(range(1, 700, 7)) + [639, 640, 641, 1000, 4000, 20000]:
        info, _ = nb.plan(n, n // 2, 1, 148)
        rows = -(-n // 32)
        if n > 640:
            assert info["resident_ctas"] == 0, n
            continue
        c, per, warps = info["resident_ctas"], info["resident_rows"], info["resident_warps"]
        assert 1 <= c <= 16 and c * per >= rows > (c - 1) * per, (n, c, per)
        assert warps in (8, 16, 32) and per <= warps, (n, per, warps)
        assert warps == 32 or per * info["slabs"] <= warps, (n, per, info["slabs"], warps)
        assert info["variant"] == "w32x1" and info["tile_targets"] == 32
    assert nb.plan(500, 250, 2, 148)[0]["resident_ctas"] == 0          # a World split over GPUs: launch per step
    monkeypatch.setenv("NBODY_NO_RESIDENT", "1")
    assert nb.plan(500, 250, 1, 148)[0]["resident_ctas"] == 0
    monkeypatch.delenv("NBODY_NO_RESIDENT")
    monkeypatch.setenv("NBODY_RESIDENT_MAX_N", "1536")
    assert nb.plan(1536, 700, 1, 148)[0]["resident_ctas"] == 16


def test_small_worlds_with_few_tiles_split_their_tiles_over_a_cluster(nb, monkeypatch):
    """Host-only: the T = 1 small-World plan shares a tile's slabs over a cluster of min(4, SMs / tiles) CTAs when that
    is at least 2 and the World has enough sources for the shorter pair loop to pay for the cluster barrier.  It is a
    launch detail: variant, slab length and slab count -- the summation order -- are the same with and without."""
    for n, m, want in [(250, 120, 1), (700, 337, 1), (800, 386, 1), (1000, 485, 4), (1200, 586, 3), (1600, 779, 2),
                       (2368, 1151, 2), (2400, 1182, 1), (4736, 2300, 1), (6000, 2951, 1), (20000, 9956, 1)]:
        info, _ = nb.plan(n, m, 1, 148)
        assert info["cluster_split"] == want, (n, info)
        assert info["cluster_split"] == 1 or info["cluster_split"] * info["tiles"] <= 148
        monkeypatch.setenv("NBODY_NO_SPLIT", "1")
        plain, _ = nb.plan(n, m, 1, 148)
        monkeypatch.delenv("NBODY_NO_SPLIT")
        assert plain["cluster_split"] == 1
        assert {k: v for k, v in plain.items() if k != "cluster_split"} == {k: v for k, v in info.items() if k != "cluster_split"}
