"""Reference-exact verification mode (exact_kernel.cuh): with NBODY_EXACT=avx|sse|scalar (or NB_SIM_FLAG_EXACT) the
CUDA path reproduces the reference's CPU update of that SIMD build BIT FOR BIT -- IEEE sqrt and division, no FMA, one
accumulator per SIMD lane in the reference's lane order, zero-padded tail pack, lanes added left to right.

This is the strongest parity statement the path admits: the golden fixtures were produced by running the reference
itself (tools/make_golden.py), the oracle is pinned to them, and here the GPU matches both to the last bit, after
one step and after ten chained ones.  The literal north_star tolerance (every particle within 1e-5 of the reference)
is met with an error of exactly zero in this mode, at every size -- at roughly a quarter of the fast kernels' speed.
"""
import numpy as np
import pytest

from _oracle import PARTICLE, acc_of, rel_acc_error

pytestmark = pytest.mark.gpu

BUILDS = {"avx": 8, "sse": 4, "scalar": 1}


def _same_bits(a, b):
    """Bitwise equality of two particle arrays (NaNs included)."""
    return a.tobytes() == b.tobytes()


@pytest.mark.parametrize("name", ["galaxy_1000_2.npz", "galaxy_4003_2.npz", "galaxy_6000_3.npz"])
def test_exact_mode_reproduces_the_reference_fixtures_bit_for_bit(nb, golden, monkeypatch, name):
    g = golden(name)
    dt = float(g["dt"])
    monkeypatch.setenv("NBODY_EXACT", "avx")
    with nb.World(g["input"]) as w:
        w.update_gpu(dt, 1)
        one = w.particles()
        assert w.pipeline_info()["variant"] == "ref8"
        w.update_gpu(dt, 9)                      # chained: 9 launches, ping-pong source snapshots
        ten = w.particles()
    assert _same_bits(one, g["step1"]), "one step differs from the reference's own UpdateWorld_CPU output"
    assert _same_bits(ten, g["step10"]), "ten steps differ from the reference's own UpdateWorld_CPU output"
    for build in ("sse", "scalar"):
        monkeypatch.setenv("NBODY_EXACT", build)
        with nb.World(g["input"]) as w:
            w.update_gpu(dt, 1)
            assert _same_bits(w.particles(), g["step1_" + build]), build


@pytest.mark.parametrize("build", ["avx", "sse", "scalar"])
@pytest.mark.parametrize("n,gal", [(16384, 2), (20000, 2), (50000, 2)])
def test_exact_mode_equals_the_oracle_bit_for_bit(nb, oracle, monkeypatch, n, gal, build):
    w0, m = oracle.partition(nb.make_galaxies(n, gal))
    ref = oracle.update(w0, m, 0.01, 2, BUILDS[build])
    monkeypatch.setenv("NBODY_EXACT", build)
    with nb.World(w0) as w:
        w.update_gpu(0.01, 2)
        got = w.particles()
    assert _same_bits(got, ref)
    # the literal tolerance of north_star, met with room to spare: the error is exactly zero
    assert rel_acc_error(acc_of(got), acc_of(ref)).max() == 0.0


@pytest.mark.parametrize("n", [1, 2, 7, 8, 9, 255, 256, 257, 511, 512, 513, 1025])
def test_exact_mode_ragged_sizes_and_tail_packs(nb, oracle, n):
    """Tail packs (mass_len not a multiple of 8), a last CTA that is not full, massless tails -- through the flag
    instead of the environment."""
    from nbody_b200.api import FLAG_EXACT
    rng = np.random.default_rng(7000 + n)
    ps = np.zeros(n, dtype=PARTICLE)
    ps["px"] = rng.normal(0, 1000, n)
    ps["py"] = rng.normal(0, 1000, n)
    ps["vx"] = rng.normal(0, 1, n)
    ps["mass"] = np.where(rng.random(n) < 0.6, rng.uniform(10, 1000, n), 0.0)
    ps["radius"] = rng.uniform(0.5, 9.5, n)
    w0, m = oracle.partition(ps)
    ref = oracle.update(w0, m, 0.01, 3, 8)
    with nb.SimPipeline(n, m, flags=FLAG_EXACT) as sim:
        sim.set_data(w0)
        sim.update(3, 0.01)
        got = sim.get_data()
        assert sim.info()["variant"] == "ref8"
    assert _same_bits(got, ref)


def test_exact_mode_on_several_gpus(nb, oracle):
    """Target sharding and the position exchange do not touch the arithmetic: still the reference's bits."""
    from nbody_b200.api import FLAG_EXACT, FLAG_NO_PEER_STORE
    gpus = min(nb.device_count(), 4)
    if gpus < 2:
        pytest.skip("needs 2 GPUs")
    w0, m = oracle.partition(nb.make_galaxies(20000, 2))
    ref = oracle.update(w0, m, 0.01, 3, 8)
    for flags in (FLAG_EXACT, FLAG_EXACT | FLAG_NO_PEER_STORE):
        with nb.SimPipeline(len(w0), m, local_gpus=gpus, flags=flags) as sim:
            sim.set_data(w0)
            sim.update(3, 0.01)
            assert _same_bits(sim.get_data(), ref)
