"""Pin the oracle (oracle/nbody_oracle.c) before anything is checked against it.

1. against fixtures produced by running the reference (tests/golden, tools/make_golden.py);
2. against the compiled reference itself where oracle/_ref is present;
3. against closed-form cases (two bodies, massless target, self term).
"""
import numpy as np
import pytest

from _oracle import PARTICLE, RefLib, acc_of, mass_len_of, rel_acc_error

GOLDEN = ["galaxy_1000_2.npz", "galaxy_4003_2.npz", "galaxy_6000_3.npz"]


@pytest.mark.parametrize("name", GOLDEN)
def test_partition_matches_reference_fixture(oracle, golden, name):
    g = golden(name)
    mine, m = oracle.partition(g["input"])
    assert mine.tobytes() == g["world0"].tobytes()
    assert m == mass_len_of(g["world0"])


@pytest.mark.parametrize("name", GOLDEN)
@pytest.mark.parametrize("key,lanes,steps", [("step1", 8, 1), ("step10", 8, 10), ("step1_sse", 4, 1),
                                              ("step1_scalar", 1, 1)])
def test_oracle_bit_exact_against_reference_fixture(oracle, golden, name, key, lanes, steps):
    g = golden(name)
    w0 = g["world0"]
    out = oracle.update(w0, mass_len_of(w0), float(g["dt"]), steps, lanes)
    assert out.tobytes() == g[key].tobytes()


@pytest.mark.parametrize("flavour,lanes", [("avx", 8), ("sse", 4), ("scalar", 1)])
@pytest.mark.parametrize("n,gal", [(250, 2), (1237, 2), (3000, 3)])
def test_oracle_bit_exact_against_compiled_reference(oracle, flavour, lanes, n, gal):
    if not RefLib.available(flavour):
        pytest.skip("oracle/_ref not built")
    r = RefLib(flavour)
    ps = r.make_galaxies(n, gal, seed=4242)
    w0 = r.world_state(ps)
    ref = r.update_cpu(ps, 0.02, 2)
    out = oracle.update(w0, mass_len_of(w0), 0.02, 2, lanes)
    assert out.tobytes() == ref.tobytes()


def test_range_step_matches_full_step(oracle, golden):
    g = golden("galaxy_1000_2.npz")
    w0 = g["world0"]
    m = mass_len_of(w0)
    part = oracle.step_range(w0, m, 100, 300, 0.01)
    full = g["step1"]
    assert part[100:300].tobytes() == full[100:300].tobytes()
    assert part[:100].tobytes() == w0[:100].tobytes()
    assert part[300:].tobytes() == w0[300:].tobytes()


def _two_body(d=100.0, m0=5.0, m1=7.0, r0=2.0, r1=3.0):
    ps = np.zeros(2, dtype=PARTICLE)
    ps["px"] = [0.0, d]
    ps["mass"] = [m0, m1]
    ps["radius"] = [r0, r1]
    return ps


def test_two_body_closed_form(oracle):
    d, m0, m1, r0, r1 = 100.0, 5.0, 7.0, 2.0, 3.0
    out = oracle.update(_two_body(d, m0, m1, r0, r1), 2, 0.5, 1, 8)
    a0 = 10.0 * m1 * d / (d * d + r0) ** 1.5   # softening = TARGET radius, added linearly
    a1 = -10.0 * m0 * d / (d * d + r1) ** 1.5
    assert out["ax"][0] == pytest.approx(a0, rel=1e-6)
    assert out["ax"][1] == pytest.approx(a1, rel=1e-6)
    assert out["ay"][0] == 0.0 and out["ay"][1] == 0.0
    # semi-implicit Euler: v += a dt, then x += v_new dt
    assert out["vx"][0] == pytest.approx(a0 * 0.5, rel=1e-6)
    assert out["px"][0] == pytest.approx(a0 * 0.25, rel=1e-6)


def test_massless_target_feels_but_does_not_pull(oracle):
    ps = np.zeros(3, dtype=PARTICLE)
    ps["px"] = [0.0, 50.0, -30.0]
    ps["mass"] = [100.0, 0.0, 0.0]
    ps["radius"] = [1.0, 0.5, 0.5]
    arr, m = oracle.partition(ps)
    assert m == 1
    out = oracle.update(arr, m, 0.1, 1, 8)
    assert out["ax"][0] == 0.0            # self term is exactly zero, nobody else has mass
    assert out["ax"][1] < 0 and out["ax"][2] > 0


def test_fp64_truth_close_to_fp32_oracle(oracle, golden):
    g = golden("galaxy_1000_2.npz")
    w0 = g["world0"]
    m = mass_len_of(w0)
    truth, _ = oracle.acc_f64(w0, m)
    err = rel_acc_error(acc_of(g["step1"]), truth)
    assert np.median(err) < 1e-6 and err.max() < 1e-4


def test_diagnostics_two_body(oracle):
    ps = _two_body(100.0, 5.0, 7.0, 2.0, 3.0)
    ps["vx"] = [1.0, -2.0]
    d = oracle.diagnostics(ps, 2)
    assert d["P"][0] == pytest.approx(5.0 - 14.0)
    assert d["K"] == pytest.approx(0.5 * 5.0 * 1.0 + 0.5 * 7.0 * 4.0)
    u = -0.5 * 10.0 * 5.0 * 7.0 * (1 / np.sqrt(1e4 + 2.0) + 1 / np.sqrt(1e4 + 3.0))
    assert d["U"] == pytest.approx(u)
