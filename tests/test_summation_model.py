"""CPU model of the CUDA kernel's summation order (no GPU needed).

The plan (NbSimPlan: tile size, slab length) fixes the order in which the step kernel adds the pair terms of a target:
inside a slab, even-numbered sources go to one FP32 accumulator and odd-numbered ones to another, in source order
(FMA); the slab partial is their sum; slab partials are added in slab order (step_kernel.cuh, force_kernel.cuh).
Here that order is replayed in NumPy FP32 (exact 1/sqrt instead of MUFU.RSQ, which changes single terms by ~1 ulp but
not the growth of the rounding error with the chain length) on the reference's golden fixtures, so that a planner
change that lengthens the chains shows up as lost accuracy on the build machine, before any GPU run.
"""
import numpy as np
import pytest

from _oracle import acc_of, mass_len_of, rel_acc_error

F32 = np.float32


def _fma(a, b, c):
    # exact for FP32 inputs: the product is exact in FP64, one rounding to FP64 and one to FP32
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(F32)


def replay_step_kernel_acc(w0, m, slab_len, nslab):
    n = len(w0)
    x, y, rho = w0["px"], w0["py"], w0["radius"]
    gm = (w0["mass"][:m] * F32(10.0)).astype(F32)          # NB_G folded once per source (layout_kernels.cuh)
    total = None
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
        total = part if total is None else (total + part).astype(F32)
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
