"""NbStepNoAlloc (include/nbody_sim.h, nbody_b200/csrc/noalloc.cu): one step of the hot path that allocates nothing --
its scratch is a static __device__ array -- checked through the C ABI against the existing entry point
(SetSimulationData / PerformSimUpdate / GetSimulationData), within the tolerance the parity tests use, on calls in a
row (every call shares the scratch array, which keeps the last call's contents).

Collected last on purpose (the file name).  Passes alone in a fresh process and after
the other tests.  The caller's buffers are torch tensors here -- the test may allocate, the entry point may not.
"""
import ctypes as C

import numpy as np
import pytest

from _oracle import PARTICLE, acc_of, rel_acc_error

pytestmark = pytest.mark.gpu

TOL_ACC = 1e-5        # per-particle acceleration, 99th percentile (tests/test_gpu_parity.py)
TOL_ACC_WORST = 5e-5  # worst single particle
MAX_PARTICLES = 16384  # NB_NOALLOC_MAX_PARTICLES


def _torch():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


def _to_device(torch, arr):
    host = torch.from_numpy(np.ascontiguousarray(arr).view(np.uint8).copy())
    return host.to("cuda")


def _to_host(torch, dev, n):
    torch.cuda.synchronize()
    return dev.cpu().numpy().view(PARTICLE)[:n].copy()


def _step_no_alloc(nb, torch, d_in, d_out, n, m, dt):
    stream = torch.cuda.current_stream().cuda_stream
    rc = nb.api.lib().NbStepNoAlloc(C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), n, m, dt, C.c_void_p(stream))
    assert rc == 0
    return rc


def _compare(got, ref, tag):
    err = rel_acc_error(acc_of(got), acc_of(ref))
    p99, worst = float(np.percentile(err, 99)), float(err.max())
    print(f"\n[no-alloc] {tag}: acc rel err vs the existing entry point p99 {p99:.2e} max {worst:.2e}")
    assert p99 <= TOL_ACC and worst <= TOL_ACC_WORST, (tag, p99, worst)
    scale = max(float(np.abs(ref["px"]).max()), float(np.abs(ref["py"]).max()))
    for k in ("px", "py"):
        assert np.abs(got[k] - ref[k]).max() <= 1e-6 * scale, (tag, k)
    vs = float(np.hypot(ref["vx"], ref["vy"]).max())
    for k in ("vx", "vy"):
        assert np.abs(got[k] - ref[k]).max() <= 1e-5 * vs, (tag, k)
    assert np.array_equal(got["mass"], ref["mass"]) and np.array_equal(got["radius"], ref["radius"])


@pytest.mark.parametrize("n,gal", [(6000, 3), (MAX_PARTICLES, 2), (1000, 2)])
def test_no_alloc_step_matches_the_existing_entry_point_on_two_calls_in_a_row(nb, oracle, n, gal):
    torch = _torch()
    torch.cuda.set_device(0)
    arr, m = oracle.partition(nb.make_galaxies(n, gal))
    dt = 0.01
    # the existing entry point: one step, then one more
    with nb.SimPipeline(n, m) as sim:
        sim.set_data(arr)
        sim.update(1, dt)
        ref1 = sim.get_data()
        sim.update(1, dt)
        ref2 = sim.get_data()
    # the entry point that allocates nothing: two calls in a row, the second on the first one's output
    d0 = _to_device(torch, arr)
    d1 = torch.empty_like(d0)
    d2 = torch.empty_like(d0)
    before = torch.cuda.memory_allocated()
    _step_no_alloc(nb, torch, d0, d1, n, m, dt)
    _step_no_alloc(nb, torch, d1, d2, n, m, dt)
    got1, got2 = _to_host(torch, d1, n), _to_host(torch, d2, n)
    assert torch.cuda.memory_allocated() == before
    _compare(got1, ref1, f"N={n} call 1")
    _compare(got2, ref2, f"N={n} call 2")
    # the input was not touched
    assert _to_host(torch, d0, n).tobytes() == arr.tobytes()


def test_no_alloc_calls_of_different_sizes_share_the_scratch_array(nb, oracle):
    """A large call, then a small one with fewer slabs and tiles, then the large one again: stale partials and
    tickets of an earlier call must not show, and a repeated call gives the same bits."""
    torch = _torch()
    torch.cuda.set_device(0)
    big, mb = oracle.partition(nb.make_galaxies(12000, 2))
    small, ms = oracle.partition(nb.make_galaxies(700, 2))
    outs = []
    for arr, m in ((big, mb), (small, ms), (big, mb)):
        d_in = _to_device(torch, arr)
        d_out = torch.empty_like(d_in)
        _step_no_alloc(nb, torch, d_in, d_out, len(arr), m, 0.01)
        outs.append(_to_host(torch, d_out, len(arr)))
    assert outs[0].tobytes() == outs[2].tobytes()
    for arr, m, got in ((big, mb, outs[0]), (small, ms, outs[1])):
        with nb.SimPipeline(len(arr), m) as sim:
            sim.set_data(arr)
            sim.update(1, 0.01)
            ref = sim.get_data()
        _compare(got, ref, f"N={len(arr)} after a call of another size")


def test_no_alloc_rejects_what_does_not_fit_and_accepts_the_edges(nb, oracle, capfd):
    torch = _torch()
    torch.cuda.set_device(0)
    lib = nb.api.lib()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    buf = torch.zeros((MAX_PARTICLES + 8) * 32, dtype=torch.uint8, device="cuda")
    out = torch.zeros_like(buf)
    a, b = C.c_void_p(buf.data_ptr()), C.c_void_p(out.data_ptr())
    assert lib.NbStepNoAlloc(a, b, MAX_PARTICLES + 1, 10, 0.01, stream) == 1      # beyond the static scratch array
    assert "NB_NOALLOC_MAX_PARTICLES" in capfd.readouterr().err
    assert lib.NbStepNoAlloc(a, b, 100, 101, 0.01, stream) == 1                    # mass_len > total_len
    assert lib.NbStepNoAlloc(a, a, 100, 50, 0.01, stream) == 1                     # in place
    assert lib.NbStepNoAlloc(a, b, 0, 0, 0.01, stream) == 0                        # empty: nothing to do
    torch.cuda.synchronize()
    assert not out.any()
    # no massive particles: accelerations exactly zero, particles drift with their velocity
    ps = np.zeros(300, dtype=PARTICLE)
    rng = np.random.default_rng(5)
    ps["px"], ps["py"] = rng.normal(0, 100, 300), rng.normal(0, 100, 300)
    ps["vx"], ps["radius"] = rng.normal(0, 1, 300), 1.0
    d_in = _to_device(torch, ps)
    d_out = torch.empty_like(d_in)
    assert lib.NbStepNoAlloc(C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), 300, 0, 0.5, stream) == 0
    got = _to_host(torch, d_out, 300)
    assert not got["ax"].any() and not got["ay"].any()
    assert np.array_equal(got["px"], (ps["px"] + ps["vx"] * np.float32(0.5)).astype(np.float32))
