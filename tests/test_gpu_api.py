"""Section 2 of include/nbody_sim.h on a B200: asynchronous stepping, caller's stream, introspection,
the slab override, FP64 accelerations on the device."""
import numpy as np
import pytest

from _oracle import acc_of, rel_acc_error

pytestmark = pytest.mark.gpu


def _setup(nb, oracle, n=6000, gal=3):
    return oracle.partition(nb.make_galaxies(n, gal))


def test_step_async_then_synchronize_equals_blocking_update(nb, oracle):
    arr, m = _setup(nb, oracle)
    with nb.SimPipeline(len(arr), m) as a, nb.SimPipeline(len(arr), m) as b:
        a.set_data(arr)
        b.set_data(arr)
        a.update(6, 0.01)
        b.step_async(2, 0.01)
        b.step_async(4, 0.01)            # enqueued behind the first call, no host wait in between
        b.synchronize()
        assert b.last_step_ms() > 0
        assert a.get_data().tobytes() == b.get_data().tobytes()


def test_runs_on_the_callers_stream(nb, oracle):
    import torch
    arr, m = _setup(nb, oracle)
    stream = torch.cuda.Stream()
    with nb.SimPipeline(len(arr), m, stream=stream.cuda_stream) as a, nb.SimPipeline(len(arr), m) as b:
        a.set_data(arr)
        b.set_data(arr)
        a.step_async(3, 0.01)
        stream.synchronize()             # the work went to OUR stream: its sync is enough
        b.update(3, 0.01)
        assert a.get_data().tobytes() == b.get_data().tobytes()


def test_info_and_plan_agree(nb, oracle):
    arr, m = _setup(nb, oracle, 20000, 2)
    with nb.SimPipeline(len(arr), m) as sim:
        sim.set_data(arr)
        before = sim.info()["kernels_launched"]
        sim.update(5, 0.01)
        info = sim.info()
    plan, ranges = nb.plan(len(arr), m, 1, info["sm_count"])
    assert info["kernels_launched"] - before == 5          # one launch per step on one GPU
    assert (info["tile_targets"], info["slab_len"], info["slabs"]) == (plan["tile_targets"], plan["slab_len"], plan["slabs"])
    assert info["tiles_local"] == plan["tiles"] and info["pairs_per_step"] == len(arr) * m
    assert info["cc_major"] >= 10 and info["sm_count"] > 0 and "B200" in info["device_name"]


def test_slab_override_changes_rounding_only(nb, oracle):
    arr, m = _setup(nb, oracle)
    with nb.SimPipeline(len(arr), m) as a, nb.SimPipeline(len(arr), m, slab_len=512) as b:
        a.set_data(arr)
        b.set_data(arr)
        a.update(1, 0.01)
        b.update(1, 0.01)
        assert b.info()["slab_len"] == 512
        ga, gb = a.get_data(), b.get_data()
    assert ga.tobytes() != gb.tobytes()                     # a different summation order...
    assert np.percentile(rel_acc_error(acc_of(ga), acc_of(gb)), 99) < 5e-6   # ...and nothing else


def test_device_fp64_accelerations_match_oracle_fp64(nb, oracle):
    arr, m = _setup(nb, oracle)
    with nb.SimPipeline(len(arr), m) as sim:
        sim.set_data(arr)
        dev = sim.acc_f64(100, 1100)
    truth, _ = oracle.acc_f64(arr, m, 100, 1100)
    # the device folds G into the mass in FP32 first (one extra FP32 rounding per source)
    assert np.percentile(rel_acc_error(dev, truth), 99) < 2e-7
