"""Multi-GPU parity (needs >= 2 B200s; skipped otherwise): sharding targets over GPUs and
re-replicating the massive positions each step must not change a single bit of the result,
whichever exchange mechanism carries the positions."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(nb, arr, m, steps, dt, **kw):
    with nb.SimPipeline(len(arr), m, **kw) as sim:
        sim.set_data(arr)
        sim.update(steps, dt)
        out = sim.get_data()
        info = sim.info()
    return out, info


@pytest.mark.parametrize("n", [6000, 50000, 300000])
@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_single_process_multi_gpu_bitwise_equal_to_one_gpu(nb, oracle, n, gpus):
    if nb.device_count() < gpus:
        pytest.skip(f"needs {gpus} GPUs")
    from nbody_b200.api import FLAG_NO_OVERLAP, FLAG_NO_PEER_STORE
    arr, m = oracle.partition(nb.make_galaxies(n, 2))
    one, _ = _run(nb, arr, m, 5, 0.01, local_gpus=1)
    peer, info = _run(nb, arr, m, 5, 0.01, local_gpus=gpus)                       # peer stores from the epilogue
    assert info["local_gpus"] == gpus
    assert peer.tobytes() == one.tobytes()
    nccl, _ = _run(nb, arr, m, 5, 0.01, local_gpus=gpus, flags=FLAG_NO_PEER_STORE)  # NCCL, overlapped
    assert nccl.tobytes() == one.tobytes()
    plain, _ = _run(nb, arr, m, 5, 0.01, local_gpus=gpus, flags=FLAG_NO_PEER_STORE | FLAG_NO_OVERLAP)
    assert plain.tobytes() == one.tobytes()


def test_world_api_with_nbody_ngpus_env(nb, oracle, monkeypatch):
    if nb.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ps = nb.make_galaxies(20000, 2)
    with nb.World(ps) as w:
        w.update_gpu(0.01, 3)
        one = w.particles()
    monkeypatch.setenv("NBODY_NGPUS", "2")
    with nb.World(ps) as w:
        w.update_gpu(0.01, 3)
        two = w.particles()
        assert w.pipeline_info()["local_gpus"] == 2
    assert two.tobytes() == one.tobytes()
