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


@pytest.mark.parametrize("n", [6000, 20000, 50000, 300000])   # 20 000: the balanced plan
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


@pytest.mark.parametrize("gpus", [2, 8])
def test_slab_groups_bitwise_equal_across_gpu_counts(nb, oracle, monkeypatch, gpus):
    """The GROUPED plan (large Worlds; forced here with NBODY_PART_BUDGET) is a function of (N, M) only."""
    if nb.device_count() < gpus:
        pytest.skip(f"needs {gpus} GPUs")
    n = 300000
    arr, m = oracle.partition(nb.make_galaxies(n, 2))
    base, _ = nb.plan(n, m, 1)
    monkeypatch.setenv("NBODY_PART_BUDGET", str(4 * base["tiles"] * base["tile_targets"] * 8 + 8))
    one, info1 = _run(nb, arr, m, 3, 0.01, local_gpus=1)
    many, info = _run(nb, arr, m, 3, 0.01, local_gpus=gpus)
    assert info1["group_slabs"] > 1 and info["group_slabs"] == info1["group_slabs"]
    assert many.tobytes() == one.tobytes()


@pytest.mark.parametrize("gpus", [2, 4])
@pytest.mark.parametrize("n", [6000, 16384])
def test_balanced_kernel_bitwise_equal_across_gpu_counts(nb, oracle, monkeypatch, n, gpus):
    """balanced_kernel.cuh shares its (row, slab) units out per launch range; sharding changes the ranges and with
    them which rows straddle a CTA boundary, never the summation order."""
    if nb.device_count() < gpus:
        pytest.skip(f"needs {gpus} GPUs")
    from nbody_b200.api import FLAG_NO_PEER_STORE
    monkeypatch.setenv("NBODY_VARIANT", "b16x2")
    arr, m = oracle.partition(nb.make_galaxies(n, 2))
    one, info1 = _run(nb, arr, m, 4, 0.01, local_gpus=1)
    assert info1["variant"] == "b16x2"
    many, _ = _run(nb, arr, m, 4, 0.01, local_gpus=gpus)
    assert many.tobytes() == one.tobytes()
    nccl, _ = _run(nb, arr, m, 4, 0.01, local_gpus=gpus, flags=FLAG_NO_PEER_STORE)
    assert nccl.tobytes() == one.tobytes()


def test_caller_device_is_restored(nb, oracle):
    """Advisor finding (round 1): the entry points used to leave the CUDA current device changed."""
    torch = pytest.importorskip("torch")
    if nb.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    arr, m = oracle.partition(nb.make_galaxies(5000, 2))
    torch.cuda.set_device(1)
    x = torch.zeros(4, device="cuda")                       # context on device 1, current device 1
    with nb.SimPipeline(len(arr), m, device=0, local_gpus=1) as sim:
        sim.set_data(arr)
        sim.update(2, 0.01)
        sim.get_data()
        assert torch._C._cuda_getDevice() == 1
    assert torch._C._cuda_getDevice() == 1 and x.device.index == 1
    torch.cuda.set_device(0)


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


@pytest.mark.parametrize("ranks,n", [(2, 50000), (2, 300000), (4, 100000), (8, 300000)])
def test_one_rank_per_gpu_nccl_bitwise_equal_to_one_gpu(nb, ranks, n):
    """torchrun, one process per GPU, NCCL position exchange, sharded upload (each rank is handed an array whose
    foreign ranges are poisoned with NaN) and sharded read-back: tools/check_multirank.py."""
    import os
    import subprocess
    import sys
    if nb.device_count() < ranks:
        pytest.skip(f"needs {ranks} GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={ranks}",
           "--master-addr", "127.0.0.1", "--master-port", str(29540 + ranks),
           os.path.join(root, "tools", "check_multirank.py"), str(n), "3"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=root)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "bit-identical to 1 GPU" in out.stdout
