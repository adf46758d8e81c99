"""Boids over spatial bands with a one-row halo exchange (include/helios_gpu.h "Spatial bands"; BASELINE.json north_star:
"position allgather or halo exchange").  Shard g owns a band of grid rows; boids migrate between bands through the message
router and the ECS command path (Destroy + Spawn/Add).  What is checked:

  * the result per boid id equals the single-world run of the id-range design (tests/test_boids.py pins that one to an
    O(n^2) numpy statement) bit for bit, at any shard count — the fixed-point sums make it independent of ownership;
  * no boid is lost or duplicated by the migration, the first step moves ~(G-1)/G of the flock, later steps a trickle;
  * the N>1 host path under gloo (world_size 2) on CPU, the CUDA path with all bands on one device, and two processes over
    NCCL + CUDA IPC (peer loads of the neighbours' edge rows) when two devices are present."""
import os
import sys

import numpy as np
import pytest
import torch

from heliosengine_b200 import boids, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_local(backend, n, world_size, steps, params=None, vel_scale=1.0):
    s = boids.BoidsShard(backend, n, world_size, params=params, vel_scale=vel_scale)
    for _ in range(steps):
        s.step_local()
    out = s.state()
    s.close()
    return out


@pytest.mark.parametrize("n_shards", [1, 2, 3, 4, 8])
def test_oracle_bands_match_single_world(oracle, n_shards):
    n, ws, steps = 4096, 24, 4
    ref = run_local(oracle, n, ws, steps)
    pos, vel, moved = boids.emulate_bands(oracle, torch, "cpu", n, ws, n_shards, steps)
    assert np.array_equal(ref[0], pos) and np.array_equal(ref[1], vel)
    if n_shards > 1:
        assert moved[0][0] == moved[0][1] and moved[0][0] > n * (n_shards - 1) // n_shards * 0.8  # the initial sort into bands
        assert all(out == arrived for out, arrived in moved)
        assert sum(m[0] for m in moved[1:]) < n // 4  # afterwards only boids crossing a band edge


def test_oracle_bands_fast_boids_cross_many_rows(oracle):
    """dt = 1/2 and max_speed 8: boids jump up to 4 grid rows per step — more than a band (ws 12 / 4 shards = 3 rows), so
    migrations go beyond the adjacent band; the halo is still one row because every boid is re-homed before the next bin."""
    params = dict(boids.DEFAULT_PARAMS, max_speed=8.0, dt=0.5)
    n, ws, steps = 1500, 12, 4
    ref = run_local(oracle, n, ws, steps, params=params, vel_scale=6.0)
    pos, vel, moved = boids.emulate_bands(oracle, torch, "cpu", n, ws, 4, steps, params=params, vel_scale=6.0)
    assert np.array_equal(ref[0], pos) and np.array_equal(ref[1], vel)
    assert all(m[0] > 0 for m in moved)


# (boids, world size, bands): one-row bands (both edge rows are the same row), a last band of one row, bands that hold no
# boid at all, a flock smaller than the band count
SMALL_GEOMETRIES = [(3000, 8, 8), (2000, 10, 4), (5, 16, 4), (3, 4, 2), (64, 6, 3)]


@pytest.mark.parametrize("n,ws,g", SMALL_GEOMETRIES)
def test_oracle_small_geometries(oracle, n, ws, g):
    ref = run_local(oracle, n, ws, 5)
    pos, vel, moved = boids.emulate_bands(oracle, torch, "cpu", n, ws, g, 5)
    assert np.array_equal(ref[0], pos) and np.array_equal(ref[1], vel)


@pytest.mark.gpu
@pytest.mark.parametrize("n,ws,g", SMALL_GEOMETRIES)
def test_gpu_small_geometries(gpu, oracle, n, ws, g):
    ref = run_local(oracle, n, ws, 5)
    pos, vel, moved = boids.emulate_bands(gpu, torch, "cuda:0", n, ws, g, 5)
    assert np.array_equal(ref[0], pos) and np.array_equal(ref[1], vel)


def test_band_geometry_and_errors(oracle):
    s = boids.BandShard(oracle, 64, 10, 4, 3)
    assert (s.rows_per, s.row_lo, s.row_hi) == (3, 9, 10)  # R = ceil(10 / 4); the last band is what is left
    with pytest.raises(capi.HeliosError):
        s.bin()  # id-range placement: boids outside the band -> migrate first
    s.close()
    with pytest.raises(capi.HeliosError):
        boids.BandShard(oracle, 64, 4, 8, 0)  # 8 shards over 4 grid rows


def _worker(rank, world_size, port, kind, n, ws, steps, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    if kind == "nccl":
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=world_size, device_id=dev)
        shard = boids.BandShard(capi.load_gpu(), n, ws, world_size, rank, device=rank)
        stream = torch.cuda.Stream(device=dev)
        shard.world.set_stream(stream.cuda_stream)
        ctx = torch.cuda.stream(stream)
    else:
        import contextlib
        dev = "cpu"
        dist.init_process_group("gloo", rank=rank, world_size=world_size)
        shard = boids.BandShard(capi.load_oracle(), n, ws, world_size, rank)
        ctx = contextlib.nullcontext()
    with ctx:
        # a small cap: the initial sort into bands overflows the light exchange and takes the general router, later steps fit
        x = boids.BandExchange(shard, torch, dist, dev, peer=(kind == "nccl"), light_cap=256)
        for _ in range(steps):
            x.step()
        assert 1 <= x.general_routes < steps, x.general_routes
        ids, p, v = shard.state()
        x.close()
    np.save(os.path.join(out_dir, f"i{rank}.npy"), ids)
    np.save(os.path.join(out_dir, f"p{rank}.npy"), p)
    np.save(os.path.join(out_dir, f"v{rank}.npy"), v)
    np.save(os.path.join(out_dir, f"m{rank}.npy"), np.array(x.migrated, dtype=np.int64))
    shard.close()
    dist.barrier()
    dist.destroy_process_group()


def _collect(tmp_path, world_size, n):
    states = [(np.load(tmp_path / f"i{r}.npy"), np.load(tmp_path / f"p{r}.npy"), np.load(tmp_path / f"v{r}.npy")) for r in range(world_size)]
    return boids.gather_band_state(states, n)


def test_two_ranks_over_gloo_match_single_world(oracle, tmp_path):
    """The N>1 host path (leavers -> route (all_to_all) -> arrivals -> bin -> all_gather of the edge rows -> step) on CPU."""
    import torch.multiprocessing as mp
    n, ws, steps = 2048, 16, 4
    port = 30500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, "gloo", n, ws, steps, str(tmp_path)), nprocs=2, join=True)
    ref = run_local(oracle, n, ws, steps)
    pos, vel = _collect(tmp_path, 2, n)
    assert np.array_equal(ref[0], pos) and np.array_equal(ref[1], vel)
    m = [np.load(tmp_path / f"m{r}.npy") for r in range(2)]
    assert m[0][:, 0].sum() == m[1][:, 1].sum() and m[1][:, 0].sum() == m[0][:, 1].sum()  # what one rank sends the other receives


@pytest.mark.gpu
@pytest.mark.parametrize("n_shards,n,ws", [(1, 4096, 24), (2, 65536, 128), (3, 100000, 17), (4, 65536, 128), (8, 262144, 256)])
def test_gpu_bands_match_single_world_oracle(gpu, oracle, n_shards, n, ws):
    steps = 4
    ref = run_local(oracle, n, ws, steps)
    pos, vel, moved = boids.emulate_bands(gpu, torch, "cuda:0", n, ws, n_shards, steps)
    assert np.array_equal(ref[0], pos) and np.array_equal(ref[1], vel)
    assert all(out == arrived for out, arrived in moved)


@pytest.mark.gpu
def test_gpu_bands_fast_boids_and_crowded_cells(gpu, oracle):
    params = dict(boids.DEFAULT_PARAMS, max_speed=8.0, dt=0.5)
    for n, ws, g, scale in [(3000, 12, 4, 6.0), (40000, 6, 2, 1.0)]:
        ref = run_local(oracle, n, ws, 3, params=params, vel_scale=scale)
        pos, vel, _ = boids.emulate_bands(gpu, torch, "cuda:0", n, ws, g, 3, params=params, vel_scale=scale)
        assert np.array_equal(ref[0], pos) and np.array_equal(ref[1], vel)


@pytest.mark.gpu
def test_gpu_bands_full_density_vs_id_range_design(gpu):
    """cfg5 density (4 boids per cell) at 4M boids over 8 bands == the single-shard GPU run, per boid id."""
    n, ws, steps = 1 << 22, 1024, 3
    ref = run_local(gpu, n, ws, steps)
    pos, vel, moved = boids.emulate_bands(gpu, torch, "cuda:0", n, ws, 8, steps)
    assert np.array_equal(ref[0], pos) and np.array_equal(ref[1], vel)
    assert moved[0][0] > n * 7 // 8 * 0.9 and moved[-1][0] < n // 100


@pytest.mark.gpu
def test_two_gpus_bands_over_nccl_and_ipc(oracle, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    n, ws, steps = 262144, 256, 4
    port = 30900 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, "nccl", n, ws, steps, str(tmp_path)), nprocs=2, join=True)
    ref = run_local(oracle, n, ws, steps)
    pos, vel = _collect(tmp_path, 2, n)
    assert np.array_equal(ref[0], pos) and np.array_equal(ref[1], vel)
