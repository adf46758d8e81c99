"""Boids (BASELINE configs[4]): sharded uniform-grid neighbour system.

The reference has no such system (parity unpinned): the oracle's CPU statement is validated here against an
independent brute-force numpy evaluation, the result is shown to be independent of the shard count, the multi-rank host
path runs under gloo (world_size 2) on CPU, and the CUDA path must match the oracle bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch

from heliosengine_b200 import boids, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def brute_force_step(pos, vel, world_size, p):
    """O(n^2) float32 numpy statement of one step (same per-pair arithmetic, int64 sums)."""
    f = np.float32
    n = pos.shape[0] // 2
    P, V = pos.reshape(n, 2).astype(f), np.clip(vel.reshape(n, 2).astype(f), f(-4.0), f(4.0))  # saturated when binned
    newP, newV = P.copy(), V.copy()
    dt, ws, wa, wc, vmax = (f(p[k]) for k in ("dt", "w_separation", "w_alignment", "w_cohesion", "max_speed"))
    cell = np.clip(P.astype(np.int32), 0, world_size - 1)
    for i in range(n):
        d = (P - P[i]).astype(f)
        d2 = (d[:, 0] * d[:, 0]).astype(f) + (d[:, 1] * d[:, 1]).astype(f)
        near = (np.abs(cell - cell[i]).max(axis=1) <= 1) & (d2 < f(1.0))
        near[i] = False
        cnt = int(near.sum())
        v = V[i].copy()
        if cnt:
            dn, vn, d2n = d[near], V[near], d2[near]
            coh = np.rint((dn * f(1048576.0)).astype(f)).astype(np.int64).sum(axis=0)
            ali = np.rint((vn * f(1048576.0)).astype(f)).astype(np.int64).sum(axis=0)
            inv = (f(1.0) / (d2n + f(0.01)).astype(f)).astype(f)
            sep = np.rint((((-dn) * inv[:, None]).astype(f) * f(65536.0)).astype(f)).astype(np.int64).sum(axis=0)
            nn = f(cnt)
            cohf = ((coh.astype(f) / f(1048576.0)).astype(f) / nn).astype(f)
            alif = (((ali.astype(f) / f(1048576.0)).astype(f) / nn).astype(f) - v).astype(f)
            sepf = (sep.astype(f) / f(65536.0)).astype(f)
            a = (((wc * cohf).astype(f) + (wa * alif).astype(f)).astype(f) + (ws * sepf).astype(f)).astype(f)
            v = (v + (a * dt).astype(f)).astype(f)
        s2 = f(f(v[0] * v[0]) + f(v[1] * v[1]))
        if s2 > f(vmax * vmax):
            v = (v * f(vmax / np.sqrt(s2, dtype=f))).astype(f)
        q = (P[i] + (v * dt).astype(f)).astype(f)
        q = np.clip(q, f(0.0), f(f(world_size) - f(0.00390625)))
        newP[i], newV[i] = q, v
    return newP.reshape(-1), newV.reshape(-1)


def run_local(backend, n, world_size, steps, params=None, vel_scale=1.0):
    s = boids.BoidsShard(backend, n, world_size, params=params, vel_scale=vel_scale)
    for _ in range(steps):
        s.step_local()
    out = s.state()
    s.close()
    return out


def test_oracle_matches_brute_force(oracle):
    n, ws = 600, 8  # dense: ~9 boids per cell, plenty of neighbours and speed clamps
    pos, vel = boids.initial_state(n, ws, 0, n)
    p1, v1 = run_local(oracle, n, ws, 1)
    bp, bv = brute_force_step(pos, vel, ws, boids.DEFAULT_PARAMS)
    assert np.array_equal(p1, bp) and np.array_equal(v1, bv)
    assert not np.array_equal(p1, pos)


def test_oracle_saturates_fast_boids(oracle):
    """Velocities beyond +-4 (user-written; max_speed 8 keeps them alive) saturate component-wise when binned."""
    n, ws = 300, 6
    params = dict(boids.DEFAULT_PARAMS, max_speed=8.0)
    pos, vel = boids.initial_state(n, ws, 0, n, vel_scale=6.0)
    assert np.abs(vel).max() > 4.0
    p1, v1 = run_local(oracle, n, ws, 1, params=params, vel_scale=6.0)
    bp, bv = brute_force_step(pos, vel, ws, params)
    assert np.array_equal(p1, bp) and np.array_equal(v1, bv)


def test_shard_ranges():
    assert boids.shard_range(64, 4, 0) == (0, 16) and boids.shard_range(64, 4, 3) == (48, 64)
    with pytest.raises(ValueError):
        boids.shard_range(10, 4, 0)
    a = boids.initial_state(1000, 16, 0, 1000)[0]
    b = np.concatenate([boids.initial_state(1000, 16, lo, lo + 250)[0] for lo in range(0, 1000, 250)])
    assert np.array_equal(a, b)  # shards start from slices of the same global world


@pytest.mark.parametrize("n_shards", [2, 4, 8])
def test_result_is_independent_of_shard_count(oracle, n_shards):
    n, ws, steps = 4096, 24, 3
    ref = run_local(oracle, n, ws, steps)
    got = boids.emulate_shards(oracle, torch, "cpu", n, ws, n_shards, steps)
    assert np.array_equal(ref[0], got[0]) and np.array_equal(ref[1], got[1])


def _gloo_worker(rank, world_size, port, n, ws, steps, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    o = capi.load_oracle()
    shard = boids.BoidsShard(o, n, ws, world_size, rank)
    buf = boids.ExchangeBuffers(torch, "cpu", world_size, shard.n_local, shard.n_cells)
    for _ in range(steps):
        boids.distributed_step(shard, buf, dist)
    p, v = shard.state()
    np.save(os.path.join(out_dir, f"p{rank}.npy"), p)
    np.save(os.path.join(out_dir, f"v{rank}.npy"), v)
    shard.close()
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_over_gloo_match_single_rank(oracle, tmp_path):
    """The N>1 host path (bin -> all_gather_into_tensor x2 -> step) under world_size 2 on CPU."""
    import torch.multiprocessing as mp
    n, ws, steps = 2048, 16, 3
    port = 29500 + os.getpid() % 2000
    mp.spawn(_gloo_worker, args=(2, port, n, ws, steps, str(tmp_path)), nprocs=2, join=True)
    ref = run_local(oracle, n, ws, steps)
    p = np.concatenate([np.load(tmp_path / f"p{r}.npy") for r in range(2)])
    v = np.concatenate([np.load(tmp_path / f"v{r}.npy") for r in range(2)])
    assert np.array_equal(ref[0], p) and np.array_equal(ref[1], v)


@pytest.mark.gpu
@pytest.mark.parametrize("n,ws", [(1, 4), (600, 8), (4096, 24), (100000, 160), (100000, 17)])
def test_gpu_single_shard_matches_oracle(gpu, oracle, n, ws):
    steps = 3
    g, o = run_local(gpu, n, ws, steps), run_local(oracle, n, ws, steps)
    assert np.array_equal(g[0], o[0]) and np.array_equal(g[1], o[1])


@pytest.mark.gpu
def test_gpu_fast_boids_and_crowded_cells(gpu, oracle):
    """Saturating velocities, and cells crowded enough (~28 boids each, ~250 candidates) for several hit-queue rounds."""
    params = dict(boids.DEFAULT_PARAMS, max_speed=8.0)
    for n, ws, scale in [(3000, 6, 6.0), (7000, 16, 1.0), (40000, 5, 1.0)]:
        g = run_local(gpu, n, ws, 3, params=params, vel_scale=scale)
        o = run_local(oracle, n, ws, 3, params=params, vel_scale=scale)
        assert np.array_equal(g[0], o[0]) and np.array_equal(g[1], o[1])


@pytest.mark.gpu
@pytest.mark.parametrize("n_shards", [2, 4, 8])
def test_gpu_shards_match_single_shard_oracle(gpu, oracle, n_shards):
    """Shards emulated on one GPU (sequential calls over shared exchange buffers) == the oracle's single-shard world."""
    n, ws, steps = 65536, 128, 3
    ref = run_local(oracle, n, ws, steps)
    got = boids.emulate_shards(gpu, torch, "cuda:0", n, ws, n_shards, steps)
    assert np.array_equal(ref[0], got[0]) and np.array_equal(ref[1], got[1])


@pytest.mark.gpu
@pytest.mark.parametrize("n_shards", [2, 3, 8, 16])
def test_gpu_peer_memory_exchange_matches_oracle(gpu, oracle, n_shards):
    """helios_gpu_boids_exchange_*: shards read each other's buffers in place (peer pointers wired inside one process)."""
    n, ws, steps = 49152, 96, 4  # 4 steps: both buffer generations are reused
    ref = run_local(oracle, n, ws, steps)
    got = boids.emulate_shards_peer(gpu, n, ws, n_shards, steps)
    assert np.array_equal(ref[0], got[0]) and np.array_equal(ref[1], got[1])


@pytest.mark.gpu
def test_gpu_full_density_properties(gpu):
    """cfg5 density (4 boids per cell) at 4M boids: in bounds, speed-limited, deterministic across two runs."""
    n, ws = 1 << 22, 1024
    a = run_local(gpu, n, ws, 2)
    b = run_local(gpu, n, ws, 2)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])  # atomics order does not leak into the result
    assert a[0].min() >= 0 and a[0].max() < ws
    v = a[1].reshape(-1, 2).astype(np.float64)
    assert np.sqrt((v * v).sum(axis=1)).max() <= 2.0 * (1 + 1e-6)


def _nccl_peer_worker(rank, world_size, port, n, ws, steps, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world_size, device_id=dev)
    g = capi.load_gpu()
    shard = boids.BoidsShard(g, n, ws, world_size, rank, device=rank)
    stream = torch.cuda.Stream(device=dev)
    shard.world.set_stream(stream.cuda_stream)
    with torch.cuda.stream(stream):
        peer = boids.PeerExchange(shard, torch, dist, dev)
        for _ in range(steps):
            peer.step()
        p, v = shard.state()
        np.save(os.path.join(out_dir, f"p{rank}.npy"), p)
        np.save(os.path.join(out_dir, f"v{rank}.npy"), v)
        peer.close()
    shard.close()
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (CUDA IPC + NVLink peer access between processes)")
def test_two_gpus_peer_memory_exchange_matches_oracle(oracle, tmp_path):
    """One process per GPU: IPC-exported buffers, merge kernels pulling the other GPU's slab over NVLink, NCCL barrier."""
    import torch.multiprocessing as mp
    n, ws, steps = 262144, 256, 4
    port = 29600 + os.getpid() % 2000
    mp.spawn(_nccl_peer_worker, args=(2, port, n, ws, steps, str(tmp_path)), nprocs=2, join=True)
    ref = run_local(oracle, n, ws, steps)
    p = np.concatenate([np.load(tmp_path / f"p{r}.npy") for r in range(2)])
    v = np.concatenate([np.load(tmp_path / f"v{r}.npy") for r in range(2)])
    assert np.array_equal(ref[0], p) and np.array_equal(ref[1], v)
