"""Cross-shard message routing: stable split by owner shard -> all-to-all -> merge with re-based keys.

CPU: the oracle's split / merge against a numpy statement, and the whole N>1 host path under gloo (world_size 2).
GPU: the CUDA split (histogram + scan + stable scatter) against the oracle's, byte for byte, at ragged sizes; a 2-GPU NCCL
run is tests/test_routing.py::test_two_gpus_over_nccl (skipped with fewer than two devices)."""
import os
import sys

import numpy as np
import pytest
import torch

from heliosengine_b200 import capi, routing, worlds

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EVENT = worlds.COLLISION_EVENT
SIZE = 24


def make_events(n, n_global, seed):
    """n CollisionEvents {a, b: Entity; impulse f32; pad u32} with a.index uniform over the GLOBAL index space."""
    rng = np.random.default_rng(seed)
    ev = np.zeros(n, dtype=np.dtype([("a", "<u8"), ("b", "<u8"), ("impulse", "<f4"), ("pad", "<u4")]))
    ev["a"] = rng.integers(0, n_global, n).astype(np.uint64) | (np.uint64(1) << np.uint64(32))
    ev["b"] = rng.integers(0, n_global, n).astype(np.uint64) | (np.uint64(1) << np.uint64(32))
    ev["impulse"] = rng.random(n, dtype=np.float32) * 4
    ev["pad"] = np.arange(n, dtype=np.uint32)  # emission position: makes order mistakes visible
    return ev.view(np.uint8).reshape(-1)


def world_with_events(backend, events):
    w = capi.World(backend)
    w.message_register(EVENT, SIZE)
    if events.size:
        w.message_write(EVENT, events)
    return w


def split_on(backend, events, per_shard, g, device):
    w = world_with_events(backend, events)
    slab = torch.zeros(max(1, events.size), dtype=torch.uint8, device=device)
    counts = routing.split(w, EVENT, 1, 0, per_shard, g, slab)
    out = slab.cpu().numpy()[: events.size].copy()
    w.close()
    return counts, out


@pytest.mark.parametrize("n,g", [(0, 2), (1, 2), (255, 3), (256, 2), (1000, 8), (5003, 4)])
def test_oracle_split_matches_numpy(oracle, n, g):
    per = 1000
    ev = make_events(n, per * g + 500, seed=n + g)  # indices past the last range clamp to the last shard
    counts, slab = split_on(oracle, ev, per, g, "cpu")
    want = routing.reference_routing([ev] + [np.zeros(0, np.uint8)] * (g - 1), SIZE, 0, per)
    assert counts == [w.size // SIZE for w in want]
    # the slab holds destination 0's messages, then 1's ... with their GLOBAL keys (re-basing happens at the merge)
    m = ev.reshape(-1, SIZE)
    idx = m[:, 0:4].copy().view("<u4").reshape(-1)
    owner = np.minimum(idx // per, g - 1)
    expect = np.concatenate([m[owner == d] for d in range(g)]).reshape(-1) if n else np.zeros(0, np.uint8)
    assert np.array_equal(slab, expect)


def _gloo_worker(rank, world_size, port, per, n, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    o = capi.load_oracle()
    w = world_with_events(o, make_events(n + 37 * rank, per * world_size, seed=100 + rank))
    info = routing.route(w, EVENT, torch, "cpu", dist, rank, world_size, per)
    assert sum(info["sent"]) == n + 37 * rank
    np.save(os.path.join(out_dir, f"q{rank}.npy"), w.message_read(EVENT, 1))
    w.close()
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_over_gloo_route_to_owners(oracle, tmp_path):
    """The N>1 host path (split -> all_to_all sizes -> all_to_all slabs -> merge) under world_size 2 on CPU."""
    import torch.multiprocessing as mp
    per, n = 4096, 3000
    port = 31500 + os.getpid() % 2000
    mp.spawn(_gloo_worker, args=(2, port, per, n, str(tmp_path)), nprocs=2, join=True)
    want = routing.reference_routing([make_events(n + 37 * r, per * 2, seed=100 + r) for r in range(2)], SIZE, 0, per)
    for r in range(2):
        got = np.load(tmp_path / f"q{r}.npy")
        assert np.array_equal(got, want[r]), r


@pytest.mark.gpu
@pytest.mark.parametrize("n,g", [(0, 2), (1, 2), (255, 3), (257, 2), (100003, 8), (1 << 20, 8), (70001, 64)])
def test_gpu_split_matches_oracle(gpu, oracle, n, g):
    per = 1 << 14
    ev = make_events(n, per * g + 777, seed=n + g)
    assert split_on(gpu, ev, per, g, "cuda:0")[0] == split_on(oracle, ev, per, g, "cpu")[0]
    a, b = split_on(gpu, ev, per, g, "cuda:0")[1], split_on(oracle, ev, per, g, "cpu")[1]
    assert np.array_equal(a, b)


@pytest.mark.gpu
def test_gpu_single_rank_route_rebases_and_feeds_consumers(gpu, oracle):
    """One rank owning everything: route() = identity order, keys re-based by 0; then merge with a rebase and read back."""
    ev = make_events(5000, 9000, seed=5)
    for backend, dev in ((gpu, "cuda:0"), (oracle, "cpu")):
        w = world_with_events(backend, ev)
        routing.route(w, EVENT, torch, dev, None, 0, 1, 9000)
        assert np.array_equal(w.message_read(EVENT, 1), ev)
        t = torch.from_numpy(ev.copy()).to(dev)
        routing.merge(w, EVENT, 1, t, 5000, 0, 0)
        routing.merge(w, EVENT, 0, t, 1234, 8, 3)  # PreviousMessages, key = evt.b, re-based by 3
        got = w.message_read(EVENT, 0).reshape(-1, SIZE)
        want = ev.reshape(-1, SIZE)[:1234].copy()
        k = want[:, 8:12].copy().view("<u4").reshape(-1) - np.uint32(3)
        want[:, 8:12] = k.view(np.uint8).reshape(-1, 4)
        assert np.array_equal(got, want)
        w.close()


@pytest.mark.gpu
@pytest.mark.parametrize("g,n", [(1, 3000), (2, 100003), (5, 70001), (8, 0)])
def test_gpu_peer_router_matches_numpy(gpu, g, n):
    """The peer-memory form with every shard on ONE device (the peers' pointers are ordinary device pointers): count on
    all shards, scatter on all shards, deliver on all shards — two steps, so both buffer generations are used."""
    per = 1 << 14
    ws = [capi.World(gpu) for _ in range(g)]
    for w in ws:
        w.message_register(EVENT, SIZE)
    routers = [routing.Router(w, EVENT, g, r, per, inbox_capacity=max(1, n) * g + 64) for r, w in enumerate(ws)]
    for gen in range(2):
        local = [rt.local(gen) for rt in routers]
        for rt in routers:
            for s2 in range(g):
                rt.set_peer(s2, gen, *local[s2])
    for step in range(3):
        queues = [make_events(n + 11 * r, per * g + 99, seed=1000 * step + r) for r in range(g)]
        for w, q in zip(ws, queues):
            t = torch.from_numpy(q.copy()).to("cuda:0")
            routing.merge(w, EVENT, 1, t, q.size // SIZE, 0, 0)
        for rt in routers:
            rt.count(1)
        for w in ws:
            w.synchronize()  # stands in for the barrier
        for rt in routers:
            rt.scatter()
        for w in ws:
            w.synchronize()
        want = routing.reference_routing(queues, SIZE, 0, per)
        for r, (w, rt) in enumerate(zip(ws, routers)):
            assert rt.deliver(1, True) == want[r].size // SIZE
            assert np.array_equal(w.message_read(EVENT, 1), want[r]), (step, r)
    for rt in routers:
        rt.close()
    for w in ws:
        w.close()


def _peer_worker(rank, world_size, port, per, n, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world_size, device_id=dev)
    g = capi.load_gpu()
    w = capi.World(g, device=rank)
    w.set_stream(torch.cuda.current_stream(dev).cuda_stream)  # the all-reduce barriers and the kernels share one stream
    w.message_register(EVENT, SIZE)
    pr = routing.PeerRouter(w, EVENT, torch, dist, dev, rank, world_size, per, inbox_capacity=2 * n + 4096)
    for step in range(3):
        q = make_events(n + 37 * rank, per * world_size, seed=100 * (step + 1) + rank)
        routing.merge(w, EVENT, 1, torch.from_numpy(q.copy()).to(dev), q.size // SIZE, 0, 0)
        pr.route()
        np.save(os.path.join(out_dir, f"p{step}_{rank}.npy"), w.message_read(EVENT, 1))
    pr.close()
    w.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_two_gpus_over_peer_memory(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    per, n = 1 << 18, 400000
    port = 35500 + os.getpid() % 2000
    mp.spawn(_peer_worker, args=(2, port, per, n, str(tmp_path)), nprocs=2, join=True)
    for step in range(3):
        want = routing.reference_routing([make_events(n + 37 * r, per * 2, seed=100 * (step + 1) + r) for r in range(2)], SIZE, 0, per)
        for r in range(2):
            assert np.array_equal(np.load(tmp_path / f"p{step}_{r}.npy"), want[r]), (step, r)


def _nccl_worker(rank, world_size, port, per, n, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world_size, device_id=dev)
    g = capi.load_gpu()
    w = capi.World(g, device=rank)
    w.message_register(EVENT, SIZE)
    w.message_write(EVENT, make_events(n + 37 * rank, per * world_size, seed=100 + rank))
    routing.route(w, EVENT, torch, dev, dist, rank, world_size, per)
    np.save(os.path.join(out_dir, f"q{rank}.npy"), w.message_read(EVENT, 1))
    w.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_two_gpus_over_nccl(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    per, n = 1 << 18, 400000
    port = 33500 + os.getpid() % 2000
    mp.spawn(_nccl_worker, args=(2, port, per, n, str(tmp_path)), nprocs=2, join=True)
    want = routing.reference_routing([make_events(n + 37 * r, per * 2, seed=100 + r) for r in range(2)], SIZE, 0, per)
    for r in range(2):
        assert np.array_equal(np.load(tmp_path / f"q{r}.npy"), want[r]), r
