"""helios_gpu_message_route_split_static — the allocation-free variant of the cross-shard message split (scratch in a
file-scope static __device__ array; no counterpart in the reference) against
the existing entry point helios_gpu_message_route_split, through the C ABI.  Bar: bit-exact slab and counts (a stable
partition of bytes), on two calls in a row (the second finds the first one's leftovers in the shared scratch).
Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

from heliosengine_b200 import capi, routing, worlds

pytestmark = pytest.mark.gpu

EVENT = np.dtype([("a", "<u8"), ("b", "<u8"), ("impulse", "<f4"), ("pad", "<u4")])  # CollisionEvent, 24 B, key = a's index


def _events(n, n_shards, per_shard, seed):
    rng = np.random.default_rng(seed)
    ev = np.zeros(n, dtype=EVENT)
    ev["a"] = rng.integers(0, per_shard * n_shards + 3, n).astype(np.uint64) | (np.uint64(1) << np.uint64(32))
    ev["b"] = rng.integers(0, 1 << 40, n).astype(np.uint64)
    ev["pad"] = np.arange(n, dtype=np.uint32)  # makes every message unique: the comparison checks the ORDER too
    return ev


def test_static_split_matches_route_split_twice_in_a_row(gpu):
    import torch
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    w = capi.World(gpu, device=0)
    w.set_stream(stream.cuda_stream)
    w.message_register(worlds.COLLISION_EVENT, EVENT.itemsize)
    # two calls in a row; the second is smaller, with another shard count: stale histogram cells beyond it must not matter
    for call, (n, n_shards, per_shard) in enumerate([(200_003, 8, 4096), (70_001, 5, 1000), (1, 64, 7)]):
        ev = _events(n, n_shards, per_shard, seed=call)
        raw = torch.from_numpy(ev.view(np.uint8).reshape(-1).copy()).to(dev)
        # existing entry point: the world's queue -> slab + host counts
        with torch.cuda.stream(stream):
            routing.merge(w, worlds.COLLISION_EVENT, 1, raw, n, 0, 0)
            slab_ref = torch.zeros(n * EVENT.itemsize, dtype=torch.uint8, device=dev)
            counts_ref = routing.split(w, worlds.COLLISION_EVENT, 1, 0, per_shard, n_shards, slab_ref)
            # new entry point: caller-owned device buffers only
            slab = torch.zeros(n * EVENT.itemsize, dtype=torch.uint8, device=dev)
            counts = torch.full((n_shards,), -1, dtype=torch.int64, device=dev)
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(dev)[0]
            routing.split_static(gpu, stream.cuda_stream, raw, n, EVENT.itemsize, 0, per_shard, n_shards, slab, counts)
            torch.cuda.synchronize()
            if call:  # (the first launch of a kernel may load its module lazily)
                assert torch.cuda.mem_get_info(dev)[0] >= free_before, "the allocation-free entry point allocated device memory"
        assert counts.cpu().tolist() == counts_ref and sum(counts_ref) == n, call
        assert torch.equal(slab, slab_ref), call
        # and it is the stable partition it claims to be
        owner = np.minimum((ev["a"] & np.uint64(0xFFFFFFFF)) // np.uint64(per_shard), n_shards - 1)
        expect = np.concatenate([ev[owner == g] for g in range(n_shards)])
        assert np.array_equal(slab.cpu().numpy().view(EVENT), expect), call
    w.close()


def test_static_split_limits_and_empty_input(gpu):
    import torch
    dev = torch.device("cuda", 0)
    counts = torch.full((4,), -1, dtype=torch.int64, device=dev)
    slab = torch.zeros(24, dtype=torch.uint8, device=dev)
    routing.split_static(gpu, 0, slab, 0, 24, 0, 10, 4, slab, counts)  # n = 0: counts become zero, nothing else is touched
    torch.cuda.synchronize()
    assert counts.cpu().tolist() == [0, 0, 0, 0]
    too_many = routing.STATIC_SPLIT_MAX_MESSAGES + 1
    big = torch.empty(too_many * 24, dtype=torch.uint8, device=dev)
    with pytest.raises(capi.HeliosError) as ei:
        routing.split_static(gpu, 0, big, too_many, 24, 0, 10, 4, big, counts)
    assert ei.value.status == capi.ERR_CAPACITY and str(routing.STATIC_SPLIT_MAX_MESSAGES) in str(ei.value)
    with pytest.raises(capi.HeliosError) as ei:
        routing.split_static(gpu, 0, slab, 1, 24, 0, 10, 65, slab, counts)
    assert ei.value.status == capi.ERR_INVALID_ARGUMENT
