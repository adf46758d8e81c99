"""Cross-shard message routing over torch.distributed (SURVEY.md §8e "Messages").

Entities are partitioned by global entity-index range, `entities_per_shard` per rank.  A message is keyed by the Entity at
`key_offset`; `route()` delivers every message of a queue to the rank that owns its key:

    split      stable G-way partition on the device (helios_gpu_message_route_split, csrc/routing.cu)
    sizes      all_to_all_single of G int64 counts
    slabs      all_to_all_single with split sizes — NCCL over NVLink on the GPU path, gloo in the CPU tests
    merge      the queue becomes what arrived, keys re-based to local indices (helios_gpu_message_route_merge)

After the call rank g's queue holds exactly the messages it owns, ordered by (source rank, position in the source queue).
The same function drives the oracle backend on CPU tensors (tests, gloo) and the CUDA backend on device tensors."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi


def split(world: capi.World, name: str, which: int, key_offset: int, entities_per_shard: int, n_shards: int, slab) -> list[int]:
    """slab: a uint8 torch tensor (device memory for the CUDA backend) with room for every queued message."""
    size = world.sizes["msg:" + name]
    counts = (C.c_uint64 * n_shards)()
    world._check(world.b.fn("message_route_split")(world.h, C.c_uint64(capi.fnv1a(name)), C.c_uint32(which), C.c_uint32(key_offset),
                                                   C.c_uint64(entities_per_shard), C.c_uint32(n_shards), C.c_void_p(slab.data_ptr()),
                                                   C.c_uint64(slab.numel() // size), counts))
    return [int(c) for c in counts]


def merge(world: capi.World, name: str, which: int, src, n: int, key_offset: int, index_rebase: int):
    world._check(world.b.fn("message_route_merge")(world.h, C.c_uint64(capi.fnv1a(name)), C.c_uint32(which),
                                                   C.c_void_p(src.data_ptr()) if n else None, C.c_uint64(n), C.c_uint32(key_offset),
                                                   C.c_uint64(index_rebase)))


def route(world: capi.World, name: str, torch, device, dist, rank: int, n_shards: int, entities_per_shard: int,
          key_offset: int = 0, which: int = 1, rebase: bool = True) -> dict:
    """Routes queue `which` (1 = current) of message type `name`; returns the send / receive counts per rank."""
    size = world.sizes["msg:" + name]
    queued = world.message_counts(name)[which]
    send = torch.empty(max(1, queued) * size, dtype=torch.uint8, device=device)
    send_counts = split(world, name, which, key_offset, entities_per_shard, n_shards, send)
    sc = torch.tensor(send_counts, dtype=torch.int64, device=device)
    rc = torch.empty_like(sc)
    if dist is not None and n_shards > 1:
        dist.all_to_all_single(rc, sc)
    else:
        rc.copy_(sc)
    recv_counts = [int(x) for x in rc.cpu().tolist()]
    total = sum(recv_counts)
    recv = torch.empty(max(1, total) * size, dtype=torch.uint8, device=device)
    if dist is not None and n_shards > 1:
        dist.all_to_all_single(recv[: total * size], send[: queued * size], [c * size for c in recv_counts],
                               [c * size for c in send_counts])
    else:
        recv[: total * size].copy_(send[: queued * size])
    if device != "cpu" and str(device) != "cpu":
        torch.cuda.current_stream(device).synchronize()  # the world's stream reads what the collective wrote
    merge(world, name, which, recv, total, key_offset, rank * entities_per_shard if rebase else 0)
    return {"sent": send_counts, "received": recv_counts, "bytes_sent": queued * size, "bytes_received": total * size}


def reference_routing(queues: list[np.ndarray], size: int, key_offset: int, entities_per_shard: int) -> list[np.ndarray]:
    """Plain numpy statement used by the tests: queues[r] = rank r's messages (bytes); returns what every rank must hold."""
    g = len(queues)
    out = []
    for dst in range(g):
        parts = []
        for src in range(g):
            m = np.ascontiguousarray(queues[src]).reshape(-1, size)
            idx = m[:, key_offset:key_offset + 4].copy().view("<u4").reshape(-1)
            owner = np.minimum(idx // entities_per_shard, g - 1)
            sel = m[owner == dst].copy()
            if sel.size:
                k = sel[:, key_offset:key_offset + 4].copy().view("<u4").reshape(-1) - np.uint32(dst * entities_per_shard)
                sel[:, key_offset:key_offset + 4] = k.astype("<u4").view(np.uint8).reshape(-1, 4)
            parts.append(sel)
        out.append(np.concatenate(parts).reshape(-1) if parts else np.zeros(0, np.uint8))
    return out
