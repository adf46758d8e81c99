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



STATIC_SPLIT_MAX_MESSAGES = 262144  # HELIOS_GPU_STATIC_SPLIT_MAX_MESSAGES


def split_static(backend: capi.Backend, stream_ptr: int, msgs, n: int, msg_size: int, key_offset: int, entities_per_shard: int,
                 n_shards: int, slab, counts):
    """helios_gpu_message_route_split_static: the allocation-free split.  msgs / slab: uint8 device tensors, counts: a uint64
    (int64) device tensor of n_shards elements; asynchronous on `stream_ptr`.  Raises HeliosError with the library's message."""
    st = backend.fn("message_route_split_static")(C.c_void_p(stream_ptr), C.c_void_p(msgs.data_ptr()) if n else None, C.c_uint64(n),
                                                  C.c_uint32(msg_size), C.c_uint32(key_offset), C.c_uint64(entities_per_shard),
                                                  C.c_uint32(n_shards), C.c_void_p(slab.data_ptr()), C.c_uint64(slab.numel() // msg_size),
                                                  C.c_void_p(counts.data_ptr()))
    if st != capi.OK:
        msg = backend.fn("static_split_last_error", C.c_char_p)()
        raise capi.HeliosError(st, msg.decode() if msg else "")

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


ROUTED_COMMAND = "helios::gpu::RoutedCommand"  # the 64-byte message type commands travel as
ROUTED_KEY_OFFSET = 8                          # offsetof(helios_gpu_routed_command, entity)


def route_commands(world: capi.World, cmds: np.ndarray, payload, n_spawns: int, torch, device, dist, rank: int, n_shards: int,
                   entities_per_shard: int) -> dict:
    """Sharded ApplyDeferred (include/helios_gpu.h "sharded deferred Commands"): this rank's command queue `cmds` / `payload`
    (targets are GLOBAL entity ids, or capi.spawn_placeholder(j) for the rank's j-th Commands::Spawn, j < n_spawns) is
    delivered to the owners of the targets and applied there in (source rank, position) order.

        ids      all_gather of the spawn counts; the last shard (open-ended range) publishes the global next index and reserves
                 the ids of every rank's spawns — the reference's single next_index_ sequence in (rank, emission) order
        pack     helios_gpu_commands_route_pack -> 64-byte records -> the ROUTED_COMMAND queue of this world
        route    split / all_to_all / merge of that queue (route(), key = the entity)
        apply    helios_gpu_commands_apply_routed: reserved-entity flush + batch apply on the device"""
    if "msg:" + ROUTED_COMMAND not in world.sizes:
        world.message_register(ROUTED_COMMAND, capi.ROUTED_COMMAND_DTYPE.itemsize, capi.MSG_MANUAL)
    last = n_shards - 1
    mine = torch.tensor([n_spawns, (last * entities_per_shard + world.next_entity_index()) if rank == last else 0],
                        dtype=torch.int64, device=device)
    if dist is not None and n_shards > 1:
        everyone = torch.empty(2 * n_shards, dtype=torch.int64, device=device)
        dist.all_gather_into_tensor(everyone, mine)
        everyone = everyone.cpu().view(n_shards, 2)
    else:
        everyone = mine.cpu().view(1, 2)
    spawns = [int(x) for x in everyone[:, 0].tolist()]
    first = int(everyone[last, 1]) + sum(spawns[:rank])
    total_spawns = sum(spawns)
    if first + n_spawns > 0xFFFFFFFF:
        raise capi.HeliosError(capi.ERR_CAPACITY, "entity indices exhausted")
    records = world.commands_route_pack(cmds, payload, first)
    if records.size:
        world.message_write(ROUTED_COMMAND, records)
    if rank == last and total_spawns:
        world.reserve_entities(total_spawns)  # fresh indices [next, next + total): what the records name after the re-base
    info = route(world, ROUTED_COMMAND, torch, device, dist, rank, n_shards, entities_per_shard, ROUTED_KEY_OFFSET, 1, True)
    world.commands_apply_routed(ROUTED_COMMAND, 1)
    info["spawn_first_index"] = first
    info["total_spawns"] = total_spawns
    return info


def reference_command_routing(queues: list[np.ndarray], entities_per_shard: int) -> list[np.ndarray]:
    """Plain numpy statement used by the tests: queues[r] = rank r's routed records (GLOBAL ids); returns, per shard, the
    records it must apply — (source rank, position) order, ids re-based to the shard's local indices."""
    size = capi.ROUTED_COMMAND_DTYPE.itemsize
    out = reference_routing([np.ascontiguousarray(q).view(np.uint8).reshape(-1) for q in queues], size, ROUTED_KEY_OFFSET,
                            entities_per_shard)
    return [o.view(capi.ROUTED_COMMAND_DTYPE) for o in out]


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


class Router:
    """helios_gpu_router: the peer-memory form (csrc/routing.cu).  count -> barrier -> scatter -> barrier -> deliver."""

    def __init__(self, world: capi.World, name: str, n_shards: int, my_shard: int, entities_per_shard: int,
                 inbox_capacity: int, key_offset: int = 0):
        self.w, self.name, self.n_shards, self.me = world, name, n_shards, my_shard
        self.h = C.c_void_p()
        world._check(world.b.fn("router_create")(world.h, C.c_uint64(capi.fnv1a(name)), C.c_uint32(key_offset),
                                                 C.c_uint64(entities_per_shard), C.c_uint32(n_shards), C.c_uint32(my_shard),
                                                 C.c_uint64(inbox_capacity), C.byref(self.h)))

    def local(self, generation: int) -> tuple[int, int]:
        inbox, matrix = C.c_void_p(), C.c_void_p()
        self.w._check(self.w.b.fn("router_local")(self.h, C.c_uint32(generation), C.byref(inbox), C.byref(matrix)))
        return inbox.value, matrix.value

    def set_peer(self, shard: int, generation: int, inbox: int, matrix: int):
        self.w._check(self.w.b.fn("router_set_peer")(self.h, C.c_uint32(shard), C.c_uint32(generation), C.c_void_p(inbox),
                                                     C.c_void_p(matrix)))

    def count(self, which: int = 1):
        self.w._check(self.w.b.fn("router_count")(self.h, C.c_uint32(which)))

    def scatter(self):
        self.w._check(self.w.b.fn("router_scatter")(self.h))

    def deliver(self, which: int = 1, rebase: bool = True) -> int:
        n = C.c_uint64()
        self.w._check(self.w.b.fn("router_deliver")(self.h, C.c_uint32(which), C.c_uint32(1 if rebase else 0), C.byref(n)))
        return n.value

    def close(self):
        if self.h:
            self.w.b.fn("router_destroy")(self.h)
            self.h = C.c_void_p()


IPC_HANDLE_BYTES = 64


class PeerRouter:
    """One process per GPU: the inboxes and count matrices are exchanged once as CUDA IPC handles; afterwards a step is
    count -> 4-byte all-reduce -> scatter (P2P stores over NVLink) -> 4-byte all-reduce -> deliver."""

    def __init__(self, world: capi.World, name: str, torch, dist, device, rank: int, n_shards: int, entities_per_shard: int,
                 inbox_capacity: int, key_offset: int = 0):
        self.dist, self.w, self.torch, self.device = dist, world, torch, device
        self.router = Router(world, name, n_shards, rank, entities_per_shard, inbox_capacity, key_offset)
        b = world.b
        mine = torch.zeros(4 * IPC_HANDLE_BYTES, dtype=torch.uint8)
        for gen in range(2):
            for k, ptr in enumerate(self.router.local(gen)):
                h = (C.c_uint8 * IPC_HANDLE_BYTES)()
                world._check(b.fn("ipc_export")(world.h, C.c_void_p(ptr), h))
                off = (gen * 2 + k) * IPC_HANDLE_BYTES
                mine[off:off + IPC_HANDLE_BYTES] = torch.frombuffer(bytearray(h), dtype=torch.uint8)
        everyone = torch.empty(n_shards * mine.numel(), dtype=torch.uint8, device=device)
        dist.all_gather_into_tensor(everyone, mine.to(device))
        everyone = everyone.cpu().numpy().reshape(n_shards, 2, 2, IPC_HANDLE_BYTES)
        self.opened = []
        for g in range(n_shards):
            if g == rank:
                continue
            for gen in range(2):
                ptrs = []
                for k in range(2):
                    h = (C.c_uint8 * IPC_HANDLE_BYTES).from_buffer_copy(everyone[g, gen, k].tobytes())
                    p = C.c_void_p()
                    world._check(b.fn("ipc_open")(world.h, h, C.byref(p)))
                    ptrs.append(p.value)
                    self.opened.append(p.value)
                self.router.set_peer(g, gen, ptrs[0], ptrs[1])
        self.token = torch.zeros(1, dtype=torch.int32, device=device)

    def route(self, which: int = 1, rebase: bool = True) -> int:
        self.w.require_stream(self.torch, self.device)  # the all-reduces below order the kernels only on a shared stream
        self.router.count(which)
        self.dist.all_reduce(self.token)  # every shard's count matrix is complete
        self.router.scatter()
        self.dist.all_reduce(self.token)  # every shard's inbox is complete
        return self.router.deliver(which, rebase)

    def close(self):
        self.w.synchronize()
        self.dist.barrier()  # nobody closes a mapping a peer may still be writing through
        for p in self.opened:
            self.w.b.fn("ipc_close")(self.w.h, C.c_void_p(p))
        self.opened = []
        self.router.close()
