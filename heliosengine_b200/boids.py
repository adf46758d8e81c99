"""Sharded boids (BASELINE configs[4]): entity-ID-range shards, one per rank, exchanged with one all-gather per step.

Host-side driver over the C ABI (`helios_gpu_boids_*`, include/helios_gpu.h).  The collective is `torch.distributed`
plumbing: NCCL on GPUs (slabs and counts live in torch CUDA tensors whose pointers are handed to the library), gloo on
CPU with the oracle backend for the world_size>1 tests.  Shard g owns global boid ids [g*n/G, (g+1)*n/G); its initial
state is the slice of the global seeded streams, so any shard count starts from the same world.
"""
from __future__ import annotations

import ctypes as C
import struct

import numpy as np

from . import capi, worlds

DEFAULT_PARAMS = dict(dt=2.0 ** -6, w_separation=1.5, w_alignment=1.0, w_cohesion=1.0, max_speed=2.0)
SLAB_BYTES = 16   # {x, y, dx, dy}
COUNT_BYTES = 2   # u16 per cell


class BoidsParams(C.Structure):
    _fields_ = [("world_size", C.c_float), ("dt", C.c_float), ("w_separation", C.c_float), ("w_alignment", C.c_float),
                ("w_cohesion", C.c_float), ("max_speed", C.c_float)]


def shard_range(n_total: int, n_shards: int, shard: int) -> tuple[int, int]:
    """Entity-ID range of one shard (equal shards: the all-gather is a fixed-size collective)."""
    if n_total % n_shards:
        raise ValueError(f"{n_total} boids do not split evenly over {n_shards} shards")
    per = n_total // n_shards
    return shard * per, (shard + 1) * per


def initial_state(n_total: int, world_size: int, lo: int, hi: int, seed: int = worlds.SEED, vel_scale: float = 1.0):
    """Positions uniform over [0, world_size)^2, velocities in [-vel_scale, vel_scale)^2 for global ids [lo, hi)."""
    n = hi - lo
    pos = worlds.uniform_f32(seed, 2 * n, 9101, 0.0, float(world_size), offset=2 * lo)
    vel = worlds.uniform_f32(seed, 2 * n, 9102, -float(vel_scale), float(vel_scale), offset=2 * lo)
    return pos, vel


class BoidsShard:
    """One rank's shard: a world holding {Position, Velocity} for its id range + the boids system."""

    def __init__(self, backend: capi.Backend, n_total: int, world_size: int, n_shards: int = 1, shard: int = 0, device: int = 0,
                 params: dict | None = None, vel_scale: float = 1.0):
        self.b = backend
        self.n_shards, self.shard = n_shards, shard
        lo, hi = shard_range(n_total, n_shards, shard)
        self.n_local = hi - lo
        self.world = capi.World(backend, device=device, entity_capacity=self.n_local + 1024)
        worlds.register_all(self.world, ["Position", "Velocity"])
        pos, vel = initial_state(n_total, world_size, lo, hi, vel_scale=vel_scale)
        self.world.spawn_batch([worlds.cname("Position"), worlds.cname("Velocity")], self.n_local, [pos, vel], want_entities=False)
        p = dict(DEFAULT_PARAMS)
        p.update(params or {})
        self.params = BoidsParams(float(world_size), p["dt"], p["w_separation"], p["w_alignment"], p["w_cohesion"], p["max_speed"])
        self.h = C.c_void_p()
        self.world._check(backend.fn("boids_create")(self.world.h, C.c_uint64(capi.fnv1a(worlds.cname("Position"))),
                                                     C.c_uint64(capi.fnv1a(worlds.cname("Velocity"))), C.byref(self.params),
                                                     C.byref(self.h)))
        n_local, n_cells = C.c_uint64(), C.c_uint64()
        self.world._check(backend.fn("boids_sizes")(self.h, C.byref(n_local), C.byref(n_cells)))
        assert n_local.value == self.n_local
        self.n_cells = n_cells.value

    # -- C ABI ----------------------------------------------------------------------------------------------------------
    def bin(self, slab_ptr: int, counts_ptr: int):
        self.world._check(self.b.fn("boids_bin")(self.h, C.c_void_p(slab_ptr), C.c_void_p(counts_ptr)))

    def step(self, slabs_ptr: int, counts_ptr: int):
        self.world._check(self.b.fn("boids_step")(self.h, C.c_void_p(slabs_ptr), C.c_void_p(counts_ptr),
                                                  C.c_uint32(self.n_shards), C.c_uint32(self.shard)))

    def step_local(self):
        self.world._check(self.b.fn("boids_step_local")(self.h))

    # -- peer-memory exchange (GPU library only) ---------------------------------------------------------------------------
    def exchange_create(self):
        self.world._check(self.b.fn("boids_exchange_create")(self.h, C.c_uint32(self.n_shards), C.c_uint32(self.shard)))

    def exchange_local(self, generation: int) -> tuple[int, int]:
        slab, counts = C.c_void_p(), C.c_void_p()
        self.world._check(self.b.fn("boids_exchange_local")(self.h, C.c_uint32(generation), C.byref(slab), C.byref(counts)))
        return slab.value, counts.value

    def exchange_set_peer(self, shard: int, generation: int, slab_ptr: int, counts_ptr: int):
        self.world._check(self.b.fn("boids_exchange_set_peer")(self.h, C.c_uint32(shard), C.c_uint32(generation),
                                                               C.c_void_p(slab_ptr), C.c_void_p(counts_ptr)))

    def exchange_bin(self):
        self.world._check(self.b.fn("boids_exchange_bin")(self.h))

    def exchange_step(self):
        self.world._check(self.b.fn("boids_exchange_step")(self.h))

    def state(self) -> tuple[np.ndarray, np.ndarray]:
        """(positions, velocities) of this shard in entity-id order (the table's row order never changes)."""
        snap = self.world.snapshot()
        arch = [a for a in snap["archetypes"] if len(a["types"]) == 2][0]
        P, V = capi.fnv1a(worlds.cname("Position")), capi.fnv1a(worlds.cname("Velocity"))
        return arch["columns"][P].view(np.float32).copy(), arch["columns"][V].view(np.float32).copy()

    def close(self):
        if self.h:
            self.b.fn("boids_destroy")(self.h)
            self.h = C.c_void_p()
        self.world.close()


class ExchangeBuffers:
    """All-gather buffers: slabs[g] / counts[g] = shard g's cell-ordered slab and per-cell counts."""

    def __init__(self, torch, device, n_shards: int, n_local: int, n_cells: int):
        self.torch = torch
        self.n_shards, self.n_local, self.n_cells = n_shards, n_local, n_cells
        self.slabs = torch.empty(n_shards * n_local * SLAB_BYTES, dtype=torch.uint8, device=device)
        self.counts = torch.empty(n_shards * n_cells * COUNT_BYTES, dtype=torch.uint8, device=device)

    def slab_slot(self, g: int):
        return self.slabs[g * self.n_local * SLAB_BYTES:(g + 1) * self.n_local * SLAB_BYTES]

    def counts_slot(self, g: int):
        return self.counts[g * self.n_cells * COUNT_BYTES:(g + 1) * self.n_cells * COUNT_BYTES]

    @property
    def exchange_bytes_per_rank(self) -> int:
        """Bytes one rank receives per step."""
        return (self.n_shards - 1) * (self.n_local * SLAB_BYTES + self.n_cells * COUNT_BYTES)


def distributed_step(shard: BoidsShard, buf: ExchangeBuffers, dist=None, sync=None):
    """One step of one rank: bin into the rank's slot, all-gather slabs and counts in place, neighbour pass + integrate.

    `sync` (GPU): called between the library's stream and the collective's stream so that each sees the other's work."""
    g = shard.shard
    shard.bin(buf.slab_slot(g).data_ptr(), buf.counts_slot(g).data_ptr())
    if shard.n_shards > 1:
        if sync:
            sync()
        dist.all_gather_into_tensor(buf.slabs, buf.slab_slot(g))
        dist.all_gather_into_tensor(buf.counts, buf.counts_slot(g))
        if sync:
            sync()
    shard.step(buf.slabs.data_ptr(), buf.counts.data_ptr())


IPC_HANDLE_BYTES = 64


class PeerExchange:
    """NVLink form of the exchange, one process per GPU: every rank exports its two generations of {slab, counts} with
    CUDA IPC, opens the other ranks' handles, and from then on a step is  bin -> 4-byte all-reduce (the barrier, on the
    world's stream) -> merge-reading-peers + neighbour pass.  No slab ever goes through a collective."""

    def __init__(self, shard: BoidsShard, torch, dist, device):
        self.shard, self.dist = shard, dist
        b, w = shard.b, shard.world
        shard.exchange_create()
        mine = torch.zeros(4 * IPC_HANDLE_BYTES, dtype=torch.uint8)
        for gen in range(2):
            for k, ptr in enumerate(shard.exchange_local(gen)):
                h = (C.c_uint8 * IPC_HANDLE_BYTES)()
                w._check(b.fn("ipc_export")(w.h, C.c_void_p(ptr), h))
                off = (gen * 2 + k) * IPC_HANDLE_BYTES
                mine[off:off + IPC_HANDLE_BYTES] = torch.frombuffer(bytearray(h), dtype=torch.uint8)
        everyone = torch.empty(shard.n_shards * mine.numel(), dtype=torch.uint8, device=device)
        dist.all_gather_into_tensor(everyone, mine.to(device))
        everyone = everyone.cpu().numpy().reshape(shard.n_shards, 2, 2, IPC_HANDLE_BYTES)
        self.opened = []
        for g in range(shard.n_shards):
            if g == shard.shard:
                continue
            for gen in range(2):
                ptrs = []
                for k in range(2):
                    h = (C.c_uint8 * IPC_HANDLE_BYTES).from_buffer_copy(everyone[g, gen, k].tobytes())
                    p = C.c_void_p()
                    w._check(b.fn("ipc_open")(w.h, h, C.byref(p)))
                    ptrs.append(p.value)
                    self.opened.append(p.value)
                shard.exchange_set_peer(g, gen, ptrs[0], ptrs[1])
        self.token = torch.zeros(1, dtype=torch.int32, device=device)
        # per peer: its slab, its u8 cell counts and its per-tile sums (1024-cell tiles)
        self.bytes_pulled_per_step = (shard.n_shards - 1) * (shard.n_local * SLAB_BYTES + shard.n_cells + 4 * -(-shard.n_cells // 1024))

    def step(self):
        self.shard.exchange_bin()
        self.dist.all_reduce(self.token)  # device-side barrier: every shard's buffers of this generation are complete
        self.shard.exchange_step()

    def close(self):
        w = self.shard.world
        w.synchronize()
        self.dist.barrier()  # nobody closes a mapping a peer may still be reading
        for p in self.opened:
            self.shard.b.fn("ipc_close")(w.h, C.c_void_p(p))
        self.opened = []


def emulate_shards_peer(backend: capi.Backend, n_total: int, world_size: int, n_shards: int, steps: int,
                        params: dict | None = None):
    """Peer-memory exchange with all shards in ONE process on one GPU: the "peer" pointers are the other shards' own
    buffers (no IPC needed inside a process).  Returns the global (positions, velocities)."""
    shards = [BoidsShard(backend, n_total, world_size, n_shards, g, params=params) for g in range(n_shards)]
    for s in shards:
        s.exchange_create()
    for s in shards:
        for o in shards:
            if o is not s:
                for gen in range(2):
                    s.exchange_set_peer(o.shard, gen, *o.exchange_local(gen))
    for _ in range(steps):
        for s in shards:
            s.exchange_bin()
        for s in shards:  # the barrier
            s.world.synchronize()
        for s in shards:
            s.exchange_step()
        for s in shards:
            s.world.synchronize()
    states = [s.state() for s in shards]
    for s in shards:
        s.close()
    return np.concatenate([p for p, _ in states]), np.concatenate([v for _, v in states])


def emulate_shards(backend: capi.Backend, torch, device, n_total: int, world_size: int, n_shards: int, steps: int,
                   params: dict | None = None):
    """All shards in ONE process on one device (tests; ranks emulated as sequential calls over shared buffers):
    returns the global (positions, velocities) after `steps` steps."""
    shards = [BoidsShard(backend, n_total, world_size, n_shards, g, params=params) for g in range(n_shards)]
    buf = ExchangeBuffers(torch, device, n_shards, shards[0].n_local, shards[0].n_cells)
    for _ in range(steps):
        for s in shards:  # every shard bins from the pre-step state ...
            s.bin(buf.slab_slot(s.shard).data_ptr(), buf.counts_slot(s.shard).data_ptr())
        for s in shards:  # ... then steps over the complete exchange buffers
            s.world.synchronize()
        for s in shards:
            s.step(buf.slabs.data_ptr(), buf.counts.data_ptr())
        for s in shards:
            s.world.synchronize()
    states = [s.state() for s in shards]
    for s in shards:
        s.close()
    return np.concatenate([p for p, _ in states]), np.concatenate([v for _, v in states])
