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
        self.shard, self.dist, self.torch, self.device = shard, dist, torch, device
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
        self.shard.world.require_stream(self.torch, self.device)
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


# =====================================================================================================================
# Spatial bands: the halo-exchange form (include/helios_gpu.h "Spatial bands").  Shard g owns the grid rows
# [g * R, (g + 1) * R) and the boids whose cell lies in them; boids that cross a band edge migrate through the message
# router and the ECS command path (Destroy here, Spawn + Add there); per step each shard reads ONE grid row from each
# neighbour.  The initial placement is still the entity-ID range (every shard spawns its id slice wherever those boids are):
# the first step's migration sorts the flock into bands.
# =====================================================================================================================
BOID_ID = worlds.NS + "BoidId"
MIGRATION = worlds.NS + "BoidMigration"
LIGHT_MIGRATION_CAP = 4096  # leavers per rank and step that travel in the fixed-size all-gather
MIGRATION_DTYPE = np.dtype([("key", "<u8"), ("id", "<u4"), ("pad", "<u4"), ("x", "<f4"), ("y", "<f4"), ("dx", "<f4"), ("dy", "<f4")])
assert MIGRATION_DTYPE.itemsize == 32


class BandShard:
    """One rank's band: a world holding {Position, Velocity, BoidId} + the boids system in band mode."""

    def __init__(self, backend: capi.Backend, n_total: int, world_size: int, n_shards: int = 1, shard: int = 0, device: int = 0,
                 params: dict | None = None, vel_scale: float = 1.0):
        self.b = backend
        self.n_shards, self.shard, self.n_total = n_shards, shard, n_total
        lo, hi = shard * n_total // n_shards, (shard + 1) * n_total // n_shards  # initial placement: id ranges (any remainder)
        self.world = capi.World(backend, device=device, entity_capacity=(hi - lo) + (hi - lo) // 4 + 1024,
                                flags=capi.WORLD_NO_BUILTIN_MESSAGES)
        P, V = worlds.cname("Position"), worlds.cname("Velocity")
        worlds.register_all(self.world, ["Position", "Velocity"])
        self.world.register_component(BOID_ID, 4, 4)
        self.world.message_register(MIGRATION, MIGRATION_DTYPE.itemsize, capi.MSG_MANUAL)
        pos, vel = initial_state(n_total, world_size, lo, hi, vel_scale=vel_scale)
        if hi > lo:
            self.world.spawn_batch([P, V, BOID_ID], hi - lo, [pos, vel, np.arange(lo, hi, dtype=np.uint32)], want_entities=False)
        p = dict(DEFAULT_PARAMS)
        p.update(params or {})
        self.params = BoidsParams(float(world_size), p["dt"], p["w_separation"], p["w_alignment"], p["w_cohesion"], p["max_speed"])
        self.h = C.c_void_p()
        self.world._check(backend.fn("boids_band_create")(self.world.h, C.c_uint64(capi.fnv1a(P)), C.c_uint64(capi.fnv1a(V)),
                                                          C.c_uint64(capi.fnv1a(BOID_ID)), C.byref(self.params),
                                                          C.c_uint32(n_shards), C.c_uint32(shard), C.byref(self.h)))
        n, rp, lo_r, hi_r, eb = C.c_uint64(), C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint64()
        self.world._check(backend.fn("boids_band_info")(self.h, C.byref(n), C.byref(rp), C.byref(lo_r), C.byref(hi_r), C.byref(eb)))
        self.rows_per, self.row_lo, self.row_hi, self.export_bytes = rp.value, lo_r.value, hi_r.value, eb.value

    # -- C ABI ----------------------------------------------------------------------------------------------------------
    def n_local(self) -> int:
        n = C.c_uint64()
        self.world._check(self.b.fn("boids_band_info")(self.h, C.byref(n), None, None, None, None))
        return n.value

    def export_ptr(self, generation: int, side: int) -> int:
        p = C.c_void_p()
        self.world._check(self.b.fn("boids_band_export")(self.h, C.c_uint32(generation), C.c_uint32(side), C.byref(p)))
        return p.value

    def set_halo(self, generation: int, side: int, ptr: int | None):
        self.world._check(self.b.fn("boids_band_set_halo")(self.h, C.c_uint32(generation), C.c_uint32(side), C.c_void_p(ptr)))

    def leavers(self) -> int:
        n = C.c_uint64()
        self.world._check(self.b.fn("boids_band_leavers")(self.h, C.c_uint64(capi.fnv1a(MIGRATION)), C.byref(n)))
        return n.value

    def arrivals(self) -> int:
        n = C.c_uint64()
        self.world._check(self.b.fn("boids_band_arrivals")(self.h, C.c_uint64(capi.fnv1a(MIGRATION)), C.byref(n)))
        return n.value

    def pack(self, dst_ptr: int, cap: int):
        self.world._check(self.b.fn("boids_band_pack")(self.h, C.c_uint64(capi.fnv1a(MIGRATION)), C.c_void_p(dst_ptr), C.c_uint32(cap)))

    def unpack(self, gathered_ptr: int, cap: int) -> bool:
        """True: the queue now holds this shard's arrivals.  False: some rank overflowed `cap` — route the general way."""
        ov = C.c_uint32()
        self.world._check(self.b.fn("boids_band_unpack")(self.h, C.c_uint64(capi.fnv1a(MIGRATION)), C.c_void_p(gathered_ptr),
                                                         C.c_uint32(self.n_shards), C.c_uint32(cap), C.byref(ov)))
        return ov.value == 0

    def bin(self) -> int:
        g = C.c_uint32()
        self.world._check(self.b.fn("boids_band_bin")(self.h, C.byref(g)))
        return g.value

    def step(self):
        self.world._check(self.b.fn("boids_band_step")(self.h))

    def state(self):
        """(ids, positions, velocities) of the boids this shard holds now, in table row order."""
        snap = self.world.snapshot()
        P, V, I = (capi.fnv1a(x) for x in (worlds.cname("Position"), worlds.cname("Velocity"), BOID_ID))
        arch = [a for a in snap["archetypes"] if len(a["types"]) == 3]
        if not arch or arch[0]["entities"].size == 0:
            return np.zeros(0, np.uint32), np.zeros(0, np.float32), np.zeros(0, np.float32)
        a = arch[0]
        return a["columns"][I].view(np.uint32).copy(), a["columns"][P].view(np.float32).copy(), a["columns"][V].view(np.float32).copy()

    def close(self):
        if self.h:
            self.b.fn("boids_destroy")(self.h)
            self.h = C.c_void_p()
        self.world.close()


def gather_band_state(states, n_total):
    """[(ids, pos, vel)] of all shards -> global (positions, velocities) in boid-id order; every id exactly once."""
    ids = np.concatenate([s[0] for s in states])
    assert ids.size == n_total and np.array_equal(np.sort(ids), np.arange(n_total, dtype=np.uint32)), "boids lost or duplicated"
    order = np.argsort(ids, kind="stable")
    pos = np.concatenate([s[1].reshape(-1, 2) for s in states])[order].reshape(-1)
    vel = np.concatenate([s[2].reshape(-1, 2) for s in states])[order].reshape(-1)
    return pos, vel


class BandExchange:
    """One rank of a multi-process run.  Migration goes through routing.route (split -> all_to_all -> merge); the halos are
    either read in place from the neighbours' export buffers (CUDA IPC, `peer=True`: NVLink loads inside the assemble kernel,
    one 4-byte all-reduce as the barrier) or shipped with one all_gather of the export buffers (gloo / CPU tests)."""

    def __init__(self, shard: BandShard, torch, dist, device, peer: bool, light_cap: int = LIGHT_MIGRATION_CAP,
                 barrier: str = "flags"):
        """barrier (peer only): "flags" = neighbours wait on each other's published step numbers on the device
        (helios_gpu_boids_band_set_barrier), "nccl" = one 4-byte all-reduce per step."""
        self.s, self.torch, self.dist, self.device, self.peer, self.cap = shard, torch, dist, device, peer, light_cap
        self.flags = peer and barrier == "flags" and shard.n_shards > 1
        self.opened = []
        self.migrated = []  # (leavers, arrivals) of this rank per step
        self.general_routes = 0
        g, G = shard.shard, shard.n_shards
        if G == 1:
            return
        self.mig_send = torch.zeros(16 + light_cap * 32, dtype=torch.uint8, device=device)
        self.mig_recv = torch.zeros(G * self.mig_send.numel(), dtype=torch.uint8, device=device)
        if peer:
            w, b = shard.world, shard.b
            mine = torch.zeros(4 * IPC_HANDLE_BYTES, dtype=torch.uint8)
            for gen in range(2):
                for side in range(2):
                    h = (C.c_uint8 * IPC_HANDLE_BYTES)()
                    w._check(b.fn("ipc_export")(w.h, C.c_void_p(shard.export_ptr(gen, side)), h))
                    off = (gen * 2 + side) * IPC_HANDLE_BYTES
                    mine[off:off + IPC_HANDLE_BYTES] = torch.frombuffer(bytearray(h), dtype=torch.uint8)
            everyone = torch.empty(G * mine.numel(), dtype=torch.uint8, device=device)
            dist.all_gather_into_tensor(everyone, mine.to(device))
            everyone = everyone.cpu().numpy().reshape(G, 2, 2, IPC_HANDLE_BYTES)
            for nb, my_side, their_side in ((g - 1, 0, 1), (g + 1, 1, 0)):  # lower halo = lower neighbour's LAST row
                if nb < 0 or nb >= G:
                    continue
                for gen in range(2):
                    h = (C.c_uint8 * IPC_HANDLE_BYTES).from_buffer_copy(everyone[nb, gen, their_side].tobytes())
                    p = C.c_void_p()
                    w._check(b.fn("ipc_open")(w.h, h, C.byref(p)))
                    self.opened.append(p.value)
                    shard.set_halo(gen, my_side, p.value)
            self.token = torch.zeros(1, dtype=torch.int32, device=device)
            if self.flags:
                shard.world._check(shard.b.fn("boids_band_set_barrier")(shard.h, C.c_uint32(1)))
                dist.barrier()  # every rank has opened its neighbours' buffers before anybody publishes into them
        else:
            if shard.b.is_gpu:
                raise capi.HeliosError(capi.ERR_INVALID_ARGUMENT, "BandExchange(peer=False) ships the edge rows through host memory: it is the "
                                       "CPU (oracle / gloo) form; on GPUs use peer=True (the neighbours' exports are read in place)")
            self.recv = torch.empty(G * 2 * shard.export_bytes, dtype=torch.uint8, device=device)
            self.send = torch.empty(2 * shard.export_bytes, dtype=torch.uint8, device=device)

    def step(self):
        from . import routing
        s, G, g = self.s, self.s.n_shards, self.s.shard
        if G > 1 and self.peer:
            s.world.require_stream(self.torch, self.device)  # collectives and kernels are ordered by ONE stream
        if G > 1:
            out = s.leavers()
            # light exchange: one fixed-size all-gather of everybody's (few) leavers; the general router when they do not fit
            s.pack(self.mig_send.data_ptr(), self.cap)
            self.dist.all_gather_into_tensor(self.mig_recv, self.mig_send)
            if not s.unpack(self.mig_recv.data_ptr(), self.cap):
                self.general_routes += 1
                routing.route(s.world, MIGRATION, self.torch, self.device, self.dist, g, G, s.rows_per, 0, 1, False)
            self.migrated.append((out, s.arrivals()))
        gen = s.bin()
        if G > 1:
            if self.peer and self.flags:
                pass  # band_step waits for the neighbours' published step on the device
            elif self.peer:
                self.dist.all_reduce(self.token)  # every shard's exports of this generation are complete
            else:
                eb = s.export_bytes
                for side in range(2):
                    src = (C.c_uint8 * eb).from_address(s.export_ptr(gen, side))
                    self.send[side * eb:(side + 1) * eb] = self.torch.frombuffer(src, dtype=self.torch.uint8)
                self.dist.all_gather_into_tensor(self.recv, self.send)
                if g > 0:
                    s.set_halo(gen, 0, self.recv[((g - 1) * 2 + 1) * eb:].data_ptr())
                if g + 1 < G:
                    s.set_halo(gen, 1, self.recv[((g + 1) * 2 + 0) * eb:].data_ptr())
        s.step()

    def close(self):
        if self.opened:
            w = self.s.world
            w.synchronize()
            self.dist.barrier()
            for p in self.opened:
                self.s.b.fn("ipc_close")(w.h, C.c_void_p(p))
            self.opened = []


def emulate_bands(backend: capi.Backend, torch, device, n_total: int, world_size: int, n_shards: int, steps: int,
                  params: dict | None = None, vel_scale: float = 1.0):
    """All bands in ONE process (tests): ranks emulated as sequential calls; halos are the other shards' export buffers, the
    migration exchange is the numpy statement of the router.  Returns (positions, velocities, migrations per step)."""
    from . import routing
    shards = [BandShard(backend, n_total, world_size, n_shards, g, params=params, vel_scale=vel_scale) for g in range(n_shards)]
    for s in shards:
        for gen in range(2):
            if s.shard > 0:
                s.set_halo(gen, 0, shards[s.shard - 1].export_ptr(gen, 1))
            if s.shard + 1 < n_shards:
                s.set_halo(gen, 1, shards[s.shard + 1].export_ptr(gen, 0))
    moved = []
    for _ in range(steps):
        if n_shards > 1:
            out = [s.leavers() for s in shards]
            queues = [s.world.message_read(MIGRATION, 1) for s in shards]
            want = routing.reference_routing(queues, MIGRATION_DTYPE.itemsize, 0, shards[0].rows_per)
            for s, q in zip(shards, want):
                # reference_routing re-bases keys; the migration exchange does not (the key is a grid row, not an entity)
                k = q.reshape(-1, 32)[:, 0:4].copy().view("<u4").reshape(-1) + np.uint32(s.shard * s.rows_per)
                q = q.copy()
                q.reshape(-1, 32)[:, 0:4] = k.view(np.uint8).reshape(-1, 4)
                t = torch.from_numpy(q).to(device)
                if str(device) != "cpu":
                    torch.cuda.synchronize()  # the copy ran on torch's stream, the merge reads on the world's
                routing.merge(s.world, MIGRATION, 1, t, q.size // 32, 0, 0)
                s.world.synchronize()         # `t` may be recycled by torch's allocator after this iteration
            moved.append((sum(out), sum(s.arrivals() for s in shards)))
        for s in shards:
            s.bin()
        for s in shards:
            s.world.synchronize()  # the barrier
        for s in shards:
            s.step()
        for s in shards:
            s.world.synchronize()
    states = [s.state() for s in shards]
    for s in shards:
        s.close()
    pos, vel = gather_band_state(states, n_total)
    return pos, vel, moved
