"""Synthetic worlds of the BASELINE.json configs — identical seeded inputs for any backend.

Generator = SplitMix64, seed 20261104, one stream per (config, column);
floats = lo + (hi - lo) * (u >> 40) * 2^-24.
"""
from __future__ import annotations

import struct

import numpy as np

from . import capi

SEED = 20261104
NS = "helios::bench::"

# name -> (size, numpy dtype of the flat storage)
COMPONENTS = {
    "Position": (8, np.float32), "Velocity": (8, np.float32), "Acceleration": (8, np.float32), "Mass": (4, np.float32),
    "Health": (4, np.int32), "Team": (4, np.uint32), "Damping": (4, np.float32), "Frozen": (4, np.uint32),
    "Shield": (4, np.uint32), "Poison": (4, np.uint32), "Tint": (4, np.uint32),
}
OPTIONAL = ["Health", "Team", "Damping", "Frozen", "Shield", "Poison"]
BASE4 = ["Position", "Velocity", "Acceleration", "Mass"]
COLLISION_EVENT = NS + "CollisionEvent"  # {Entity a; Entity b; f32 impulse; u32 pad} = 24 B
DT = 2.0 ** -6


def cname(short: str) -> str:
    return NS + short


def splitmix64(seed: int, n: int, stream: int = 0, offset: int = 0) -> np.ndarray:
    """Outputs offset+1 .. offset+n of SplitMix64 started at seed + stream * 0x632BE59BD9B4E019 (random access: a shard
    generates exactly its slice of the global sequence)."""
    with np.errstate(over="ignore"):
        s0 = np.uint64((seed + stream * 0x632BE59BD9B4E019) & 0xFFFFFFFFFFFFFFFF)
        z = s0 + (np.arange(offset + 1, offset + n + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def uniform_f32(seed: int, n: int, stream: int, lo: float, hi: float, offset: int = 0) -> np.ndarray:
    u = (splitmix64(seed, n, stream, offset) >> np.uint64(40)).astype(np.float64) * (2.0 ** -24)
    return (lo + (hi - lo) * u).astype(np.float32)


def column(short: str, n: int, stream: int, seed: int = SEED) -> np.ndarray:
    """Values of one component column (flat numpy array of n * size bytes)."""
    if short == "Position":
        return uniform_f32(seed, 2 * n, stream, 0.0, 1024.0)
    if short == "Velocity":
        return uniform_f32(seed, 2 * n, stream, -1.0, 1.0)
    if short == "Acceleration":
        return uniform_f32(seed, 2 * n, stream, -0.125, 0.125)
    if short == "Mass":
        return uniform_f32(seed, n, stream, 0.5, 2.0)
    if short == "Damping":
        return uniform_f32(seed, n, stream, 0.9, 1.0)
    if short == "Health":
        return (1 + (splitmix64(seed, n, stream) % np.uint64(100))).astype(np.int32)
    if short == "Team":
        return (splitmix64(seed, n, stream) % np.uint64(8)).astype(np.uint32)
    if short == "Frozen":
        return np.ones(n, dtype=np.uint32)
    if short in ("Shield", "Poison", "Tint"):
        return (splitmix64(seed, n, stream) % np.uint64(1000)).astype(np.uint32)
    raise KeyError(short)


def register_all(world: capi.World, names=None):
    for short in (names or COMPONENTS):
        world.register_component(cname(short), COMPONENTS[short][0], 4)


def build_cfg1(world: capi.World, n: int = 1 << 20):
    """configs[0]: n entities {Position, Velocity}, one archetype, README ChangePosition system."""
    register_all(world, ["Position", "Velocity"])
    world.spawn_batch([cname("Position"), cname("Velocity")], n, [column("Position", n, 1), column("Velocity", n, 2)],
                      want_entities=False)
    q = world.query([(cname("Position"), True), (cname("Velocity"), False)])
    s = world.schedule("Update")
    s.add_system("ChangePosition", "integrate", q)
    s.build()
    return s, q


def build_lattice(world: capi.World, n: int, optional: list[str], stream_base: int = 100):
    """n entities over the 2^k archetype lattice {BASE4} x subsets(optional); archetype created in subset-mask order,
    archetype m holds ids [m*rows, (m+1)*rows)."""
    register_all(world, BASE4 + OPTIONAL)
    k = len(optional)
    n_arch = 1 << k
    rows = n // n_arch
    for m in range(n_arch):
        shorts = BASE4 + [optional[b] for b in range(k) if m >> b & 1]
        cols = [column(sh, rows, stream_base + m * 16 + i) for i, sh in enumerate(shorts)]
        world.spawn_batch([cname(sh) for sh in shorts], rows, cols, want_entities=False)
    return rows


def build_cfg2(world: capi.World, n: int = 1 << 24, filtered: bool = False, flags: int = 0):
    """configs[1]: n entities across 64 archetypes (4 mandatory + 6 optional 4-byte components);
    Query<Position&, Velocity&, const Acceleration&, const Mass&[, With<Health>, Without<Frozen>]>,
    body v += a*(dt/m); p += v*dt."""
    build_lattice(world, n, OPTIONAL)
    q = world.query([(cname("Position"), True), (cname("Velocity"), True), (cname("Acceleration"), False), (cname("Mass"), False)],
                    with_=[cname("Health")] if filtered else [], without=[cname("Frozen")] if filtered else [])
    s = world.schedule("Update")
    s.add_system("AccelIntegrate", "accel_integrate", q, params=struct.pack("<f", DT), flags=flags)
    s.build()
    return s, q


CFG2_BYTES_PER_ENTITY = 44  # read 8+8+8+4, write 8+8
CFG1_BYTES_PER_ENTITY = 24


def build_cfg3(world: capi.World, n: int = 1 << 23):
    """configs[2] world: n entities over the 16-table lattice with optional {Health, Damping, Frozen, Team}."""
    build_lattice(world, n, ["Health", "Damping", "Frozen", "Team"], stream_base=5000)


def cfg3_commands(world: capi.World, snap_entities: list[np.ndarray], snap_types: list[list[int]], frame: int,
                  n_cmds: int, seed: int = SEED):
    """One frame of structural churn: 25 % Spawn().AddComponents(P,V,A,M), 25 % Destroy, 25 % AddComponents(Health) on
    entities lacking it, 25 % RemoveComponents<Damping> on entities having it; targets are DISTINCT live entities.
    Returns (commands, payload, n_spawn).  The caller reserves n_spawn entities first (Commands::Spawn)."""
    q = n_cmds // 4
    h_health, h_damp = capi.fnv1a(cname("Health")), capi.fnv1a(cname("Damping"))
    have_h = [e for e, t in zip(snap_entities, snap_types) if h_health in t]
    lack_h = [e for e, t in zip(snap_entities, snap_types) if h_health not in t and len(t) > 0]
    have_d = [e for e, t in zip(snap_entities, snap_types) if h_damp in t]
    all_e = np.concatenate([e for e in snap_entities if e.size]) if snap_entities else np.empty(0, np.uint64)
    rng = np.random.Generator(np.random.PCG64(seed * 1000003 + frame))

    def pick(pool_list, k, exclude):
        pool = np.concatenate(pool_list) if pool_list else np.empty(0, np.uint64)
        if exclude is not None and exclude.size:
            pool = np.setdiff1d(pool, exclude, assume_unique=False)
        k = min(k, pool.size)
        return rng.choice(pool, size=k, replace=False) if k else np.empty(0, np.uint64)

    despawn = pick([all_e], q, None)
    insert = pick(lack_h, q, despawn)
    remove = pick(have_d, q, np.concatenate([despawn, insert]))
    n_spawn = q
    return despawn, insert, remove, n_spawn


def assemble_cfg3(world: capi.World, spawned: np.ndarray, despawn, insert, remove, frame: int, seed: int = SEED):
    """Interleave the four command streams in a seeded random order and build records + payload."""
    m = world.mask
    base_mask = m([cname(x) for x in BASE4])
    n = spawned.size + despawn.size + insert.size + remove.size
    cmds = np.zeros(n, dtype=capi.COMMAND_DTYPE)
    kinds = np.concatenate([np.full(spawned.size, 0), np.full(despawn.size, 1), np.full(insert.size, 2), np.full(remove.size, 3)])
    ents = np.concatenate([spawned, despawn, insert, remove]).astype(np.uint64)
    rng = np.random.Generator(np.random.PCG64(seed * 7919 + frame))
    perm = rng.permutation(n)
    kinds, ents = kinds[perm], ents[perm]
    cmds["entity"] = ents
    cmds["kind"] = np.choose(kinds, [capi.CMD_ADD, capi.CMD_DESTROY, capi.CMD_ADD, capi.CMD_REMOVE])
    cmds["component_mask"] = np.choose(kinds, [base_mask, 0, m([cname("Health")]), m([cname("Damping")])]).astype(np.uint64)
    # payload: spawn = P,V,A,M in ascending slot order (28 B), insert = Health (4 B)
    sizes = np.choose(kinds, [28, 0, 4, 0]).astype(np.uint32)
    offs = np.zeros(n, dtype=np.uint64)
    np.cumsum(sizes[:-1], out=offs[1:])
    cmds["payload_offset"] = offs.astype(np.uint32)
    total = int(sizes.sum())
    payload = np.zeros(total // 4 + 4, dtype=np.uint32)
    vals = splitmix64(seed, total // 4 + 4, 9000 + frame)
    fl = ((vals >> np.uint64(40)).astype(np.float64) * (2.0 ** -24)).astype(np.float32)
    payload_f = fl.view(np.uint32).copy()
    # Health payloads are ints in [1,100]
    ins_idx = np.nonzero(kinds == 2)[0]
    payload_f[(offs[ins_idx] // 4).astype(np.int64)] = (1 + (vals[: ins_idx.size] % np.uint64(100))).astype(np.uint32)
    # Mass (last 4 bytes of a spawn payload) must be non-zero: shift into [0.5, 1.5)
    sp_idx = np.nonzero(kinds == 0)[0]
    mpos = (offs[sp_idx] // 4 + 6).astype(np.int64)
    payload_f[mpos] = (payload_f[mpos].view(np.float32) + np.float32(0.5)).view(np.uint32)
    payload[:] = payload_f[: payload.size]
    return cmds, payload.view(np.uint8)[:total + 16]


def build_cfg4(world: capi.World, n: int = 1 << 22, sparse_emit: bool = False):
    """configs[3]: n entities {Position, Velocity, Health, Team, Tint}; 8-system schedule with a CollisionEvent channel.

    The emitter is REGISTERED LAST on purpose: every system entry's MergeMessages moves
    world-current into world-previous before appending its own messages (system_local_data.hpp:122-132), and the
    per-frame World::Update clears automatic-policy `previous` (message/manager.hpp:400-419).  Only messages merged by
    the last entry are still in `current` at World::Update, become `previous` there and are what next frame's reader
    sees — the documented "write frame N -> read frame N+1" flow."""
    register_all(world, ["Position", "Velocity", "Health", "Team", "Tint"])
    shorts = ["Position", "Velocity", "Health", "Team", "Tint"]
    world.spawn_batch([cname(s) for s in shorts], n, [column(s, n, 7000 + i) for i, s in enumerate(shorts)], want_entities=False)
    world.message_register(COLLISION_EVENT, 24)
    P, V, H, T, TI = (cname(x) for x in shorts)
    s = world.schedule("Update")
    s.add_system("S0_integrate", "integrate", world.query([(P, True), (V, False)]))
    s.add_system("S1_damp", "damp", world.query([(V, True)]), params=struct.pack("<f", 0.99))
    s.add_system("S2_regen", "regen", world.query([(H, True)]))
    s.add_system("S3_team_tint", "team_tint", world.query([(T, False), (TI, True)]))
    s.add_system("S4_consume", "consume", None, params=struct.pack("<I", world.slots[H]), reads_message=COLLISION_EVENT,
                 extra_writes=[H])
    s.add_system("S5_stats", "stats", world.query([(H, False)]), write_resources=["helios::bench::Stats"])
    s.add_system("S6_exit_check", "exit_check", None, read_resources=["helios::bench::Stats"],
                 write_resources=["helios::bench::Exit"])
    s.add_system("S7_detect", "detect_sparse" if sparse_emit else "detect", world.query([(P, False), (T, False)]),
                 writes_message=COLLISION_EVENT, writer_capacity=n if sparse_emit else 0)
    s.build()
    return s


# per entity: S0 24, S1 16, S2 8, S3 8, S5 4, S7 12 (+ the 8-byte entity id); per event: 24 written + 24 read + 8 Health rmw
CFG4_COMPONENT_BYTES_PER_ENTITY = 24 + 16 + 8 + 8 + 4 + 12 + 8
CFG4_BYTES_PER_EVENT = 24 + 24 + 8


def snapshot_digest(snap: dict) -> str:
    """SHA-256 of a canonical snapshot: equal digests == bit-exact world state."""
    import hashlib
    h = hashlib.sha256()
    h.update(struct.pack("<IIqQ", len(snap["archetypes"]), snap["next_index"], snap["free_cursor"], snap["entity_count"]))
    for a in snap["archetypes"]:
        h.update(struct.pack("<II", len(a["types"]), a["entities"].size))
        h.update(np.asarray(a["types"], dtype=np.uint64).tobytes())
        h.update(np.ascontiguousarray(a["entities"]).tobytes())
        for t in a["types"]:
            h.update(np.ascontiguousarray(a["columns"][t]).tobytes())
    h.update(np.ascontiguousarray(snap["generations"]).tobytes())
    h.update(np.ascontiguousarray(snap["free_indices"]).tobytes())
    for t in sorted(snap.get("messages", {})):  # every message queue, the builtin structural messages included
        prev, cur = snap["messages"][t]
        h.update(struct.pack("<QQQ", t, prev.size, cur.size))
        h.update(np.ascontiguousarray(prev).tobytes())
        h.update(np.ascontiguousarray(cur).tobytes())
    return h.hexdigest()


def snapshots_equal(a: dict, b: dict, float_cols: dict | None = None) -> list[str]:
    """Bit-exact comparison of two canonical snapshots; returns the list of differences (empty = equal)."""
    diffs = []
    if len(a["archetypes"]) != len(b["archetypes"]):
        diffs.append(f"archetype count {len(a['archetypes'])} != {len(b['archetypes'])}")
        return diffs
    for i, (x, y) in enumerate(zip(a["archetypes"], b["archetypes"])):
        if x["types"] != y["types"]:
            diffs.append(f"archetype {i}: type lists differ")
            continue
        if x["entities"].shape != y["entities"].shape or not np.array_equal(x["entities"], y["entities"]):
            diffs.append(f"archetype {i}: entity rows differ ({x['entities'].size} vs {y['entities'].size})")
            continue
        for t in x["types"]:
            if t in x["columns"] and t in y["columns"] and not np.array_equal(x["columns"][t], y["columns"][t]):
                diffs.append(f"archetype {i}: column {t:016x} differs")
    for key in ("next_index", "free_cursor", "entity_count"):
        if a[key] != b[key]:
            diffs.append(f"{key}: {a[key]} != {b[key]}")
    for key in ("generations", "free_indices"):
        if a[key].shape != b[key].shape or not np.array_equal(a[key], b[key]):
            diffs.append(f"{key} differ")
    ma, mb = a.get("messages", {}), b.get("messages", {})
    if sorted(ma) != sorted(mb):
        diffs.append(f"registered message types differ: {sorted(f'{t:016x}' for t in set(ma) ^ set(mb))}")
    for t in sorted(set(ma) & set(mb)):
        for which, (x, y) in enumerate(zip(ma[t], mb[t])):
            if x.shape != y.shape or not np.array_equal(x, y):
                diffs.append(f"message {t:016x} {'current' if which else 'previous'} queue differs ({x.size} vs {y.size} bytes)")
    return diffs
