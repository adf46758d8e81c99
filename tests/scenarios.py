"""Seeded parity scenarios shared by the golden-fixture generator, the CPU tests (oracle) and the GPU tests.

Every scenario builds a world on the given backend, drives it through the C ABI and returns a dict of SHA-256 digests of
the canonical world snapshot (archetype order, type lists, entity rows, raw column bytes, generation table, free list) plus
whatever else the scenario observes (match sets, message buffers, schedule resources).  Identical digests == bit-exact
parity; tests/golden/golden.json holds the digests the oracle produced when the fixtures were generated."""
import hashlib
import struct

import numpy as np

from heliosengine_b200 import capi, worlds


snapshot_digest = worlds.snapshot_digest


def _bytes_digest(*arrays) -> str:
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def cfg1(backend, n=4099, frames=3, mode=capi.EXEC_SERIAL):
    w = capi.World(backend)
    s, q = worlds.build_cfg1(w, n)
    out = {"count": q.count(), "match": q.match()}
    for _ in range(frames):
        s.run_and_wait(mode)
    out["world"] = snapshot_digest(w.snapshot())
    w.close()
    return out


def cfg2(backend, n=64 * 1500, filtered=True, frames=3, mode=capi.EXEC_SERIAL, flags=0):
    w = capi.World(backend)
    s, q = worlds.build_cfg2(w, n, filtered=filtered, flags=flags)
    out = {"count": q.count(), "match": q.match()}
    for _ in range(frames):
        s.run_and_wait(mode)
    out["world"] = snapshot_digest(w.snapshot())
    w.close()
    return out


def cfg3(backend, n=16 * 2048, frac=0.10, frames=2):
    """Structural churn.  Targets are drawn from the backend's own live lists — identical on both sides as long as parity
    holds, so a divergence in frame k shows up in frame k's digest."""
    w = capi.World(backend)
    worlds.build_cfg3(w, n)
    out = {}
    for frame in range(frames):
        snap = w.snapshot(columns=False)
        ents = [a["entities"] for a in snap["archetypes"]]
        types = [a["types"] for a in snap["archetypes"]]
        despawn, insert, remove, n_spawn = worlds.cfg3_commands(w, ents, types, frame, int(n * frac))
        spawned = w.reserve_entities(n_spawn)
        cmds, payload = worlds.assemble_cfg3(w, spawned, despawn, insert, remove, frame)
        w.enqueue(cmds, payload)
        w.flush()
        out[f"frame{frame}"] = snapshot_digest(w.snapshot())
    w.close()
    return out


def cfg4(backend, n=8192, frames=3, mode=capi.EXEC_SERIAL):
    """8-system schedule + CollisionEvent channel: RunAndWait + World::Update per frame."""
    w = capi.World(backend)
    s = worlds.build_cfg4(w, n)
    out = {"order": s.execution_order(), "conflicts": s.conflicts()}
    for frame in range(frames):
        s.run_and_wait(mode)
        prev_n, cur_n = w.message_counts(worlds.COLLISION_EVENT)
        out[f"frame{frame}_counts_after_run"] = [prev_n, cur_n]
        out[f"frame{frame}_messages_after_run"] = _bytes_digest(w.message_read(worlds.COLLISION_EVENT, 0),
                                                                  w.message_read(worlds.COLLISION_EVENT, 1))
        w.update()
        out[f"frame{frame}_counts_after_update"] = list(w.message_counts(worlds.COLLISION_EVENT))
        out[f"frame{frame}_resources"] = s.read_resources(16).tobytes().hex()
    out["world"] = snapshot_digest(w.snapshot())
    w.close()
    return out


def consumable(backend, n=4099, frames=4, mode=capi.EXEC_SERIAL):
    """Consumable messages (manager.hpp:376-398,421-496): a MANUAL-policy CollisionEvent channel, so events live until a
    reader consumes them.  S0 consumes the strong ones (impulse > 2.5: Health damage + MarkConsumed), S1 reads whatever
    is left, S2 emits this frame's events (registered last, as in cfg4).  Counts and bytes of both queues are recorded
    after every run and every World::Update."""
    w = capi.World(backend)
    shorts = ["Position", "Health", "Team"]
    worlds.register_all(w, shorts)
    w.spawn_batch([worlds.cname(x) for x in shorts], n, [worlds.column(x, n, 9000 + i) for i, x in enumerate(shorts)],
                  want_entities=False)
    w.message_register(worlds.COLLISION_EVENT, 24, capi.MSG_MANUAL)
    P, H, T = (worlds.cname(x) for x in shorts)
    s = w.schedule("Update")
    s.add_system("S0_consume_strong", "consume_strong", None, params=struct.pack("<If", w.slots[H], 2.5),
                 reads_message=worlds.COLLISION_EVENT, extra_writes=[H], flags=capi.SYSTEM_CONSUMES_MESSAGES)
    s.add_system("S1_consume_rest", "consume", None, params=struct.pack("<I", w.slots[H]), reads_message=worlds.COLLISION_EVENT,
                 extra_writes=[H])
    s.add_system("S2_detect", "detect", w.query([(P, False), (T, False)]), writes_message=worlds.COLLISION_EVENT)
    s.build()
    out = {"order": s.execution_order()}
    for frame in range(frames):
        s.run_and_wait(mode)
        out[f"frame{frame}_counts_after_run"] = list(w.message_counts(worlds.COLLISION_EVENT))
        out[f"frame{frame}_messages_after_run"] = _bytes_digest(w.message_read(worlds.COLLISION_EVENT, 0),
                                                                  w.message_read(worlds.COLLISION_EVENT, 1))
        w.update()
        out[f"frame{frame}_counts_after_update"] = list(w.message_counts(worlds.COLLISION_EVENT))
    out["world"] = snapshot_digest(w.snapshot())
    w.close()
    return out


def churn_small(backend, seed=3, frames=4):
    """Many commands per entity on tiny tables: long swap-and-pop chains, tables emptied and refilled, new archetypes,
    stale ids, Try-variants and Clear."""
    shorts = ["Position", "Velocity", "Health", "Team", "Damping"]
    w = capi.World(backend)
    worlds.register_all(w, shorts)
    tables = [(["Position", "Velocity"], 7), (["Position", "Velocity", "Health"], 5), (["Position"], 3), (["Health", "Team"], 6)]
    for i, (sh, rows) in enumerate(tables):
        w.spawn_batch([worlds.cname(x) for x in sh], rows, [worlds.column(x, rows, 40 + 8 * i + k, seed=seed) for k, x in enumerate(sh)])
    out = {}
    stale = []
    for f in range(frames):
        rng = np.random.default_rng(seed * 100 + f)
        snap = w.snapshot(columns=False)
        live = np.concatenate([a["entities"] for a in snap["archetypes"]])
        n_cmd = 60
        spawned = w.reserve_entities(6)
        pool = np.concatenate([live, spawned, np.asarray(stale, dtype=np.uint64)])
        cmds = np.zeros(n_cmd, dtype=capi.COMMAND_DTYPE)
        payload = bytearray()
        try_kinds = {capi.CMD_ADD: capi.CMD_TRY_ADD, capi.CMD_REMOVE: capi.CMD_TRY_REMOVE, capi.CMD_DESTROY: capi.CMD_TRY_DESTROY}
        for i in range(n_cmd):
            ent = int(pool[rng.integers(0, pool.size)])
            kind = int(rng.choice([capi.CMD_ADD, capi.CMD_REMOVE, capi.CMD_DESTROY, capi.CMD_CLEAR, capi.CMD_ADD, capi.CMD_REMOVE]))
            if kind != capi.CMD_CLEAR:
                kind = try_kinds[kind]  # every record tolerates dead targets (the reference asserts on the non-Try forms)
            k = int(rng.integers(1, 3))
            sh = list(rng.choice(shorts, size=k, replace=False)) if kind in (capi.CMD_TRY_ADD, capi.CMD_TRY_REMOVE) else []
            cmds[i]["kind"] = kind
            cmds[i]["entity"] = ent
            cmds[i]["component_mask"] = w.mask(worlds.cname(x) for x in sh)
            cmds[i]["payload_offset"] = len(payload)
            if kind == capi.CMD_TRY_ADD:
                for x in sorted(sh, key=lambda y: w.slots[worlds.cname(y)]):
                    payload += rng.integers(1, 2 ** 31 - 1, size=worlds.COMPONENTS[x][0] // 4, dtype=np.uint32).tobytes()
            if kind == capi.CMD_TRY_DESTROY:
                stale.append(ent)
        # Clear on a dead entity is an assert in the reference: retarget those records at live entities
        dead = set(stale)
        for i in range(n_cmd):
            if cmds[i]["kind"] == capi.CMD_CLEAR and (int(cmds[i]["entity"]) in dead):
                cmds[i]["kind"] = capi.CMD_TRY_REMOVE
        w.enqueue(cmds, bytes(payload))
        w.flush()
        out[f"frame{f}"] = snapshot_digest(w.snapshot())
    w.close()
    return out


SPARSE_TYPES = {"PlayerTag": 0, "FrozenTag": 0, "SparseBuff": 4, "SparseVelocity": 8}


def sparse_world(backend, rows=(37, 21, 12), entity_capacity=0):
    """Archetype tables {Position, Health}, {Position, Health, Team}, {Health} plus four sparse-set types (two tags)."""
    w = capi.World(backend, entity_capacity=entity_capacity)
    worlds.register_all(w, ["Position", "Health", "Team"])
    for name, size in SPARSE_TYPES.items():
        w.register_sparse_component(worlds.cname(name), size)
    tables = [["Position", "Health"], ["Position", "Health", "Team"], ["Health"]]
    for i, (sh, n) in enumerate(zip(tables, rows)):
        w.spawn_batch([worlds.cname(x) for x in sh], n, [worlds.column(x, n, 700 + 8 * i + k) for k, x in enumerate(sh)])
    return w


def sparse_commands(w, rng, pool, n_cmd, stale, arch_shorts=("Team", "Position"), allow_destroy=True):
    """Random records over archetype AND sparse components: Add / TryAdd / Remove / TryRemove / Clear / Destroy."""
    cmds = np.zeros(n_cmd, dtype=capi.COMMAND_DTYPE)
    payload = bytearray()
    names = list(SPARSE_TYPES) + list(arch_shorts)
    sizes = {**SPARSE_TYPES, **{x: worlds.COMPONENTS[x][0] for x in arch_shorts}}
    dead = set(stale)
    for i in range(n_cmd):
        ent = int(pool[rng.integers(0, pool.size)])
        kinds = [capi.CMD_TRY_ADD, capi.CMD_TRY_ADD, capi.CMD_ADD, capi.CMD_TRY_REMOVE, capi.CMD_REMOVE, capi.CMD_CLEAR]
        if allow_destroy:
            kinds.append(capi.CMD_TRY_DESTROY)
        kind = int(rng.choice(kinds))
        if ent in dead and kind in (capi.CMD_ADD, capi.CMD_REMOVE, capi.CMD_CLEAR):
            kind = capi.CMD_TRY_REMOVE  # the non-Try forms assert on dead entities in the reference
        k = int(rng.integers(1, 4))
        sh = list(rng.choice(names, size=k, replace=False)) if kind not in (capi.CMD_CLEAR, capi.CMD_TRY_DESTROY) else []
        cmds[i]["kind"] = kind
        cmds[i]["entity"] = ent
        cmds[i]["component_mask"] = w.mask(worlds.cname(x) for x in sh)
        cmds[i]["payload_offset"] = len(payload)
        if kind in (capi.CMD_ADD, capi.CMD_TRY_ADD):
            for x in sorted(sh, key=lambda y: w.slots[worlds.cname(y)]):
                if sizes[x]:
                    payload += rng.integers(1, 90, size=sizes[x] // 4, dtype=np.uint32).tobytes()
        if kind == capi.CMD_TRY_DESTROY:
            stale.append(ent)
            dead.add(ent)
    return cmds, bytes(payload)


def sparse_state(w):
    """Digest of the archetype snapshot plus every sparse storage in dense order (reverse map + value bytes)."""
    parts = [snapshot_digest(w.snapshot())]
    for name in SPARSE_TYPES:
        sp = w.snapshot_sparse(worlds.cname(name))
        parts.append(_bytes_digest(sp["indices"], sp["data"]))
    return parts


def sparse_sets(backend, seed=11, frames=4, n_cmd=90, mode=capi.EXEC_SERIAL):
    """Sparse-set components end to end: deferred commands over tags and sparse data (dense order after swap-and-pop),
    Destroy / Clear purging every storage, queries with sparse With / Without / data terms under the three rules of
    helios_gpu_query_desc::sparse_rule, a nullable `T*` fetch, and systems running behind those filters."""
    w = sparse_world(backend)
    P, H, T = (worlds.cname(x) for x in ("Position", "Health", "Team"))
    TAG, FROZEN, BUFF = (worlds.cname(x) for x in ("PlayerTag", "FrozenTag", "SparseBuff"))
    queries = {}
    for rule in (capi.SPARSE_STRICT, capi.SPARSE_FOREACH, capi.SPARSE_WITH_ENTITY):
        queries[f"with_tag_r{rule}"] = w.query([(H, False)], with_=[TAG], sparse_rule=rule)
        queries[f"without_frozen_r{rule}"] = w.query([(H, False)], with_=[P], without=[FROZEN], sparse_rule=rule)
        queries[f"tag_not_frozen_buff_r{rule}"] = w.query([(H, False)], with_=[TAG], without=[FROZEN, T], sparse_data=[BUFF],
                                                          sparse_rule=rule)
    s = w.schedule("Update")
    s.add_system("regen_players", "regen", w.query([(H, True)], with_=[TAG]))
    s.add_system("buff_heal", "buff_heal", w.query([(H, True)], without=[FROZEN], sparse_data=[BUFF]))
    ordinal = list(SPARSE_TYPES).index("SparseBuff")
    s.add_system("buff_optional", "buff_optional", w.query([(H, True)], with_=[P]),
                 params=struct.pack("<QII", capi.fnv1a(BUFF), ordinal, 0))
    s.build()
    out = {"order": s.execution_order()}
    stale = []
    for f in range(frames):
        rng = np.random.default_rng(seed * 1000 + f)
        live = np.concatenate([a["entities"] for a in w.snapshot(columns=False)["archetypes"]])
        spawned = w.reserve_entities(5)
        pool = np.concatenate([live, spawned, np.asarray(stale[-8:], dtype=np.uint64)])
        cmds, payload = sparse_commands(w, rng, pool, n_cmd, stale)
        w.enqueue(cmds, payload)
        w.flush()
        out[f"frame{f}_state"] = sparse_state(w)
        out[f"frame{f}_counts"] = {k: q.count() for k, q in queries.items()}
        s.run_and_wait(mode)
        out[f"frame{f}_after_systems"] = snapshot_digest(w.snapshot())
    w.close()
    return out


def run_conditions(backend, n=4099, frames=6, mode=capi.EXEC_SERIAL):
    """Run conditions (run_condition.hpp; schedule.cpp:427-461; executor/main_thread.cpp:38-52) of every kind, evaluated on
    the device: verdicts flip from frame to frame because of what the guarded systems themselves do.
      S1 damp     — own: CountIf(pos.x > 512) > its initial value (a query terminal as the predicate); set MOVE: constant true
      S2 regen    — Res: last frame's Σhp (written by S5) below a threshold -> stops once the world has healed
      S3 consume  — MessageReader non-empty
      S4 despawn  — WorldView: a sentinel entity is alive (the host destroys it before frame 3)
      S6 ghost    — member of set OFF (constant false): never runs, its integrate never moves anything
      S7 detect   — emits only while the channel is empty -> runs every other frame
    Recorded per frame: which systems ran, message counts, resources, world digest."""
    w = capi.World(backend)
    shorts = ["Position", "Velocity", "Health", "Team"]
    worlds.register_all(w, shorts)
    w.spawn_batch([worlds.cname(x) for x in shorts], n, [worlds.column(x, n, 6100 + i) for i, x in enumerate(shorts)],
                  want_entities=False)
    T_ = worlds.cname("Team")
    sentinel = w.spawn_batch([T_], 1, [worlds.column("Team", 1, 6200)])
    sentinel = int(np.asarray(sentinel).reshape(-1)[0])
    w.message_register(worlds.COLLISION_EVENT, 24)
    P, V, H, T = (worlds.cname(x) for x in shorts)
    hp0 = int(worlds.column("Health", n, 6102).astype(np.int64).sum())
    probe = w.query([(P, False)])
    count0 = probe.count_if_pos_x_gt(512.0)
    MOVE, OFF = 11, 12
    s = w.schedule("Update")
    s0 = s.add_system("S0_integrate", "integrate", w.query([(P, True), (V, False)]))
    s1 = s.add_system("S1_damp", "damp", w.query([(V, True)]), params=struct.pack("<f", 0.5))
    s2 = s.add_system("S2_regen", "regen", w.query([(H, True)]), read_resources=["helios::bench::Stats"])
    s3 = s.add_system("S3_consume", "consume", None, params=struct.pack("<I", w.slots[H]), reads_message=worlds.COLLISION_EVENT,
                      extra_writes=[H])
    s4 = s.add_system("S4_despawn_dead", "despawn_dead", w.query([(H, False)]), flags=capi.SYSTEM_EMITS_COMMANDS)
    s.add_system("S5_stats", "stats", w.query([(H, False)]), write_resources=["helios::bench::Stats"])
    s6 = s.add_system("S6_ghost", "integrate", w.query([(P, True), (V, False)]))
    s7 = s.add_system("S7_detect", "detect", w.query([(P, False), (T, False)]), writes_message=worlds.COLLISION_EVENT)
    s.set_run_if(MOVE, "AlwaysTrue", capi.COND_CONSTANT, rhs=1)
    s.set_run_if(OFF, "AlwaysFalse", capi.COND_CONSTANT, rhs=0)
    s.in_set(s0, MOVE)
    s.in_set(s1, MOVE)
    s.in_set(s6, MOVE)
    s.in_set(s6, OFF)
    s.run_if(s1, "MoreOnTheRight", capi.COND_TERMINAL, builtin="count_if_pos_x_gt", query=w.query([(P, False)]),
             params=struct.pack("<f", 512.0), field=capi.COND_FIELD_COUNT, cmp=capi.CMP_GT, rhs=count0)
    s.run_if(s2, "NotHealedYet", capi.COND_RESOURCE, offset=0, cmp=capi.CMP_LT, rhs=hp0 - hp0 // 8)
    s.run_if(s3, "HasEvents", capi.COND_MESSAGES, reads_message=worlds.COLLISION_EVENT, cmp=capi.CMP_GT, rhs=0)
    s.run_if(s4, "SentinelAlive", capi.COND_CUSTOM, builtin="entity_alive", params=struct.pack("<Q", sentinel))
    s.run_if(s7, "ChannelEmpty", capi.COND_MESSAGES, reads_message=worlds.COLLISION_EVENT, cmp=capi.CMP_EQ, rhs=0)
    s.build()
    out = {"order": s.execution_order(), "count0": count0}
    for frame in range(frames):
        if frame == 3:
            cmds = np.zeros(1, dtype=capi.COMMAND_DTYPE)
            cmds[0]["kind"] = capi.CMD_DESTROY
            cmds[0]["entity"] = sentinel
            w.enqueue(cmds)
            w.flush()
        s.run_and_wait(mode)
        out[f"frame{frame}_ran"] = s.last_verdicts()
        out[f"frame{frame}_counts_after_run"] = list(w.message_counts(worlds.COLLISION_EVENT))
        w.update()
        out[f"frame{frame}_resources"] = s.read_resources(16).tobytes().hex()
        out[f"frame{frame}_world"] = snapshot_digest(w.snapshot())
    w.close()
    return out


SCORE_EVENT = worlds.NS + "ScoreEvent"  # {Entity who; i32 points; u32 pad} = 16 B


def multi_messages(backend, n=4099, frames=3, mode=capi.EXEC_SERIAL, tight=False):
    """Several MessageWriter / MessageReader parameters in one system (system/param_traits.hpp:193-240): S0 reads both
    types, S1 writes two CollisionEvents and one ScoreEvent per row (3 messages per row of two types).  `tight` leaves
    writer 0 at its default capacity (one message per row): the run must REPORT the overflow."""
    w = capi.World(backend)
    shorts = ["Position", "Health", "Team"]
    worlds.register_all(w, shorts)
    w.spawn_batch([worlds.cname(x) for x in shorts], n, [worlds.column(x, n, 9500 + i) for i, x in enumerate(shorts)],
                  want_entities=False)
    w.message_register(worlds.COLLISION_EVENT, 24)
    w.message_register(SCORE_EVENT, 16)
    P, H, T = (worlds.cname(x) for x in shorts)
    s = w.schedule("Update")
    s.add_system("S0_consume_two", "consume_two", None, params=struct.pack("<I", w.slots[H]),
                 reads_message=worlds.COLLISION_EVENT, more_reads=[SCORE_EVENT], extra_writes=[H])
    s.add_system("S1_emit_multi", "emit_multi", w.query([(P, False), (T, False)]), writes_message=worlds.COLLISION_EVENT,
                 writer_capacity=0 if tight else 2 * n, more_writes=[(SCORE_EVENT, 0)])
    s.build()
    out = {}
    for frame in range(frames):
        s.run_and_wait(mode)
        out[f"frame{frame}_counts"] = [list(w.message_counts(worlds.COLLISION_EVENT)), list(w.message_counts(SCORE_EVENT))]
        out[f"frame{frame}_messages"] = _bytes_digest(w.message_read(worlds.COLLISION_EVENT, 1), w.message_read(SCORE_EVENT, 1))
        w.update()
    out["world"] = snapshot_digest(w.snapshot())
    w.close()
    return out


SCENARIOS = {
    "cfg1_n4099": lambda b: cfg1(b),
    "cfg1_n1": lambda b: cfg1(b, n=1),
    "cfg2_filtered": lambda b: cfg2(b, filtered=True),
    "cfg2_all": lambda b: cfg2(b, filtered=False),
    "cfg3_churn": lambda b: cfg3(b),
    "cfg4_schedule_messages": lambda b: cfg4(b),
    "consumable_messages": lambda b: consumable(b),
    "churn_small_s3": lambda b: churn_small(b, seed=3),
    "sparse_sets": lambda b: sparse_sets(b),
    "churn_small_s11": lambda b: churn_small(b, seed=11),
    "run_conditions": lambda b: run_conditions(b),
    "multi_messages": lambda b: multi_messages(b),
}
