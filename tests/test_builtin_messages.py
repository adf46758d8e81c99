"""Builtin structural messages (reference builtin_messages.hpp; written by every structural World call, world.hpp:979-1036,
1124-1196; the reference's own cases are src/ecs/tests/builtin_messages.cpp:67-701), driven through the C ABI on both
backends.  The CPU suite pins the oracle to the expected answers; the GPU suite pins the CUDA path to the same answers."""
import struct

import numpy as np
import pytest

from heliosengine_b200 import capi, worlds

P, V, H = (worlds.cname(x) for x in ("Position", "Velocity", "Health"))
SCORE = "helios::bench::Score"


def _cmds(w, specs):
    """specs: (kind, entity, [component names]); Add payloads are zero bytes."""
    cmds = np.zeros(len(specs), dtype=capi.COMMAND_DTYPE)
    payload = bytearray()
    for i, (kind, ent, names) in enumerate(specs):
        cmds[i]["kind"] = kind
        cmds[i]["entity"] = ent
        cmds[i]["component_mask"] = w.mask(names)
        cmds[i]["payload_offset"] = len(payload)
        if kind in (capi.CMD_ADD, capi.CMD_TRY_ADD):
            for nm in sorted(names, key=lambda x: w.slots[x]):
                payload += bytes(w.sizes[nm])
    return cmds, bytes(payload)


def _counts(w, kind, comp=None):
    """[previous, current] entity lists of one builtin message kind."""
    return [w.builtin_messages(kind, comp, 0).tolist(), w.builtin_messages(kind, comp, 1).tolist()]


def cases(backend):
    out = {}
    # builtin_messages.cpp:67-116 EntityAddedMsg: one per CreateEntity, in `previous` after Update, gone after the next
    w = capi.World(backend)
    worlds.register_all(w, ["Position", "Velocity", "Health"])
    kinds = [capi.BUILTIN_ENTITY_ADDED, capi.BUILTIN_ENTITY_DESTROYED, capi.BUILTIN_COMPONENTS_CLEARED]
    registered = dict(w.message_types())
    out["builtin_registered_at_creation"] = [w.builtin_message_type(k) in registered for k in kinds]  # world.hpp:42,1343-1347
    e = w.create_entities(2)
    out["created_current"] = _counts(w, capi.BUILTIN_ENTITY_ADDED) == [[], e.tolist()]
    w.update()
    out["created_previous"] = _counts(w, capi.BUILTIN_ENTITY_ADDED) == [e.tolist(), []]
    w.update()
    out["created_cleared"] = _counts(w, capi.BUILTIN_ENTITY_ADDED)
    # Commands::Spawn = reserve + flush: no EntityAddedMsg (world.hpp:957-963 vs :979-984)
    r = w.reserve_entities(3)
    w.flush()
    out["spawn_no_entity_added"] = [w.entity_count(), _counts(w, capi.BUILTIN_ENTITY_ADDED)]
    # :262-305 ComponentAddedMsg<T>; registration happens with the first valid Add (world.hpp:1033)
    added_p = w.builtin_message_type(capi.BUILTIN_COMPONENT_ADDED, P)
    out["added_unregistered_before"] = added_p in dict(w.message_types())
    c, pay = _cmds(w, [(capi.CMD_ADD, int(r[2]), [P]), (capi.CMD_ADD, int(e[0]), [P, V]), (capi.CMD_TRY_ADD, int(r[2]), [P, H]),
                       (capi.CMD_ADD, int(r[2]), [P]), (capi.CMD_TRY_ADD, int(e[1]), [V])])
    w.enqueue(c, pay)
    w.flush()
    out["added_registered_after"] = added_p in dict(w.message_types())
    # per type, command order; Add writes for every listed type (overwrite included), TryAdd only for what it added
    out["added_P"] = _counts(w, capi.BUILTIN_COMPONENT_ADDED, P) == [[], [int(r[2]), int(e[0]), int(r[2])]]
    out["added_V"] = _counts(w, capi.BUILTIN_COMPONENT_ADDED, V) == [[], [int(e[0]), int(e[1])]]
    out["added_H"] = _counts(w, capi.BUILTIN_COMPONENT_ADDED, H) == [[], [int(r[2])]]
    w.update()
    w.update()
    out["added_cleared"] = _counts(w, capi.BUILTIN_COMPONENT_ADDED, P)
    # :346-390 ComponentRemovedMsg<T>: Remove = every listed type, TryRemove = only what was removed (world.hpp:1148-1161)
    c, pay = _cmds(w, [(capi.CMD_REMOVE, int(e[0]), [P]), (capi.CMD_TRY_REMOVE, int(e[0]), [P, V]),
                       (capi.CMD_TRY_REMOVE, int(e[1]), [H]), (capi.CMD_REMOVE, int(r[2]), [P, H])])
    w.enqueue(c, pay)
    w.flush()
    out["removed_P"] = _counts(w, capi.BUILTIN_COMPONENT_REMOVED, P) == [[], [int(e[0]), int(r[2])]]
    out["removed_V"] = _counts(w, capi.BUILTIN_COMPONENT_REMOVED, V) == [[], [int(e[0])]]
    out["removed_H"] = _counts(w, capi.BUILTIN_COMPONENT_REMOVED, H) == [[], [int(r[2])]]  # registered by e[1]'s Try, written by r[2]
    # :420-470 ComponentsClearedMsg: also for an entity without components; :168-210 EntityDestroyedMsg; Try on a dead id
    c, pay = _cmds(w, [(capi.CMD_CLEAR, int(e[1]), []), (capi.CMD_CLEAR, int(r[0]), []), (capi.CMD_DESTROY, int(r[1]), []),
                       (capi.CMD_TRY_DESTROY, int(r[1]), []), (capi.CMD_TRY_DESTROY, int(e[1]), []), (capi.CMD_TRY_ADD, int(r[1]), [P])])
    w.enqueue(c, pay)
    w.flush()
    out["cleared"] = _counts(w, capi.BUILTIN_COMPONENTS_CLEARED) == [[], [int(e[1]), int(r[0])]]
    out["destroyed"] = _counts(w, capi.BUILTIN_ENTITY_DESTROYED) == [[], [int(r[1]), int(e[1])]]
    w.update()
    out["destroyed_previous"] = _counts(w, capi.BUILTIN_ENTITY_DESTROYED) == [[int(r[1]), int(e[1])], []]
    w.update()
    out["destroyed_cleared"] = _counts(w, capi.BUILTIN_ENTITY_DESTROYED)
    # the bulk loader is CreateEntity + AddComponents per row: EntityAddedMsg and ComponentAddedMsg<T> per type
    s = w.spawn_batch([P, H], 4, [np.zeros(8, np.float32), np.arange(4, dtype=np.int32)])
    out["spawn_batch"] = [_counts(w, capi.BUILTIN_ENTITY_ADDED) == [[], s.tolist()],
                          _counts(w, capi.BUILTIN_COMPONENT_ADDED, P) == [[], s.tolist()],
                          _counts(w, capi.BUILTIN_COMPONENT_ADDED, H) == [[], s.tolist()],
                          _counts(w, capi.BUILTIN_COMPONENT_ADDED, V)]
    # CreateEntity recycles the two destroyed indices first (LIFO): the messages carry the bumped generation
    out["recycled_generation"] = [[int(x) >> 32 for x in s], [int(x) & 0xFFFFFFFF for x in s[:2]] == [int(e[1]) & 0xFFFFFFFF, int(r[1]) & 0xFFFFFFFF]]
    # :475-640 ResourceInsertedMsg<T> / ResourceRemovedMsg<T>: one byte per message
    ins = capi.fnv1a("helios::ecs::ResourceInsertedMsg") ^ ((capi.fnv1a(SCORE) * capi._FNV_PRIME) & 0xFFFFFFFFFFFFFFFF)
    rem = capi.fnv1a("helios::ecs::ResourceRemovedMsg") ^ ((capi.fnv1a(SCORE) * capi._FNV_PRIME) & 0xFFFFFFFFFFFFFFFF)
    w.resource_insert(SCORE, struct.pack("<i", 42))
    w.resource_insert(SCORE, struct.pack("<i", 43))           # "Replacing an existing resource also emits a message"
    tried = w.resource_try_insert(SCORE, struct.pack("<i", 1))  # present: registered, nothing written
    out["resource_inserted"] = [tried, w.message_read_hash(ins, 1, 1).size]
    out["resource_removed"] = [w.resource_try_remove(SCORE), w.resource_try_remove(SCORE), w.message_read_hash(rem, 1, 1).size]
    w.close()
    # throughput mode: HELIOS_GPU_WORLD_NO_BUILTIN_MESSAGES — only the three types of World() are registered, nothing is written
    w = capi.World(backend, flags=capi.WORLD_NO_BUILTIN_MESSAGES)
    worlds.register_all(w, ["Position"])
    e = w.create_entities(2)
    c, pay = _cmds(w, [(capi.CMD_ADD, int(e[0]), [P]), (capi.CMD_DESTROY, int(e[1]), [])])
    w.enqueue(c, pay)
    w.flush()
    out["switch_off"] = [len(w.message_types()), sum(p.size + c.size for p, c in w.snapshot_messages().values())]
    w.close()
    return out


EXPECTED = {
    "builtin_registered_at_creation": [True, True, True], "created_current": True, "created_previous": True,
    "created_cleared": [[], []], "spawn_no_entity_added": [5, [[], []]], "added_unregistered_before": False,
    "added_registered_after": True, "added_P": True, "added_V": True, "added_H": True, "added_cleared": [[], []],
    "removed_P": True, "removed_V": True, "removed_H": True, "cleared": True, "destroyed": True, "destroyed_previous": True,
    "destroyed_cleared": [[], []], "spawn_batch": [True, True, True, [[], []]], "recycled_generation": [[2, 2, 1, 1], True],
    "resource_inserted": [False, 2], "resource_removed": [True, False, 1], "switch_off": [3, 0],
}


def test_oracle_builtin_messages(oracle):
    assert cases(oracle) == EXPECTED


@pytest.mark.gpu
def test_gpu_builtin_messages(gpu):
    assert cases(gpu) == EXPECTED


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 255, 256, 257, 5000, 70001])
def test_gpu_builtin_messages_tile_boundaries(gpu, oracle, n):
    """Ordered per-type compaction across tile / warp boundaries: n mixed commands over n + 64 entities, every kind,
    several types per command; every queue must equal the oracle's byte for byte (snapshots include the message queues)."""
    snaps = []
    for b in (gpu, oracle):
        w = capi.World(b)
        shorts = ["Position", "Velocity", "Health", "Team", "Damping"]
        worlds.register_all(w, shorts)
        w.spawn_batch([P], n + 64, [np.zeros(2 * (n + 64), np.float32)], want_entities=False)
        w.update()
        w.update()
        rng = np.random.default_rng(n)
        ents = capi.make_entity(rng.permutation(n + 64)[:n])  # distinct live targets
        kinds = rng.choice([capi.CMD_ADD, capi.CMD_TRY_ADD, capi.CMD_REMOVE, capi.CMD_TRY_REMOVE, capi.CMD_CLEAR, capi.CMD_DESTROY,
                            capi.CMD_TRY_DESTROY], size=n)
        specs = []
        for k, e in zip(kinds, ents):
            names = [worlds.cname(x) for x in rng.choice(shorts, size=int(rng.integers(1, 4)), replace=False)]
            if k == capi.CMD_REMOVE:
                names = [P]  # the non-Try form asserts on absent components in the reference
            specs.append((int(k), int(e), names if k in (capi.CMD_ADD, capi.CMD_TRY_ADD, capi.CMD_REMOVE, capi.CMD_TRY_REMOVE) else []))
        c, pay = _cmds(w, specs)
        w.enqueue(c, pay)
        w.flush()
        snaps.append(w.snapshot())
        w.close()
    assert len(snaps[0]["messages"]) >= 4 and (n < 5000 or len(snaps[0]["messages"]) >= 8)
    assert not worlds.snapshots_equal(snaps[0], snaps[1])
