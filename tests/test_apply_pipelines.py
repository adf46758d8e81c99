"""The two command pipelines of the CUDA library — the steady-state one (csrc/apply_fast.cu: one command per entity, every
destination archetype exists, ten launches and a single read-back) and the general one (csrc/apply.cu) — held against each
other and against the oracle's sequential FIFO replay (reference command/queue.hpp:158-163, component/manager.hpp:1893-1952,
component/archetype.hpp:408-442).  Bar: bit-exact snapshots (ids, generations, free list, archetype and row order, column
bytes, every message queue)."""
import numpy as np
import pytest

from heliosengine_b200 import capi, worlds

pytestmark = pytest.mark.gpu

SHORTS = ["Position", "Velocity", "Acceleration", "Mass", "Health", "Team", "Damping", "Frozen"]


def _world(backend, tables, flags=0):
    w = capi.World(backend, flags=flags)
    worlds.register_all(w, SHORTS)
    for i, (shorts, rows) in enumerate(tables):
        cols = [worlds.column(s, rows, 800 + 16 * i + k) for k, s in enumerate(shorts)]
        w.spawn_batch([worlds.cname(s) for s in shorts], rows, cols, want_entities=False)
    return w


def _frame(w, live_by_table, types_by_table, rng, n_cmd, n_spawn, spawned):
    """Distinct targets (one command per entity): Destroy / Add Health / Remove Damping / TryAdd Team / Clear / Remove on a
    missing component (TryRemove) + spawns with components.  Every destination archetype exists after frame 0."""
    all_e = np.concatenate([e for e in live_by_table if e.size])
    targets = rng.choice(all_e, size=min(n_cmd, all_e.size), replace=False)
    kinds = rng.choice(6, size=targets.size)
    specs = [(capi.CMD_ADD, int(e), ["Position", "Velocity", "Acceleration", "Mass"]) for e in spawned]
    for k, e in zip(kinds, targets):
        specs.append([(capi.CMD_DESTROY, int(e), []), (capi.CMD_ADD, int(e), ["Health"]), (capi.CMD_TRY_REMOVE, int(e), ["Damping"]),
                      (capi.CMD_TRY_ADD, int(e), ["Team", "Health"]), (capi.CMD_TRY_DESTROY, int(e), []),
                      (capi.CMD_TRY_REMOVE, int(e), ["Frozen", "Team"])][k])
    order = rng.permutation(len(specs))
    specs = [specs[i] for i in order]
    cmds = np.zeros(len(specs), dtype=capi.COMMAND_DTYPE)
    payload = bytearray()
    for i, (kind, ent, shorts) in enumerate(specs):
        cmds[i]["kind"] = kind
        cmds[i]["entity"] = ent
        cmds[i]["component_mask"] = w.mask(worlds.cname(s) for s in shorts)
        cmds[i]["payload_offset"] = len(payload)
        if kind in (capi.CMD_ADD, capi.CMD_TRY_ADD):
            for s in sorted(shorts, key=lambda x: w.slots[worlds.cname(x)]):
                payload += rng.integers(1, 2 ** 31 - 1, size=worlds.COMPONENTS[s][0] // 4, dtype=np.uint32).tobytes()
    return cmds, bytes(payload)


def _run(gpu, oracle, tables, frames, n_cmd, n_spawn, seed, expect_fast_from=1):
    ws = [_world(gpu, tables), _world(gpu, tables, flags=capi.WORLD_GENERAL_APPLY), _world(oracle, tables)]
    pipelines = []
    for f in range(frames):
        snap = ws[2].snapshot(columns=False)
        live = [a["entities"] for a in snap["archetypes"]]
        types = [a["types"] for a in snap["archetypes"]]
        snaps = []
        for w in ws:
            spawned = w.reserve_entities(n_spawn)
            cmds, payload = _frame(w, live, types, np.random.default_rng(seed * 100 + f), n_cmd, n_spawn, spawned)
            w.enqueue(cmds, payload)
            w.flush()
            snaps.append(w.snapshot())
            w.update()
        pipelines.append((ws[0].apply_stats().pipeline, ws[1].apply_stats().pipeline))
        d_fast = worlds.snapshots_equal(snaps[0], snaps[2])
        d_gen = worlds.snapshots_equal(snaps[1], snaps[2])
        assert not d_gen, (f, "general pipeline vs oracle", d_gen[:4])
        assert not d_fast, (f, "steady-state pipeline vs oracle", d_fast[:4], pipelines)
    assert all(p[1] == 0 for p in pipelines)
    assert all(p[0] == 1 for p in pipelines[expect_fast_from:]), pipelines  # frame 0 may create archetypes / grow queues
    for w in ws:
        w.close()


@pytest.mark.parametrize("rows,n_cmd,n_spawn", [(7, 9, 2), (300, 200, 40), (5000, 3000, 700), (40000, 30000, 5000)])
def test_pipelines_agree_small_to_medium(gpu, oracle, rows, n_cmd, n_spawn):
    tables = [(["Position", "Velocity", "Acceleration", "Mass"], rows), (["Position", "Velocity", "Acceleration", "Mass", "Health"], rows),
              (["Position", "Velocity", "Acceleration", "Mass", "Damping"], rows // 2 + 1),
              (["Position", "Velocity", "Acceleration", "Mass", "Health", "Damping", "Team"], rows // 3 + 1),
              (["Position", "Health", "Frozen", "Team"], rows // 4 + 1)]
    _run(gpu, oracle, tables, frames=5, n_cmd=n_cmd, n_spawn=n_spawn, seed=rows, expect_fast_from=3)


def test_pipelines_agree_tail_heavy(gpu, oracle):
    """Removals concentrated on the tail of one table (long swap-and-pop chains: the removed entity is itself a row that was
    moved earlier in the batch) while another table only grows."""
    tables = [(["Position", "Velocity"], 4000), (["Position", "Velocity", "Health"], 50)]
    ws = [_world(gpu, tables), _world(gpu, tables, flags=capi.WORLD_GENERAL_APPLY), _world(oracle, tables)]
    for f in range(4):
        snap = ws[2].snapshot(columns=False)
        t0 = snap["archetypes"][1]["entities"]  # archetype 0 is `{}`
        rng = np.random.default_rng(77 + f)
        tail = t0[-600:]
        victims = rng.permutation(tail)[:450]
        movers = rng.permutation(t0[:-600])[:300]
        snaps = []
        for w in ws:
            cmds = np.zeros(victims.size + movers.size, dtype=capi.COMMAND_DTYPE)
            cmds["entity"] = np.concatenate([victims, movers])
            cmds["kind"] = np.concatenate([np.full(victims.size, capi.CMD_DESTROY), np.full(movers.size, capi.CMD_ADD)])
            cmds["component_mask"] = np.concatenate([np.zeros(victims.size, np.uint64),
                                                     np.full(movers.size, w.mask([worlds.cname("Health")]), np.uint64)])
            perm = np.random.default_rng(5 + f).permutation(cmds.size)
            cmds = cmds[perm]
            cmds["payload_offset"] = 0
            w.enqueue(cmds, np.arange(4, dtype=np.uint32))
            w.flush()
            snaps.append(w.snapshot())
        assert not worlds.snapshots_equal(snaps[1], snaps[2]), f
        assert not worlds.snapshots_equal(snaps[0], snaps[2]), f
        assert ws[0].apply_stats().pipeline == 1 or f == 0
    for w in ws:
        w.close()


def test_duplicates_and_new_archetypes_fall_back(gpu, oracle):
    """Two commands on one entity, or a destination archetype that does not exist yet: the steady-state pipeline aborts before
    it writes anything and the general one applies the batch."""
    tables = [(["Position", "Velocity"], 100)]
    wg, wo = _world(gpu, tables), _world(oracle, tables)
    ents = wo.snapshot(columns=False)["archetypes"][1]["entities"]
    for w in (wg, wo):
        cmds = np.zeros(3, dtype=capi.COMMAND_DTYPE)
        cmds["entity"] = [ents[3], ents[3], ents[9]]
        cmds["kind"] = [capi.CMD_ADD, capi.CMD_REMOVE, capi.CMD_ADD]
        cmds["component_mask"] = [w.mask([worlds.cname("Health")]), w.mask([worlds.cname("Velocity")]), w.mask([worlds.cname("Team")])]
        cmds["payload_offset"] = [0, 0, 4]
        w.enqueue(cmds, np.arange(2, dtype=np.uint32))
        w.flush()
    assert wg.apply_stats().pipeline == 0 and wg.apply_stats().fast_fallbacks == 1
    assert not worlds.snapshots_equal(wg.snapshot(), wo.snapshot())
    wg.close()
    wo.close()
