"""GPU parity: deferred Commands (spawn / despawn / insert / remove / clear, with archetype migration) applied as one
parallel batch on the device vs the oracle's strictly sequential FIFO execution.

Bar: post-Commands world state bit-exact — entity ids and generations, free list, archetype creation order, table row
order (append + swap-and-pop), every column byte.
"""
import numpy as np
import pytest

from heliosengine_b200 import capi, worlds

pytestmark = pytest.mark.gpu

SHORTS = ["Position", "Velocity", "Acceleration", "Mass", "Health", "Team", "Damping", "Frozen"]


def make_world(backend, tables, rng_seed=1):
    """tables: list of (component short names, rows)."""
    w = capi.World(backend)
    worlds.register_all(w, SHORTS)
    ents = []
    for i, (shorts, rows) in enumerate(tables):
        cols = [worlds.column(s, rows, 300 + 16 * i + k, seed=rng_seed) for k, s in enumerate(shorts)]
        ents.append(w.spawn_batch([worlds.cname(s) for s in shorts], rows, cols))
    return w, (np.concatenate(ents) if ents else np.empty(0, np.uint64))


def build_commands(w, specs, rng):
    """specs: list of (kind, entity, [component shorts]); returns (records, payload bytes)."""
    cmds = np.zeros(len(specs), dtype=capi.COMMAND_DTYPE)
    payload = bytearray()
    for i, (kind, ent, shorts) in enumerate(specs):
        cmds[i]["kind"] = kind
        cmds[i]["entity"] = ent
        cmds[i]["component_mask"] = w.mask(worlds.cname(s) for s in shorts)
        cmds[i]["payload_offset"] = len(payload)
        if kind in (capi.CMD_ADD, capi.CMD_TRY_ADD):
            for s in sorted(shorts, key=lambda x: w.slots[worlds.cname(x)]):  # ascending slot order, 4-byte padded
                size = worlds.COMPONENTS[s][0]
                payload += rng.integers(1, 2 ** 31 - 1, size=size // 4, dtype=np.uint32).tobytes()
    return cmds, bytes(payload)


def run_both(gpu, oracle, tables, make_specs, frames=1, seed=7):
    """Builds the same world on both backends and applies the same command frames; returns both snapshots per frame."""
    out = []
    ws = []
    for b in (gpu, oracle):
        w, ents = make_world(b, tables)
        ws.append((w, ents))
    for f in range(frames):
        rng = np.random.default_rng(seed * 1000 + f)
        # entities are identical on both sides; generate specs once from the oracle-side snapshot of live entities
        snap = ws[1][0].snapshot(columns=False)
        live = np.concatenate([a["entities"] for a in snap["archetypes"]]) if snap["archetypes"] else np.empty(0, np.uint64)
        n_spawn, spec_fn = make_specs(rng, live, f)
        specs = None
        frame_snaps = []
        for w, _ in ws:
            spawned = w.reserve_entities(n_spawn)
            if specs is None:
                specs = spec_fn(spawned)
            cmds, payload = build_commands(w, specs, np.random.default_rng(seed * 77 + f))
            w.enqueue(cmds, payload)
            w.flush()
            frame_snaps.append(w.snapshot())
        diffs = worlds.snapshots_equal(frame_snaps[0], frame_snaps[1])
        assert not diffs, (f, diffs[:5])
        out.append(frame_snaps)
    stats = ws[0][0].apply_stats()
    return out, stats, ws


def test_basic_verbs(gpu, oracle):
    tables = [(["Position", "Velocity"], 5), (["Position", "Velocity", "Health"], 4), (["Position"], 3)]

    def make_specs(rng, live, f):
        def fn(sp):
            e = [int(x) for x in live]
            return [
                (capi.CMD_ADD, sp[0], ["Position", "Velocity"]),          # Spawn().AddComponents(...)
                (capi.CMD_DESTROY, e[1], []),                             # swap-and-pop in table 1
                (capi.CMD_ADD, e[0], ["Health"]),                         # migrate {P,V} -> {P,V,H}
                (capi.CMD_REMOVE, e[5], ["Health"]),                      # migrate back
                (capi.CMD_TRY_ADD, e[6], ["Health", "Team"]),             # Health kept, Team added -> new archetype
                (capi.CMD_ADD, e[7], ["Health"]),                         # overwrite only
                (capi.CMD_CLEAR, e[9], []),                               # -> {} table
                (capi.CMD_TRY_DESTROY, e[1], []),                         # already dead: skipped silently
                (capi.CMD_DESTROY, e[1], []),                             # already dead: counted invalid
                (capi.CMD_TRY_REMOVE, e[10], ["Mass"]),                   # absent component: no-op
                (capi.CMD_ADD, sp[1], ["Position"]),
                (capi.CMD_DESTROY, sp[1], []),                            # spawned and destroyed in one batch
                (capi.CMD_ADD, sp[2], ["Mass"]),
                (capi.CMD_ADD, sp[2], ["Position"]),                      # two migrations of one new entity
                (capi.CMD_REMOVE, sp[2], ["Mass"]),
            ]
        return 4, fn                                                      # sp[3] stays in {} (spawn without components)

    out, stats, ws = run_both(gpu, oracle, tables, make_specs)
    assert stats.invalid == 1
    assert ws[0][0].entity_count() == ws[1][0].entity_count()


def test_tail_removal_orders(gpu, oracle):
    """Adversarial orders for the swap-and-pop replay: remove the tail forwards, backwards, and holes then tail."""
    for order in ("forward", "backward", "holes_then_tail", "interleaved_append"):
        tables = [(["Position"], 12), (["Position", "Health"], 3)]

        def make_specs(rng, live, f, order=order):
            def fn(sp):
                e = [int(x) for x in live[:12]]
                if order == "forward":
                    return [(capi.CMD_DESTROY, x, []) for x in e[6:]]
                if order == "backward":
                    return [(capi.CMD_DESTROY, x, []) for x in reversed(e[6:])]
                if order == "holes_then_tail":
                    return [(capi.CMD_DESTROY, e[i], []) for i in (0, 1, 2, 11, 10, 3, 9, 4)]
                specs = []
                for i in range(6):  # append (migration in) and remove (migration out) interleaved on the same table
                    specs.append((capi.CMD_ADD, sp[i], ["Position"]))
                    specs.append((capi.CMD_ADD, e[11 - i], ["Health"]))
                    specs.append((capi.CMD_DESTROY, e[i], []))
                return specs
            return 6, fn

        run_both(gpu, oracle, tables, make_specs)


@pytest.mark.parametrize("seed", range(6))
def test_random_churn_small_tables(gpu, oracle, seed):
    """Many commands per entity, tiny tables: long move chains, tables emptied and refilled, archetypes created."""
    tables = [(["Position"], 17), (["Position", "Velocity"], 9), (["Position", "Health"], 30), (["Mass"], 1)]
    comps = ["Position", "Velocity", "Mass", "Health", "Team"]

    def make_specs(rng, live, f):
        n_spawn = int(rng.integers(0, 12))

        def fn(sp):
            pool = [int(x) for x in live] + [int(x) for x in sp]
            specs = []
            for _ in range(int(rng.integers(20, 160))):
                kind = int(rng.choice([capi.CMD_DESTROY, capi.CMD_TRY_DESTROY, capi.CMD_ADD, capi.CMD_TRY_ADD, capi.CMD_REMOVE,
                                       capi.CMD_TRY_REMOVE, capi.CMD_CLEAR], p=[0.12, 0.08, 0.3, 0.15, 0.2, 0.1, 0.05]))
                ent = pool[int(rng.integers(0, len(pool)))]
                k = int(rng.integers(1, 4))
                shorts = list(rng.choice(comps, size=k, replace=False)) if kind in (2, 3, 4, 5) else []
                specs.append((kind, ent, shorts))
            return specs
        return n_spawn, fn

    run_both(gpu, oracle, tables, make_specs, frames=4, seed=seed)


@pytest.mark.parametrize("n,frac", [(16 * 4096, 0.10), (16 * 20000, 0.10), (16 * 4096, 0.6)])
def test_cfg3_shape(gpu, oracle, n, frac):
    """configs[2] at reduced scale: 16-table lattice, 25 % spawn / despawn / insert Health / remove Damping on distinct
    live entities, randomly interleaved, several frames."""
    res = []
    for b in (gpu, oracle):
        w = capi.World(b)
        worlds.build_cfg3(w, n)
        res.append(w)
    for frame in range(3):
        snap = res[1].snapshot(columns=False)
        ents = [a["entities"] for a in snap["archetypes"]]
        types = [a["types"] for a in snap["archetypes"]]
        despawn, insert, remove, n_spawn = worlds.cfg3_commands(res[1], ents, types, frame, int(n * frac))
        snaps = []
        for w in res:
            spawned = w.reserve_entities(n_spawn)
            cmds, payload = worlds.assemble_cfg3(w, spawned, despawn, insert, remove, frame)
            w.enqueue(cmds, payload)
            w.flush()
            snaps.append(w.snapshot())
        diffs = worlds.snapshots_equal(snaps[0], snaps[1])
        assert not diffs, (frame, diffs[:5])
    st = res[0].apply_stats()
    assert st.used_sequential_fallback == 0


def test_device_emitted_commands(gpu, oracle):
    """A GPU system emits Destroy commands from its body (despawn_dead): same ids, same order as the serial ForEach."""
    res = []
    for b in (gpu, oracle):
        w = capi.World(b)
        worlds.register_all(w, ["Position", "Health"])
        n = 5000
        hp = (np.arange(n) % 7 - 2).astype(np.int32)  # some <= 0
        w.spawn_batch([worlds.cname("Position"), worlds.cname("Health")], n, [worlds.column("Position", n, 1), hp])
        w.spawn_batch([worlds.cname("Health")], 3000, [(np.arange(3000) % 5 - 1).astype(np.int32)])
        s = w.schedule("Update")
        s.add_system("DespawnDead", "despawn_dead", w.query([(worlds.cname("Health"), False)]), flags=capi.SYSTEM_EMITS_COMMANDS)
        s.build()
        s.run_and_wait()
        s.run_and_wait()  # second frame: nothing left to despawn
        res.append(w)
    diffs = worlds.snapshots_equal(res[0].snapshot(), res[1].snapshot())
    assert not diffs, diffs[:5]
