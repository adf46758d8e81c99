"""Optional access to archetype-stored components — `T*` / `const T*` query arguments (reference query/iterator.hpp:49-68;
the reference's own case src/ecs/tests/query/query.cpp:1035-1075): the optional type is no match term, tables with and
without it are visited, the body gets a null pointer where the column is absent, and the access still registers in the
system's policy (`T*` = write, `const T*` = read, query/details/traits.hpp:73-97)."""
import numpy as np
import pytest

from heliosengine_b200 import capi, worlds

P, V, H = (worlds.cname(x) for x in ("Position", "Velocity", "Health"))


def build(backend, rows=(1, 1, 0)):
    """Tables {Position}, {Position, Velocity}, {Velocity} with rows[i] entities each."""
    w = capi.World(backend)
    worlds.register_all(w, ["Position", "Velocity", "Health"])
    for i, (shorts, n) in enumerate(zip((["Position"], ["Position", "Velocity"], ["Velocity"]), rows)):
        w.spawn_batch([worlds.cname(s) for s in shorts], n, [worlds.column(s, n, 60 + 4 * i + k) for k, s in enumerate(shorts)],
                      want_entities=False)
    return w


def scenario(backend, rows, frames=3, mode=capi.EXEC_SERIAL):
    w = build(backend, rows)
    q = w.query([(P, True)], optional=[(V, False)])
    out = {"count": q.count(), "match": q.match()}
    s = w.schedule("Update")
    s.add_system("DriftOptional", "drift_optional", q)                               # Query<Position&, const Velocity*>
    s.add_system("Damp", "damp", w.query([(V, True)]), params=np.float32(0.5).tobytes())  # Query<Velocity&>
    s.add_system("Regen", "regen", w.query([(H, True)]))                             # unrelated
    s.build()
    out["conflicts"] = s.conflicts()
    for _ in range(frames):
        s.run_and_wait(mode)
    out["snapshot"] = w.snapshot()
    w.close()
    return out


def test_reference_case_on_the_oracle(oracle):
    """query.cpp:1065-1074: one entity with Position only, one with Position + Velocity -> Count() == 2."""
    r = scenario(oracle, (1, 1, 0), frames=1)
    assert r["count"] == 2 and len(r["match"]) == 2
    # the const Velocity* reader conflicts with the Velocity writer, not with the Health system
    assert r["conflicts"] == [0b010, 0b001, 0]
    tables = r["snapshot"]["archetypes"]
    pos_only = np.frombuffer(tables[1]["columns"][capi.fnv1a(P)].tobytes(), np.float32)
    assert np.array_equal(pos_only, worlds.column("Position", 1, 60) + np.array([0.25, -0.5], np.float32))  # null pointer branch
    pos_vel = np.frombuffer(tables[2]["columns"][capi.fnv1a(P)].tobytes(), np.float32)
    assert np.array_equal(pos_vel, worlds.column("Position", 1, 64) + worlds.column("Velocity", 1, 65))


@pytest.mark.gpu
@pytest.mark.parametrize("rows", [(1, 1, 0), (5, 0, 3), (0, 7, 0), (3000, 2049, 77), (100_000, 150_001, 1)])
@pytest.mark.parametrize("mode", [capi.EXEC_SERIAL, capi.EXEC_GRAPH])
def test_gpu_optional_component_matches_oracle(gpu, oracle, rows, mode):
    g, o = scenario(gpu, rows, mode=mode), scenario(oracle, rows)
    assert g["count"] == o["count"] == rows[0] + rows[1]
    assert g["match"] == o["match"] and g["conflicts"] == o["conflicts"]
    assert not worlds.snapshots_equal(g["snapshot"], o["snapshot"])


@pytest.mark.gpu
def test_gpu_optional_of_an_unregistered_type_is_always_null(gpu, oracle):
    outs = []
    for b in (gpu, oracle):
        w = build(b, (9, 4, 0))
        s = w.schedule("Update")
        s.add_system("DriftOptional", "drift_optional", w.query([(P, True)], optional=[("helios::bench::NeverRegistered", False)]))
        s.build()
        s.run_and_wait(capi.EXEC_SERIAL)
        outs.append(w.snapshot())
        w.close()
    assert not worlds.snapshots_equal(outs[0], outs[1])
