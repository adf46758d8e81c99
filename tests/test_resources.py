"""World resources (reference resource/manager.hpp, world.hpp:491-620; tests/resource/manager.cpp): HBM-resident values
read and written in place by system kernels (Res<const T> / ResMut<T>), and run conditions over them."""
import struct

import pytest

import scenarios
from heliosengine_b200 import capi, worlds

GRAVITY = "helios::bench::Gravity"
COUNTER = "helios::bench::FrameCounter"
OTHER = "helios::bench::Other"


def manager_kats(backend):
    """tests/resource/manager.cpp through the C ABI."""
    w = capi.World(backend)
    out = {"empty_count": w.resource_count(), "has_on_empty": w.resource_has(COUNTER)}           # :64, :460
    w.resource_insert(COUNTER, struct.pack("<QQ", 42, 0))                                         # :146
    out["inserted"] = struct.unpack("<QQ", w.resource_read(COUNTER))[0]
    w.resource_insert(COUNTER, struct.pack("<QQ", 7, 0))                                          # :161 Insert replaces
    out["replaced"] = struct.unpack("<QQ", w.resource_read(COUNTER))[0]
    out["try_insert_present"] = w.resource_try_insert(COUNTER, struct.pack("<QQ", 9, 9))         # :197 false, value kept
    out["kept"] = struct.unpack("<QQ", w.resource_read(COUNTER))[0]
    out["try_insert_other"] = w.resource_try_insert(OTHER, struct.pack("<f", 1.5))               # :208
    out["count_two"] = w.resource_count()                                                         # :171
    w.resource_write(COUNTER, struct.pack("<QQ", 100, 1))                                         # :370 modification persists
    out["written"] = list(struct.unpack("<QQ", w.resource_read(COUNTER)))
    w.resource_remove(COUNTER)                                                                    # :299, :309
    out["after_remove"] = [w.resource_has(COUNTER), w.resource_has(OTHER), w.resource_count()]
    out["try_remove_missing"] = w.resource_try_remove(COUNTER)                                    # :343
    try:
        w.resource_remove(COUNTER)
        out["remove_missing"] = "ok"
    except capi.HeliosError as e:
        out["remove_missing"] = e.status                                                          # the reference asserts
    w.resource_insert(COUNTER, struct.pack("<QQ", 1, 1))                                          # :321 re-insert after remove
    out["reinserted"] = w.resource_has(COUNTER)
    out["try_remove_present"] = w.resource_try_remove(OTHER)                                      # :333
    w.resource_clear()                                                                            # :122, :528
    out["after_clear"] = [w.resource_count(), w.resource_has(COUNTER)]
    w.close()
    return out


EXPECTED = {"empty_count": 0, "has_on_empty": False, "inserted": 42, "replaced": 7, "try_insert_present": False, "kept": 7,
            "try_insert_other": True, "count_two": 2, "written": [100, 1], "after_remove": [False, True, 1],
            "try_remove_missing": False, "remove_missing": capi.ERR_NOT_FOUND, "reinserted": True, "try_remove_present": True,
            "after_clear": [0, False]}


def gravity_world(backend, n=5003, frames=6, mode=capi.EXEC_SERIAL):
    """S0 tick_counter (ResMut<FrameCounter>) -> S1 apply_gravity (Res<const Gravity>, guarded: only while ticks <= 4)
    -> S2 integrate.  The host rewrites Gravity between frames; no graph is re-captured for it."""
    w = capi.World(backend)
    worlds.register_all(w, ["Position", "Velocity"])
    P, V = worlds.cname("Position"), worlds.cname("Velocity")
    w.spawn_batch([P, V], n, [worlds.column("Position", n, 4100), worlds.column("Velocity", n, 4101)], want_entities=False)
    w.resource_insert(GRAVITY, struct.pack("<ffff", 0.0, -9.8125, 1.0 / 64, 0.0))
    w.resource_insert(COUNTER, struct.pack("<QQ", 0, 0))
    s = w.schedule("Update")
    s.add_system("S0_tick", "tick_counter", None, params=struct.pack("<QQ", capi.fnv1a(COUNTER), 3), write_resources=[COUNTER])
    s1 = s.add_system("S1_gravity", "apply_gravity", w.query([(V, True)]), params=struct.pack("<QQ", capi.fnv1a(GRAVITY), 0),
                      read_resources=[GRAVITY, COUNTER])
    s.add_system("S2_integrate", "integrate", w.query([(P, True), (V, False)]))
    s.run_if(s1, "FirstFourFrames", capi.COND_WORLD_RESOURCE, resource=COUNTER, offset=0, cmp=capi.CMP_LE, rhs=4)
    s.build()
    out = {"order": s.execution_order(), "conflicts": s.conflicts()}
    for f in range(frames):
        if f == 2:
            w.resource_write(GRAVITY, struct.pack("<ffff", 0.5, 2.25, 1.0 / 32, 0.0))
        s.run_and_wait(mode)
        out[f"frame{f}_ran"] = s.last_verdicts()
        out[f"frame{f}_counter"] = list(struct.unpack("<QQ", w.resource_read(COUNTER)))
        out[f"frame{f}_world"] = scenarios.snapshot_digest(w.snapshot())
    w.close()
    return out


def test_oracle_resource_manager_kats(oracle):
    assert manager_kats(oracle) == EXPECTED


def test_oracle_gravity_world(oracle):
    r = gravity_world(oracle)
    assert [r[f"frame{f}_ran"][1] for f in range(6)] == [1, 1, 1, 1, 0, 0]
    assert r["frame5_counter"] == [6, 18]
    assert r["conflicts"] == [0b010, 0b101, 0b010]  # tick x gravity (FrameCounter W/R), gravity x integrate (Velocity W/R)


@pytest.mark.gpu
def test_gpu_resource_manager_kats(gpu):
    assert manager_kats(gpu) == EXPECTED


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [capi.EXEC_SERIAL, capi.EXEC_STREAMS, capi.EXEC_GRAPH])
def test_gpu_resources_in_systems_and_conditions_match_oracle(gpu, oracle, mode):
    assert gravity_world(gpu, mode=mode) == gravity_world(oracle)


@pytest.mark.gpu
def test_gpu_missing_resource_fails_loudly(gpu):
    w = capi.World(gpu)
    worlds.register_all(w, ["Velocity"])
    s = w.schedule("Update")
    s.add_system("gravity", "apply_gravity", w.query([(worlds.cname("Velocity"), True)]),
                 params=struct.pack("<QQ", capi.fnv1a(GRAVITY), 0))
    s.build()
    with pytest.raises(capi.HeliosError) as e:
        s.run_and_wait()
    assert e.value.status == capi.ERR_NOT_FOUND
    w.close()
