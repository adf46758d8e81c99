"""Host-side logic that needs no GPU: type hashing, synthetic-world generators, schedule order / conflict matrix,
message lifecycle and command records — exercised through the oracle's mirror of the C ABI."""
import numpy as np
import pytest

from heliosengine_b200 import capi, worlds


def test_fnv1a_type_index_vectors():
    # FNV-1a-64 published vectors (utils/hash.hpp:26-32)
    assert capi.fnv1a("") == 0xCBF29CE484222325
    assert capi.fnv1a("a") == 0xAF63DC4C8601EC8C
    assert capi.fnv1a("foobar") == 0x85944171F73967E8


def test_splitmix64_reference_values():
    # SplitMix64 from seed 0: first outputs of the published generator
    v = worlds.splitmix64(0, 3)
    assert [int(x) for x in v] == [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4, 0x06C45D188009454F]


def test_command_record_layout_matches_header():
    assert capi.COMMAND_DTYPE.itemsize == 24
    assert capi.COMMAND_DTYPE.fields["entity"][1] == 8 and capi.COMMAND_DTYPE.fields["component_mask"][1] == 16


def test_generators_are_deterministic_and_in_range():
    p = worlds.column("Position", 1000, 5)
    assert p.dtype == np.float32 and p.size == 2000 and p.min() >= 0 and p.max() < 1024
    assert np.array_equal(p, worlds.column("Position", 1000, 5))
    assert not np.array_equal(p, worlds.column("Position", 1000, 6))
    m = worlds.column("Mass", 1000, 1)
    assert m.min() >= 0.5 and m.max() < 2.0
    h = worlds.column("Health", 1000, 1)
    assert h.min() >= 1 and h.max() <= 100


def test_cfg4_conflict_matrix_and_order(oracle):
    """AccessPolicy::ConflictsWith (system/access_policy.hpp:172-196) over the 8-system schedule; conflict rows are
    indexed by position in the execution order (schedule.cpp:371-395)."""
    w = capi.World(oracle)
    s = worlds.build_cfg4(w, 64)
    assert s.execution_order() == list(range(8))
    c = s.conflicts()
    expect = {
        0: {7},          # integrate (W Pos)  x detect (R Pos)
        1: {0},          # damp (W Vel)       x integrate (R Vel)
        2: {4, 5},       # regen (W Health)   x consume (W Health), stats (R Health)
        3: set(),        # team_tint (R Team, W Tint): detect only READS Team
        4: {2, 5},
        5: {2, 4, 6},    # stats also writes the Stats resource exit_check reads
        6: {5},
        7: {0},
    }
    expect[0].add(1)
    for i in range(8):
        assert {j for j in range(8) if c[i] >> j & 1} == expect[i], i
    w.close()


def test_explicit_order_and_cycle(oracle):
    w = capi.World(oracle)
    worlds.register_all(w, ["Position", "Velocity"])
    s = w.schedule("Update")
    a = s.add_system("A", "integrate", w.query([(worlds.cname("Position"), True), (worlds.cname("Velocity"), False)]))
    b = s.add_system("B", "damp", w.query([(worlds.cname("Velocity"), True)]), params=b"\x00\x00\x80\x3f")
    s.add_order(b, a)
    s.build()
    assert s.execution_order() == [1, 0]
    s.add_order(a, b)
    with pytest.raises(capi.HeliosError) as e:
        s.build()
    assert e.value.status == capi.ERR_CYCLE
    w.close()


def test_message_double_buffer_lifecycle(oracle):
    """message/manager.cpp:148-213 through the ABI: written → current; Update → previous; second Update clears automatic."""
    w = capi.World(oracle)
    w.message_register("Msg", 8)
    w.message_register("Manual", 4, capi.MSG_MANUAL)
    w.message_write("Msg", np.arange(6, dtype=np.uint32))
    w.message_write("Manual", np.array([42], dtype=np.uint32))
    assert w.message_counts("Msg") == (0, 3)
    w.update()
    assert w.message_counts("Msg") == (3, 0)
    assert np.array_equal(w.message_read("Msg", 0).view(np.uint32), np.arange(6, dtype=np.uint32))
    w.update()
    assert w.message_counts("Msg") == (0, 0)
    assert w.message_counts("Manual") == (1, 0)
    w.close()


def test_commands_spawn_ids_never_recycle(oracle):
    """Commands::Spawn reserves {next_index, gen 1} (entity/manager.hpp:294-305) even when the free list is non-empty,
    while World::CreateEntity recycles LIFO with a bumped generation (:428-448)."""
    w = capi.World(oracle)
    worlds.register_all(w, ["Position"])
    e = w.create_entities(4)
    cmds = np.zeros(2, dtype=capi.COMMAND_DTYPE)
    cmds["kind"] = capi.CMD_DESTROY
    cmds["entity"] = e[[1, 2]]
    w.enqueue(cmds)
    w.flush()
    r = w.reserve_entities(2)
    assert [int(x) & 0xFFFFFFFF for x in r] == [4, 5] and all(int(x) >> 32 == 1 for x in r)
    w.flush()
    c = w.create_entities(2)
    assert [(int(x) & 0xFFFFFFFF, int(x) >> 32) for x in c] == [(2, 2), (1, 2)]
    assert w.entity_count() == 6
    w.close()


def _run_condition_host_rules(backend):
    """RunIf / Set RunIf / InSet mark the schedule dirty (tests/schedule/system_set_handle.cpp:190-224): Run before Build is
    an error (schedule.cpp:56-58); malformed condition descriptors are rejected; set conditions reach every member."""
    from heliosengine_b200 import capi, worlds
    w = capi.World(backend)
    worlds.register_all(w, ["Health"])
    H = worlds.cname("Health")
    w.spawn_batch([H], 10, [np.full(10, 5, dtype=np.int32)], want_entities=False)
    s = w.schedule("Update")
    a = s.add_system("a", "regen", w.query([(H, True)]))
    b = s.add_system("b", "regen", w.query([(H, True)]))
    s.build()
    s.run_and_wait()
    assert s.last_verdicts() == [1, 1]
    s.set_run_if(7, "Off", capi.COND_CONSTANT, rhs=0)
    s.in_set(b, 7)
    with pytest.raises(capi.HeliosError) as e:
        s.run_and_wait()
    assert e.value.status == capi.ERR_STATE
    s.build()
    s.build()  # conditions survive a double Build (tests/schedule/schedule.cpp:900-945)
    s.run_and_wait()
    assert s.last_verdicts() == [1, 0]
    for bad in (dict(kind=capi.COND_RESOURCE, offset=4), dict(kind=capi.COND_RESOURCE, offset=4096),
                dict(kind=capi.COND_MESSAGES), dict(kind=capi.COND_TERMINAL), dict(kind=capi.COND_WORLD_RESOURCE, offset=0),
                dict(kind=capi.COND_CONSTANT, cmp=9), dict(kind=17)):
        with pytest.raises(capi.HeliosError) as e:
            s.run_if(a, "bad", **bad)
        assert e.value.status == capi.ERR_INVALID_ARGUMENT, bad
    with pytest.raises(capi.HeliosError) as e:
        s.run_if(5, "nobody", capi.COND_CONSTANT, rhs=1)
    assert e.value.status == capi.ERR_NOT_FOUND
    w.close()


def test_run_condition_host_rules_oracle(oracle):
    _run_condition_host_rules(oracle)


@pytest.mark.gpu
def test_run_condition_host_rules_gpu(gpu):
    _run_condition_host_rules(gpu)
