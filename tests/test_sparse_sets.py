"""Sparse-set components: component/sparse_storage.hpp, container/sparse_set.hpp, the sparse paths of
component/manager.hpp:1314-1590 and the per-entity sparse terms of queries (query/iterator.hpp:539-577, quirk #7).

CPU leg: the oracle against hand-derived dense orders (the reference's own expectations are in oracle/kat_tests.cpp,
TestSparseSets).  GPU leg: the CUDA path against the oracle on the same seeded command streams — archetype snapshot, every
storage's dense order and value bytes, query counts under the three sparse rules, systems behind sparse filters."""
import numpy as np
import pytest

import scenarios
from heliosengine_b200 import capi, worlds


def _small(backend):
    w = capi.World(backend)
    w.register_component("Position", 8)
    w.register_component("Health", 4)
    w.register_sparse_component("PlayerTag", 0)
    w.register_sparse_component("SparseBuff", 4)
    ents = w.spawn_batch(["Position", "Health"], 8, [np.zeros((8, 2), np.float32), np.full(8, 50, np.int32)])
    return w, ents


def _check_small(backend):
    w, e = _small(backend)
    tag, buff = w.mask(["PlayerTag"]), w.mask(["SparseBuff"])
    pay = np.array([7, 9, 11, 13, 15], dtype=np.int32)
    cmds = np.zeros(10, dtype=capi.COMMAND_DTYPE)
    cmds[0] = (capi.CMD_ADD, 0, e[1], tag)
    cmds[1] = (capi.CMD_ADD, 0, e[2], buff)            # buff: [2]=7
    cmds[2] = (capi.CMD_ADD, 4, e[3], buff | tag)      # buff: [2,3]=7,9   tag: [1,3]
    cmds[3] = (capi.CMD_ADD, 8, e[5], buff | tag)      # buff: [2,3,5]=7,9,11   tag: [1,3,5]
    cmds[4] = (capi.CMD_DESTROY, 0, e[1], 0)           # tag: [5,3]
    cmds[5] = (capi.CMD_REMOVE, 0, e[2], buff)         # buff: [5,3]=11,9
    cmds[6] = (capi.CMD_TRY_ADD, 12, e[3], buff)       # present: stays 9
    cmds[7] = (capi.CMD_ADD, 16, e[5], buff)           # overwrite: 15
    cmds[8] = (capi.CMD_TRY_REMOVE, 0, e[7], buff)     # absent: nothing
    cmds[9] = (capi.CMD_CLEAR, 0, e[3], 0)             # leaves every storage: buff [5]=15, tag [5]
    w.enqueue(cmds, pay)
    w.flush()
    t, b = w.snapshot_sparse("PlayerTag"), w.snapshot_sparse("SparseBuff")
    assert t["indices"].tolist() == [5] and t["data"].size == 0
    assert b["indices"].tolist() == [5] and b["data"].view(np.int32).tolist() == [15]
    assert w.query([("Health", False)], with_=["PlayerTag"]).count() == 1
    assert w.query([("Health", False)], with_=["PlayerTag"], sparse_rule=capi.SPARSE_FOREACH).count() == 6
    assert w.query([("Health", False)], without=["PlayerTag"]).count() == 5  # 6 rows have Health (e1 destroyed, e3 cleared), e5 is tagged
    assert w.query([("Health", False)], without=["PlayerTag"], sparse_rule=capi.SPARSE_WITH_ENTITY).count() == 6
    assert w.query([("Health", False)], sparse_data=["SparseBuff"]).count() == 1
    # second batch: re-add after removal lands at the tail; removing the middle moves the last element in
    cmds = np.zeros(5, dtype=capi.COMMAND_DTYPE)
    for i, k in enumerate((0, 2, 4, 6)):
        cmds[i] = (capi.CMD_ADD, 4 * i, e[k], buff)    # buff: [5,0,2,4,6]
    cmds[4] = (capi.CMD_REMOVE, 0, e[0], buff)         # buff: [5,6,2,4]
    w.enqueue(cmds, np.array([1, 2, 3, 4], dtype=np.int32))
    w.flush()
    b = w.snapshot_sparse("SparseBuff")
    assert b["indices"].tolist() == [5, 6, 2, 4] and b["data"].view(np.int32).tolist() == [15, 4, 2, 3]
    w.close()


def test_oracle_sparse_commands_hand_checked(oracle):
    _check_small(oracle)


def test_oracle_sparse_scenario_is_deterministic(oracle):
    assert scenarios.sparse_sets(oracle, seed=5, frames=2) == scenarios.sparse_sets(oracle, seed=5, frames=2)


@pytest.mark.gpu
def test_gpu_sparse_commands_hand_checked(gpu):
    _check_small(gpu)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [11, 12, 13])
def test_gpu_sparse_scenario_matches_oracle(gpu, oracle, seed):
    got, want = scenarios.sparse_sets(gpu, seed=seed), scenarios.sparse_sets(oracle, seed=seed)
    assert got == want, {k: (got.get(k), want[k]) for k in want if got.get(k) != want[k]}


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [capi.EXEC_STREAMS, capi.EXEC_GRAPH])
def test_gpu_sparse_scenario_other_executors(gpu, oracle, mode):
    assert scenarios.sparse_sets(gpu, seed=21, mode=mode) == scenarios.sparse_sets(oracle, seed=21)


def _big(backend, n_cmd, frames, seed):
    """Thousands of records per batch over a few thousand entities, entity table growing while sparse types exist."""
    w = scenarios.sparse_world(backend, rows=(3000, 2000, 1000), entity_capacity=1024)
    out = []
    stale = []
    for f in range(frames):
        rng = np.random.default_rng(seed * 77 + f)
        live = np.concatenate([a["entities"] for a in w.snapshot(columns=False)["archetypes"]])
        spawned = w.reserve_entities(2500)
        pool = np.concatenate([live, spawned, np.asarray(stale[-50:], dtype=np.uint64)])
        cmds, payload = scenarios.sparse_commands(w, rng, pool, n_cmd, stale)
        w.enqueue(cmds, payload)
        w.flush()
        out.append(scenarios.sparse_state(w))
        H = worlds.cname("Health")
        out.append([w.query([(H, False)], with_=[worlds.cname("PlayerTag")], without=[worlds.cname("FrozenTag")], sparse_rule=r).count()
                    for r in (capi.SPARSE_STRICT, capi.SPARSE_FOREACH, capi.SPARSE_WITH_ENTITY)])
    w.close()
    return out


@pytest.mark.gpu
def test_gpu_sparse_large_batches_match_oracle(gpu, oracle):
    assert _big(gpu, 6000, 3, 4) == _big(oracle, 6000, 3, 4)


@pytest.mark.gpu
def test_gpu_sparse_registration_rules(gpu):
    w = capi.World(gpu)
    w.register_component("Position", 8)
    tag = w.register_sparse_component("Tag", 0)
    assert w.register_sparse_component("Tag", 0) == tag  # idempotent
    with pytest.raises(capi.HeliosError):
        w.register_sparse_component("Position", 8)        # already archetype-stored
    with pytest.raises(capi.HeliosError):
        w.spawn_batch(["Position", "Tag"], 1, [np.zeros(2, np.float32), np.zeros(0, np.uint8)])
    with pytest.raises(capi.HeliosError):
        w.query([("Tag", False)])                          # sparse data goes in sparse_data
    w.close()


def _bulk_tagging(backend):
    """Append-only batches (tags and values attached to thousands of entities at once: the parallel placement path),
    then a batch with removals over the same storages (ordered replay), then append-only again."""
    w = scenarios.sparse_world(backend, rows=(6000, 3000, 1000))
    TAG, BUFF, SV = (worlds.cname(x) for x in ("PlayerTag", "SparseBuff", "SparseVelocity"))
    ents = np.concatenate([a["entities"] for a in w.snapshot(columns=False)["archetypes"]])
    rng = np.random.default_rng(99)
    out = []

    def add(es, names, try_=False):
        names = sorted(names, key=lambda y: w.slots[y])
        words = sum(w.sizes[x] // 4 for x in names)
        cmds = np.zeros(es.size, dtype=capi.COMMAND_DTYPE)
        cmds["kind"] = capi.CMD_TRY_ADD if try_ else capi.CMD_ADD
        cmds["entity"] = es
        cmds["component_mask"] = w.mask(names)
        cmds["payload_offset"] = np.arange(es.size, dtype=np.uint32) * (4 * words)
        w.enqueue(cmds, rng.integers(1, 1 << 20, size=max(1, es.size * words), dtype=np.uint32))
        w.flush()
        out.append(scenarios.sparse_state(w))

    add(rng.permutation(ents)[:7000], [TAG, BUFF])             # both storages append-only, >= 1024 appends
    add(rng.permutation(ents)[:5000], [SV, TAG], try_=True)    # TAG: partly present (no event), SV: all new
    cmds = np.zeros(3000, dtype=capi.COMMAND_DTYPE)            # removals: BUFF goes through the ordered replay
    cmds["kind"] = capi.CMD_TRY_REMOVE
    cmds["entity"] = rng.permutation(ents)[:3000]
    cmds["component_mask"] = w.mask([BUFF])
    w.enqueue(cmds)
    w.flush()
    out.append(scenarios.sparse_state(w))
    add(rng.permutation(ents)[:4000], [BUFF])                  # overwrites + inserts mixed: replay again
    add(w.reserve_entities(2000), [TAG, SV])                   # freshly spawned entities, append-only
    w.close()
    return out


@pytest.mark.gpu
def test_gpu_sparse_bulk_tagging_matches_oracle(gpu, oracle):
    assert _bulk_tagging(gpu) == _bulk_tagging(oracle)


def _rare_terms(backend, mode=capi.EXEC_SERIAL):
    """A required sparse term carried by ~1 % of the entities: on the GPU these launches are driven from the storage's dense
    array instead of the tables (for_each_sparse_driven_kernel); results must not depend on the driver."""
    import struct
    w = scenarios.sparse_world(backend, rows=(12000, 6000, 2000))
    P, H, T = (worlds.cname(x) for x in ("Position", "Health", "Team"))
    TAG, FROZEN, BUFF = (worlds.cname(x) for x in ("PlayerTag", "FrozenTag", "SparseBuff"))
    ents = np.concatenate([a["entities"] for a in w.snapshot(columns=False)["archetypes"]])
    rng = np.random.default_rng(5)

    def add(es, names, values=None):
        cmds = np.zeros(es.size, dtype=capi.COMMAND_DTYPE)
        cmds["kind"] = capi.CMD_ADD
        cmds["entity"] = es
        cmds["component_mask"] = w.mask(names)
        if values is not None:
            cmds["payload_offset"] = np.arange(es.size, dtype=np.uint32) * 4
        w.enqueue(cmds, values if values is not None else b"")
        w.flush()

    add(rng.permutation(ents)[:200], [TAG])
    add(rng.permutation(ents)[:300], [BUFF], rng.integers(1, 50, size=300, dtype=np.uint32))
    add(rng.permutation(ents)[:5000], [FROZEN])
    s = w.schedule("Update")
    s.add_system("regen_players", "regen", w.query([(H, True)], with_=[TAG]))
    s.add_system("buff_heal", "buff_heal", w.query([(H, True)], without=[FROZEN], sparse_data=[BUFF]))
    s.build()
    out = []
    for _ in range(3):
        s.run_and_wait(mode)
        out.append(scenarios.snapshot_digest(w.snapshot()))
    q = w.query([(P, False)], with_=[TAG])
    out.append(q.count())
    for name, params in (("count_if_pos_x_gt", struct.pack("<f", 500.0)), ("find_pos_x_gt", struct.pack("<f", 900.0)),
                         ("max_by_pos_x", b""), ("min_by_pos_x", b"")):
        r = q.terminal(name, params)
        out.append((name, r.count, r.entity))
    w.close()
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [capi.EXEC_SERIAL, capi.EXEC_GRAPH])
def test_gpu_rare_sparse_terms_match_oracle(gpu, oracle, mode):
    assert _rare_terms(gpu, mode) == _rare_terms(oracle)
