"""Query terminals: Find / Any / All / None / CountIf / Fold / MinBy / MaxBy / Partition / Filter +
CollectEntities as device reductions and ordered compactions.

CPU part: the oracle's sequential definitions against the expectations of the reference's own tests
(/root/reference/src/ecs/tests/query/query.cpp:453-731, cited per case).  GPU part: the CUDA terminals against the oracle on
the same seeded worlds — counts, positions in iteration order, entities and integer folds bit-exact; the floating-point Fold
within the tolerance written in the test (the device sum is a fixed f64 tree, the reference's a left fold in f32)."""
import struct

import numpy as np
import pytest

from heliosengine_b200 import capi, worlds

P = worlds.cname("Position")


def pos_world(backend, xs):
    """ComponentManager with one Position per entry of xs (AddPos of the reference tests), in creation order."""
    w = capi.World(backend)
    worlds.register_all(w, ["Position", "Health"])
    if len(xs):
        col = np.zeros((len(xs), 2), dtype=np.float32)
        col[:, 0] = xs
        w.spawn_batch([P], len(xs), [col.reshape(-1)], want_entities=False)
    return w, w.query([(P, False)])


def thr(t):
    return struct.pack("<f", t)


# ---- reference expectations on the oracle --------------------------------------------------------------------------------
def test_find_reference_cases(oracle):
    w, q = pos_world(oracle, [])                      # query.cpp:454-460 nullopt when no entities match
    assert q.terminal("find_pos_x_gt", thr(-1e30)).index == capi.NO_ROW
    w.close()
    w, q = pos_world(oracle, [1.0])                   # :462-470 predicate never matches
    assert q.terminal("find_pos_x_gt", thr(100.0)).index == capi.NO_ROW
    w.close()
    w, q = pos_world(oracle, [1.0, 50.0])             # :472-481 first matching element (x == 50)
    assert q.terminal("find_pos_x_gt", thr(10.0)).index == 1
    w.close()
    w, q = pos_world(oracle, [5.0, 10.0])             # :483-495 first match, not all matches
    assert q.terminal("find_pos_x_gt", thr(0.0)).index == 0
    w.close()


def test_any_all_none_reference_cases(oracle):
    w, q = pos_world(oracle, [])
    assert not q.any_pos_x_gt(-1e30)                  # :499-503 Any false on empty
    assert q.all_pos_x_gt(1e30)                       # :523-527 All vacuously true
    assert q.none_pos_x_gt(-1e30)                     # :547-551 None true on empty
    w.close()
    w, q = pos_world(oracle, [1.0, 200.0])
    assert q.any_pos_x_gt(100.0)                      # :513-519
    assert not q.none_pos_x_gt(100.0)                 # :561-567
    assert not q.all_pos_x_gt(100.0)
    w.close()
    w, q = pos_world(oracle, [5.0, 10.0])
    assert q.all_pos_x_gt(0.0)                        # :529-535
    w.close()
    w, q = pos_world(oracle, [5.0, -1.0])
    assert not q.all_pos_x_gt(0.0)                    # :537-543
    w.close()


def test_count_if_and_fold_reference_cases(oracle):
    w, q = pos_world(oracle, [])
    assert q.count_if_pos_x_gt(-1e30) == 0            # :570-574
    assert q.terminal("fold_sum_pos_x").f64 == 0.0    # :604-610 Fold returns init on empty
    w.close()
    w, q = pos_world(oracle, [1.0, 200.0, 300.0])
    assert q.count_if_pos_x_gt(100.0) == 2            # :584-592
    assert q.count_if_pos_x_gt(-1e30) == 3            # :594-600 all satisfy
    w.close()
    w, q = pos_world(oracle, [1.0, 2.0, 3.0])
    r = q.terminal("fold_sum_pos_x")                  # :612-622 accumulates 6
    assert r.f64 == 6.0 and struct.unpack("<f", struct.pack("<I", r.i64))[0] == 6.0
    w.close()


def test_min_max_by_reference_cases(oracle):
    w, q = pos_world(oracle, [])
    assert q.terminal("max_by_pos_x").index == capi.NO_ROW   # :682-687
    assert q.terminal("min_by_pos_x").index == capi.NO_ROW   # :713-718
    w.close()
    w, q = pos_world(oracle, [7.0])
    assert q.terminal("max_by_pos_x").f64 == 7.0             # :689-697
    w.close()
    w, q = pos_world(oracle, [1.0, 99.0, 50.0])
    assert q.terminal("max_by_pos_x").index == 1             # :699-709
    w.close()
    w, q = pos_world(oracle, [10.0, 1.0, 50.0])
    assert q.terminal("min_by_pos_x").index == 1             # :720-730
    w.close()
    w, q = pos_world(oracle, [3.0, 9.0, 9.0, 1.0, 1.0])      # strict comparison keeps the FIRST extremum
    assert q.terminal("max_by_pos_x").index == 1 and q.terminal("min_by_pos_x").index == 3
    w.close()


def test_partition_reference_cases(oracle):
    w, q = pos_world(oracle, [])
    r = q.terminal("partition_pos_x_gt", thr(0.0))           # :636-644 both empty
    assert r.count == 0 and q.terminal_entities(0).size == 0 and q.terminal_entities(1).size == 0
    w.close()
    w, q = pos_world(oracle, [5.0, 15.0, 25.0])
    r = q.terminal("partition_pos_x_gt", thr(10.0))          # :668-678 mixed data: 2 / 1
    yes, no = q.terminal_entities(0), q.terminal_entities(1)
    assert (r.count, yes.size, no.size) == (2, 2, 1)
    assert list(yes & 0xFFFFFFFF) == [1, 2] and list(no & 0xFFFFFFFF) == [0]   # iteration order kept
    r = q.terminal("partition_pos_x_gt", thr(-1e30))         # :646-655 all in the first partition = CollectEntities
    assert r.count == 3 and q.terminal_entities(1).size == 0
    w.close()


# ---- CUDA terminals against the oracle -------------------------------------------------------------------------------------
def lattice_worlds(gpu, oracle, n):
    out = []
    for b in (gpu, oracle):
        w = capi.World(b)
        worlds.build_lattice(w, n, worlds.OPTIONAL)
        out.append(w)
    return out


QUERIES = [
    dict(components=[(P, False)], with_=[], without=[]),
    dict(components=[(P, False)], with_=[worlds.cname("Health")], without=[worlds.cname("Frozen")]),
    dict(components=[(P, False)], with_=[worlds.cname("Tint"), worlds.cname("Team")], without=[]),
]


@pytest.mark.gpu
@pytest.mark.parametrize("n", [64, 64 * 1000, 64 * 4099])
def test_gpu_terminals_match_oracle(gpu, oracle, n):
    wg, wo = lattice_worlds(gpu, oracle, n)
    for spec in QUERIES:
        qg, qo = wg.query(**spec), wo.query(**spec)
        xs = None
        for t in (-1e30, -0.5, 0.0, 0.731, 1e30):
            for name in ("count_if_pos_x_gt", "find_pos_x_gt"):
                g, o = qg.terminal(name, thr(t)), qo.terminal(name, thr(t))
                assert (g.visited, g.count, g.index, g.entity) == (o.visited, o.count, o.index, o.entity), (name, t)
            g, o = qg.terminal("partition_pos_x_gt", thr(t)), qo.terminal("partition_pos_x_gt", thr(t))
            assert (g.visited, g.count) == (o.visited, o.count)
            for grp in (0, 1):
                assert np.array_equal(qg.terminal_entities(grp), qo.terminal_entities(grp))
        for name in ("max_by_pos_x", "min_by_pos_x"):
            g, o = qg.terminal(name), qo.terminal(name)
            assert (g.visited, g.index, g.entity, g.f64) == (o.visited, o.index, o.entity, o.f64), name
        g, o = qg.terminal("fold_sum_pos_x"), qo.terminal("fold_sum_pos_x")
        assert g.visited == o.visited
        # TOLERANCE (floating-point Fold; positions are non-negative, so sum|x| = the sum): the device's f64 tree and the
        # oracle's f64 left fold agree to 1e-10 relative (bound n * 2^-53 each) ...
        scale = max(1.0, abs(o.f64))
        assert abs(g.f64 - o.f64) <= 1e-10 * scale
        # ... and the reference's literal result, the f32 left fold (bits in o.i64), lies within ITS rounding bound
        # n * 2^-24 * sum|x| of the device value
        f32_fold = struct.unpack("<f", struct.pack("<I", o.i64))[0]
        assert abs(g.f64 - f32_fold) <= 2.0 ** -24 * max(1.0, float(o.visited)) * scale
    # integer fold: exact
    spec = dict(components=[(worlds.cname("Health"), False)], with_=[], without=[])
    g, o = wg.query(**spec).terminal("fold_sum_health_hp"), wo.query(**spec).terminal("fold_sum_health_hp")
    assert (g.visited, g.i64) == (o.visited, o.i64) and g.visited == n // 2
    wg.close()
    wo.close()


@pytest.mark.gpu
def test_gpu_first_extremum_and_nan_keys(gpu, oracle):
    """Ties resolve to the first row in iteration order; NaN keys never win unless they are the very first row."""
    nan = float("nan")
    for xs in ([3.0, 9.0, 9.0, 1.0, 1.0], [nan, 5.0, 7.0], [5.0, nan, 7.0, nan, 1.0], [2.0] * 5000 + [2.5] * 3 + [2.0] * 5000):
        wg, qg = pos_world(gpu, xs)
        wo, qo = pos_world(oracle, xs)
        for name in ("max_by_pos_x", "min_by_pos_x"):
            g, o = qg.terminal(name), qo.terminal(name)
            assert (g.index, g.entity) == (o.index, o.entity), (name, xs[:6])
        wg.close()
        wo.close()


@pytest.mark.gpu
def test_gpu_terminal_empty_and_repeated(gpu):
    w, q = pos_world(gpu, [])
    r = q.terminal("count_if_pos_x_gt", thr(0.0))
    assert (r.visited, r.count) == (0, 0)
    assert q.terminal("find_pos_x_gt", thr(0.0)).index == capi.NO_ROW and q.all_pos_x_gt(0.0)
    w.close()
    w, q = pos_world(gpu, list(np.linspace(-1, 1, 10001, dtype=np.float32)))
    a = [q.terminal("fold_sum_pos_x").f64 for _ in range(3)]   # the ticket is left at zero: repeated launches agree
    assert a[0] == a[1] == a[2] and abs(a[0]) < 1e-3
    assert q.count_if_pos_x_gt(0.0) == 5000
    w.close()
