"""GPU parity: archetype-table query matching + flattened iteration vs the CPU oracle, through the C ABI.

Bar (north_star): entity ids, archetype membership, match sets and row order bit-exact; float component updates
bit-exact here as well, because both sides are built without FMA contraction (-fmad=false / -ffp-contract=off);
the contractual tolerance is 1e-6 relative and is asserted separately.
"""
import struct

import numpy as np
import pytest

from heliosengine_b200 import capi, worlds

pytestmark = pytest.mark.gpu

P, V, A, M = (worlds.cname(x) for x in worlds.BASE4)


def both(gpu, oracle):
    return capi.World(gpu), capi.World(oracle)


def assert_same_world(wg, wo):
    diffs = worlds.snapshots_equal(wg.snapshot(), wo.snapshot())
    assert not diffs, diffs


def rel_err_ok(sg, so, tol=1e-6):
    for x, y in zip(sg["archetypes"], so["archetypes"]):
        for t in x["types"]:
            a = x["columns"][t].view(np.float32).astype(np.float64)
            b = y["columns"][t].view(np.float32).astype(np.float64)
            ok = np.abs(a - b) <= tol * np.maximum(np.abs(b), 1e-30)
            assert ok.all()


@pytest.mark.parametrize("n", [0, 1, 2, 1023, 1024, 1025, 4096 + 7, 1 << 16])
def test_cfg1_integrate_ragged_sizes(gpu, oracle, n):
    """configs[0] shape at ragged sizes (empty world, single row, tile boundary +-1)."""
    worlds_ = []
    for b in (gpu, oracle):
        w = capi.World(b)
        s, q = worlds.build_cfg1(w, n)
        assert q.count() == n
        for _ in range(3):
            s.run_and_wait(capi.EXEC_SERIAL)
        worlds_.append(w)
    assert_same_world(*worlds_)


@pytest.mark.parametrize("filtered", [False, True])
@pytest.mark.parametrize("n", [64 * 37, 64 * 1024, 64 * 3000])
def test_cfg2_match_and_iterate(gpu, oracle, n, filtered):
    res = []
    for b in (gpu, oracle):
        w = capi.World(b)
        s, q = worlds.build_cfg2(w, n, filtered)
        m, c = q.match(), q.count()
        for _ in range(2):
            s.run_and_wait(capi.EXEC_SERIAL)
        res.append((w, m, c))
    (wg, mg, cg), (wo, mo, co) = res
    assert mg == mo                      # match set, creation order, empty tables skipped
    assert cg == co == (n // 4 if filtered else n)
    assert len(mg) == (16 if filtered else 64)
    assert_same_world(wg, wo)            # bit-exact incl. untouched tables
    rel_err_ok(wg.snapshot(), wo.snapshot())


def test_all_exec_modes_agree(gpu, oracle):
    """Serial stream / multi-stream / CUDA-graph execution and the materialised-view kernel path give one answer."""
    snaps = []
    for mode, flags in ((capi.EXEC_SERIAL, 0), (capi.EXEC_STREAMS, 0), (capi.EXEC_GRAPH, 0),
                        (capi.EXEC_SERIAL, capi.SYSTEM_NEEDS_VIEW), (capi.EXEC_GRAPH, capi.SYSTEM_NEEDS_VIEW)):
        w = capi.World(gpu)
        s, _ = worlds.build_cfg2(w, 64 * 2048, True, flags=flags)
        for _ in range(3):
            s.run_and_wait(mode)
        snaps.append(w.snapshot())
    wo = capi.World(oracle)
    s, _ = worlds.build_cfg2(wo, 64 * 2048, True)
    for _ in range(3):
        s.run_and_wait()
    so = wo.snapshot()
    for sg in snaps:
        assert not worlds.snapshots_equal(sg, so)


def test_query_semantics_known_answers(gpu):
    """Reference KATs re-expressed through the C ABI (src/ecs/tests/query/query.cpp:2109-2125, 1217-1242):
    With<Velocity>, Without<Health> => Count 1, sum x = 2.0; dynamic refresh after adds."""
    w = capi.World(gpu)
    H = worlds.cname("Health")
    worlds.register_all(w, ["Position", "Velocity", "Health"])
    f32 = np.float32
    w.spawn_batch([P], 1, [np.array([1, 0], f32)])
    w.spawn_batch([P, V], 1, [np.array([2, 0], f32), np.zeros(2, f32)])
    w.spawn_batch([P, V, H], 1, [np.array([3, 0], f32), np.zeros(2, f32), np.zeros(1, np.int32)])
    q = w.query([(P, False)], with_=[V], without=[H])
    assert q.count() == 1
    snap = w.snapshot()
    (tbl,) = q.match()
    x = snap["archetypes"][tbl]["columns"][capi.fnv1a(P)].view(np.float32)[0]
    assert x == 2.0
    # dynamic refresh: the same query object sees later additions
    q2 = w.query([(P, False)])
    assert q2.count() == 3
    w.spawn_batch([P], 2, [np.zeros(4, f32)])
    assert q2.count() == 5
    # unregistered With<T> matches nothing; unregistered Without<T> passes (query.hpp:1754-1757,1797-1800)
    assert w.query([(P, False)], with_=["nope::T"]).count() == 0
    assert w.query([(P, False)], without=["nope::T"]).count() == 5


def test_full_size_properties(gpu):
    """BASELINE size (16M entities, 64 archetypes): size-independent properties instead of an oracle replay.
    Linearity of the integrate body: after k steps of accel_integrate with zero acceleration every matched
    Position equals p0 + k*dt*v0 evaluated with the same float operation order; untouched tables are unchanged."""
    n = 1 << 24
    w = capi.World(gpu)
    s, q = worlds.build_cfg2(w, n, filtered=True)
    assert q.count() == n // 4 and len(q.match()) == 16
    matched = set(q.match())
    before = w.snapshot()
    s.run_and_wait(capi.EXEC_SERIAL)
    after = w.snapshot()
    hp, hv, ha, hm = (capi.fnv1a(x) for x in (P, V, A, M))
    dt = np.float32(worlds.DT)
    for i, (x, y) in enumerate(zip(before["archetypes"], after["archetypes"])):
        assert np.array_equal(x["entities"], y["entities"])
        if i not in matched:
            for t in x["types"]:
                assert np.array_equal(x["columns"][t], y["columns"][t])
            continue
        p0 = x["columns"][hp].view(np.float32); v0 = x["columns"][hv].view(np.float32)
        a0 = x["columns"][ha].view(np.float32); m0 = np.repeat(x["columns"][hm].view(np.float32), 2)
        sfac = (dt / m0).astype(np.float32)
        v1 = (v0 + (a0 * sfac).astype(np.float32)).astype(np.float32)
        p1 = (p0 + (v1 * dt).astype(np.float32)).astype(np.float32)
        assert np.array_equal(y["columns"][hv].view(np.float32), v1)
        assert np.array_equal(y["columns"][hp].view(np.float32), p1)
