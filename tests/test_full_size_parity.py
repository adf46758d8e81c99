"""BASELINE-size parity: the CUDA path against the ORACLE (the CPU restatement of the reference, oracle/helios_oracle.hpp)
on the full worlds of BASELINE.json configs[1..3] — the same seeded inputs on both sides, world state compared bit for bit
(ids, generations, free list, archetype and row order, every column byte, every message queue).  The smaller parity tests
cover the edge cases; these cover the sizes the numbers are quoted on."""
import numpy as np
import pytest

from heliosengine_b200 import capi, worlds

pytestmark = pytest.mark.gpu


def test_cfg2_full_size_vs_oracle(gpu, oracle):
    """16 777 216 entities / 64 archetypes, the headline query (16 tables, 4 194 304 rows), two steps.  The worlds are built
    without the bulk load's builtin messages (1.3 GB of them otherwise): the path under test makes no structural call."""
    n = 1 << 24
    snaps = []
    for b in (gpu, oracle):
        w = capi.World(b, entity_capacity=n + 1024, flags=capi.WORLD_NO_BUILTIN_MESSAGES)
        s, q = worlds.build_cfg2(w, n, filtered=True)
        if not b.is_gpu:
            b.fn("set_access_mode")(w.h, 0, 8)  # direct column loops over 8 threads: same arithmetic, row-independent body
        match, count = q.match(), q.count()
        for _ in range(2):
            s.run_and_wait(capi.EXEC_SERIAL)
        snaps.append((match, count, w.snapshot()))
        w.close()
    assert snaps[0][0] == snaps[1][0] and snaps[0][1] == snaps[1][1] == n // 4
    assert not worlds.snapshots_equal(snaps[0][2], snaps[1][2])


def test_cfg3_full_size_vs_oracle(gpu, oracle):
    """8 388 608 entities / 16 tables, two frames of 838 860 deferred commands (25 % spawn, despawn, insert, remove), builtin
    structural messages on; frame 2 runs on the steady-state pipeline (frame 1 sizes the message queues)."""
    n = 1 << 23
    wg, wo = capi.World(gpu, entity_capacity=n + n // 2), capi.World(oracle)
    for w in (wg, wo):
        worlds.build_cfg3(w, n)
        w.update()
        w.update()
    pipelines = []
    for frame in range(3):
        snap = wo.snapshot(columns=False)
        ents = [a["entities"] for a in snap["archetypes"]]
        types = [a["types"] for a in snap["archetypes"]]
        despawn, insert, remove, n_spawn = worlds.cfg3_commands(wo, ents, types, frame, n // 10)
        for w in (wg, wo):
            spawned = w.reserve_entities(n_spawn)
            cmds, payload = worlds.assemble_cfg3(w, spawned, despawn, insert, remove, frame)
            w.enqueue(cmds, payload)
            w.flush()
        pipelines.append(wg.apply_stats().pipeline)
        sg, so = wg.snapshot(columns=(frame == 2)), wo.snapshot(columns=(frame == 2))
        assert not worlds.snapshots_equal(sg, so), frame
        for w in (wg, wo):
            w.update()
    assert pipelines[-1] == 1, pipelines
    wg.close()
    wo.close()


def test_cfg4_full_size_vs_oracle(gpu, oracle):
    """4 194 304 entities, the 8-system schedule with 4 194 304 CollisionEvents per frame: one frame (RunAndWait + Update) in
    the CUDA-graph executor against the oracle's serial executor."""
    n = 1 << 22
    out = []
    for b, mode in ((gpu, capi.EXEC_GRAPH), (oracle, capi.EXEC_SERIAL)):
        w = capi.World(b, entity_capacity=n + 1024)
        s = worlds.build_cfg4(w, n)
        if not b.is_gpu:
            b.fn("set_access_mode")(w.h, 0, 8)
        w.update()
        w.update()
        for _ in range(2):
            s.run_and_wait(mode)
            w.update()
        out.append((w.snapshot(), s.read_resources(16).tobytes(), w.message_counts(worlds.COLLISION_EVENT)))
        w.close()
    assert out[0][1] == out[1][1] and out[0][2] == out[1][2] == (n, 0)
    assert not worlds.snapshots_equal(out[0][0], out[1][0])
