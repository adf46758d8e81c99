"""SubApp::Extract (src/app/include/helios/app/sub_app.hpp:213-220,699-..., src/app/src/scheduler.cpp:141-151):
main-world query rows copied into a sub-world, and World::ClearEntities (world.hpp:974-977) between frames.

The extract function is user code in the reference; what is checked here is the device form of its usual body
(`for row in query: e = sub.CreateEntity(); sub.AddComponents(e, row...)`) against the oracle running exactly that."""
import numpy as np
import pytest

import scenarios
from heliosengine_b200 import capi, worlds


def _extract_frames(backend, frames=3, n=64 * 40):
    main = capi.World(backend)
    sched, _ = worlds.build_cfg2(main, n, filtered=True)
    P, V, H, F = (worlds.cname(x) for x in ("Position", "Velocity", "Health", "Frozen"))
    render_q = main.query([(P, False), (V, False)], with_=[H], without=[F])   # "render extraction": 16 of the 64 tables
    health_q = main.query([(H, False)])
    sub = capi.World(backend)
    out = []
    for f in range(frames):
        sched.run_and_wait(capi.EXEC_SERIAL)          # main world frame
        sub.clear_entities()                           # the sub-world is rebuilt from scratch every frame
        n1 = sub.extract_from(render_q)
        n2 = sub.extract_from(health_q)                # a second table in the sub-world, ids continue
        out.append((n1, n2, sub.entity_count(), scenarios.snapshot_digest(sub.snapshot())))
        # the main world is untouched by the extraction
        out.append(scenarios.snapshot_digest(main.snapshot()))
        if f == 0:                                     # sub-world systems run on the extracted copy
            s2 = sub.schedule("Render")
            s2.add_system("integrate", "integrate", sub.query([(P, True), (V, False)]))
            s2.build()
        s2.run_and_wait(capi.EXEC_SERIAL)
        out.append(scenarios.snapshot_digest(sub.snapshot()))
    # extraction appends when the sub-world is not cleared, recycling nothing (fresh ids)
    n3 = sub.extract_from(render_q)
    out.append((n3, sub.entity_count(), scenarios.snapshot_digest(sub.snapshot())))
    empty = main.query([(P, False)], with_=[worlds.cname("Tint")])
    out.append(sub.extract_from(empty))
    sub.close()
    main.close()
    return out


def test_oracle_extract_counts(oracle):
    out = _extract_frames(oracle, frames=1)
    n1, n2, total, _ = out[0]
    assert n1 == 16 * 40 and n2 == 32 * 40 and total == n1 + n2
    assert out[-1] == 0


def _clear_then_reuse(backend):
    """ClearEntities restarts ids at 0, drops every table and sparse element, and the world stays usable."""
    w = scenarios.sparse_world(backend)
    rng = np.random.default_rng(3)
    ents = np.concatenate([a["entities"] for a in w.snapshot(columns=False)["archetypes"]])
    cmds, payload = scenarios.sparse_commands(w, rng, ents, 80, [])
    w.enqueue(cmds, payload)
    w.flush()
    w.clear_entities()
    out = [scenarios.sparse_state(w), w.entity_count()]
    P = worlds.cname("Position")
    e = w.spawn_batch([P], 5, [worlds.column("Position", 5, 1)])
    out.append(e.tolist())
    cmds, payload = scenarios.sparse_commands(w, rng, e, 30, [])
    w.enqueue(cmds, payload)
    w.flush()
    out.append(scenarios.sparse_state(w))
    w.close()
    return out


def test_oracle_clear_entities_restarts_ids(oracle):
    out = _clear_then_reuse(oracle)
    assert out[1] == 0 and out[2] == [capi.make_entity(i, 1) for i in range(5)]


@pytest.mark.gpu
def test_gpu_extract_matches_oracle(gpu, oracle):
    got, want = _extract_frames(gpu), _extract_frames(oracle)
    assert got == want


@pytest.mark.gpu
def test_gpu_clear_entities_matches_oracle(gpu, oracle):
    assert _clear_then_reuse(gpu) == _clear_then_reuse(oracle)
