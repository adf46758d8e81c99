"""Secondary benchmark legs of bench.py (`--config cfg1|cfg3|cfg4`): the other BASELINE.json configurations, reported in
the same JSON-line shape as the headline (cfg2) leg.  Each function returns the dict bench.py prints.

cfg3 — structural churn: `ApplyDeferred` of 10 % commands per frame (spawn / despawn / insert / remove), timed with CUDA
       events on the world stream with the frame's command records already staged in HBM; `e2e` includes the host→device
       copy of the records.  The same frames are replayed on the CPU oracle: bit-exact world state is asserted (full size)
       and the oracle's replay log gives the algorithmic bytes and the CPU baseline.
cfg4 — 8-system schedule + CollisionEvent channel: one RunAndWait + World::Update per frame, in each executor mode.
cfg1 — 1M-entity single-archetype integrate (launch-bound; reported, no roofline claim).
"""
from __future__ import annotations

import struct
import ctypes
import time

import numpy as np

from heliosengine_b200 import capi, worlds


def _events(torch, n):
    return [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]


def run_cfg3(args, torch, dev, stream, peak):
    n = args.entities or (1 << 23)
    frac = 0.10
    gpu, oracle = capi.load_gpu(), capi.load_oracle()
    wg = capi.World(gpu, device=dev.index, entity_capacity=n + (n // 2))
    wg.set_stream(stream.cuda_stream)
    worlds.build_cfg3(wg, n)
    verify = not args.no_cpu_baseline
    wo = None
    if verify:
        wo = capi.World(oracle)
        worlds.build_cfg3(wo, n)

    steps, warm = args.steps, max(3, args.warmup)
    dev_ms, e2e_s, cpu_s, alg_bytes, n_cmds, launches, stats_l = [], [], [], [], [], [], []
    with torch.cuda.stream(stream):
        for frame in range(warm + steps):
            src = wo if verify else wg
            snap = src.snapshot(columns=False)
            ents = [a["entities"] for a in snap["archetypes"]]
            types = [a["types"] for a in snap["archetypes"]]
            despawn, insert, remove, n_spawn = worlds.cfg3_commands(src, ents, types, frame, int(n * frac))
            spawned = wg.reserve_entities(n_spawn)
            cmds, payload = worlds.assemble_cfg3(wg, spawned, despawn, insert, remove, frame)
            timed = frame >= warm
            # e2e on even timed frames (host records -> device inside the timed region), device-resident otherwise
            e2e_frame = timed and ((frame - warm) % 2 == 1)
            wg.enqueue(cmds, payload)
            l0 = wg.launch_count()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if not e2e_frame:
                wg.stage_commands()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            a.record(stream)
            wg.flush()
            b.record(stream)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            st = wg.apply_stats()
            if verify:
                wo.reserve_entities(n_spawn)
                b0 = wo.apply_stats().moved_bytes
                wo.enqueue(cmds, payload)
                b0 = b0  # record bytes are added at enqueue
                c0 = time.perf_counter()
                wo.flush()
                c1 = time.perf_counter()
                if timed:
                    cpu_s.append(c1 - c0)
                    alg_bytes.append(wo.apply_stats().moved_bytes - b0)
            if timed:
                n_cmds.append(cmds.size)
                stats_l.append(st)
                launches.append(wg.launch_count() - l0)
                if e2e_frame:
                    e2e_s.append(t1 - t0)
                else:
                    dev_ms.append(a.elapsed_time(b))
            # end of frame: World::Update (world.hpp:950-955) ages the message queues — the builtin structural messages the
            # commands wrote (1.47M per frame here) are dropped two frames later, as in a reference App loop
            wg.update()
            if verify:
                wo.update()
    parity = None
    if verify:
        dg, do = worlds.snapshot_digest(wg.snapshot()), worlds.snapshot_digest(wo.snapshot())
        parity = dg == do
        assert parity, "cfg3 full-size world state differs from the oracle"
    cmds_per_frame = float(np.mean(n_cmds))
    ms = float(np.mean(dev_ms))
    value = cmds_per_frame / (ms * 1e-3)
    bytes_per_frame = float(np.mean(alg_bytes)) if alg_bytes else float(np.mean([s.moved_bytes for s in stats_l]))
    achieved = bytes_per_frame / (ms * 1e-3) / 1e9
    line = {
        "metric": "commands_applied_per_sec", "value": value, "unit": "commands/s", "ms_per_step": ms,
        "dtype": "u8/u64 (row bytes, entity ids)",
        "config": {"workload": f"cfg3: {n} entities / 16 tables, {int(cmds_per_frame)} deferred commands per frame "
                               "(25% spawn, 25% despawn, 25% insert Health, 25% remove Damping), ApplyDeferred only",
                   "l2": "frame working set (tables 8M x ~40 B) exceeds L2; no flush",
                   "parity": "post-Commands world state bit-exact vs oracle at full size" if parity else "not verified in this run"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak[0], "unit": "GB/s", "frac": achieved / peak[0],
                     "traffic": None, "peak_kind": peak[1], "algorithmic_bytes_per_launch": bytes_per_frame,
                     "bytes_per_command": bytes_per_frame / cmds_per_frame,
                     "kernel": "whole ApplyDeferred pipeline (sort + scan + resolve + row mover)"},
        "e2e": {"value": cmds_per_frame / float(np.mean(e2e_s)) if e2e_s else None, "unit": "commands/s",
                "h2d_bytes_per_step": int(cmds_per_frame * 24 + cmds_per_frame * 8), "d2h_bytes_per_step": 64,
                "note": "host command records + payload copied to HBM inside the timed region; wall clock around World::Flush"},
        "cpu_baseline": ({"value": cmds_per_frame / float(np.mean(cpu_s)), "unit": "commands/s", "cores": 1, "kind": "port",
                          "sample": f"the same {len(cpu_s)} frames replayed by the oracle (sequential FIFO apply, dense index "
                                    "structures — cheaper than the reference's flat_map inserts)"} if cpu_s else None),
        "gpu_launches": int(np.sum(launches)),
        "apply": {"resolve_rounds": int(max(s.resolve_rounds for s in stats_l)),
                  "sequential_fallback": int(max(s.used_sequential_fallback for s in stats_l)),
                  "moved_rows_per_frame": float(np.mean([s.moved_rows for s in stats_l])),
                  "launches_per_frame": float(np.mean(launches)), "step_ms": dev_ms},
    }
    wg.close()
    if wo:
        wo.close()
    return line


def run_cfg4(args, torch, dev, stream, peak):
    n = args.entities or (1 << 22)
    gpu = capi.load_gpu()
    steps, warm = args.steps, max(3, args.warmup)
    out_modes = {}
    launches = 0
    for mode, mname in ((capi.EXEC_SERIAL, "serial"), (capi.EXEC_STREAMS, "streams"), (capi.EXEC_GRAPH, "graph")):
        w = capi.World(gpu, device=dev.index, entity_capacity=n + 1024)
        w.set_stream(stream.cuda_stream)
        s = worlds.build_cfg4(w, n)
        with torch.cuda.stream(stream):
            for _ in range(warm):
                s.run_and_wait(mode)
                w.update()
            torch.cuda.synchronize()
            evs = _events(torch, steps)
            l0 = w.launch_count()
            t0 = time.perf_counter()
            for a, b in evs:
                a.record(stream)
                s.run(mode)
                b.record(stream)
                s.apply_deferred()
                w.update()
            torch.cuda.synchronize()
            wall = time.perf_counter() - t0
            launches += w.launch_count() - l0
        ms = [a.elapsed_time(b) for a, b in evs]
        prev_n, _ = w.message_counts(worlds.COLLISION_EVENT)
        out_modes[mname] = {"run_ms": float(np.mean(ms)), "frame_wall_ms": wall / steps * 1e3, "events_per_frame": prev_n,
                            "streams": s.stream_count()}
        w.close()
    best = min(out_modes, key=lambda k: out_modes[k]["run_ms"])
    ms = out_modes[best]["run_ms"]
    ev = out_modes[best]["events_per_frame"]
    bytes_per_frame = n * worlds.CFG4_COMPONENT_BYTES_PER_ENTITY + ev * worlds.CFG4_BYTES_PER_EVENT
    achieved = bytes_per_frame / (ms * 1e-3) / 1e9
    cpu = None
    if not args.no_cpu_baseline:  # the same schedule on the oracle, 1/8-scale sample, serial executor, direct column loops
        o = capi.load_oracle()
        sn = n // 8
        wo = capi.World(o)
        so = worlds.build_cfg4(wo, sn)
        for _ in range(2):
            so.run_and_wait(capi.EXEC_SERIAL)
            wo.update()
        t0 = time.perf_counter()
        for _ in range(3):
            so.run_and_wait(capi.EXEC_SERIAL)
            wo.update()
        dt = (time.perf_counter() - t0) / 3
        wo.close()
        cpu = {"value": sn / dt, "unit": "events/s", "cores": 1, "kind": "port",
               "sample": f"1/8-scale world ({sn} entities, {sn} events per frame), 3 frames of RunAndWait + World::Update on the "
                         "oracle (serial executor, direct column loops: cheaper than the reference's per-entity lookups)"}
    return {
        "metric": "events_per_sec", "value": ev / (ms * 1e-3), "unit": "events/s", "ms_per_step": ms, "dtype": "f32/i32",
        "config": {"workload": f"cfg4: {n} entities, 8-system schedule (conflict DAG), CollisionEvent channel {ev} events/frame",
                   "exec_mode": best, "l2": "frame working set ~500 MB exceeds L2; no flush"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak[0], "unit": "GB/s", "frac": achieved / peak[0],
                     "traffic": None, "peak_kind": peak[1], "algorithmic_bytes_per_launch": bytes_per_frame,
                     "kernel": "all 8 systems of one Schedule::Run"},
        "entity_updates_per_sec": n * 7 / (ms * 1e-3),
        "modes": out_modes, "gpu_launches": int(launches), "e2e": None, "cpu_baseline": cpu,
    }


def run_cfg1(args, torch, dev, stream, peak):
    n = args.entities or (1 << 20)
    gpu = capi.load_gpu()
    w = capi.World(gpu, device=dev.index, entity_capacity=n + 1024)
    w.set_stream(stream.cuda_stream)
    s, q = worlds.build_cfg1(w, n)
    steps, warm = args.steps, max(3, args.warmup)
    res = {}
    with torch.cuda.stream(stream):
        for mode, mname in ((capi.EXEC_SERIAL, "serial"), (capi.EXEC_GRAPH, "graph")):
            for _ in range(warm):
                s.run_and_wait(mode)
            torch.cuda.synchronize()
            evs = _events(torch, steps)
            for a, b in evs:
                a.record(stream)
                s.run(mode)
                b.record(stream)
                s.apply_deferred()
            torch.cuda.synchronize()
            res[mname] = float(np.median([a.elapsed_time(b) for a, b in evs]))
    ms = min(res.values())
    achieved = n * worlds.CFG1_BYTES_PER_ENTITY / (ms * 1e-3) / 1e9
    w.close()
    cpu = None
    if not args.no_cpu_baseline:  # configs[0] is the reference's own CPU-runnable case: the full 1M-entity world on the oracle
        o = capi.load_oracle()
        wo = capi.World(o, flags=capi.WORLD_NO_BUILTIN_MESSAGES)
        so, _ = worlds.build_cfg1(wo, n)
        o.fn("set_access_mode")(wo.h, 1, 1)
        timer = o.fn("time_schedule_run", ctypes.c_double)
        timer(so.h, 1)
        dt = timer(so.h, 3) / 3
        wo.close()
        cpu = {"value": n / dt, "unit": "updates/s", "cores": 1, "kind": "port",
               "sample": f"the full config ({n} entities), 3 steps, reference-faithful per-entity access path (B-ref), one core"}
    return {
        "metric": "entity_component_updates_per_sec", "value": n / (ms * 1e-3), "unit": "updates/s", "ms_per_step": ms,
        "dtype": "f32", "config": {"workload": f"cfg1: {n} entities {{Position, Velocity}}, one archetype, ChangePosition",
                                   "l2": "25 MB working set is L2-resident by construction (launch-bound config; no roofline claim)"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak[0], "unit": "GB/s", "frac": achieved / peak[0],
                     "traffic": None, "peak_kind": peak[1], "note": "L2-resident: not an HBM measurement"},
        "modes_ms": res, "gpu_launches": steps * 2, "e2e": None, "cpu_baseline": cpu,
    }


def run_terminals(args, torch, dev, stream, peak):
    """Query terminals over the cfg2 world, `Query<const Position&>` on all 64 tables: every call goes
    through the public immediate API (helios_gpu_query_terminal: launch, 48-byte result back, synchronise), timed with
    CUDA events around the call; L2 flushed (write 256 MiB, read 256 MiB) between calls."""
    n = args.entities or (1 << 24)
    gpu = capi.load_gpu()
    w = capi.World(gpu, device=dev.index, entity_capacity=n + 1024)
    w.set_stream(stream.cuda_stream)
    worlds.build_lattice(w, n, worlds.OPTIONAL)
    q = w.query([(worlds.cname("Position"), False)])
    steps, warm = args.steps, max(3, args.warmup)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    flush_rd = torch.ones(256 << 20, dtype=torch.uint8, device=dev)
    sink = torch.zeros((), dtype=torch.int64, device=dev)
    thr = __import__("struct").pack("<f", 500.0)
    # algorithmic bytes per row: Position (8) [+ entity id (8) where the result names an entity];
    # partition: Position + flag w/r/r + position w(scan)/r(add)/w/r + entity r/w
    legs = [("count_if_pos_x_gt", thr, 8), ("find_pos_x_gt", thr, 16), ("max_by_pos_x", b"", 16), ("fold_sum_pos_x", b"", 8),
            ("partition_pos_x_gt", thr, 8 + 4 * 3 + 4 * 4 + 8 + 8 + 8)]
    out, launches = {}, 0
    with torch.cuda.stream(stream):
        for name, params, bpr in legs:
            for _ in range(warm):
                q.terminal(name, params)
            ms = []
            l0 = w.launch_count()
            for _ in range(steps):
                flush.fill_(1)
                torch.sum(flush_rd, dim=(0,), dtype=torch.int64, out=sink)
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                r = q.terminal(name, params)
                b.record(stream)
                torch.cuda.synchronize()
                ms.append(a.elapsed_time(b))
            launches += w.launch_count() - l0
            med = float(np.median(ms))
            out[name] = {"ms": med, "rows_per_s": n / (med * 1e-3), "bytes_per_row": bpr,
                         "GBps": n * bpr / (med * 1e-3) / 1e9, "frac": n * bpr / (med * 1e-3) / 1e9 / peak[0],
                         "result": {"visited": r.visited, "count": r.count, "index": r.index}}
        # the same reductions as systems of a schedule: no host round trip inside the timed region (device-resident `value`)
        for name, params, bpr in legs[:4]:
            s = w.schedule("T_" + name)
            s.add_system(name, name, q, params=params)
            s.build()
            for _ in range(warm):
                s.run_and_wait(capi.EXEC_SERIAL)
            ms = []
            for _ in range(steps):
                flush.fill_(1)
                torch.sum(flush_rd, dim=(0,), dtype=torch.int64, out=sink)
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                s.run(capi.EXEC_SERIAL)
                b.record(stream)
                s.apply_deferred()
                ms.append(a.elapsed_time(b))
            launches += steps
            res = s.read_resources(48).view(np.uint64)
            imm = out[name]["result"]
            assert int(res[0]) == imm["visited"] and int(res[1]) == imm["count"] and int(res[2]) == imm["index"]
            med = float(np.median(ms))
            out[name].update({"system_ms": med, "system_GBps": n * bpr / (med * 1e-3) / 1e9,
                              "system_frac": n * bpr / (med * 1e-3) / 1e9 / peak[0]})
            s.close()
    cpu = None
    if not args.no_cpu_baseline:
        o = capi.load_oracle()
        sn = n // 8
        wo = capi.World(o)
        worlds.build_lattice(wo, sn, worlds.OPTIONAL)
        qo = wo.query([(worlds.cname("Position"), False)])
        qo.terminal("count_if_pos_x_gt", thr)
        t0 = time.perf_counter()
        for _ in range(3):
            qo.terminal("count_if_pos_x_gt", thr)
        dt = (time.perf_counter() - t0) / 3
        cpu = {"value": sn / dt, "unit": "rows/s", "cores": 1, "kind": "port",
               "sample": f"CountIf over {sn} rows / 64 tables (1/8 scale), sequential ordered walk, 3 repetitions"}
        wo.close()
    w.close()
    head = out["count_if_pos_x_gt"]
    return {
        "metric": "query_terminal_rows_per_sec", "value": n / (head["system_ms"] * 1e-3), "unit": "rows/s",
        "ms_per_step": head["system_ms"],
        "dtype": "f32 keys, u64 counts / positions, f64 sums",
        "config": {"workload": f"terminals: CountIf(Position.x > 500) over {n} rows / 64 tables through helios_gpu_query_terminal "
                               "(headline); Find, MaxBy, Fold, Partition in `terminals`",
                   "l2": "flushed between calls: 256 MiB written, then 256 MiB read"},
        "roofline": {"bound": "hbm", "achieved": head["system_GBps"], "peak": peak[0], "unit": "GB/s", "frac": head["system_frac"],
                     "traffic": None, "peak_kind": peak[1], "algorithmic_bytes_per_launch": n * 8,
                     "kernel": "helios_gpu::for_each_kernel<ReduceBody<CountIfReducer<...>>, Position> as a schedule system"},
        "terminals": out,
        "e2e": {"value": head["rows_per_s"], "unit": "rows/s", "h2d_bytes_per_step": 4, "d2h_bytes_per_step": 48,
                "note": "helios_gpu_query_terminal: predicate parameter in, launch, 48-byte result back, synchronise"},
        "cpu_baseline": cpu, "gpu_launches": int(launches),
    }


def run_sparse(args, torch, dev, stream, peak):
    """Sparse-set components: `Query<Position&, const Velocity&, With<PlayerTag>, Without<FrozenTag>>`
    (README-style tag filters; tags are sparse-set stored) over ONE table of `--entities` rows (default 16.8M).  Tags are
    attached through deferred commands (timed: the sparse apply path), then the filtered integrate system is timed at tag
    densities 25 % and 100 %, next to the same system without sparse terms.  L2 flushed between steps."""
    n = args.entities or (1 << 24)
    gpu = capi.load_gpu()
    w = capi.World(gpu, device=dev.index, entity_capacity=n + 1024)
    w.set_stream(stream.cuda_stream)
    P, V = worlds.cname("Position"), worlds.cname("Velocity")
    worlds.register_all(w, ["Position", "Velocity"])
    TAG, FROZEN = worlds.cname("PlayerTag"), worlds.cname("FrozenTag")
    w.register_sparse_component(TAG, 0)
    w.register_sparse_component(FROZEN, 0)
    ents = w.spawn_batch([P, V], n, [worlds.column("Position", n, 1), worlds.column("Velocity", n, 2)])
    steps, warm = args.steps, max(3, args.warmup)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    flush_rd = torch.ones(256 << 20, dtype=torch.uint8, device=dev)
    sink = torch.zeros((), dtype=torch.int64, device=dev)

    def time_system(q):
        s = w.schedule("S")
        s.add_system("integrate", "integrate", q)
        s.build()
        for _ in range(warm):
            s.run_and_wait(capi.EXEC_SERIAL)
        ms = []
        for _ in range(steps):
            flush.fill_(1)
            torch.sum(flush_rd, dim=(0,), dtype=torch.int64, out=sink)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            s.run(capi.EXEC_SERIAL)
            b.record(stream)
            s.apply_deferred()
            ms.append(a.elapsed_time(b))
        s.close()
        return float(np.median(ms))

    def tag(entities, name):
        cmds = np.zeros(entities.size, dtype=capi.COMMAND_DTYPE)
        cmds["kind"] = capi.CMD_ADD
        cmds["entity"] = entities
        cmds["component_mask"] = w.mask([name])
        w.enqueue(cmds)
        w.stage_commands()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        w.flush()
        torch.cuda.synchronize()
        return time.perf_counter() - t0

    out, launches = {}, 0
    l0 = w.launch_count()
    with torch.cuda.stream(stream):
        plain = time_system(w.query([(P, True), (V, False)]))
        out["no_sparse_terms"] = {"ms": plain, "GBps": n * 24 / (plain * 1e-3) / 1e9}
        q = w.query([(P, True), (V, False)], with_=[TAG], without=[FROZEN])

        def leg(key, tagged):
            visited = q.count()
            ms = time_system(q)
            # algorithmic bytes: entity id 8 + membership word 4 per row; Position r/w + Velocity r only for rows that pass
            nbytes = n * 12 + visited * 24
            out[key] = {"ms": ms, "visited": visited, "GBps": nbytes / (ms * 1e-3) / 1e9, "frac": nbytes / (ms * 1e-3) / 1e9 / peak[0]}
            if tagged * 32 < n:
                # rare required term: the launch walks the storage's dense array (for_each_sparse_driven_kernel), so the per-row
                # bytes above are not touched at all — five dependent gathers per member, latency/launch-bound, no roofline claim
                out[key].update({"GBps": None, "frac": None, "driver": "sparse set dense array",
                                 "table_walk_bytes_avoided": int(nbytes)})

        # 0.1 %: one entity in 1024 tagged (a rare tag: the selective kernel variant); 1/64 of all entities frozen
        tag(ents[::1024], TAG)
        tag(ents[32::64], FROZEN)
        leg("tagged_0.1pct", n // 1024)
        # 25 %: every fourth entity
        dt_tag = tag(ents[::4], TAG)
        out["apply_tag_commands"] = {"commands": int(ents[::4].size), "ms": dt_tag * 1e3, "commands_per_s": ents[::4].size / dt_tag,
                                     "note": "ADD PlayerTag through helios_gpu_world_flush (records already staged in HBM): command "
                                             "chains + ranked parallel placement of the appends"}
        leg("tagged_25pct", n // 4)
        tag(np.delete(ents, np.arange(0, n, 4)), TAG)
        leg("tagged_100pct", n)
    launches = w.launch_count() - l0
    cpu = None
    if not args.no_cpu_baseline:
        o = capi.load_oracle()
        sn = n // 16
        wo = capi.World(o)
        worlds.register_all(wo, ["Position", "Velocity"])
        wo.register_sparse_component(TAG, 0)
        wo.register_sparse_component(FROZEN, 0)
        eo = wo.spawn_batch([P, V], sn, [worlds.column("Position", sn, 1), worlds.column("Velocity", sn, 2)])
        cm = np.zeros(eo.size, dtype=capi.COMMAND_DTYPE)
        cm["kind"] = capi.CMD_ADD
        cm["entity"] = eo
        cm["component_mask"] = wo.mask([TAG])
        wo.enqueue(cm)
        wo.flush()
        so = wo.schedule("S")
        so.add_system("integrate", "integrate", wo.query([(P, True), (V, False)], with_=[TAG], without=[FROZEN]))
        so.build()
        so.run_and_wait(capi.EXEC_SERIAL)
        t0 = time.perf_counter()
        for _ in range(3):
            so.run_and_wait(capi.EXEC_SERIAL)
        dt = (time.perf_counter() - t0) / 3
        cpu = {"value": sn / dt, "unit": "rows/s", "cores": 1, "kind": "port",
               "sample": f"tag-filtered integrate over {sn} rows (1/16 scale), all tagged, 3 repetitions"}
        wo.close()
    w.close()
    head = out["tagged_100pct"]
    return {
        "metric": "sparse_filtered_rows_per_sec", "value": n / (head["ms"] * 1e-3), "unit": "rows/s", "ms_per_step": head["ms"],
        "dtype": "f32 bodies, u32 membership words",
        "config": {"workload": f"sparse: integrate With<PlayerTag> Without<FrozenTag> over {n} rows, one table; tags are sparse-set "
                               "components attached by deferred commands",
                   "l2": "flushed between steps: 256 MiB written, then 256 MiB read"},
        "roofline": {"bound": "hbm", "achieved": head["GBps"], "peak": peak[0], "unit": "GB/s", "frac": head["frac"], "traffic": None,
                     "peak_kind": peak[1], "algorithmic_bytes_per_launch": int(n * 12 + head["visited"] * 24),
                     "kernel": "helios_gpu::for_each_kernel<..., kSparse = true, IntegrateBody, Position, Velocity>"},
        "sparse": out, "e2e": None, "cpu_baseline": cpu, "gpu_launches": int(launches),
    }


def run_extract(args, torch, dev, stream, peak):
    """SubApp::Extract: the cfg2 main world (16.8M entities / 64 tables); every frame the render
    sub-world is cleared and refilled with {Position, Velocity} of the rows `With<Health>, Without<Frozen>` matches
    (16 tables, 4.19M rows) — id placement + one device-to-device gather launch on the sub-world's stream."""
    n = args.entities or (1 << 24)
    gpu = capi.load_gpu()
    main = capi.World(gpu, device=dev.index, entity_capacity=n + 1024)
    main.set_stream(stream.cuda_stream)
    worlds.build_lattice(main, n, worlds.OPTIONAL)
    P, V, H, F = (worlds.cname(x) for x in ("Position", "Velocity", "Health", "Frozen"))
    q = main.query([(P, False), (V, False)], with_=[H], without=[F])
    rows = q.count()
    steps, warm = args.steps, max(3, args.warmup)

    def measure(flags):
        sub = capi.World(gpu, device=dev.index, entity_capacity=rows + 1024, flags=flags)
        sub.set_stream(stream.cuda_stream)
        ms, wall = [], []
        l0 = sub.launch_count()
        with torch.cuda.stream(stream):
            for i in range(warm + steps):
                sub.clear_entities()
                sub.update()
                sub.update()  # the previous frame's builtin messages age out (a sub-app's World::Update every frame)
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0 = time.perf_counter()
                a.record(stream)
                got = sub.extract_from(q)
                b.record(stream)
                torch.cuda.synchronize()
                if i >= warm:
                    ms.append(a.elapsed_time(b))
                    wall.append((time.perf_counter() - t0) * 1e3)
                assert got == rows
        launches = sub.launch_count() - l0
        sub.close()
        return float(np.median(ms)), float(np.median(wall)), int(launches)

    # per row: Position + Velocity read and written (32 B), entity id written (8), location + generation written (12);
    # with the builtin structural messages on (the default world): EntityAddedMsg + ComponentAddedMsg<Position> +
    # ComponentAddedMsg<Velocity>, 8 B each, written by the destination's CreateEntity + AddComponents (world.hpp:982,1035)
    med, wall, launches = measure(0)
    med_q, wall_q, launches_q = measure(capi.WORLD_NO_BUILTIN_MESSAGES)
    bpr, bpr_q = 16 * 2 + 8 + 12 + 24, 16 * 2 + 8 + 12
    main.close()
    return {
        "metric": "extracted_rows_per_sec", "value": rows / (med * 1e-3), "unit": "rows/s", "ms_per_step": med,
        "dtype": "bytes", "config": {"workload": f"extract: {rows} rows x (Position, Velocity) from 16 of 64 main-world tables into a "
                                                 "cleared sub-world (builtin structural messages written, as the reference does)",
                                     "l2": "working set 320 MB exceeds L2; no flush"},
        "roofline": {"bound": "hbm", "achieved": rows * bpr / (med * 1e-3) / 1e9, "peak": peak[0], "unit": "GB/s",
                     "frac": rows * bpr / (med * 1e-3) / 1e9 / peak[0], "traffic": None, "peak_kind": peak[1],
                     "algorithmic_bytes_per_launch": rows * bpr,
                     "kernel": "k_match + k_place_entities + k_extract_rows (one launch for every table and column) + 3 message writers"},
        "without_builtin_messages": {"ms_per_step": med_q, "value": rows / (med_q * 1e-3), "bytes_per_row": bpr_q,
                                     "frac": rows * bpr_q / (med_q * 1e-3) / 1e9 / peak[0], "gpu_launches": launches_q,
                                     "host_wall_ms": wall_q},
        "host_wall_ms": wall, "e2e": None, "cpu_baseline": None, "gpu_launches": launches,
    }


def run_conditions(args, torch, dev, stream, peak):
    """Run conditions: what guarding the cfg2 headline system costs.  Same world and body as the
    headline; the system is (a) unguarded, (b) guarded by two conditions that pass — a query terminal `Any(pos.x > t)` over
    the same 4.19M rows plus a constant — and (c) guarded by a condition that fails.  Device time of Schedule::Run per mode."""
    n = args.entities or (1 << 24)
    gpu = capi.load_gpu()
    w = capi.World(gpu, device=dev.index, entity_capacity=n + 1024)
    w.set_stream(stream.cuda_stream)
    worlds.build_lattice(w, n, worlds.OPTIONAL)
    P, V, A, M, H, F = (worlds.cname(x) for x in ("Position", "Velocity", "Acceleration", "Mass", "Health", "Frozen"))

    def make(kind):
        q = w.query([(P, True), (V, True), (A, False), (M, False)], with_=[H], without=[F])
        s = w.schedule("Update_" + kind)
        i = s.add_system("AccelIntegrate", "accel_integrate", q, params=struct.pack("<f", worlds.DT))
        if kind == "pass":
            s.run_if(i, "AnyOnTheRight", capi.COND_TERMINAL, builtin="count_if_pos_x_gt", query=w.query([(P, False)], with_=[H], without=[F]),
                     params=struct.pack("<f", 512.0), field=capi.COND_FIELD_COUNT, cmp=capi.CMP_GT, rhs=0)
            s.run_if(i, "AlwaysTrue", capi.COND_CONSTANT, rhs=1)
        elif kind == "pass_cheap":
            s.run_if(i, "AlwaysTrue", capi.COND_CONSTANT, rhs=1)
        elif kind == "skip":
            s.run_if(i, "AlwaysFalse", capi.COND_CONSTANT, rhs=0)
        s.build()
        return s, q

    steps, warm = args.steps, max(3, args.warmup)
    res, launches = {}, {}
    rows = None
    with torch.cuda.stream(stream):
        for kind in ("unguarded", "pass_cheap", "pass", "skip"):
            s, q = make(kind)
            rows = q.count()
            for mode, mname in ((capi.EXEC_SERIAL, "serial"), (capi.EXEC_GRAPH, "graph")):
                for _ in range(warm):
                    s.run_and_wait(mode)
                torch.cuda.synchronize()
                l0 = w.launch_count()
                evs = _events(torch, steps)
                for a, b in evs:
                    a.record(stream)
                    s.run(mode)
                    b.record(stream)
                    s.apply_deferred()
                torch.cuda.synchronize()
                res[f"{kind}_{mname}"] = float(np.median([a.elapsed_time(b) for a, b in evs]))
                launches[f"{kind}_{mname}"] = (w.launch_count() - l0) / steps
                assert s.last_verdicts() == [0 if kind == "skip" else 1]
    ms = res["pass_cheap_graph"]
    bytes_per_launch = rows * worlds.CFG2_BYTES_PER_ENTITY
    achieved = bytes_per_launch / (ms * 1e-3) / 1e9
    w.close()
    return {
        "metric": "entity_component_updates_per_sec", "value": rows / (ms * 1e-3), "unit": "updates/s", "ms_per_step": ms,
        "dtype": "f32",
        "config": {"workload": f"conditions: cfg2 filtered query ({rows} rows of {n} entities) behind device run conditions; "
                               "value = guarded-and-passing (constant condition), CUDA-graph executor",
                   "l2": "184 MB touched per launch exceeds L2; no flush"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak[0], "unit": "GB/s", "frac": achieved / peak[0], "traffic": None,
                     "peak_kind": peak[1], "algorithmic_bytes_per_launch": bytes_per_launch,
                     "kernel": "k_cond_begin + k_cond_test + k_cond_commit + IF{for_each_kernel<AccelBody>}"},
        "modes_ms": res, "launches_per_run": launches, "e2e": None, "cpu_baseline": None,
        "gpu_launches": int(sum(launches.values()) * steps),
    }


def run_async(args, torch, dev, stream, peak):
    """Async messages: n entities {Position, Health}; `alert_low_health` pushes a 16-byte alert for
    every hp < 40 (~39 % of the rows) through warp-aggregated appends to the device ring, `heal_alerted` drains the ring and
    heals the entities by id (random access through the location table)."""
    n = args.entities or (1 << 24)
    gpu = capi.load_gpu()
    w = capi.World(gpu, device=dev.index, entity_capacity=n + 1024)
    w.set_stream(stream.cuda_stream)
    worlds.register_all(w, ["Position", "Health"])
    H = worlds.cname("Health")
    w.spawn_batch([worlds.cname("Position"), H], n, [worlds.column("Position", n, 5100), worlds.column("Health", n, 5101)],
                  want_entities=False)
    ALERT = "helios::bench::LowHealthAlert"
    w.async_message_register(ALERT, 16, capacity=n)
    params = struct.pack("<QiIQ", capi.fnv1a(ALERT), 40, w.slots[H], 0)
    emit = w.schedule("Emit")
    emit.add_system("S1_alert", "alert_low_health", w.query([(H, False)]), params=params)
    emit.build()
    heal = w.schedule("Heal")
    heal.add_system("S2_heal", "heal_alerted", None, params=params, extra_writes=[H])
    heal.build()
    steps, warm = args.steps, max(3, args.warmup)
    t_push, t_drain, pushed = [], [], 0
    hp0 = np.ascontiguousarray(worlds.column("Health", n, 5101))
    arch = [i for i, a in enumerate(w.snapshot(columns=False)["archetypes"]) if a["entities"].size == n][0]
    l0 = w.launch_count()
    with torch.cuda.stream(stream):
        for i in range(warm + steps):
            a, b, c, d = (torch.cuda.Event(enable_timing=True) for _ in range(4))
            a.record(stream)
            emit.run(capi.EXEC_SERIAL)
            b.record(stream)
            emit.apply_deferred()
            pushed = w.async_message_size(ALERT)
            c.record(stream)
            heal.run(capi.EXEC_SERIAL)
            d.record(stream)
            heal.apply_deferred()
            torch.cuda.synchronize()
            assert w.async_message_size(ALERT) == 0
            if i >= warm:
                t_push.append(a.elapsed_time(b))
                t_drain.append(c.elapsed_time(d))
            # undo the healing so every step pushes the same number of alerts
            w.async_message_clear(ALERT)
            w.column_write(arch, H, 0, n, hp0.ctypes.data)
    launches = w.launch_count() - l0
    push_ms, drain_ms = float(np.median(t_push)), float(np.median(t_drain))
    push_bytes = n * (8 + 4) + pushed * 16          # entity id + Health read per row, one 16-byte alert written per push
    drain_bytes = pushed * (16 + 8 + 4 + 4 + 4)      # alert read, location + generation read, Health read-modify-write
    w.close()
    return {
        "metric": "async_messages_per_sec", "value": pushed / (push_ms * 1e-3), "unit": "messages/s", "ms_per_step": push_ms,
        "dtype": "bytes",
        "config": {"workload": f"async: {n} entities, {pushed} alerts of 16 B pushed per step by for_each_kernel<AlertBody> "
                               "(one tail atomic per warp), then drained by async_drain_kernel",
                   "l2": "201 MB scanned per push launch exceeds L2; no flush"},
        "roofline": {"bound": "hbm", "achieved": push_bytes / (push_ms * 1e-3) / 1e9, "peak": peak[0], "unit": "GB/s",
                     "frac": push_bytes / (push_ms * 1e-3) / 1e9 / peak[0], "traffic": None, "peak_kind": peak[1],
                     "algorithmic_bytes_per_launch": push_bytes, "kernel": "for_each_kernel<AlertBody, Health> (push)"},
        "drain": {"ms": drain_ms, "messages_per_sec": pushed / (drain_ms * 1e-3), "algorithmic_bytes": drain_bytes,
                  "achieved_GBps": drain_bytes / (drain_ms * 1e-3) / 1e9,
                  "note": "random 4-byte read-modify-writes by entity id: sector-bound, not stream-bound"},
        "e2e": None, "cpu_baseline": None, "gpu_launches": int(launches),
    }


def run_routing(args, torch, dev, stream, peak, dist=None, rank=0, world_size=1):
    """Cross-shard message routing: every rank holds 4 194 304 CollisionEvents (cfg 4's per-frame volume)
    keyed uniformly over the entity-index ranges of all ranks; a step = stable split by owner on the device, all-to-all of
    counts and slabs (NCCL over NVLink), merge.  value = messages routed per second over all ranks (max-over-ranks time)."""
    from heliosengine_b200 import routing
    n = args.entities or (1 << 22)
    per = 1 << 22
    gpu = capi.load_gpu()
    w = capi.World(gpu, device=dev.index)
    w.set_stream(stream.cuda_stream)
    name = worlds.COLLISION_EVENT
    w.message_register(name, 24)
    rng = np.random.default_rng(77 + rank)
    ev = np.zeros(n, dtype=np.dtype([("a", "<u8"), ("b", "<u8"), ("impulse", "<f4"), ("pad", "<u4")]))
    ev["a"] = rng.integers(0, per * world_size, n).astype(np.uint64) | (np.uint64(1) << np.uint64(32))
    ev["pad"] = np.arange(n, dtype=np.uint32)
    ev_dev = torch.from_numpy(ev.view(np.uint8).reshape(-1).copy()).to(dev)
    slab = torch.empty(n * 24, dtype=torch.uint8, device=dev)
    steps, warm = args.steps, max(3, args.warmup)
    t_split, t_total = [], []
    use_peer = dist is not None and world_size > 1 and getattr(args, "exchange", "p2p") == "p2p"
    peer = routing.PeerRouter(w, name, torch, dist, dev, rank, world_size, per, inbox_capacity=2 * n) if use_peer else None
    info = {"received": [0]}
    l0 = w.launch_count()
    with torch.cuda.stream(stream):
        for i in range(warm + steps):
            routing.merge(w, name, 1, ev_dev, n, 0, 0)  # reload this rank's outgoing queue
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            a.record(stream)
            routing.split(w, name, 1, 0, per, world_size, slab)
            b.record(stream)
            if peer is not None:
                info = {"received": [peer.route()]}
            else:
                info = routing.route(w, name, torch, dev, dist, rank, world_size, per)
            c.record(stream)
            torch.cuda.synchronize()
            if i >= warm:
                t_split.append(a.elapsed_time(b))
                t_total.append(b.elapsed_time(c))
    if peer is not None:
        peer.close()
    launches = w.launch_count() - l0
    split_ms, route_ms = float(np.median(t_split)), float(np.median(t_total))
    if dist is not None:
        t = torch.tensor([split_ms, route_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        split_ms, route_ms = float(t[0]), float(t[1])
    received = sum(info["received"])
    w.close()
    split_bytes = n * (24 + 24 + 24)  # key pass reads every sector of the 24-byte records, scatter reads + writes them
    return {
        "metric": "routed_messages_per_sec", "value": n * world_size / (route_ms * 1e-3), "unit": "messages/s",
        "ms_per_step": route_ms, "dtype": "bytes",
        "config": {"workload": f"routing: {n} CollisionEvents (24 B) per rank keyed uniformly over {world_size} entity-index shards; "
                               "step = split + all_to_all(counts) + all_to_all(slabs) + merge",
                   "l2": "100 MB of records per rank per pass; slabs and queue exceed L2 together; no flush",
                   "parallelism": (f"entity-range shards x{world_size}, " + ("peer-memory scatter over NVLink (2 four-byte all-reduce barriers)"
                                   if use_peer else "NCCL all-to-all")) if world_size > 1 else "1 rank (no exchange)"},
        "roofline": {"bound": "hbm", "achieved": split_bytes / (split_ms * 1e-3) / 1e9, "peak": peak[0], "unit": "GB/s",
                     "frac": split_bytes / (split_ms * 1e-3) / 1e9 / peak[0], "traffic": None, "peak_kind": peak[1],
                     "algorithmic_bytes_per_launch": split_bytes,
                     "kernel": "k_route_hist + scan + k_route_scatter (the split alone, incl. the count read-back)"},
        "split_ms": split_ms, "received_last_step": received, "scaling": "weak",
        "e2e": None, "cpu_baseline": None, "gpu_launches": int(launches),
    }


def run_cfg5(args, torch, dev, stream, peak, dist=None, rank=0, world_size=1):
    """Boids, STRONG scaling: `--entities` boids in total (default 64M, density 4 per cell) split by entity-id range over
    the ranks; per step: bin own shard, all-gather slabs + counts (NCCL), neighbour pass + integrate."""
    from heliosengine_b200 import boids
    if getattr(args, "boids", "bands") == "bands":
        return run_cfg5_bands(args, torch, dev, stream, peak, dist, rank, world_size)
    n_total = args.entities or (1 << 26)
    ws = int(round((n_total / 4) ** 0.5))
    gpu = capi.load_gpu()
    shard = boids.BoidsShard(gpu, n_total, ws, world_size, rank, device=dev.index)
    shard.world.set_stream(stream.cuda_stream)
    buf = boids.ExchangeBuffers(torch, dev, world_size, shard.n_local, shard.n_cells)
    steps, warm = args.steps, max(3, args.warmup)
    # N > 1: peer-memory exchange (merge kernels read the other GPUs' slabs over NVLink) unless --exchange nccl
    peer = None
    if world_size > 1 and getattr(args, "exchange", "p2p") == "p2p":
        with torch.cuda.stream(stream):
            peer = boids.PeerExchange(shard, torch, dist, dev)

    def one_step():
        if peer:
            peer.step()
        else:
            boids.distributed_step(shard, buf, dist)

    def barrier():
        torch.cuda.synchronize()
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(warm):
            one_step()
        barrier()
        l0 = shard.world.launch_count()
        evs = _events(torch, steps)
        for a, b in evs:
            a.record(stream)
            one_step()
            b.record(stream)
        barrier()
        launches = shard.world.launch_count() - l0
        step_ms = [a.elapsed_time(b) for a, b in evs]
        # phase split (one extra untimed-in-`value` step): bin / exchange / neighbour pass
        ph = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        g = shard.shard
        ph[0].record(stream)
        if peer:
            shard.exchange_bin()
            ph[1].record(stream)
            dist.all_reduce(peer.token)
            ph[2].record(stream)
            shard.exchange_step()
        else:
            shard.bin(buf.slab_slot(g).data_ptr(), buf.counts_slot(g).data_ptr())
            ph[1].record(stream)
            if world_size > 1:
                dist.all_gather_into_tensor(buf.slabs, buf.slab_slot(g))
                dist.all_gather_into_tensor(buf.counts, buf.counts_slot(g))
            ph[2].record(stream)
            shard.step(buf.slabs.data_ptr(), buf.counts.data_ptr())
        ph[3].record(stream)
        barrier()
        phases = {"bin_ms": ph[0].elapsed_time(ph[1]), "exchange_ms": ph[1].elapsed_time(ph[2]),
                  "neighbour_ms": ph[2].elapsed_time(ph[3])}

        # e2e: host state -> device, step, positions -> host (rank-local slices; PCIe-bound by construction)
        e2e = None
        if not args.no_e2e:
            P, V = worlds.cname("Position"), worlds.cname("Velocity")
            hp = torch.empty(shard.n_local * 2, dtype=torch.float32).pin_memory()
            hv = torch.empty(shard.n_local * 2, dtype=torch.float32).pin_memory()
            pos0, vel0 = boids.initial_state(n_total, ws, *boids.shard_range(n_total, world_size, rank))
            hp.copy_(torch.from_numpy(pos0))
            hv.copy_(torch.from_numpy(vel0))
            ho = torch.empty(shard.n_local * 2, dtype=torch.float32).pin_memory()
            tbl = 1  # {} is table 0, {Position, Velocity} table 1
            e2e_steps = 3
            barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                shard.world.column_write(tbl, P, 0, shard.n_local, hp.data_ptr())
                shard.world.column_write(tbl, V, 0, shard.n_local, hv.data_ptr())
                one_step()
                shard.world.column_read(tbl, P, 0, shard.n_local, ho.data_ptr())
                shard.world.synchronize()
            barrier()
            e2e = (time.perf_counter() - t0) / e2e_steps

    tm = torch.tensor([sum(step_ms), e2e or 0.0, phases["bin_ms"], phases["exchange_ms"], phases["neighbour_ms"]],
                      dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    tot_ms, e2e_s, bin_ms, ex_ms, nb_ms = (float(x) for x in tm.tolist())
    ms = tot_ms / steps
    local_bytes = 64 * shard.n_local  # compulsory HBM bytes per boid per step on the owning GPU
    achieved = local_bytes / (ms * 1e-3) / 1e9

    cpu = None
    if rank == 0 and world_size == 1 and not args.no_cpu_baseline:
        o = capi.load_oracle()
        sample_n, sample_ws = 1 << 20, 512
        so = boids.BoidsShard(o, sample_n, sample_ws)
        nthreads = int(o.fn("hardware_threads", ctypes.c_uint32)())
        o.fn("set_access_mode")(so.world.h, 0, nthreads)
        so.step_local()
        t0 = time.perf_counter()
        for _ in range(2):
            so.step_local()
        dt = (time.perf_counter() - t0) / 2
        so.close()
        cpu = {"value": sample_n / dt, "unit": "boid updates/s", "cores": nthreads, "kind": "port",
               "sample": f"{sample_n} boids in a {sample_ws}^2 world (same density, 4 per cell), 2 steps; binning single-threaded, "
                         f"neighbour pass on {nthreads} threads; no reference implementation exists for this config"}
    line = {
        "metric": "boid_updates_per_sec", "value": n_total / (ms * 1e-3), "unit": "boid updates/s", "ms_per_step": ms,
        "dtype": "f32 pair terms, int64 fixed-point sums", "scaling": "strong",
        "config": {"workload": f"cfg5: {n_total} boids, world {ws}^2, cell = radius = 1 (density 4/cell), 3x3 stencil, "
                               f"entity-id range shards x{world_size}",
                   "parallelism": (f"{world_size} shard(s); peer-memory exchange: the merge kernels read the other GPUs' slabs "
                                   f"(16 B/boid) and u16 cell counts in place over NVLink, {peer.bytes_pulled_per_step} B pulled "
                                   f"per rank and step; one 4-byte NCCL all-reduce per step as the barrier" if peer else
                                   f"{world_size} shard(s); per step 2 NCCL all-gathers (slabs 16 B/boid + u16 cell counts), "
                                   f"{buf.exchange_bytes_per_rank} B received per rank"),
                   "l2": "per-step working set (slabs + tables) exceeds L2; no flush",
                   "parity": "bit-exact vs the CPU statement at test sizes, any shard count; no reference implementation (unpinned)"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak[0], "unit": "GB/s", "frac": achieved / peak[0],
                     "traffic": None, "peak_kind": peak[1], "algorithmic_bytes_per_launch": local_bytes,
                     "kernel": "bin (count/scan/scatter) + k_boids_neighbours, per GPU; exchange excluded from the bytes",
                     "exchange_bytes_per_rank": peer.bytes_pulled_per_step if peer else buf.exchange_bytes_per_rank},
        "phases_ms": {"bin": bin_ms, "exchange": ex_ms, "neighbour": nb_ms},
        "e2e": ({"value": n_total / e2e_s, "unit": "boid updates/s", "h2d_bytes_per_step": shard.n_local * 16,
                 "d2h_bytes_per_step": shard.n_local * 8,
                 "note": "per rank and step: Position+Velocity columns from pinned host memory, step, Position back"}
                if e2e_s else None),
        "cpu_baseline": cpu, "gpu_launches": int(launches),
    }
    if peer:
        peer.close()
    shard.close()
    return line


def run_cfg5_bands(args, torch, dev, stream, peak, dist=None, rank=0, world_size=1):
    """Boids over spatial bands, STRONG scaling: `--entities` boids in total (default 64M, density 4 per cell).  Rank g owns
    the grid rows [g * R, (g + 1) * R); per step: migrate the boids that left the band (router + Destroy / Spawn commands),
    bin the own boids, publish the two edge rows, barrier, read ONE row from each neighbour in place over NVLink, neighbour
    pass + integrate.  The initial placement is the entity-id range: the first step sorts the flock into bands and is timed
    and reported apart (`first_step_ms`)."""
    from heliosengine_b200 import boids, routing
    n_total = args.entities or (1 << 26)
    ws = int(round((n_total / 4) ** 0.5))
    gpu = capi.load_gpu()
    shard = boids.BandShard(gpu, n_total, ws, world_size, rank, device=dev.index)
    shard.world.set_stream(stream.cuda_stream)
    steps, warm = args.steps, max(3, args.warmup)

    def barrier():
        torch.cuda.synchronize()
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        x = boids.BandExchange(shard, torch, dist, dev, peer=world_size > 1, barrier=getattr(args, "band_barrier", "flags"))
        barrier()
        a, b = _events(torch, 1)[0]
        a.record(stream)
        x.step()  # the initial sort into bands: (G-1)/G of the flock migrates
        b.record(stream)
        barrier()
        first_ms = a.elapsed_time(b)
        first_moved = x.migrated[0] if x.migrated else (0, 0)
        for _ in range(warm):
            x.step()
        import gc
        gc.collect()
        gc.disable()  # a step is ~1 ms with three host read-backs on each of N ranks: a collector pause on ONE rank stalls all
        l0 = shard.world.launch_count()
        evs = _events(torch, steps)
        barrier()     # last thing before the timed region: the ranks enter it together (a collection takes a different time
                      # on every rank, and a rank that starts late is waited for by its neighbours inside step 1)
        for a, b in evs:
            a.record(stream)
            x.step()
            b.record(stream)
        barrier()
        gc.enable()
        launches = shard.world.launch_count() - l0
        step_ms = [a.elapsed_time(b) for a, b in evs]
        moved = x.migrated[-steps:] if x.migrated else []
        # phase split (one extra step, not part of `value`)
        ph = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        ph[0].record(stream)
        if world_size > 1:
            shard.leavers()
            routing.route(shard.world, boids.MIGRATION, torch, dev, dist, rank, world_size, shard.rows_per, 0, 1, False)
            shard.arrivals()
        ph[1].record(stream)
        shard.bin()
        ph[2].record(stream)
        if world_size > 1 and not x.flags:
            dist.all_reduce(x.token)
        ph[3].record(stream)
        shard.step()
        ph[4].record(stream)
        barrier()
        phases = [ph[i].elapsed_time(ph[i + 1]) for i in range(4)]
        n_local = shard.n_local()

        e2e = None
        if not args.no_e2e:
            P, V = worlds.cname("Position"), worlds.cname("Velocity")
            snap = shard.world.snapshot(columns=False)
            tbl = [i for i, a_ in enumerate(snap["archetypes"]) if len(a_["types"]) == 3][0]
            _, p0, v0 = shard.state()
            hp, hv = torch.from_numpy(p0).pin_memory(), torch.from_numpy(v0).pin_memory()
            ho = torch.empty(n_local * 2, dtype=torch.float32).pin_memory()
            e2e_steps = 3
            barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):  # the same in-band state every step: the row count stays put
                shard.world.column_write(tbl, P, 0, n_local, hp.data_ptr())
                shard.world.column_write(tbl, V, 0, n_local, hv.data_ptr())
                x.step()
                shard.world.column_read(tbl, P, 0, n_local, ho.data_ptr())
                shard.world.synchronize()
            barrier()
            e2e = (time.perf_counter() - t0) / e2e_steps

    per_step = torch.tensor(step_ms, dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(per_step, op=dist.ReduceOp.MAX)  # a step is as slow as its slowest rank
    step_ms_max = [float(v) for v in per_step.tolist()]
    tm = torch.tensor([sum(step_ms), e2e or 0.0, first_ms] + phases, dtype=torch.float64, device=dev)
    cnt = torch.tensor([n_local, first_moved[0], sum(m[0] for m in moved)], dtype=torch.float64, device=dev)
    nmax = torch.tensor([float(n_local)], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        dist.all_reduce(nmax, op=dist.ReduceOp.MAX)
    tot_ms, e2e_s, first_ms, mig_ms, bin_ms, bar_ms, nb_ms = (float(v) for v in tm.tolist())
    assert int(cnt[0]) == n_total, "boids lost or duplicated by the migration"
    ms = sum(step_ms_max) / steps
    local_bytes = 64 * float(nmax[0])  # compulsory HBM bytes per boid per step on the owning GPU (largest band)
    achieved = local_bytes / (ms * 1e-3) / 1e9
    halo_bytes = 2 * (4 * ws * 16 + (ws + 1) * 4 + 16) if world_size > 1 else 0

    cpu = None
    if rank == 0 and world_size == 1 and not args.no_cpu_baseline:
        o = capi.load_oracle()
        sample_n, sample_ws = 1 << 20, 512
        so = boids.BoidsShard(o, sample_n, sample_ws)
        nthreads = int(o.fn("hardware_threads", ctypes.c_uint32)())
        o.fn("set_access_mode")(so.world.h, 0, nthreads)
        so.step_local()
        t0 = time.perf_counter()
        for _ in range(2):
            so.step_local()
        dt = (time.perf_counter() - t0) / 2
        so.close()
        cpu = {"value": sample_n / dt, "unit": "boid updates/s", "cores": nthreads, "kind": "port",
               "sample": f"{sample_n} boids in a {sample_ws}^2 world (same density, 4 per cell), 2 steps; binning single-threaded, "
                         f"neighbour pass on {nthreads} threads; no reference implementation exists for this config"}
    line = {
        "metric": "boid_updates_per_sec", "value": n_total / (ms * 1e-3), "unit": "boid updates/s", "ms_per_step": ms,
        "dtype": "f32 pair terms, int64 fixed-point sums", "scaling": "strong",
        "config": {"workload": f"cfg5: {n_total} boids, world {ws}^2, cell = radius = 1 (density 4/cell), 3x3 stencil, "
                               f"spatial bands x{world_size} ({shard.rows_per} grid rows each), one-row halo exchange",
                   "parallelism": (f"{world_size} band(s); per step each rank reads its two neighbours' edge rows in place over NVLink "
                                   f"(~{halo_bytes} B), band-crossing boids migrate through the message router (NCCL all-to-all of "
                                   "32-byte records, one fixed-size all-gather in the steady state) and Destroy / Spawn commands; barrier: "
                                   + ("neighbours wait on each other's published step numbers on the device" if x.flags else "one 4-byte NCCL all-reduce")
                                   if world_size > 1 else "1 band = the whole grid (no exchange, no migration)"),
                   "l2": "per-step working set (slab + tables) exceeds L2; no flush",
                   "parity": "per boid id bit-exact vs the single-world CPU statement at test sizes, any band count "
                             "(tests/test_boids_bands.py); no reference implementation (unpinned)"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak[0], "unit": "GB/s", "frac": achieved / peak[0],
                     "traffic": None, "peak_kind": peak[1], "algorithmic_bytes_per_launch": local_bytes,
                     "kernel": "bin (count/scan/scatter) + k_boids_neighbours on the largest band; exchange excluded from the bytes",
                     "exchange_bytes_per_rank": halo_bytes},
        "phases_ms": {"migrate": mig_ms, "bin": bin_ms, "barrier": bar_ms, "neighbour": nb_ms},
        "first_step_ms": first_ms, "first_step_migrated": int(cnt[1]), "migrated_per_step": float(cnt[2]) / max(1, steps),
        "largest_band": int(nmax[0]), "step_ms_rank0": step_ms, "step_ms_max_over_ranks": step_ms_max,
        # `value` uses the MEAN step (stalls included); the median is the steady state between stalls
        "median_ms_per_step": float(np.median(step_ms_max)), "value_at_median": n_total / (float(np.median(step_ms_max)) * 1e-3),
        "e2e": ({"value": n_total / e2e_s, "unit": "boid updates/s", "h2d_bytes_per_step": int(nmax[0]) * 16,
                 "d2h_bytes_per_step": int(nmax[0]) * 8,
                 "note": "per rank and step: Position+Velocity columns from pinned host memory, step, Position back"}
                if e2e_s else None),
        "cpu_baseline": cpu, "gpu_launches": int(launches),
    }
    x.close()
    shard.close()
    return line


def run_cfg3_sharded(args, torch, dev, stream, peak, dist=None, rank=0, world_size=1):
    """cfg3 over entity-range shards (SURVEY.md §8e "Commands apply"), WEAK scaling: every rank holds cfg3's world (`--entities`
    per rank, default 8 388 608 in 16 tables) and emits cfg3's per-frame command mix against entities of EVERY shard.
    A step = stable split of the rank's 64-byte command records by owner + all-to-all (NCCL over NVLink) + merge + device
    unpack + ApplyDeferred on the owner.  value = commands applied per second over all ranks (max-over-ranks step time).

    Untimed per frame: each rank draws cfg3's valid, distinct targets from its own tables, and the records are dealt out
    round-robin to the other ranks so that every rank's outgoing queue names entities of all shards; the spawn ids come from
    the last shard's next index (one sequence in (rank, emission) order), every rank emitting 1/G of cfg3's spawn count."""
    from heliosengine_b200 import routing
    n = args.entities or (1 << 23)
    g = world_size
    gpu = capi.load_gpu()
    w = capi.World(gpu, device=dev.index, entity_capacity=n + n // 2, flags=capi.WORLD_NO_BUILTIN_MESSAGES)
    w.set_stream(stream.cuda_stream)
    worlds.build_cfg3(w, n)
    name = routing.ROUTED_COMMAND
    w.message_register(name, 64, capi.MSG_MANUAL)
    steps, warm = args.steps, max(3, args.warmup)
    dev_ms, apply_ms, e2e_s, n_applied, launches, invalid, trace = [], [], [], [], [], 0, []

    def exchange(t_send, send_counts):
        sc = torch.tensor(send_counts, dtype=torch.int64, device=dev)
        rc = torch.empty_like(sc)
        if dist is not None:
            dist.all_to_all_single(rc, sc)
        else:
            rc.copy_(sc)
        rcl = [int(x) for x in rc.cpu().tolist()]
        out = torch.empty(max(1, sum(rcl)) * 64, dtype=torch.uint8, device=dev)
        if dist is not None:
            dist.all_to_all_single(out[: sum(rcl) * 64], t_send, [c * 64 for c in rcl], [c * 64 for c in send_counts])
        else:
            out[: sum(rcl) * 64].copy_(t_send)
        return out, sum(rcl)

    with torch.cuda.stream(stream):
        for frame in range(warm + steps):
            snap = w.snapshot(columns=False)
            ents = [a["entities"] for a in snap["archetypes"]]
            types = [a["types"] for a in snap["archetypes"]]
            despawn, insert, remove, n_spawn = worlds.cfg3_commands(w, ents, types, frame * 64 + rank, int(n * 0.10))
            n_spawn //= g
            cmds, payload = worlds.assemble_cfg3(w, capi.spawn_placeholder(np.arange(n_spawn)), despawn, insert, remove, frame * 64 + rank)
            live = (cmds["entity"] >> np.uint64(32)) != capi.SPAWN_PLACEHOLDER_GENERATION
            cmds["entity"][live] += np.uint64(rank * n)  # local -> global index
            # ids: the routine of routing.route_commands, spelled out so that the records can be dealt before the timed region
            mine = torch.tensor([n_spawn, ((g - 1) * n + w.next_entity_index()) if rank == g - 1 else 0], dtype=torch.int64, device=dev)
            every = torch.empty(2 * g, dtype=torch.int64, device=dev)
            if dist is not None:
                dist.all_gather_into_tensor(every, mine)
            else:
                every.copy_(mine)
            every = every.cpu().view(g, 2)
            sp = [int(x) for x in every[:, 0].tolist()]
            first = int(every[g - 1, 1]) + sum(sp[:rank])
            timed = frame >= warm
            e2e_frame = timed and ((frame - warm) % 2 == 1)
            records = w.commands_route_pack(cmds, payload, first)
            # deal: record i of this rank is EMITTED by rank (rank + i) % g
            order = np.argsort((np.arange(records.size) % g), kind="stable")
            dealt = records[order]
            counts = np.bincount((np.arange(records.size) % g), minlength=g)
            send_counts = [0] * g
            for j in range(g):
                send_counts[(rank + j) % g] = int(counts[j])
            seg = np.concatenate([[0], np.cumsum(counts)])
            parts = [None] * g
            for j in range(g):
                parts[(rank + j) % g] = dealt[seg[j]:seg[j + 1]]
            t_send = torch.from_numpy(np.concatenate(parts).view(np.uint8).reshape(-1).copy()).to(dev)
            emitted, n_emit = exchange(t_send, send_counts)
            host_records = emitted[: n_emit * 64].cpu().numpy() if e2e_frame else None
            if rank == g - 1 and sum(sp):
                w.reserve_entities(sum(sp))
            if not e2e_frame:
                routing.merge(w, name, 1, emitted, n_emit, routing.ROUTED_KEY_OFFSET, 0)  # this rank's outgoing queue, device-resident
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            l0 = w.launch_count()
            a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            t0 = time.perf_counter()
            a.record(stream)
            if e2e_frame:
                w.message_write(name, host_records)
            info = routing.route(w, name, torch, dev, dist, rank, g, n, routing.ROUTED_KEY_OFFSET, 1, True)
            b.record(stream)
            w.commands_apply_routed(name, 1)
            c.record(stream)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            st = w.apply_stats()
            invalid += st.invalid
            trace.append((frame, "e2e" if e2e_frame else "dev", round(a.elapsed_time(b), 3), round(b.elapsed_time(c), 3),
                          int(st.pipeline), int(st.fast_fallbacks), int(st.resolve_rounds), round((t1 - t0) * 1e3, 3)))
            if timed:
                launches.append(w.launch_count() - l0)
                n_applied.append(sum(info["received"]))
                if e2e_frame:
                    e2e_s.append(t1 - t0)
                else:
                    dev_ms.append(a.elapsed_time(c))
                    apply_ms.append(b.elapsed_time(c))
            w.update()
    t = torch.tensor([float(np.mean(dev_ms)), float(np.mean(apply_ms)), float(np.mean(e2e_s)) if e2e_s else 0.0], device=dev, dtype=torch.float64)
    tot = torch.tensor([float(np.mean(n_applied)), float(invalid), float(w.entity_count())], device=dev, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms, ap_ms, e2e = float(t[0]), float(t[1]), float(t[2])
    total_cmds = float(tot[0])
    stats = w.apply_stats()
    w.close()
    assert int(tot[1]) == 0, f"{int(tot[1])} routed commands were invalid on their owner"
    return {
        "metric": "commands_applied_per_sec", "value": total_cmds / (ms * 1e-3), "unit": "commands/s", "ms_per_step": ms,
        "dtype": "u8/u64 (row bytes, entity ids)", "scaling": "weak",
        "config": {"workload": f"cfg3 sharded: {n} entities / 16 tables per rank x {g} ranks, {int(total_cmds)} commands per frame in "
                               "total, targets uniform over all shards; step = split by owner + all_to_all + merge + unpack + ApplyDeferred",
                   "l2": "per-rank working set (tables 8M x ~40 B) exceeds L2; no flush",
                   "parallelism": f"entity-range shards x{g}, NCCL all-to-all of 64-byte command records",
                   "parity": "tests/test_sharded_commands.py (per-shard oracle replay + single-world entity states)"},
        "apply_ms": ap_ms, "route_ms": ms - ap_ms, "median_ms_rank0": float(np.median(dev_ms)), "pipeline": int(stats.pipeline),
        "entities_after": int(tot[2]),
        "frames": trace,  # (frame, kind, route ms, apply ms, pipeline, fallbacks so far, resolve rounds, wall ms) on rank 0
        "roofline": None,
        "e2e": {"value": total_cmds / e2e if e2e else None, "unit": "commands/s", "h2d_bytes_per_step": int(total_cmds / g * 64),
                "d2h_bytes_per_step": 64 + 8 * g,
                "note": "host command records written to the queue inside the timed region; wall clock around route + apply"},
        "cpu_baseline": None, "gpu_launches": int(np.sum(launches)),
    }
