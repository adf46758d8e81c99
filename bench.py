#!/usr/bin/env python
"""bench.py — headline benchmark of the ECS hot path on B200 (contract in the task statement / DESIGN.md §6).

Workload (BASELINE.json configs[1]): 16 777 216 entities across 64 archetypes (4 mandatory + 6 optional 4-byte
components), Query<Position&, Velocity&, const Acceleration&, const Mass&, With<Health>, Without<Frozen>>
(16 matching tables, 4 194 304 rows), body v += a*(dt/m); p += v*dt.  One "step" = one Schedule::Run of that system
(device-side archetype-mask match + one flattened ranged launch).  Metric = matched entity-component updates / s.

  value     device-resident throughput: CUDA events around each step on the launching stream, L2 flushed between steps
  e2e       same metric through the C ABI with HOST buffers: per step the frame's Acceleration columns are copied
            host->device from pinned memory, the schedule runs, the Position columns are read back device->host
  roofline  algorithmic bytes (44 B per matched entity) / kernel time vs MEASURED_PEAKS.json hbm_gbs
  --impl reference   the CPU oracle's reference-faithful access path (B-ref) on a bounded sample
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "entity_component_updates_per_sec"
UNIT = "updates/s"
N_ENTITIES = 1 << 24
DEFAULT_ENTITIES = N_ENTITIES
# dram__bytes_read.sum + dram__bytes_write.sum of one launch of the default workload (ncu --set full, see traffic_source)
NCU_TRAFFIC_DEFAULT = 117461760 + 21580800
WORKLOAD = ("cfg2: 16777216 entities / 64 archetypes, 4-component Query with With<Health> Without<Frozen> "
            "(16 tables, 4194304 rows), accel-integrate body")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--entities", type=int, default=0, help="entities per GPU (default: the config's BASELINE size)")
    ap.add_argument("--config", default="cfg2", choices=["cfg1", "cfg2", "cfg3", "cfg4", "cfg5", "terminals", "sparse", "extract", "conditions", "async", "routing", "cfg3_sharded"],
                    help="cfg2 (default) is the headline workload; the others are secondary legs (bench_configs.py)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--all-tables", action="store_true", help="drop the With/Without filters (cfg 2-all)")
    ap.add_argument("--no-secondary", action="store_true", help="headline line only (skip the cfg2-all / cfg3 / cfg4 / cfg5 / routing legs)")
    ap.add_argument("--boids", default="bands", choices=["bands", "idrange"],
                    help="cfg5 ownership: spatial bands with a one-row halo exchange (default), or entity-id ranges with a "
                         "whole-flock exchange (the A/B: --exchange p2p = peer-memory merge, nccl = all-gather)")
    ap.add_argument("--band-barrier", default="flags", choices=["flags", "nccl"],
                    help="cfg5 bands at N > 1: neighbours wait on published step numbers on the device, or one 4-byte all-reduce per step")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="cfg5 / routing at N>1: peer-memory exchange fused into the kernels (default) or NCCL collectives")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device, enabled=True, n_gpus=1, period_ms=200):
        """One sampler per BOX: under torchrun only rank 0 starts it (`enabled`), and it watches every GPU — eight nvidia-smi
        loops on an eight-thread host, each taking the driver's lock every 100 ms, showed up as 2-8 ms stalls of single
        ranks inside ~1 ms multi-GPU steps.  Period 200 ms, as in the profiling recipe."""
        self.device, self.rows, self.proc, self.enabled, self.n_gpus, self.period_ms = device, [], None, enabled, n_gpus, period_ms

    def start(self):
        if not self.enabled:
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.enabled:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "sampled by rank 0 for every GPU of the box"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        first = self.device if self.n_gpus == 1 else 0  # the GPUs this job runs on: [first, first + n_gpus)
        self.rows = [r for r in self.rows if len(r) >= 9 and r[0].isdigit() and first <= int(r[0]) < first + self.n_gpus]
        sm = sorted(float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit())
        mx = max([float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()] or [0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 9 for i in range(4) if r[5 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": reasons, "samples": len(sm)}


def cpu_reference_leg(entities, steps, warmup, threads=1, mode=1, budget_s=None):
    """Times the oracle's reference-faithful path (B-ref: per-entity unordered_map probe + 2 binary searches + column-map
    find per component, serial ForEach — SURVEY §3.3).  With `budget_s` the number of timed steps is capped so that the
    timed region stays inside the budget (the step count actually run is returned)."""
    import ctypes
    from heliosengine_b200 import capi, worlds
    o = capi.load_oracle()
    # the world is only iterated here (no structural call inside the timed region): skip the builtin messages of the bulk load
    w = capi.World(o, flags=capi.WORLD_NO_BUILTIN_MESSAGES)
    sched, q = worlds.build_cfg2(w, entities, filtered=True)
    o.fn("set_access_mode")(w.h, mode, threads)
    matched = q.count()
    timer = o.fn("time_schedule_run", ctypes.c_double)
    est = None
    for i in range(max(1, warmup)):
        t = timer(sched.h, 1)
        est = t if i else None  # the first pass also builds the lookup index
    if budget_s is not None:
        est = est if est is not None else timer(sched.h, 1)
        steps = max(1, min(steps, int(budget_s / max(est, 1e-9))))
    t = timer(sched.h, steps)
    w.close()
    return matched * steps / t, t / steps, matched, steps


def run_reference(args):
    """The reference arm: the reference's own CPU algorithm for the path (the oracle restatement — the reference cannot be
    built in this image) on the FULL headline world, its per-entity access path, one core (the reference runs
    one system on one thread: no intra-system parallelism).  Steps are capped by a time budget, never the size."""
    import ctypes
    from heliosengine_b200 import capi
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    entities = args.entities or N_ENTITIES
    warm = max(1, min(args.warmup, 3))
    v, spt, matched, steps = cpu_reference_leg(entities, args.steps, warm, threads=1, mode=1, budget_s=120.0)
    host_threads = capi.load_oracle().fn("hardware_threads", ctypes.c_uint32)()
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": spt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "entities_per_gpu": entities, "matched_rows_per_gpu": matched,
                   "steps_note": f"{steps} timed steps of the full-size world (requested {args.steps}; capped by a 120 s budget, "
                                 "never by shrinking the world)"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
                         "sample": f"full-size world ({entities} entities, {matched} matched rows per step), {steps} steps; "
                                   "reference-faithful access path (B-ref); the reference runs one system on one core "
                                   f"(no intra-system parallelism); host has {host_threads} threads"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def run_secondary(args):
    """cfg1 / cfg3 / cfg4 legs (single GPU) and cfg5 (boids, sharded over the ranks): same JSON-line shape."""
    import torch
    import bench_configs
    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.config not in ("cfg5", "routing", "cfg3_sharded") and rank != 0:
        return
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    stream = torch.cuda.Stream(device=dev)
    clocks = ClockSampler(local_rank, enabled=rank == 0, n_gpus=world_size)
    clocks.start()
    if args.config in ("cfg5", "routing", "cfg3_sharded"):
        dist = None
        if world_size > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=dev)
        line = getattr(bench_configs, "run_" + args.config)(args, torch, dev, stream, peaks(), dist, rank, world_size)
        n_gpus = world_size
        if world_size > 1:
            dist.destroy_process_group()
    else:
        line = getattr(bench_configs, "run_" + args.config)(args, torch, dev, stream, peaks())
        n_gpus = 1
    base = {"n_gpus": n_gpus, "steps": args.steps, "warmup": max(3, args.warmup), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "data": "synthetic", "clocks": clocks.stop()}
    base.update(line)
    if rank == 0:
        print(json.dumps(base))


def secondary_legs(args, torch, dist, dev, rank, world_size, local_rank):
    """The other BASELINE configs, each to the same bar as the headline (own timing, roofline, CPU baseline where one exists,
    clocks): cfg2 with the filters dropped (L2-insensitive: 738 MB per step), cfg3 (ApplyDeferred, full-size digest parity
    against the oracle), cfg4 (8-system schedule + messages) on rank 0 at N = 1; cfg5 (boids) and cross-shard routing —
    the two workloads with an exchange step — on every rank at every N."""
    import copy
    import bench_configs
    out = {}

    def leg(name, fn, steps, warm, collective=False, **over):
        a2 = copy.copy(args)
        a2.steps, a2.warmup, a2.entities = steps, warm, 0
        for k, v in over.items():
            setattr(a2, k, v)
        clocks = ClockSampler(local_rank, enabled=rank == 0, n_gpus=world_size)
        clocks.start()
        t0 = time.perf_counter()
        try:
            st = torch.cuda.Stream(device=dev)
            r = fn(a2, torch, dev, st, peaks(), dist if world_size > 1 else None, rank, world_size) if collective else fn(a2, torch, dev, st, peaks())
            r["steps"], r["warmup"] = steps, max(3, warm)
        except Exception as e:  # a failing leg must not take the headline line with it
            r = {"error": repr(e)}
        r["clocks"] = clocks.stop()
        r["leg_wall_s"] = time.perf_counter() - t0
        out[name] = r
        torch.cuda.synchronize()
        torch.cuda.empty_cache()

    if world_size == 1:
        leg("cfg1", bench_configs.run_cfg1, 50, 5)
        leg("cfg2_all_tables", run_cfg2_all, 20, 3)
        leg("cfg3", bench_configs.run_cfg3, 3, 3)
        leg("cfg4", bench_configs.run_cfg4, 10, 3)
    # warm-up 12: after the initial sort into bands every rank's buffers (entity tables, free list, slab) reach their steady
    # capacity within the first ten steps; each growth is a one-off of 2-5 ms on ONE rank that all ranks wait for
    leg("cfg5", bench_configs.run_cfg5, 10, 12 if world_size > 1 else 3, collective=True, no_e2e=True)
    leg("routing", bench_configs.run_routing, 10, 3, collective=True)
    if world_size > 1:
        leg("cfg3_sharded", bench_configs.run_cfg3_sharded, 4, 3, collective=True)
    return out


def run_cfg2_all(args, torch, dev, stream, peak):
    """cfg 2-all: the headline query with the filters dropped — 64 tables, 16 777 216 rows, 738 MB per step: six times the L2,
    so the number does not depend on what the previous step left in the cache."""
    from heliosengine_b200 import capi, worlds
    n = args.entities or N_ENTITIES
    w = capi.World(capi.load_gpu(), device=dev.index, entity_capacity=n + 1024, flags=capi.WORLD_NO_BUILTIN_MESSAGES)
    w.set_stream(stream.cuda_stream)
    sched, q = worlds.build_cfg2(w, n, filtered=False)
    matched = q.count()
    with torch.cuda.stream(stream):
        for _ in range(max(3, args.warmup)):
            sched.run_and_wait(capi.EXEC_SERIAL)
        torch.cuda.synchronize()
        l0 = w.launch_count()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for a, b in evs:
            a.record(stream)
            sched.run(capi.EXEC_SERIAL)
            b.record(stream)
            sched.apply_deferred()
        torch.cuda.synchronize()
        launches = w.launch_count() - l0
    ms = sorted(a.elapsed_time(b) for a, b in evs)
    mean_ms = sum(ms) / len(ms)
    nbytes = matched * worlds.CFG2_BYTES_PER_ENTITY
    achieved = nbytes / (mean_ms * 1e-3) / 1e9
    w.close()
    return {"metric": METRIC, "value": matched / (mean_ms * 1e-3), "unit": UNIT, "ms_per_step": mean_ms, "dtype": "f32",
            "config": {"workload": WORKLOAD + " [filters dropped: 64 tables, all rows]", "matched_rows_per_gpu": matched,
                       "l2": "738 MB streamed per step (6x the L2): no flush needed"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak[0], "unit": "GB/s", "frac": achieved / peak[0],
                         "traffic": None, "peak_kind": peak[1], "algorithmic_bytes_per_launch": nbytes,
                         "kernel": "helios_gpu::for_each_kernel<AccelBody,...>"},
            "step_ms_min_med_max": [ms[0], ms[len(ms) // 2], ms[-1]], "gpu_launches": int(launches), "e2e": None, "cpu_baseline": None}


def run_b200(args):
    import torch
    import torch.distributed as dist
    from heliosengine_b200 import capi, worlds

    args.entities = args.entities or N_ENTITIES

    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    # Entity-range sharding, weak scaling: every rank owns an independent shard of `entities` entities (rows are
    # independent units; SURVEY §8e: no data-path collective for query iteration).
    gpu = capi.load_gpu()
    # built without the bulk load's builtin messages: the timed region iterates, it makes no structural call
    w = capi.World(gpu, device=local_rank, entity_capacity=args.entities + 1024, flags=capi.WORLD_NO_BUILTIN_MESSAGES)
    stream = torch.cuda.Stream(device=dev)
    w.set_stream(stream.cuda_stream)
    sched, q = worlds.build_cfg2(w, args.entities, filtered=not args.all_tables)
    matched_tables = q.match()
    matched = q.count()
    rows_per_table = matched // max(1, len(matched_tables))
    bytes_per_step = matched * worlds.CFG2_BYTES_PER_ENTITY

    # L2 flush between timed steps: write 256 MiB (> 126 MB L2), then READ a second 256 MiB buffer so that the lines left
    # in L2 are clean — otherwise the timed kernel is charged the write-back of the flush's own dirty lines.
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    flush_rd = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
    flush_sink = torch.zeros((), dtype=torch.int64, device=dev)

    def l2_flush():
        flush.fill_(1)
        torch.sum(flush_rd, dim=(0,), dtype=torch.int64, out=flush_sink)

    def barrier():
        torch.cuda.synchronize()
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(max(3, args.warmup)):
            sched.run(capi.EXEC_SERIAL)
            sched.apply_deferred()
        barrier()
        clocks = ClockSampler(local_rank, enabled=rank == 0, n_gpus=world_size, period_ms=50)  # the headline's timed region is short
        clocks.start()
        launches0 = w.launch_count()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        t_wall0 = time.perf_counter()
        for a, b in evs:
            l2_flush()  # outside the event pair
            a.record(stream)
            sched.run(capi.EXEC_SERIAL)
            b.record(stream)
            sched.apply_deferred()
        barrier()
        t_wall = time.perf_counter() - t_wall0
        launches = w.launch_count() - launches0
        step_ms = [a.elapsed_time(b) for a, b in evs]
        dev_ms = sum(step_ms)

        # ---- e2e: host buffers, H2D of the frame's Acceleration, run, D2H of Position -------------------------------
        A, P = worlds.cname("Acceleration"), worlds.cname("Position")
        h_acc = [torch.empty(rows_per_table * 2, dtype=torch.float32).pin_memory() for _ in matched_tables]
        h_pos = [torch.empty(rows_per_table * 2, dtype=torch.float32).pin_memory() for _ in matched_tables]
        for t in h_acc:
            t.uniform_(-0.125, 0.125)
        e2e_steps = max(3, min(args.steps, 20))

        # Frames are pipelined the way a streaming application would: frame i+1's inputs go up on a copy stream while
        # frame i's results come down on the world's stream (PCIe is full duplex, the copy engines are separate).  Every
        # frame still uploads its own inputs, runs, and returns its own results inside the timed region.
        up_stream = torch.cuda.Stream(device=dev)
        ev_run = torch.cuda.Event()   # frame i's kernels are enqueued/finished: its Acceleration may be overwritten
        ev_up = torch.cuda.Event()    # frame i+1's inputs have arrived

        def upload():
            up_stream.wait_event(ev_run)
            for tbl, src in zip(matched_tables, h_acc):
                w.column_write_on(up_stream.cuda_stream, tbl, A, 0, rows_per_table, src.data_ptr())
            ev_up.record(up_stream)

        def e2e_step():
            stream.wait_event(ev_up)
            sched.run(capi.EXEC_SERIAL)
            ev_run.record(stream)
            for tbl, dst in zip(matched_tables, h_pos):
                w.column_read(tbl, P, 0, rows_per_table, dst.data_ptr())
            upload()                # next frame's inputs, concurrently with the read-back above
            sched.apply_deferred()  # synchronises the world's stream: this frame's results are on the host

        ev_run.record(stream)
        upload()
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        barrier()
        e2e_s = time.perf_counter() - t0
        clk = clocks.stop()

    # max over ranks
    tm = torch.tensor([dev_ms, e2e_s, t_wall], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    dev_ms, e2e_s, t_wall = (float(x) for x in tm.tolist())
    total_units = matched * world_size
    value = total_units * args.steps / (dev_ms * 1e-3)
    e2e_value = total_units * e2e_steps / e2e_s
    peak, peak_kind = peaks()
    kernel_ms = dev_ms / args.steps  # the system kernel dominates the step (the match kernel is ~2 us)
    achieved = bytes_per_step / (kernel_ms * 1e-3) / 1e9

    cpu = None
    if rank == 0 and world_size == 1 and not args.no_cpu_baseline:
        import ctypes
        # B-ref on the SAME world as the GPU arm (full size; the sample is bounded in steps, not in entities: ~2 s per step)
        v, spt, m, ran = cpu_reference_leg(args.entities, 2, 1, threads=1, mode=1, budget_s=20.0)
        nthreads = capi.load_oracle().fn("hardware_threads", ctypes.c_uint32)()
        sample_entities = 1 << 21
        vopt, _, _, _ = cpu_reference_leg(sample_entities, 5, 2, threads=nthreads, mode=0)
        cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": f"full-size world ({args.entities} entities, {m} matched rows/step), {ran} steps after 1 warm-up, B-ref = oracle "
                         "with the reference's per-entity lookup path, single core (the reference has no intra-system parallelism)",
               "b_opt": {"value": vopt, "cores": int(nthreads), "note": f"direct column loops split over all host threads, 1/8-scale world ({sample_entities} entities)"}}

    w.close()
    secondary = None
    if not args.no_secondary:
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        secondary = secondary_legs(args, torch, dist, dev, rank, world_size, local_rank)

    if rank == 0:
        default_workload = args.entities == DEFAULT_ENTITIES and not args.all_tables
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD if not args.all_tables else WORKLOAD + " [filters dropped: 64 tables]",
                       "entities_per_gpu": args.entities, "matched_rows_per_gpu": matched,
                       "parallelism": f"entity-range shards x{world_size}, no data-path collective",
                       "l2": "flushed between timed steps: 256 MiB written, then 256 MiB read (L2 left holding clean lines)", "exec_mode": "serial stream",
                       "fma": "off (-fmad=false), bit-exact vs oracle"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_TRAFFIC_DEFAULT if default_workload else None,
                         # bytes that reached HBM inside the kernel (ncu) / CUDA-event kernel time: the L2-insensitive reading of
                         # the same launch (45.5 MB of its written lines are still dirty in the 126 MB L2 when the stop event fires);
                         # the other L2-insensitive figure is secondary.cfg2_all_tables (738 MB per step)
                         "frac_dram_in_kernel": (NCU_TRAFFIC_DEFAULT / (kernel_ms * 1e-3) / 1e9 / peak) if default_workload else None,
                         "traffic_source": "profiles/r01_foreach_accel_full.md (ncu --set full, one launch of this workload): "
                                           "117.46 MB read + 21.58 MB written inside the kernel; the other 45.5 MB of written "
                                           "lines are still dirty in L2 when it ends",
                         "peak_kind": peak_kind, "algorithmic_bytes_per_launch": bytes_per_step,
                         "kernel": "helios_gpu::for_each_kernel<AccelBody,...>"},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": matched * 8, "d2h_bytes_per_step": matched * 8,
                    "steps": e2e_steps, "note": "per step: pinned-host Acceleration columns -> device (copy stream, overlapping the previous read-back), Schedule::Run, Position columns -> host"},
            "gpu_launches": int(launches), "clocks": clk, "wall_s_timed_region": t_wall,
            "step_ms_min_med_max": [min(step_ms), sorted(step_ms)[len(step_ms) // 2], max(step_ms)],
            "secondary": secondary,
        }
        print(json.dumps(line))
    if world_size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.config != "cfg2":
        run_secondary(a)
    else:
        run_b200(a)
