"""Sharded deferred Commands (SURVEY.md §8e "Commands apply"; include/helios_gpu.h "sharded deferred Commands").

Entities are partitioned by index range over G shards; every rank holds a command queue whose targets are GLOBAL ids (or
spawn placeholders).  routing.route_commands() delivers each command to the owner of its target and applies it there.

Two statements are checked:
  * per shard, bit-exact: the shard's world equals an oracle world with the same initial state that executed — through the
    ordinary enqueue + World::Flush path — the commands it owns in (source rank, position) order (snapshot equality: tables,
    row order, entity tables, message queues);
  * across shards: the per-entity state (alive, component set, values) and the ids of the spawned entities equal those of ONE
    oracle world holding all G*E entities that executed the ranks' queues back to back (command/queue.hpp:137-163 FIFO;
    entity/manager.hpp:294-305 one next_index_ sequence).

CPU: the oracle backend on one process (G shards in a loop) and under gloo with world_size 2.  GPU: the CUDA backend with
every shard on one device, and two processes over NCCL when two devices are present."""
import os
import sys

import numpy as np
import pytest
import torch

from heliosengine_b200 import capi, routing, worlds

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHORTS = ["Position", "Velocity", "Health", "Team"]
P, V, H, T = (worlds.cname(x) for x in SHORTS)
GEN1 = np.uint64(1) << np.uint64(32)


def fill(world, first_global, count, warm=False):
    """`count` entities whose values are functions of the GLOBAL index: the first half {Position}, the rest {Position, Velocity}.
    warm: a first local batch (TryRemove<Velocity> on the last eighth) — it finds the {Position} table full, so it takes the
    general pipeline and grows the table ahead; the batch under test then meets the steady-state pipeline's preconditions."""
    half = count // 2
    g = np.arange(first_global, first_global + count, dtype=np.float32)
    pos = np.stack([g, -g], axis=1).astype(np.float32)
    vel = np.stack([g * 0.5, g * 0.25], axis=1).astype(np.float32)
    if half:
        world.spawn_batch([P], half, [pos[:half].copy()], want_entities=False)
    if count - half:
        ents = world.spawn_batch([P, V], count - half, [pos[half:].copy(), vel[half:].copy()])
        if warm:
            cmds = np.zeros(count // 8, dtype=capi.COMMAND_DTYPE)
            cmds["kind"], cmds["entity"], cmds["component_mask"] = capi.CMD_TRY_REMOVE, ents[-(count // 8):], world.mask([V])
            world.enqueue(cmds, b"")
            world.flush()


def make_world(backend, device=0, first_global=0, count=0, flags=0, warm=False):
    w = capi.World(backend, device=device, flags=flags)
    worlds.register_all(w, SHORTS)
    w.message_register(routing.ROUTED_COMMAND, 64, capi.MSG_MANUAL)
    fill(w, first_global, count, warm)
    w.update()
    w.update()  # the loader's builtin messages age out: only the commands' messages are compared
    return w


def make_queue(w, rank, n_global, n, n_spawns, seed, distinct=False):
    """Rank `rank`'s queue: n commands on uniformly drawn GLOBAL targets + n_spawns spawns (Add {P, V} on a placeholder,
    every third followed by a TryAdd {Health} on the same placeholder).  distinct: one command per entity across ALL ranks
    (rank r draws from the residue class r mod 4 of a shared permutation) — the steady-state pipeline's precondition."""
    rng = np.random.default_rng(seed + 17 * rank)
    if distinct:
        perm = np.random.default_rng(seed).permutation(n_global)
        targets = perm[rank::4][:n]
        n = targets.size
    else:
        targets = rng.integers(0, n_global, n)
    kinds = rng.choice([capi.CMD_ADD, capi.CMD_TRY_ADD, capi.CMD_TRY_REMOVE, capi.CMD_CLEAR, capi.CMD_TRY_DESTROY, capi.CMD_DESTROY],
                       size=n, p=[0.3, 0, 0.4, 0, 0.3, 0] if distinct else [0.3, 0.2, 0.2, 0.05, 0.2, 0.05])
    specs = []
    for k, t in zip(kinds, targets):
        names = [worlds.cname(x) for x in rng.choice(SHORTS, size=int(rng.integers(1, 4)), replace=False)]
        if distinct:  # only moves between the two tables that exist: {P,V} -> {P} by TryRemove, Add overwrites in place
            names = [V] if k == capi.CMD_TRY_REMOVE else [P]
        specs.append((int(k), int(t) | int(GEN1), names if k in (capi.CMD_ADD, capi.CMD_TRY_ADD, capi.CMD_TRY_REMOVE) else []))
    at = sorted(rng.integers(0, len(specs) + 1, n_spawns).tolist())
    for j in reversed(range(n_spawns)):
        ins = [(capi.CMD_ADD, int(capi.spawn_placeholder(j)), [P, V])]
        if j % 3 == 0:
            ins.append((capi.CMD_TRY_ADD, int(capi.spawn_placeholder(j)), [H]))
        specs[at[j]:at[j]] = ins
    return specs_to_queue(w, specs, rank)


def specs_to_queue(w, specs, rank=0):
    """(kind, entity, [component names]) -> command records + payload (values derived from the position and the rank)."""
    cmds = np.zeros(len(specs), dtype=capi.COMMAND_DTYPE)
    payload = bytearray()
    for i, (kind, ent, names) in enumerate(specs):
        cmds[i]["kind"], cmds[i]["entity"], cmds[i]["component_mask"] = kind, ent, w.mask(names)
        cmds[i]["payload_offset"] = len(payload)
        if kind in (capi.CMD_ADD, capi.CMD_TRY_ADD):
            for nm in sorted(names, key=lambda x: w.slots[x]):
                payload += (np.arange(w.sizes[nm] // 4, dtype=np.uint32) + np.uint32(1000 * i + 7 * rank + w.slots[nm])).tobytes()
    return cmds, bytes(payload)


def count_spawns(cmds):
    ph = cmds["entity"][(cmds["entity"] >> np.uint64(32)) == capi.SPAWN_PLACEHOLDER_GENERATION]
    return int((ph & np.uint64(0xFFFFFFFF)).max()) + 1 if ph.size else 0


def records_to_commands(rec):
    cmds = np.zeros(rec.size, dtype=capi.COMMAND_DTYPE)
    cmds["kind"], cmds["entity"], cmds["component_mask"] = rec["kind"], rec["entity"], rec["component_mask"]
    cmds["payload_offset"] = np.arange(rec.size, dtype=np.uint32) * 40
    return cmds, np.ascontiguousarray(rec["payload"]).reshape(-1)


def entity_states(snap, rebase):
    """{global index: (generation, ((type hash, value bytes), ...))} of every entity stored in a table."""
    out = {}
    for a in snap["archetypes"]:
        for r, e in enumerate(a["entities"].tolist()):
            vals = []
            for t in a["types"]:
                col = a["columns"].get(t)
                size = col.size // max(1, a["entities"].size) if col is not None else 0
                vals.append((t, bytes(col[r * size:(r + 1) * size]) if size else b""))
            out[(e & 0xFFFFFFFF) + rebase] = (e >> 32, tuple(vals))
    return out


def expected(oracle, queues, per, g, last_count, flags=0, warm=False):
    """(per-shard replay snapshots, single-world entity states): the two statements of the module docstring."""
    n_global = per * (g - 1) + last_count
    # ids: one sequence over (rank, emission)
    spawns = [count_spawns(q[0]) for q in queues]
    ws = [make_world(oracle, 0, r * per, per if r < g - 1 else last_count, flags, warm) for r in range(g)]
    recs = [ws[r].commands_route_pack(queues[r][0], queues[r][1], n_global + sum(spawns[:r])) for r in range(g)]
    shard_snaps = []
    for r, rec in enumerate(routing.reference_command_routing(recs, per)):
        if r == g - 1 and sum(spawns):
            ws[r].reserve_entities(sum(spawns))
        c, pay = records_to_commands(rec)
        ws[r].enqueue(c, pay)
        ws[r].flush()
        shard_snaps.append(ws[r].snapshot())
        ws[r].close()
    single = make_world(oracle, 0, 0, 0, flags)
    for r in range(g):  # the same index -> value assignment as the shards: shard after shard
        fill(single, r * per, per if r < g - 1 else last_count, warm)
    if sum(spawns):
        single.reserve_entities(sum(spawns))
    for rec in recs:
        c, pay = records_to_commands(rec)
        single.enqueue(c, pay)
    single.flush()
    states = entity_states(single.snapshot(), 0)
    single.close()
    return shard_snaps, states


def check(shard_snaps, want_snaps, want_states, per):
    merged = {}
    for r, (got, want) in enumerate(zip(shard_snaps, want_snaps)):
        assert not worlds.snapshots_equal(got, want), r
        merged.update(entity_states(got, r * per))
    assert merged == want_states


def run_in_process(backend, oracle, g, per, last_count, n, n_spawns, seed, device="cpu", flags=0, distinct=False, warm=False,
                   specs=None):
    """All G shards in this process (`dist` = None): split on every shard, hand the slabs over, merge + apply on every shard.
    specs: per-rank hand-written command lists instead of the random queues."""
    n_global = per * (g - 1) + last_count
    ws = [make_world(backend, 0, r * per, per if r < g - 1 else last_count, flags, warm) for r in range(g)]
    if specs is not None:
        queues = [specs_to_queue(ws[r], specs[r], r) for r in range(g)]
    else:
        queues = [make_queue(ws[r], r, n_global, n + 13 * r, n_spawns + r if n_spawns else 0, seed, distinct) for r in range(g)]
    spawns = [count_spawns(q[0]) for q in queues]
    first = n_global
    slabs, counts = [], []
    for r, w in enumerate(ws):
        rec = w.commands_route_pack(queues[r][0], queues[r][1], first + sum(spawns[:r]))
        if rec.size:
            w.message_write(routing.ROUTED_COMMAND, rec)
        slab = torch.zeros(max(1, rec.size) * 64, dtype=torch.uint8, device=device)
        counts.append(routing.split(w, routing.ROUTED_COMMAND, 1, routing.ROUTED_KEY_OFFSET, per, g, slab))
        slabs.append(slab)
    if sum(spawns):
        assert ws[-1].next_entity_index() == last_count
        ws[-1].reserve_entities(sum(spawns))
    snaps = []
    for d, w in enumerate(ws):
        parts = []
        for s in range(g):
            lo = sum(counts[s][:d]) * 64
            parts.append(slabs[s][lo:lo + counts[s][d] * 64])
        recv = torch.cat(parts) if parts else torch.zeros(0, dtype=torch.uint8, device=device)
        routing.merge(w, routing.ROUTED_COMMAND, 1, recv, recv.numel() // 64, routing.ROUTED_KEY_OFFSET, d * per)
        w.commands_apply_routed(routing.ROUTED_COMMAND, 1)
        assert w.message_counts(routing.ROUTED_COMMAND) == (0, 0)
        snaps.append(w.snapshot())
    stats = [w.apply_stats() for w in ws]
    for w in ws:
        w.close()
    want_snaps, want_states = expected(oracle, queues, per, g, last_count, flags, warm)
    check(snaps, want_snaps, want_states, per)
    return stats


@pytest.mark.parametrize("g,per,last,n,spawns", [(1, 64, 64, 40, 3), (2, 128, 100, 300, 5), (3, 500, 500, 2000, 0), (4, 256, 7, 900, 11)])
def test_oracle_shards_in_one_process(oracle, g, per, last, n, spawns):
    run_in_process(oracle, oracle, g, per, last, n, spawns, seed=g * 1000 + n)


def test_oracle_shards_distinct_targets_after_a_local_batch(oracle):
    run_in_process(oracle, oracle, 4, 1 << 12, 1 << 12, 900, 0, seed=5, distinct=True, warm=True)


def E(i):
    return int(i) | int(GEN1)


def PH(j):
    return int(capi.spawn_placeholder(j))


EDGE_CASES = {
    # nobody has anything to say: an empty exchange and an empty apply on every shard
    "empty_queues": (3, 64, 64, [[], [], []]),
    # one rank speaks, everything goes to shard 0
    "all_to_first_shard": (3, 64, 64, [[], [(capi.CMD_ADD, E(5), [H]), (capi.CMD_TRY_REMOVE, E(40), [V]), (capi.CMD_DESTROY, E(63), [])], []]),
    # spawn-only batches; a placeholder used again by its own rank (TryAdd, then Destroy of the just-spawned entity), and
    # the same global entity named by two ranks (rank order decides: rank 0's Destroy, then rank 2's TryAdd is invalid)
    "spawns_and_reuse": (3, 64, 10, [[(capi.CMD_ADD, PH(0), [P]), (capi.CMD_TRY_ADD, PH(0), [H]), (capi.CMD_ADD, PH(1), [P, V]),
                                      (capi.CMD_DESTROY, PH(1), []), (capi.CMD_DESTROY, E(70), [])],
                                     [(capi.CMD_ADD, PH(0), [P, V])],
                                     [(capi.CMD_TRY_ADD, E(70), [T]), (capi.CMD_ADD, PH(0), [T])]]),
    # targets past every range land on the open-ended last shard, where they are dead ids: skipped and counted, as on one world
    "dead_targets_beyond_the_ranges": (2, 32, 32, [[(capi.CMD_TRY_DESTROY, E(5000), []), (capi.CMD_TRY_ADD, E(64), [H]), (capi.CMD_ADD, E(3), [H])],
                                                   [(capi.CMD_TRY_REMOVE, E(99999), [P]), (capi.CMD_CLEAR, E(33), [])]]),
    # every command of every rank names ONE entity: the (source rank, position) order is all that decides the outcome
    "one_hot_entity": (4, 16, 16, [[(capi.CMD_ADD, E(20), [H])], [(capi.CMD_TRY_REMOVE, E(20), [H, P])], [(capi.CMD_CLEAR, E(20), [])],
                                   [(capi.CMD_TRY_ADD, E(20), [V, T]), (capi.CMD_DESTROY, E(20), []), (capi.CMD_TRY_ADD, E(20), [P])]]),
}


@pytest.mark.parametrize("case", sorted(EDGE_CASES))
def test_oracle_edge_queues(oracle, case):
    g, per, last, specs = EDGE_CASES[case]
    run_in_process(oracle, oracle, g, per, last, 0, 0, seed=0, specs=specs)


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(EDGE_CASES))
def test_gpu_edge_queues(gpu, oracle, case):
    g, per, last, specs = EDGE_CASES[case]
    run_in_process(gpu, oracle, g, per, last, 0, 0, seed=0, device="cuda:0", specs=specs)


def test_pack_limits(oracle):
    w = make_world(oracle, 0, 0, 8)
    worlds.register_all(w, ["Acceleration", "Mass", "Damping", "Frozen", "Shield"])
    names = [worlds.cname(x) for x in ["Position", "Velocity", "Acceleration", "Mass", "Health", "Team", "Damping", "Frozen"]]  # 44 B
    cmds = np.zeros(1, dtype=capi.COMMAND_DTYPE)
    cmds[0]["kind"], cmds[0]["entity"], cmds[0]["component_mask"] = capi.CMD_ADD, int(GEN1), w.mask(names)
    with pytest.raises(capi.HeliosError) as e:
        w.commands_route_pack(cmds, bytes(44), 0)
    assert e.value.status == capi.ERR_CAPACITY
    cmds[0]["component_mask"] = w.mask(names[:7])  # 40 bytes: fits
    rec = w.commands_route_pack(cmds, bytes(range(40)), 0)
    assert rec[0]["payload_bytes"] == 40 and bytes(rec[0]["payload"]) == bytes(range(40))
    with pytest.raises(capi.HeliosError):
        w.commands_route_pack(cmds, bytes(39), 0)  # payload past the end of the buffer
    w.close()


def _worker(rank, world_size, port, kind, per, last, n, n_spawns, seed, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import pickle
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    if kind == "nccl":
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=world_size, device_id=dev)
        backend = capi.load_gpu()
    else:
        dev = "cpu"
        dist.init_process_group("gloo", rank=rank, world_size=world_size)
        backend = capi.load_oracle()
    n_global = per * (world_size - 1) + last
    w = make_world(backend, rank if kind == "nccl" else 0, rank * per, per if rank < world_size - 1 else last)
    for step in range(2):  # the second step targets entities the first one spawned
        cmds, pay = make_queue(w, rank, n_global + (step * (2 * n_spawns + 1)), n + 13 * rank, n_spawns + rank, seed + step)
        info = routing.route_commands(w, cmds, pay, n_spawns + rank, torch, dev, dist, rank, world_size, per)
        assert info["total_spawns"] == 2 * n_spawns + 1
    with open(os.path.join(out_dir, f"s{rank}.pkl"), "wb") as f:
        pickle.dump(w.snapshot(), f)
    w.close()
    dist.barrier()
    dist.destroy_process_group()


def _two_rank_expected(oracle, per, last, n, n_spawns, seed):
    """Replay of _worker's two steps on per-shard oracle worlds through the ordinary enqueue + flush path."""
    ws = [make_world(oracle, 0, r * per, per if r == 0 else last) for r in range(2)]
    n_global = per + last
    for step in range(2):
        qs = [make_queue(ws[r], r, n_global + (step * (2 * n_spawns + 1)), n + 13 * r, n_spawns + r, seed + step) for r in range(2)]
        first = per + ws[1].next_entity_index()
        recs = [ws[r].commands_route_pack(qs[r][0], qs[r][1], first + (n_spawns if r else 0)) for r in range(2)]
        ws[1].reserve_entities(2 * n_spawns + 1)
        for r, rec in enumerate(routing.reference_command_routing(recs, per)):
            c, pay = records_to_commands(rec)
            ws[r].enqueue(c, pay)
            ws[r].flush()
    snaps = [w.snapshot() for w in ws]
    for w in ws:
        w.close()
    return snaps


def test_two_ranks_over_gloo(oracle, tmp_path):
    """The N>1 host path of route_commands (all_gather of spawn counts -> pack -> route -> apply) under world_size 2 on CPU."""
    import pickle
    import torch.multiprocessing as mp
    per, last, n, n_spawns, seed = 512, 300, 1500, 4, 77
    port = 33500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, "gloo", per, last, n, n_spawns, seed, str(tmp_path)), nprocs=2, join=True)
    want = _two_rank_expected(oracle, per, last, n, n_spawns, seed)
    for r in range(2):
        with open(tmp_path / f"s{r}.pkl", "rb") as f:
            assert not worlds.snapshots_equal(pickle.load(f), want[r]), r


@pytest.mark.gpu
@pytest.mark.parametrize("g,per,last,n,spawns", [(1, 64, 64, 40, 3), (2, 4096, 1000, 9000, 5), (4, 1 << 14, 1 << 14, 70001, 0),
                                                 (8, 1 << 13, 77, 30000, 9)])
def test_gpu_shards_on_one_device(gpu, oracle, g, per, last, n, spawns):
    run_in_process(gpu, oracle, g, per, last, n, spawns, seed=g * 1000 + n, device="cuda:0")


@pytest.mark.gpu
def test_gpu_shards_steady_state_pipeline(gpu, oracle):
    """One command per entity over the whole box (cfg3's shape): every shard's batch takes the steady-state pipeline."""
    stats = run_in_process(gpu, oracle, 4, 1 << 15, 1 << 15, 9000, 0, seed=5, device="cuda:0",
                           flags=capi.WORLD_NO_BUILTIN_MESSAGES, distinct=True, warm=True)
    assert all(s.pipeline == 1 for s in stats), [s.pipeline for s in stats]


@pytest.mark.gpu
def test_two_gpus_over_nccl(oracle, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import pickle
    import torch.multiprocessing as mp
    per, last, n, n_spawns, seed = 1 << 16, 50000, 120000, 6, 91
    port = 36500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, "nccl", per, last, n, n_spawns, seed, str(tmp_path)), nprocs=2, join=True)
    want = _two_rank_expected(oracle, per, last, n, n_spawns, seed)
    for r in range(2):
        with open(tmp_path / f"s{r}.pkl", "rb") as f:
            assert not worlds.snapshots_equal(pickle.load(f), want[r]), r
