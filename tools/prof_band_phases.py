"""Per-phase times of the band-mode boids step (tools, not product): run under torchrun with N ranks on N GPUs; every phase is
bracketed by device synchronisations, so the sum overstates the pipelined step.  Used for profiles/r02_multi_gpu_scaling.md."""
import os, sys, time, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from heliosengine_b200 import boids, capi, routing
rank, G = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
n_total = 1 << 26; ws = 4096
shard = boids.BandShard(capi.load_gpu(), n_total, ws, G, rank, device=rank)
stream = torch.cuda.Stream(device=dev); shard.world.set_stream(stream.cuda_stream)
rows = []
def T(row, name, fn):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = fn(); torch.cuda.synchronize(); row[name] = round((time.perf_counter() - t0) * 1e3, 3); return r
with torch.cuda.stream(stream):
    x = boids.BandExchange(shard, torch, dist, dev, peer=True)
    x.step()  # the initial sort into bands
    NSTEPS = int(os.environ.get('PROF_STEPS', '12'))
    for _ in range(NSTEPS):
        row = {}
        T(row, "leavers", shard.leavers)
        T(row, "pack", lambda: shard.pack(x.mig_send.data_ptr(), x.cap))
        T(row, "gather", lambda: dist.all_gather_into_tensor(x.mig_recv, x.mig_send))
        ok = T(row, "unpack", lambda: shard.unpack(x.mig_recv.data_ptr(), x.cap))
        assert ok
        T(row, "arrivals", shard.arrivals)
        st = shard.world.apply_stats(); row["pipe"] = int(st.pipeline); row["fb"] = int(st.fast_fallbacks); row["cmds"] = int(st.commands)
        T(row, "bin", shard.bin)
        T(row, "barrier", lambda: dist.all_reduce(x.token))
        T(row, "step", shard.step)
        rows.append(row)
    allrows = [None] * G
    dist.all_gather_object(allrows, rows)
    if rank == 0:
        keys = ["leavers", "pack", "gather", "unpack", "arrivals", "bin", "barrier", "step"]
        med = {k: float(np.median([max(r[i][k] for r in allrows) for i in range(NSTEPS)])) for k in keys}
        print("median of max-over-ranks per phase:", med)
        for i in range(NSTEPS):
            worst = {k: max(r[i][k] for r in allrows) for k in keys}
            if not any(worst[k] > 1.5 * med[k] + 0.05 for k in keys):
                continue
            slow = {k: (round(worst[k], 2), int(np.argmax([r[i][k] for r in allrows]))) for k in keys if worst[k] > 1.5 * med[k] + 0.05}
            print("step", i + 1, "slow phases (ms, rank):", slow)
        for i in range(0):
            print(i, {k: max(r[i][k] for r in allrows) for k in keys}, "pipe", [r[i]["pipe"] for r in allrows], "fb", [r[i]["fb"] for r in allrows], "slow arrivals rank", int(np.argmax([r[i]["arrivals"] for r in allrows])))
    x.close()
shard.close(); dist.destroy_process_group()
