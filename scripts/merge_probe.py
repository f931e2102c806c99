"""Multi-GPU probe: per-rank tile-kernel time and the NCCL merge time, separately (torchrun)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np, torch, torch.distributed as dist
import mpb200 as mp
from mpb200.dist import merge_plan_across_ranks
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n, m = 1 << 20, 256
T = np.random.default_rng(1234).random(n)
ctx = mp.Context(local)
stream = torch.cuda.ExternalStream(ctx.stream, device=local)
plan = mp.Plan(ctx, n, m, mp.F64, 0); plan.set_shard(rank, world); plan.upload_series(T)
sub = n - m + 1
corr = torch.empty(sub, dtype=torch.float64, device="cuda"); idx = torch.empty(sub, dtype=torch.int32, device="cuda")
with torch.cuda.stream(stream):
    dist.all_reduce(idx)
ctx.finish()
for rep in range(4):
    dist.barrier(); torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        e[0].record(stream); plan.precompute(); plan.compute(); e[1].record(stream)
        merge_plan_across_ranks(plan, corr, idx); e[2].record(stream)
        plan.finalize(); e[3].record(stream)
    torch.cuda.synchronize()
    t = plan.timings()
    print(json.dumps({"rank": rank, "rep": rep, "compute_ms": round(t["compute_ms"], 3), "pre+compute": round(e[0].elapsed_time(e[1]), 3),
                      "merge_ms": round(e[1].elapsed_time(e[2]), 3), "fin_ms": round(e[2].elapsed_time(e[3]), 3), "items": t["items"], "wins": t["slow_updates"], "hits": t["filter_hits"], "replays": t["replays"]}), flush=True)
dist.barrier(); dist.destroy_process_group()
