"""Device-to-host ceiling of the box: N processes (torchrun), one GPU each, plain 8.3 MB cudaMemcpyAsync copies from device
memory into pinned host memory (the size of one 1080p RGBA8 frame), all ranks at once.  Prints per-rank and aggregate
GB/s: what a frame-gather over PCIe can reach at most at N GPUs, whatever the renderer does.
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/d2h_ceiling.py [MB]"""
import json, os, sys, time
import torch
import torch.distributed as dist
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nbytes = int(float(sys.argv[1]) * 1e6) if len(sys.argv) > 1 else 1920 * 1080 * 4
src = torch.empty(nbytes, dtype=torch.uint8, device="cuda").random_(0, 255)
dst = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(8)]
def run(n):
    for k in range(n):
        dst[k % 8].copy_(src, non_blocking=True)
    torch.cuda.synchronize()
run(16)
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
n = 400
t0 = time.perf_counter(); run(n); dt = time.perf_counter() - t0
gbs = n * nbytes / dt / 1e9
t = torch.tensor([gbs], dtype=torch.float64, device="cuda")
allg = [torch.zeros_like(t) for _ in range(world)]
if world > 1:
    dist.all_gather(allg, t)
else:
    allg = [t]
if rank == 0:
    per = [float(x[0]) for x in allg]
    print(json.dumps({"tool": "d2h_ceiling", "n_gpus": world, "bytes_per_copy": nbytes, "copies": n, "per_rank_gbs": [round(p, 2) for p in per],
                      "aggregate_gbs": round(sum(per), 2), "frames_per_s_ceiling_1080p": round(sum(per) * 1e9 / (1920 * 1080 * 4), 1),
                      "cpus": len(os.sched_getaffinity(0))}), flush=True)
if world > 1:
    dist.destroy_process_group()
