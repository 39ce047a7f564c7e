"""Host-side PCIe check for the e2e scaling question: device-to-host copy bandwidth of one GPU alone and of all GPUs of the
box at once (torchrun, one rank per GPU), 8 MB pinned transfers like a step's observation slabs."""
import os, time, json
import torch, torch.distributed as dist
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1: dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 8 << 20
d = torch.empty(n, dtype=torch.uint8, device="cuda"); h = torch.empty(n, dtype=torch.uint8).pin_memory()
def run(active, secs=1.0):
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    if not active: 
        time.sleep(secs + 0.2); return 0.0
    t0 = time.perf_counter(); k = 0
    while time.perf_counter() - t0 < secs:
        for _ in range(8): h.copy_(d, non_blocking=True)
        torch.cuda.synchronize(); k += 8
    return k * n / (time.perf_counter() - t0) / 1e9
for _ in range(3): h.copy_(d, non_blocking=True)
alone = run(rank == 0)
allg = run(True)
res = torch.tensor([alone, allg], dtype=torch.float64, device="cuda")
if world > 1:
    out = [torch.zeros_like(res) for _ in range(world)]; dist.all_gather(out, res)
else: out = [res]
if rank == 0:
    print(json.dumps({"gpus": world, "d2h_alone_gbs": float(out[0][0]), "d2h_all_at_once_gbs_per_gpu": [round(float(o[1]), 1) for o in out],
                      "aggregate_gbs": float(sum(o[1] for o in out))}))
if world > 1: dist.destroy_process_group()
