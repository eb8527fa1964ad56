"""Time NCCL all_gather_into_tensor of one FP32 block per rank (the PageRank exchange) under the current NCCL_* env."""
import os, sys, torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
src = torch.ones(nb, device="cuda"); dst = torch.empty(nb * world, device="cuda")
for _ in range(5): dist.all_gather_into_tensor(dst, src)
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): dist.all_gather_into_tensor(dst, src)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / 50
if rank == 0:
    tag = " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("NCCL_"))
    print(f"world={world} block={nb * 4 / 1e6:.1f}MB  {us:8.1f} us  busbw={(world - 1) * nb * 4 / us / 1e3:7.1f} GB/s  [{tag}]", flush=True)
dist.destroy_process_group()
