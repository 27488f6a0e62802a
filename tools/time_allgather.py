"""NCCL all_gather timing at the sizes the sharded kNN exchanges (run under torchrun):
torchrun --nproc-per-node N tools/time_allgather.py"""
import os
import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
for name, numel, dtype in (("hi f32 [123287,25]", 123287 * 25, torch.float32), ("dist f64 [123287,25]", 123287 * 25, torch.float64),
                           ("idx i64 [123287,25]", 123287 * 25, torch.int64)):
    t = torch.ones(numel, dtype=dtype, device=dev)
    out = torch.empty(world * numel, dtype=dtype, device=dev)
    for _ in range(3):
        dist.all_gather_into_tensor(out, t)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        dist.all_gather_into_tensor(out, t)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    if rank == 0:
        gb = out.numel() * out.element_size() / 1e9
        print(f"{name}: {ms:.3f} ms per all_gather ({gb / ms * 1e3:.1f} GB/s received per rank)")
dist.destroy_process_group()
