"""Section timing of the row-sharded kNN protocol (run under torchrun): where a step's time goes beyond the kernels."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vqa_counterexamples_b200 import knn as K, synth  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
n, k = 123287, 25
x = synth.clustered_features_cuda(n, 2048, seed=1234, device=dev)
r0, r1 = K.shard_rows(n, rank, world)


def tick(acc, name, t0):
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    acc[name] = acc.get(name, 0.0) + (t1 - t0) * 1e3
    return t1


for it in range(6):
    acc = {}
    dist.barrier()
    torch.cuda.synchronize()
    t = time.perf_counter()
    shard = K._CudaShard(x[r0:r1], r0, x, k)
    t = tick(acc, "fit+prep", t)
    hi = shard.phase1()
    t = tick(acc, "phase1", t)
    hi_all = K._all_gather(hi, world, None)
    t = tick(acc, "gather_hi", t)
    bound = K.kth_bound(hi_all)
    t = tick(acc, "kth", t)
    d, i = shard.phase2(bound)
    t = tick(acc, "phase2", t)
    d_all, i_all = K._all_gather(d, world, None), K._all_gather(i, world, None)
    t = tick(acc, "gather_out", t)
    md, mi = K.merge_topk(d_all, i_all)
    t = tick(acc, "merge", t)
    if it >= 3:
        print(f"rank {rank} it {it}: " + "  ".join(f"{a} {b:.2f}" for a, b in acc.items()) + f"  total {sum(acc.values()):.2f}", flush=True)
dist.destroy_process_group()
