"""Scratch timing of the kNN self-join at a given size (not the bench): python tools/time_knn.py N [reps]"""
import sys
import time

import torch

sys.path.insert(0, ".")
from vqa_counterexamples_b200 import synth
from vqa_counterexamples_b200.knn import NearestNeighbors

n = int(sys.argv[1]) if len(sys.argv) > 1 else 30000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
exact = len(sys.argv) > 3 and sys.argv[3] == "exact"
x = synth.clustered_features_cuda(n, 2048, seed=1234)
torch.cuda.synchronize()
t0 = time.perf_counter()
nn_ = NearestNeighbors(n_neighbors=25, force_exact=exact).fit(x)
torch.cuda.synchronize()
print(f"fit(prep) {1e3 * (time.perf_counter() - t0):.2f} ms")
for r in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    d, i = nn_.kneighbors(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"n={n} rep={r} {ms:.2f} ms  {n / ms * 1e3:.0f} q/s  {2.0 * n * n * 2048 / ms / 1e9:.1f} TFLOP/s  stats={nn_.last_stats}")
