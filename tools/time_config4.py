"""BASELINE configs[3] on ONE GPU: 1 000 000 x 2048 database, 131 072 queries, k = 24 (the sharded runs split the
database rows over G GPUs): python tools/time_config4.py [n_db] [n_q] [reps]"""
import sys

import torch

sys.path.insert(0, ".")
from vqa_counterexamples_b200 import synth
from vqa_counterexamples_b200.knn import NearestNeighbors

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 131072
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
x = synth.clustered_features_cuda(n, 2048, seed=77, n_centres=max(1, n // 60))
q = synth.clustered_features_cuda(nq, 2048, seed=78, n_centres=max(1, n // 60))
nn_ = NearestNeighbors(n_neighbors=24).fit(x)
nn_.profile = True
for r in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    d, i = nn_.kneighbors(q)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    ph = nn_.last_phase_ms
    print(f"n_db={n} n_q={nq} rep={r} {ms:.1f} ms  {nq / ms * 1e3:.0f} q/s | gemm {ph['candidates']:.1f} ms = "
          f"{2.0 * n * nq * 2048 / ph['candidates'] / 1e9:.0f} TFLOP/s  finalize {ph['finalize']:.2f}  rerank {ph['rerank']:.2f}  "
          f"stats={nn_.last_stats}")
