"""Interleaved A/B of the kNN wave overlap at bench size: python tools/ab_knn_overlap.py [N] [rounds]"""
import os
import sys
import torch
sys.path.insert(0, ".")
from vqa_counterexamples_b200 import synth
from vqa_counterexamples_b200.knn import NearestNeighbors

n = int(sys.argv[1]) if len(sys.argv) > 1 else 123287
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 6
x = synth.clustered_features_cuda(n, 2048, seed=1234)
nn_ = NearestNeighbors(n_neighbors=25).fit(x)
for _ in range(3):
    nn_.kneighbors(x)
res = {"0": [], "1": []}
for r in range(rounds):
    for mode in ("0", "1"):
        os.environ["VQACX_KNN_OVERLAP"] = mode
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            nn_.kneighbors(x)
        e1.record()
        torch.cuda.synchronize()
        res[mode].append(e0.elapsed_time(e1) / 3)
for mode, name in (("1", "overlap"), ("0", "serial ")):
    v = res[mode]
    print(name, " ".join(f"{t:.2f}" for t in v), "| median", sorted(v)[len(v) // 2])
