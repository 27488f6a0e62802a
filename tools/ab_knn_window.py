"""Interleaved A/B of the pair kernel's lock-step window (VQACX_OPT_KNN_LOCKSTEP, or with a 4th argument "sym" VQACX_OPT_KNN_SYM_LOCKSTEP) on the 123 287 x 2048 symmetric self-join."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vqa_counterexamples_b200 import _lib, synth
from vqa_counterexamples_b200.knn import NearestNeighbors

n = int(sys.argv[1]) if len(sys.argv) > 1 else 123287
windows = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [8, 0, 2, 4, 16, 32, 64]
rounds = int(sys.argv[3]) if len(sys.argv) > 3 else 4
which = _lib.OPT_KNN_SYM_LOCKSTEP if (len(sys.argv) > 4 and sys.argv[4] == "sym") else _lib.OPT_KNN_LOCKSTEP
x = synth.clustered_features_cuda(n, 2048, seed=1234)
res = {w: [] for w in windows}
for _ in range(rounds):
    for w in windows:
        with _lib.option(which, w):
            nn_ = NearestNeighbors(n_neighbors=25)
            nn_.profile = True
            nn_.fit(x)
            nn_.kneighbors(x)
            torch.cuda.synchronize()
            res[w].append(round(nn_.last_phase_ms["candidates"], 2))
for w in windows:
    print(f"window={w:3d}: candidates ms {res[w]}  median {sorted(res[w])[len(res[w]) // 2]}", flush=True)
