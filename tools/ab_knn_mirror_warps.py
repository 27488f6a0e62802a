"""Interleaved A/B of the number of column-side (mirror) warps of the mirrored launch (VQACX_OPT_KNN_MIRROR_WARPS = 4 | 8) on
the 123 287 x 2048 symmetric self-join: whole-search time, so that what the power cap gives back in one phase and takes
in another is counted."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vqa_counterexamples_b200 import _lib, synth
from vqa_counterexamples_b200.knn import NearestNeighbors

n = int(sys.argv[1]) if len(sys.argv) > 1 else 123287
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 10
x = synth.clustered_features_cuda(n, 2048, seed=1234)
res = {4: [], 8: []}
cand = {4: [], 8: []}
for _ in range(rounds):
    for w in (4, 8):
        # (4-warp mode fills half as many append logs twice as full: give them room)
        with _lib.option(_lib.OPT_KNN_MIRROR_WARPS, w), _lib.option(_lib.OPT_KNN_SYM_LOG_CAP, 120000 if w == 4 else 0):
            nn_ = NearestNeighbors(n_neighbors=25)
            nn_.profile = True
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            nn_.fit(x)
            nn_.kneighbors(x)
            e1.record()
            torch.cuda.synchronize()
            res[w].append(round(e0.elapsed_time(e1), 2))
            cand[w].append(round(nn_.last_phase_ms["candidates"], 2))
for w in (4, 8):
    print(f"mirror warps {w}: total ms {res[w]} median {sorted(res[w])[len(res[w]) // 2]} | candidates median "
          f"{sorted(cand[w])[len(cand[w]) // 2]}", flush=True)
