"""A/B: symmetric self-join (123 287 x 2048, k = 25) with different group counts / mirror capacities / lock-step windows."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vqa_counterexamples_b200 import _lib, synth
from vqa_counterexamples_b200.knn import NearestNeighbors

n = int(sys.argv[1]) if len(sys.argv) > 1 else 123287
x = synth.clustered_features_cuda(n, 2048, seed=1234)


def run(reps=4):
    best, last = 1e9, None
    for _ in range(reps):
        nn_ = NearestNeighbors(n_neighbors=25)
        nn_.profile = True
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        nn_.fit(x)
        d, i = nn_.kneighbors(x)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
        last = nn_
    return best, last.last_phase_ms, last.last_stats


for groups, cap in [(int(g), 512 if int(g) <= 7 else 2048) for g in (sys.argv[2].split(",") if len(sys.argv) > 2 else "7,9,11,13,15,7".split(","))]:
    with _lib.option(_lib.OPT_KNN_SYM_GROUPS, groups), _lib.option(_lib.OPT_KNN_SYM_MIRROR_CAP, cap):
        ms, ph, st = run()
        print(f"groups={groups} mirror_cap={cap}: {ms:.2f} ms  phases {({k: round(v, 2) for k, v in ph.items()})}  "
              f"flagged {st['flagged_rows']} mirror_keys_avg {st.get('mirror_keys_avg')}")
