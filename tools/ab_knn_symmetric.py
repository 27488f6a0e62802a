"""A/B of the symmetric self-join sweep against the plain sweep (scratch tool, not the bench):
    python tools/ab_knn_symmetric.py N D [reps]
Checks that both give identical results, prints step times, phase times and the mirror-buffer statistics."""
import sys

import torch

sys.path.insert(0, ".")
from vqa_counterexamples_b200 import synth
from vqa_counterexamples_b200.knn import NearestNeighbors

n = int(sys.argv[1]) if len(sys.argv) > 1 else 40000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
x = synth.clustered_features_cuda(n, d, seed=1234)
res = {}
for sym in (False, True):
    nn_ = NearestNeighbors(n_neighbors=25, symmetric=sym).fit(x)
    nn_.profile = True
    for r in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dist, idx = nn_.kneighbors(x)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"symmetric={sym} n={n} d={d} rep={r} {ms:.2f} ms {n / ms * 1e3:.0f} q/s phases={getattr(nn_, 'last_phase_ms', None)} "
              f"stats={nn_.last_stats}", flush=True)
    res[sym] = (dist.clone(), idx.clone())
same_i = bool((res[False][1] == res[True][1]).all())
same_d = bool((res[False][0] == res[True][0]).all())
print("identical indices:", same_i, " identical distances:", same_d)
if not same_i:
    bad = (res[False][1] != res[True][1]).any(dim=1).nonzero().flatten()
    print("rows that differ:", bad.numel(), bad[:10].tolist())
    r0 = int(bad[0])
    print(res[False][1][r0].tolist())
    print(res[True][1][r0].tolist())
    print(res[False][0][r0].tolist())
    print(res[True][0][r0].tolist())
sys.exit(0 if (same_i and same_d) else 1)
