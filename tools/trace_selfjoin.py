"""Shell timeline of the pipelined host self-join: python tools/trace_selfjoin.py [chunks...]"""
import os, sys, torch
sys.path.insert(0, ".")
from vqa_counterexamples_b200 import synth
from vqa_counterexamples_b200.knn import NearestNeighbors
n = 123287
x = synth.clustered_features_cuda(n, 2048, seed=1234)
h = torch.empty((n, 2048), dtype=torch.float32, pin_memory=True)
h.copy_(x)
torch.cuda.synchronize()
for C in [int(a) for a in sys.argv[1:]] or [1, 4]:
    for _ in range(2):
        NearestNeighbors(n_neighbors=25).fit_kneighbors(h.numpy(), n_chunks=C)
    os.environ["VQACX_KNN_WAVE_TRACE"] = "1"
    print(f"--- n_chunks={C}", file=sys.stderr)
    NearestNeighbors(n_neighbors=25).fit_kneighbors(h.numpy(), n_chunks=C)
    os.environ["VQACX_KNN_WAVE_TRACE"] = "0"
