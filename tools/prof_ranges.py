"""One profiled repetition of a hot path inside a cudaProfilerStart/Stop range (ncu --profile-from-start off):
    ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/x python tools/prof_ranges.py eval
what = eval (one NeuralCX eval batch of 4096, fp16 fast path) | train (one B = 64 train step, eager launches)
     | knn (one 123 287 x 2048 symmetric self-join) | knn_plain (the plain sweep)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from vqa_counterexamples_b200 import synth

what = sys.argv[1] if len(sys.argv) > 1 else "eval"
if what in ("eval", "train"):
    from time_ncx import build

    feats = synth.clustered_features_cuda(82783, 2048, seed=1234)
    cx, table = build(300, 2, 4096, 0.25)
    g = torch.Generator(device="cuda").manual_seed(1)
    B = 4096 if what == "eval" else 64
    b = (torch.randint(0, 82783, (B, 25), generator=g, device="cuda", dtype=torch.int32),
         table[torch.randint(0, 4096, (B,), generator=g, device="cuda")],
         torch.randint(0, 2000, (B,), generator=g, device="cuda"), torch.randint(0, 24, (B,), generator=g, device="cuda"))
    if what == "eval":
        cx.eval()
        fn = lambda: cx.eval_batch(feats, *b)
    else:
        cx.train()
        fn = lambda: cx.train_step(feats, *b, lr=1e-4)
else:
    from vqa_counterexamples_b200.knn import NearestNeighbors

    x = synth.clustered_features_cuda(123287, 2048, seed=1234)

    def fn():
        nn_ = NearestNeighbors(n_neighbors=25, symmetric=(what == "knn")).fit(x)
        nn_.kneighbors(x)

for _ in range(3):
    fn()
torch.cuda.synchronize()
torch.cuda.profiler.start()
fn()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
