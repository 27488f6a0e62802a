"""Times the CUDA question encoder (BayesianUniSkip, 620 -> 2400, 26 steps): python tools/time_gru.py [B]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vqa_counterexamples_b200.vqa_models import BayesianUniSkip  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
torch.manual_seed(0)
enc = BayesianUniSkip(5000)
g = np.random.default_rng(0)
wids = np.zeros((B, 26), dtype=np.int64)
for b in range(B):
    n = int(np.clip(g.poisson(6) + 2, 3, 26))
    wids[b, :n] = g.integers(1, 5001, n)
w = torch.from_numpy(wids).cuda()
for _ in range(2):
    enc(w)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    enc(w)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"B={B}: {ms:.3f} ms per batch, {B / ms * 1e3:.0f} questions/s (26 steps, TF32)")
