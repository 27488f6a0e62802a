"""A/B: NeuralCX train step and eval batches with / without the per-image tables (hv table, P1 table)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from time_ncx import build
from vqa_counterexamples_b200 import synth

feats = synth.clustered_features_cuda(82783, 2048, seed=1234)
cx, table = build(300, 2, 4096, 0.25)
g = torch.Generator(device="cuda").manual_seed(1)


def batch(B):
    return (torch.randint(0, 82783, (B, 25), generator=g, device="cuda", dtype=torch.int32),
            table[torch.randint(0, 4096, (B,), generator=g, device="cuda")],
            torch.randint(0, 2000, (B,), generator=g, device="cuda"), torch.randint(0, 24, (B,), generator=g, device="cuda"))


def timed(fn, n=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


b64, b4096 = batch(64), batch(4096)
for tables in (True, False, True):
    cx.use_image_tables = tables
    cx.train()
    print(f"tables={tables}: train graph {timed(lambda: cx.train_step_graph(feats, *b64, lr=1e-4)):.3f} ms, "
          f"eager {timed(lambda: cx.train_step(feats, *b64, lr=1e-4)):.3f} ms")
cx.use_image_tables = True
cx.eval()
for p1 in (True, False, True):
    cx.use_p1_table = p1
    print(f"p1={p1}: eval B=64 {timed(lambda: cx.eval_batch(feats, *b64)):.3f} ms, B=4096 {timed(lambda: cx.eval_batch(feats, *b4096), 10):.3f} ms")
cx.eval_precision = "tf32"
print(f"tf32 general path with hv table: eval B=4096 {timed(lambda: cx.eval_batch(feats, *b4096), 10):.3f} ms")
cx.use_image_tables = False
print(f"tf32 general path, no tables:    eval B=4096 {timed(lambda: cx.eval_batch(feats, *b4096), 10):.3f} ms")
