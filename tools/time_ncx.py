"""Times the fused NeuralCX train step / eval batch at BASELINE config 3 sizes (MUTAN no-att dims, H=300, L=2).
Usage: python tools/time_ncx.py [--batch 64] [--steps 50] [--eval-batch 4096] [--dim-h 300] [--layers 2]"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def build(dim_h, layers, n_q, drop_p, seed=0):
    from vqa_counterexamples_b200.cx import NeuralModel
    from vqa_counterexamples_b200.vqa_models import MutanNoAtt, QuestionTable

    torch.manual_seed(seed)
    opt = {"arch": "MutanNoAtt", "fusion": {"dim_v": 2048, "dim_q": 2400, "dim_hv": 360, "dim_hq": 360, "dim_mm": 360, "R": 10},
           "classif": {"dropout": 0.5}}
    table = torch.tanh(torch.randn(n_q, 2400) * 0.5)
    vqa = MutanNoAtt(opt, ["w"], ["a"] * 2000, seq2vec=QuestionTable(table).cuda())
    spec = {k: True for k in ("v_emb", "v_mult", "v_dist", "v_rank", "q_emb", "a_emb", "z_emb")}
    emb = (torch.randn(2000, 2400) * 0.35).numpy()
    return NeuralModel(spec, dim_h, layers, emb, drop_p, vqa_model=vqa, knn_size=24, trainable_vqa=False), table.cuda()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--eval-batch", type=int, default=4096)
    ap.add_argument("--dim-h", type=int, default=300)
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--n-img", type=int, default=82783)
    ap.add_argument("--eval-only", action="store_true", help="skip the train steps (short ncu launch lists)")
    args = ap.parse_args()
    from vqa_counterexamples_b200 import synth
    from vqa_counterexamples_b200._lib import lib

    dev = torch.device("cuda")
    feats = synth.clustered_features_cuda(args.n_img, 2048, seed=1234)
    cx, table = build(args.dim_h, args.layers, 4096, 0.25)
    g = torch.Generator(device="cuda").manual_seed(1)

    def batch(B):
        img = torch.randint(0, args.n_img, (B, 25), generator=g, device=dev, dtype=torch.int32)
        q = table[torch.randint(0, table.shape[0], (B,), generator=g, device=dev)]
        aid = torch.randint(0, 2000, (B,), generator=g, device=dev)
        comp = torch.randint(0, 24, (B,), generator=g, device=dev)
        return img, q, aid, comp

    def timed(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        l0 = lib.vqacx_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n, (lib.vqacx_launch_count() - l0) / n

    cx.train()
    b = batch(args.batch)
    if args.eval_only:
        cx.eval()
        b = batch(args.eval_batch)
        ms, launches = timed(lambda: cx.eval_batch(feats, *b), args.steps)
        print(f"eval   B={args.eval_batch}: {ms:.3f} ms/batch  {args.eval_batch / ms * 1e3:.0f} samples/s  {launches:.0f} launches")
        return
    ms, launches = timed(lambda: cx.train_step(feats, *b, lr=1e-4), args.steps)
    print(f"train  B={args.batch}: {ms:.3f} ms/step  {args.batch / ms * 1e3:.0f} samples/s  {launches:.0f} launches/step")
    ms, launches = timed(lambda: cx.train_step_graph(feats, *b, lr=1e-4), args.steps)
    print(f"train  B={args.batch} (CUDA graph): {ms:.3f} ms/step  {args.batch / ms * 1e3:.0f} samples/s  {launches:.0f} launches/step")
    cx.eval()
    for B in (args.batch, args.eval_batch):
        b = batch(B)
        ms, launches = timed(lambda: cx.eval_batch(feats, *b), max(5, args.steps // 5))
        gflop = 242.2e-3 * B
        print(f"eval   B={B}: {ms:.3f} ms/batch  {B / ms * 1e3:.0f} samples/s  {launches:.0f} launches  "
              f"{gflop / ms:.1f} TFLOP/s algorithmic")


if __name__ == "__main__":
    main()
