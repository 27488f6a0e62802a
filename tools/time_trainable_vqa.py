"""trainable_vqa=True at the reference's full dimensions (options/cx/neuralcx_trainable_vqa.yaml: MUTAN no-att 2048/2400/360,
R = 10, 2000 answers, dim_h 1024, 3 layers, batch 64; skip-thoughts GRU 620 -> 2400 over 26 steps): a few fused train
steps (forward_tv, CE, backward_tv, BPTT, three Adam passes), loss trajectory and time per step."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vqa_counterexamples_b200.cx import NeuralModel
from vqa_counterexamples_b200.vqa_models import BayesianUniSkip, MutanNoAtt
from vqa_counterexamples_b200 import synth

B, V, T, n_img = 64, 3000, 26, 20000
torch.manual_seed(0)
enc = BayesianUniSkip(V)
opt = {"arch": "MutanNoAtt", "seq2vec": {"dropout": 0.25},
       "fusion": {"dim_v": 2048, "dim_q": 2400, "dim_hv": 360, "dim_hq": 360, "dim_mm": 360, "R": 10, "dropout_v": 0.5,
                  "dropout_q": 0.5, "dropout_hv": 0, "dropout_hq": 0},
       "classif": {"dropout": 0.5}}
vqa = MutanNoAtt(opt, ["w"] * V, ["a"] * 2000, seq2vec=enc)
spec = {k: True for k in ("v_emb", "v_mult", "v_dist", "v_rank", "q_emb", "a_emb", "z_emb")}
cx = NeuralModel(spec, 1024, 3, (torch.randn(2000, 2400) * 0.35).numpy(), 0.25, vqa_model=vqa, knn_size=24, trainable_vqa=True)
assert cx.trains_seq2vec
cx.train()
feats = synth.clustered_features_cuda(n_img, 2048, seed=3)
g = torch.Generator(device="cuda").manual_seed(1)
img = torch.randint(0, n_img, (B, 25), generator=g, device="cuda", dtype=torch.int32)
wids = torch.randint(1, V + 1, (B, T), generator=g, device="cuda")
lens = torch.randint(3, T + 1, (B,), generator=g, device="cuda")
wids = wids * (torch.arange(T, device="cuda")[None, :] < lens[:, None])
aid = torch.randint(0, 2000, (B,), generator=g, device="cuda")
comp = torch.randint(0, 24, (B,), generator=g, device="cuda")
losses = []
for step in range(8):
    cx.reset_metrics()
    if step == 3:
        torch.cuda.synchronize(); t0 = time.time()
    cx.train_step(feats, img, None, aid, comp, lr=1e-4, question_wids=wids)
    losses.append(float(cx.metrics[0]) / B)
torch.cuda.synchronize()
print("loss per step:", [round(l, 4) for l in losses])
print(f"{(time.time() - t0) / 5 * 1e3:.2f} ms per trainable-VQA step (B = {B}, incl. the host sync of reading the loss)")
assert all(np.isfinite(losses)) and losses[-1] < losses[0] + 0.05
