"""Debug: per-tensor gradient error of one fused train step vs the fp64 oracle on the tiny trajectory-test config."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from oracle import ncx_oracle as O
from test_ncx_gpu import build
from vqa_counterexamples_b200.cx import unpack_state

dv, dq, df, R, na, dim_h, L, B = 64, 48, 40, 3, 37, 12, 2, 7
params = O.make_params(7, dim_v=dv, dim_q=dq, dim_h_fusion=df, R=R, n_ans=na, dim_h=dim_h, n_layers=L)
feats, q_emb, aid, comp = O.make_batch(77, B, dim_v=dv, dim_q=dq, n_ans=na)
cx = build(params, L, dim_h, q_emb); cx.train()
x = torch.as_tensor(feats).cuda().view(B * 25, dv)
img_idx = torch.arange(B * 25, dtype=torch.int32, device="cuda").view(B, 25)
cx.train_step(x, img_idx, torch.as_tensor(q_emb).cuda(), torch.as_tensor(aid).cuda(), torch.as_tensor(comp).cuda(), lr=0.0)
s, loss, grads = O.loss_and_grads(params, feats, q_emb, aid, comp, L, dtype=torch.float64)
got = unpack_state(cx.last_grads, cx.dims, cx.layout)
print("aid", aid)
for k, g in grads.items():
    ref = g.numpy(); mine = got[k].cpu().numpy().astype(np.float64)
    err = np.abs(mine - ref)
    print(f"{k:28s} max|ref| {np.abs(ref).max():.3e} maxerr {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}")
    if k == "answer_embedding.weight":
        rows = err.max(1)
        top = np.argsort(-rows)[:6]
        print("  worst rows", [(int(r), float(rows[r]), float(np.abs(ref[r]).max())) for r in top])
