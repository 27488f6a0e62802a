"""Debug: per-tensor gradient error at the BENCH shape (full dims, B = 64) for the eager step, the graph-replayed step
and the serial (no side streams) schedule, vs the fp64 oracle; answer_embedding split into ground-truth rows / others."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import ncx_oracle as O
from test_ncx_gpu import build, FULL
from vqa_counterexamples_b200 import _lib
from vqa_counterexamples_b200.cx import unpack_state

B, L = 64, FULL["n_layers"]
params = O.make_params(22, **FULL)
feats, q_emb, aid, comp = O.make_batch(2200, B, dim_v=FULL["dim_v"], dim_q=FULL["dim_q"], n_ans=FULL["n_ans"])
ref_s, ref_loss, ref_g = O.loss_and_grads(params, feats, q_emb, aid, comp, L, dtype=torch.float64)
_, _, ref_g32 = O.loss_and_grads(params, feats, q_emb, aid, comp, L, dtype=torch.float32)
x = torch.as_tensor(feats).cuda().view(B * 25, -1)
img_idx = torch.arange(B * 25, dtype=torch.int32, device="cuda").view(B, 25)
args = (x, img_idx, torch.as_tensor(q_emb).cuda(), torch.as_tensor(aid).cuda(), torch.as_tensor(comp).cuda())
gt_rows = np.unique(aid)
other = np.setdiff1d(np.arange(FULL["n_ans"]), gt_rows)


def report(tag, got):
    print("==", tag)
    for k, r in ref_g.items():
        ref = r.numpy(); mine = got[k].double().cpu().numpy()
        e = np.abs(mine - ref)
        print(f"  {k:26s} max|ref| {np.abs(ref).max():.3e}  maxerr/max {e.max() / np.abs(ref).max():.2e}  relL2 "
              f"{np.linalg.norm(mine - ref) / np.linalg.norm(ref):.2e}")
        if k == "answer_embedding.weight":
            for nm, rows in (("gt rows", gt_rows), ("other rows", other)):
                print(f"      {nm:10s} max|ref| {np.abs(ref[rows]).max():.3e} maxerr {e[rows].max():.3e} relL2 "
                      f"{np.linalg.norm((mine - ref)[rows]) / np.linalg.norm(ref[rows]):.2e}")


report("fp32 oracle (reference precision) vs fp64 oracle", {k: v for k, v in ref_g32.items()})
for tag, serial, graph in (("eager", 0, False), ("eager serial", 1, False), ("graph", 0, True), ("graph serial", 1, True)):
    with _lib.option(_lib.OPT_NCX_SERIAL, serial):
        cx = build(params, L, FULL["dim_h"], torch.zeros((1, FULL["dim_q"])), drop_p=0.0)
        cx.train()
        step = cx.train_step_graph if graph else cx.train_step
        for rep in range(2):
            step(*args, lr=0.0)
            torch.cuda.synchronize()
            report(f"{tag} rep {rep}", unpack_state(cx.last_grads, cx.dims, cx.layout))
