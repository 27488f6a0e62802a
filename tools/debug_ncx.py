"""Scratch: per-tensor error report of the CUDA NeuralCX path vs golden fixtures (python tools/debug_ncx.py)."""
import glob, os, sys
import numpy as np, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from oracle import ncx_oracle as O
import test_ncx_gpu as T

for path in sorted(glob.glob("tests/golden/ncx_*.npz")):
    z, params, (feats, q_emb, aid, comp), L, dim_h = T.load_case(path)
    cx = T.build(params, L, dim_h, q_emb)
    scores, loss, grads = T.run_reference_style_step(cx, feats, aid, comp)
    print(os.path.basename(path), "score err", np.abs(scores - z["scores"]).max(), "loss err", abs(loss - float(z["loss"])))
    for key in z.files:
        if key.startswith("gnorm:"):
            name = key[6:]; g = grads[name].astype(np.float64)
            if "grad:" + name in z.files:
                ref = z["grad:" + name].astype(np.float64); err = np.abs(g - ref).max()
            else:
                ref = z["gval:" + name].astype(np.float64); err = np.abs(g.reshape(-1)[z["gidx:" + name]] - ref).max()
            print(f"   {name:28s} max|ref| {np.abs(ref).max():.3e} maxerr {err:.3e} rel {err/ (np.abs(ref).max()+1e-30):.3e} norm {np.sqrt((g**2).sum()):.4e} vs {float(z[key]):.4e}")
    cx.eval()
    B = feats.shape[0]
    x = torch.as_tensor(feats).cuda().view(B * 25, -1)
    idx = torch.arange(B * 25, dtype=torch.int32, device="cuda").view(B, 25)
    q, a = torch.as_tensor(q_emb).cuda(), torch.as_tensor(aid).cuda()
    e = [cx.forward_indexed(x, idx, q, a) for _ in range(4)]
    print("   eval repeat max diff", max((e[0] - t).abs().max().item() for t in e[1:]), "eval err", np.abs(e[0].cpu().numpy() - z["scores_eval"]).max())
