"""Generates tests/golden/heads_*.npz by running the REFERENCE's own BlackBox / LinearContext / SemanticBaseline
(imported read-only from /root/reference: vqa.models.cx over vqa.models.noatt.MutanNoAtt) on seeded synthetic inputs.

Run in the build container:  python oracle/gen_golden_heads.py
Inputs and weights are NOT stored; they are regenerated from seeds (oracle.ncx_oracle.make_params / make_batch,
oracle.heads_oracle.make_linear). Shims: the three documented in oracle/gen_golden_ncx.py, plus
  4. torch.cuda.FloatTensor -> torch.FloatTensor (SemanticBaseline.forward builds its result with it, cx.py:203;
     no GPU in the build container).
"""
import os
import sys

import numpy as np
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True

from oracle import heads_oracle as H  # noqa: E402
from oracle import ncx_oracle as O  # noqa: E402
from oracle.gen_golden_ncx import REF, TableSeq2Vec, import_reference  # noqa: E402

CASES = [
    # name, dims (dim_v, dim_q, dim_fusion, R, n_ans), B, seed, lambda
    ("small", (64, 48, 40, 3, 37), 5, 201, 0.5),
    ("small_lam", (96, 64, 24, 2, 50), 4, 202, 0.8),
    ("full", (2048, 2400, 360, 10, 2000), 4, 203, 0.5),
]


def build_vqa(s2v, MutanNoAtt, params, dims, q_table):
    s2v.factory = lambda vocab_words, opt: TableSeq2Vec(q_table)
    opt = {"arch": "MutanNoAtt", "seq2vec": {"arch": "stub"},
           "fusion": {"dim_v": dims["dim_v"], "dim_q": dims["dim_q"], "dim_hv": dims["dim_hv"], "dim_hq": dims["dim_hq"],
                      "dim_mm": dims["dim_mm"], "R": dims["R"], "dropout_v": 0.5, "dropout_q": 0.5,
                      "activation_v": "tanh", "activation_q": "tanh", "dropout_hv": 0, "dropout_hq": 0},
           "classif": {"dropout": 0.5}}
    vqa = MutanNoAtt(opt, vocab_words=["w"] * 10, vocab_answers=["a"] * dims["n_ans"])
    sd = vqa.state_dict()
    for k, v in params.items():
        if k.startswith("vqa_model."):
            k2 = k[len("vqa_model."):]
            assert tuple(sd[k2].shape) == tuple(v.shape), k
            sd[k2] = torch.as_tensor(v)
    vqa.load_state_dict(sd)
    return vqa


def main():
    s2v, _, MutanNoAtt = import_reference()
    torch.cuda.FloatTensor = torch.FloatTensor
    from vqa.models.cx import BlackBox, LinearContext, SemanticBaseline
    out_dir = os.path.join(ROOT, "tests", "golden")
    for name, (dv, dq, df, R, na), B, seed, lam in CASES:
        params = O.make_params(seed, dim_v=dv, dim_q=dq, dim_h_fusion=df, R=R, n_ans=na, dim_h=12, n_layers=1)
        feats, q_emb, aid, comp = O.make_batch(seed + 1000, B, dim_v=dv, dim_q=dq, n_ans=na)
        dims = O.fusion_dims(params)
        wids = torch.zeros((B, 26), dtype=torch.long)
        wids[:, 0] = torch.arange(B)
        x, aid_t = torch.as_tensor(feats), torch.as_tensor(aid)
        save = {"meta": np.array([dv, dq, df, R, na, B, seed], dtype=np.int64), "lam": np.float64(lam),
                "torch_version": np.array(torch.__version__)}

        bb = BlackBox(build_vqa(s2v, MutanNoAtt, params, dims, q_emb), knn_size=24, trainable_vqa=False)
        save["blackbox"] = bb(x, wids, aid_t).detach().numpy()

        sb = SemanticBaseline(build_vqa(s2v, MutanNoAtt, params, dims, q_emb), knn_size=24, trainable_vqa=False)
        sb.set_lambda(lam)
        sb.set_answer_embedding(params["answer_embedding.weight"])
        save["semantic"] = sb(x, wids, aid_t).detach().numpy()
        rng = np.random.Generator(np.random.Philox(seed + 5))
        ij = rng.integers(0, na, size=(256, 2))
        save["cos_ij"], save["cos_val"] = ij.astype(np.int64), sb.emb_pairs[ij[:, 0], ij[:, 1]]

        lc = LinearContext(build_vqa(s2v, MutanNoAtt, params, dims, q_emb), knn_size=24, trainable_vqa=False)
        w, b = H.make_linear(seed + 9, 24, df)
        lc.linear.weight.data, lc.linear.bias.data = torch.as_tensor(w), torch.as_tensor(b)
        lc.train()
        scores = lc(x, wids, aid_t)
        loss = nn.CrossEntropyLoss(reduction="sum")(scores, torch.as_tensor(comp)) / B      # counterexamples.py:310,334
        lc.zero_grad()
        loss.backward()
        save["linear_scores"], save["linear_loss"] = scores.detach().numpy(), np.float64(loss.item())
        save["linear_gw"], save["linear_gb"] = lc.linear.weight.grad.numpy(), lc.linear.bias.grad.numpy()
        np.savez_compressed(os.path.join(out_dir, f"heads_{name}.npz"), **save)
        print(name, "blackbox[0,:3]", save["blackbox"][0, :3], "semantic[0,:3]", save["semantic"][0, :3], "lc loss",
              float(loss))


if __name__ == "__main__":
    main()
