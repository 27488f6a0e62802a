"""Generates tests/golden/ncx_*.npz by running the REFERENCE's own NeuralCX code (imported read-only from
/root/reference: vqa.models.cx.NeuralModel over vqa.models.noatt.MutanNoAtt) on seeded synthetic inputs.

Run in the build container:  python oracle/gen_golden_ncx.py
Inputs and weights are NOT stored; they are regenerated from seeds by oracle.ncx_oracle.make_params/make_batch.

Shims applied to the reference (it targets PyTorch 0.3 + a GPU), none of which touches its arithmetic:
  1. F.pairwise_distance(..., keepdim=True): modern torch returns [B], cx.py:309-320 concatenates it as [B,1];
  2. Tensor.cuda / Module.cuda -> identity (no GPU in the build container);
  3. vqa.models.seq2vec.factory -> a table lookup module: the skip-thought encoder downloads its weights
     (skipthoughts.py:44-70) and is outside the hot path; q_emb = table[question_wids[:, 0]].
"""
import os
import sys

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True

from oracle import ncx_oracle as O  # noqa: E402


class TableSeq2Vec(nn.Module):
    def __init__(self, table):
        super().__init__()
        self.register_buffer("table", torch.as_tensor(table))

    def forward(self, wids):
        return self.table[wids[:, 0]]


def import_reference():
    os.chdir(REF)  # seq2vec.py / convnets.py append relative paths to sys.path
    sys.path.insert(0, REF)
    _pd = F.pairwise_distance
    F.pairwise_distance = lambda a, b, p=2, eps=1e-6, keepdim=True: _pd(a, b, p, eps, True)
    torch.Tensor.cuda = lambda self, *a, **k: self
    nn.Module.cuda = lambda self, *a, **k: self
    import vqa.models.seq2vec as s2v
    from vqa.models.cx import NeuralModel
    from vqa.models.noatt import MutanNoAtt
    return s2v, NeuralModel, MutanNoAtt


def build_reference(s2v, NeuralModel, MutanNoAtt, params, dims, q_table, dim_h, n_layers, drop_p, lesion=()):
    s2v.factory = lambda vocab_words, opt: TableSeq2Vec(q_table)
    opt = {"arch": "MutanNoAtt", "seq2vec": {"arch": "stub"},
           "fusion": {"dim_v": dims["dim_v"], "dim_q": dims["dim_q"], "dim_hv": dims["dim_hv"], "dim_hq": dims["dim_hq"],
                      "dim_mm": dims["dim_mm"], "R": dims["R"], "dropout_v": 0.5, "dropout_q": 0.5,
                      "activation_v": "tanh", "activation_q": "tanh", "dropout_hv": 0, "dropout_hq": 0},
           "classif": {"dropout": 0.5}}
    vqa = MutanNoAtt(opt, vocab_words=["w"] * 10, vocab_answers=["a"] * dims["n_ans"])
    spec = {k: k not in lesion for k in ("v_emb", "v_mult", "v_dist", "v_rank", "q_emb", "a_emb", "z_emb")}
    cx = NeuralModel(spec, dim_h, n_layers, params["answer_embedding.weight"], drop_p, vqa_model=vqa, knn_size=24,
                     trainable_vqa=False)
    sd = cx.state_dict()
    for k, v in params.items():
        assert k in sd, k
        assert tuple(sd[k].shape) == tuple(v.shape), (k, sd[k].shape, v.shape)
        sd[k] = torch.as_tensor(v)
    cx.load_state_dict(sd)
    return cx


CASES = [
    # name, dims (dim_v, dim_q, dim_fusion, R, n_ans), dim_h, n_layers, B, seed, store full grads, lesioned flags
    ("small_l1", (64, 48, 40, 3, 37), 12, 1, 5, 101, True, ()),
    ("small_l2", (64, 48, 40, 3, 37), 12, 2, 6, 102, True, ()),
    ("small_l3", (96, 64, 24, 2, 50), 20, 3, 3, 103, True, ()),
    ("full_l2", (2048, 2400, 360, 10, 2000), 300, 2, 4, 104, False, ()),
    # options/cx/neuralcx_lesion_v_mult_dist.yaml, neuralcx_lesion_v_feat.yaml (v_rank: recorded torch.rand draws),
    # neuralcx_lesion_q_emb.yaml (changes nothing in the reference: cx.py:270 still runs vqa_forward)
    ("lesion_vmd", (64, 48, 40, 3, 37), 12, 2, 5, 105, True, ("v_mult", "v_dist")),
    ("lesion_vfeat", (64, 48, 40, 3, 37), 12, 1, 4, 106, True, ("v_mult", "v_dist", "v_rank")),
    ("lesion_q", (64, 48, 40, 3, 37), 12, 1, 4, 107, True, ("q_emb",)),
]


class RandRecorder:
    """Wraps torch.rand so the [B, 24] stand-ins the reference draws for a lesioned v_rank (cx.py:306) are kept."""

    def __init__(self):
        self.draws, self._orig = [], torch.rand

    def __enter__(self):
        def rand(*a, **k):
            t = self._orig(*a, **k)
            self.draws.append(t.clone())
            return t
        torch.rand = rand
        return self

    def __exit__(self, *exc):
        torch.rand = self._orig


def main(only=()):
    s2v, NeuralModel, MutanNoAtt = import_reference()
    out_dir = os.path.join(ROOT, "tests", "golden")
    for name, (dv, dq, df, R, na), dim_h, n_layers, B, seed, full, lesion in CASES:
        if only and name not in only:
            continue
        params = O.make_params(seed, dim_v=dv, dim_q=dq, dim_h_fusion=df, R=R, n_ans=na, dim_h=dim_h, n_layers=n_layers)
        feats, q_emb, aid, comp = O.make_batch(seed + 1000, B, dim_v=dv, dim_q=dq, n_ans=na)
        dims = O.fusion_dims(params)
        cx = build_reference(s2v, NeuralModel, MutanNoAtt, params, dims, q_emb, dim_h, n_layers, drop_p=0.0,
                             lesion=lesion)
        torch.manual_seed(seed)
        wids = torch.zeros((B, 26), dtype=torch.long)
        wids[:, 0] = torch.arange(B)
        x = torch.as_tensor(feats)
        # training-mode step exactly as counterexamples.py:325-337 (drop_p = 0 so it is deterministic)
        cx.train()
        cx.vqa_model.eval()
        with RandRecorder() as rec:
            scores = cx(x, wids, torch.as_tensor(aid))
        criterion = nn.CrossEntropyLoss(reduction="sum")  # == size_average=False
        loss = criterion(scores, torch.as_tensor(comp)) / B
        cx.zero_grad()
        loss.backward()
        # eval-mode forward (eval_model, counterexamples.py:450-465)
        cx.eval()
        with RandRecorder() as rec_eval:
            scores_eval = cx(x, wids, torch.as_tensor(aid)).detach()
        top5 = scores_eval.topk(5)[1]
        top1 = scores_eval.topk(1)[1]
        gt = torch.as_tensor(comp).view(-1, 1)
        save = {
            "meta": np.array([dv, dq, df, R, na, dim_h, n_layers, B, seed], dtype=np.int64),
            "scores": scores.detach().numpy(), "scores_eval": scores_eval.numpy(), "loss": np.float64(loss.item()),
            "recall5": (top5 == gt).sum(1).numpy().astype(np.int64), "recall1": (top1 == gt).sum(1).numpy().astype(np.int64),
            "torch_version": np.array(torch.__version__),
        }
        if lesion:
            save["lesion"] = np.array(list(lesion))
        if "v_rank" in lesion:
            save["v_rank_train"] = torch.stack(rec.draws).numpy()          # [24, B, 24]
            save["v_rank_eval"] = torch.stack(rec_eval.draws).numpy()
        rng = np.random.Generator(np.random.Philox(seed + 7))
        for k, prm in cx.named_parameters():
            if prm.grad is None:
                continue
            g = prm.grad.detach().numpy()
            save["gnorm:" + k] = np.float64(np.sqrt((g.astype(np.float64) ** 2).sum()))
            save["gsum:" + k] = np.float64(g.astype(np.float64).sum())
            if full or g.size <= 4096:
                save["grad:" + k] = g
            else:
                idx = rng.choice(g.size, size=2048, replace=False)
                save["gidx:" + k] = idx.astype(np.int64)
                save["gval:" + k] = g.reshape(-1)[idx]
        np.savez_compressed(os.path.join(out_dir, f"ncx_{name}.npz"), **save)
        print(name, "loss", float(loss), "scores[0,:4]", scores[0, :4].tolist(),
              "grads", sorted(k for k in save if k.startswith("gnorm:")))


if __name__ == "__main__":
    main(tuple(sys.argv[1:]))      # optional case names: regenerate only those fixtures
