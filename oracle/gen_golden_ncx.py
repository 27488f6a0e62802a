"""Generates tests/golden/ncx_*.npz by running the REFERENCE's own NeuralCX code (imported read-only from
/root/reference: vqa.models.cx.NeuralModel over vqa.models.noatt.MutanNoAtt) on seeded synthetic inputs.

Run in the build container:  python oracle/gen_golden_ncx.py
Inputs and weights are NOT stored; they are regenerated from seeds by oracle.ncx_oracle.make_params/make_batch.

Shims applied to the reference (it targets PyTorch 0.3 + a GPU), none of which touches its arithmetic:
  1. F.pairwise_distance(..., keepdim=True): modern torch returns [B], cx.py:309-320 concatenates it as [B,1];
  2. Tensor.cuda / Module.cuda -> identity (no GPU in the build container);
  3. vqa.models.seq2vec.factory -> a table lookup module: the skip-thought encoder downloads its weights
     (skipthoughts.py:44-70) and is outside the hot path; q_emb = table[question_wids[:, 0]].
The trainable_vqa cases (TV_CASES, cx.py:73-74) run the reference with trainable_vqa=True and the whole module in
train mode; two more shims, again outside its arithmetic:
  4. F.dropout draws its Bernoulli mask through torch.rand_like and RECORDS it (the three train-mode dropouts of the
     head: fusion.py:86,94 and noatt.py:27), so that the oracle and the CUDA path can replay the same masks;
  5. the table of shim 3 is an nn.Parameter, whose gradient is dLoss/dq_emb (what seq2vec's backward would receive).
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


class ParamTableSeq2Vec(nn.Module):
    def __init__(self, table):
        super().__init__()
        self.table = nn.Parameter(torch.as_tensor(table).clone())

    def forward(self, wids):
        return self.table[wids[:, 0]]


def build_reference(s2v, NeuralModel, MutanNoAtt, params, dims, q_table, dim_h, n_layers, drop_p, lesion=(),
                    trainable_vqa=False, head_drop=(0.5, 0.5, 0.5)):
    s2v.factory = lambda vocab_words, opt: (ParamTableSeq2Vec if trainable_vqa else TableSeq2Vec)(q_table)
    opt = {"arch": "MutanNoAtt", "seq2vec": {"arch": "stub"},
           "fusion": {"dim_v": dims["dim_v"], "dim_q": dims["dim_q"], "dim_hv": dims["dim_hv"], "dim_hq": dims["dim_hq"],
                      "dim_mm": dims["dim_mm"], "R": dims["R"], "dropout_v": head_drop[0], "dropout_q": head_drop[1],
                      "activation_v": "tanh", "activation_q": "tanh", "dropout_hv": 0, "dropout_hq": 0},
           "classif": {"dropout": head_drop[2]}}
    vqa = MutanNoAtt(opt, vocab_words=["w"] * 10, vocab_answers=["a"] * dims["n_ans"])
    spec = {k: k not in lesion for k in ("v_emb", "v_mult", "v_dist", "v_rank", "q_emb", "a_emb", "z_emb")}
    cx = NeuralModel(spec, dim_h, n_layers, params["answer_embedding.weight"], drop_p, vqa_model=vqa, knn_size=24,
                     trainable_vqa=trainable_vqa)
    sd = cx.state_dict()
    for k, v in params.items():
        if k == "vqa_model.seq2vec.table":
            continue
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


TV_CASES = [
    # name, dims, dim_h, n_layers, B, seed, head dropouts (dropout_v, dropout_q, classif.dropout)
    ("tv_small_nodrop", (64, 48, 40, 3, 37), 12, 2, 5, 201, (0.0, 0.0, 0.0)),
    ("tv_small_drop", (64, 48, 40, 3, 37), 12, 2, 6, 202, (0.5, 0.5, 0.5)),
    ("tv_small_l3", (96, 64, 24, 2, 50), 20, 3, 3, 203, (0.5, 0.5, 0.5)),
]


class DropoutRecorder:
    """F.dropout with the same semantics (inverted Bernoulli mask, identity when p = 0 or not training) whose masks are
    kept, in call order. torch.nn.Dropout modules go through the same function object."""

    def __init__(self):
        self.masks, self._orig = [], F.dropout

    def __enter__(self):
        def dropout(input, p=0.5, training=True, inplace=False):
            if not training or p == 0:
                return input
            mask = (torch.rand_like(input) >= p).to(input.dtype) / (1.0 - p)
            self.masks.append(mask.clone())
            return input * mask
        F.dropout = dropout
        return self

    def __exit__(self, *exc):
        F.dropout = self._orig


def main_tv(s2v, NeuralModel, MutanNoAtt, out_dir, only=()):
    for name, (dv, dq, df, R, na), dim_h, n_layers, B, seed, head_drop in TV_CASES:
        if only and name not in only:
            continue
        params = O.make_params(seed, dim_v=dv, dim_q=dq, dim_h_fusion=df, R=R, n_ans=na, dim_h=dim_h, n_layers=n_layers)
        feats, q_emb, aid, comp = O.make_batch(seed + 1000, B, dim_v=dv, dim_q=dq, n_ans=na)
        dims = O.fusion_dims(params)
        cx = build_reference(s2v, NeuralModel, MutanNoAtt, params, dims, q_emb, dim_h, n_layers, drop_p=0.0,
                             trainable_vqa=True, head_drop=head_drop)
        torch.manual_seed(seed)
        wids = torch.zeros((B, 26), dtype=torch.long)
        wids[:, 0] = torch.arange(B)
        x = torch.as_tensor(feats)
        cx.train()                       # trainable_vqa: the VQA model stays in train mode (cx.py:58-59 does not eval() it)
        assert cx.vqa_model.training
        with DropoutRecorder() as rec:
            scores = cx(x, wids, torch.as_tensor(aid))
        loss = nn.CrossEntropyLoss(reduction="sum")(scores, torch.as_tensor(comp)) / B
        cx.zero_grad()
        loss.backward()
        n_masks = sum(1 for p_ in head_drop if p_ > 0)
        assert len(rec.masks) == n_masks, (len(rec.masks), n_masks)
        save = {"meta": np.array([dv, dq, df, R, na, dim_h, n_layers, B, seed], dtype=np.int64),
                "head_drop": np.array(head_drop, dtype=np.float64),
                "scores": scores.detach().numpy(), "loss": np.float64(loss.item()),
                "torch_version": np.array(torch.__version__)}
        if n_masks:
            # call order: fusion.py:86 (v), fusion.py:94 (q), noatt.py:27 (classifier); rows b*25 + i
            save["mask_v"], save["mask_q"], save["mask_c"] = (m.numpy().astype(np.float32) for m in rec.masks)
        for k, prm in cx.named_parameters():
            assert prm.grad is not None, k
            key = "grad:q_emb" if k == "vqa_model.seq2vec.table" else "grad:" + k
            save[key] = prm.grad.detach().numpy()
        np.savez_compressed(os.path.join(out_dir, f"ncx_{name}.npz"), **save)
        print(name, "loss", float(loss), "scores[0,:4]", scores[0, :4].tolist(), "masks", n_masks,
              "grads", len([k for k in save if k.startswith("grad:")]))


def main(only=()):
    s2v, NeuralModel, MutanNoAtt = import_reference()
    out_dir = os.path.join(ROOT, "tests", "golden")
    main_tv(s2v, NeuralModel, MutanNoAtt, out_dir, only)
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
