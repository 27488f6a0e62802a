"""GPU tests of the re-hosted counterexamples.py driver (vqa_counterexamples_b200/counterexamples.py): the reference's
free functions (getDataFromBatch, eval_model, checkpoints) and the fused train loop, against the oracle restatement on
seeded synthetic example dicts of the reference's schema."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import ncx_oracle as O

pytestmark = pytest.mark.gpu
DV, DQ, DF, R, NA, DIM_H, L = 64, 48, 40, 3, 37, 16, 2
N_IMG, N_Q = 120, 50


def make_world(seed, n_examples):
    """(params, features [N_IMG, DV], q_table [N_Q, DQ], dataset dict of the reference's pickle schema)."""
    rng = np.random.default_rng(seed)
    params = O.make_params(seed, dim_v=DV, dim_q=DQ, dim_h_fusion=DF, R=R, n_ans=NA, dim_h=DIM_H, n_layers=L)
    feats = np.maximum(rng.normal(0.3, 0.5, size=(N_IMG, DV)), 0).astype(np.float32)
    q_table = np.tanh(rng.normal(0, 0.5, size=(N_Q, DQ))).astype(np.float32)
    names = ["COCO_val2014_%012d.jpg" % i for i in range(N_IMG)]
    perm = rng.permutation(N_IMG)
    name_to_index = {names[i]: int(perm[i]) for i in range(N_IMG)}     # row order differs from name order
    examples = []
    for e in range(n_examples):
        img = int(rng.integers(N_IMG))
        knn = rng.choice([i for i in range(N_IMG) if i != img], size=24, replace=False)
        wids = [int(rng.integers(N_Q))] + [int(w) for w in rng.integers(1, 100, size=5)] + [0] * 20
        examples.append({"image_name": names[img], "knns": [names[j] for j in knn], "question_wids": wids,
                         "answer_aid": int(rng.integers(NA)), "comp": {"knn_index": int(rng.integers(24))},
                         "question_id": e})
    ds = {"examples_list": examples, "name_to_index": name_to_index, "vocab_words": ["w"], "vocab_answers": ["a"] * NA}
    return params, feats, q_table, ds


def build_model(params, q_table, drop_p=0.0):
    from vqa_counterexamples_b200.cx import NeuralModel
    from vqa_counterexamples_b200.vqa_models import MutanNoAtt, QuestionTable

    opt = {"arch": "MutanNoAtt", "fusion": {"dim_v": DV, "dim_q": DQ, "dim_hv": DF, "dim_hq": DF, "dim_mm": DF, "R": R},
           "classif": {"dropout": 0.5}}
    vqa = MutanNoAtt(opt, ["w"], ["a"] * NA, seq2vec=QuestionTable(q_table).cuda())
    spec = {k: True for k in ("v_emb", "v_mult", "v_dist", "v_rank", "q_emb", "a_emb", "z_emb")}
    cx = NeuralModel(spec, DIM_H, L, None, drop_p, vqa_model=vqa, knn_size=24, trainable_vqa=False)
    cx.load_state_dict({k: torch.as_tensor(v) for k, v in params.items()})
    return cx


def oracle_inputs(batch, feats, q_table, name_to_index):
    idx = np.array([[name_to_index[ex["image_name"]]] + [name_to_index[n] for n in ex["knns"]] for ex in batch])
    q = q_table[[ex["question_wids"][0] for ex in batch]]
    aid = np.array([ex["answer_aid"] for ex in batch])
    comp = np.array([ex["comp"]["knn_index"] for ex in batch])
    return feats[idx], q, aid, comp


def test_getDataFromBatch_host_and_device_forms_agree():
    from vqa_counterexamples_b200.counterexamples import DeviceFeatures, getDataFromBatch

    params, feats, q_table, ds = make_world(1, 9)
    batch = ds["examples_list"][:7]
    a = getDataFromBatch(batch, feats, ds["name_to_index"])
    b = getDataFromBatch(batch, DeviceFeatures(feats), ds["name_to_index"])
    x, q, aid, comp = oracle_inputs(batch, feats, q_table, ds["name_to_index"])
    assert a[0].is_cuda and tuple(a[0].shape) == (7, 25, DV) and a[0].dtype == torch.float32
    assert torch.equal(a[0], b[0]) and np.array_equal(a[0].cpu().numpy(), x)
    for u, v in zip(a[1:], b[1:]):
        assert u.dtype == torch.int64 and torch.equal(u, v)
    assert a[1].shape == (7, 26) and np.array_equal(a[2].cpu().numpy(), aid) and np.array_equal(a[3].cpu().numpy(), comp)
    with pytest.raises(NotImplementedError):
        getDataFromBatch(batch, feats, ds["name_to_index"], pairwise=True)


def test_eval_model_matches_oracle():
    from vqa_counterexamples_b200.counterexamples import eval_model

    params, feats, q_table, ds = make_world(2, 53)
    cx = build_model(params, q_table)
    random.seed(42)
    res = eval_model(cx, ds, feats, batch_size=8, return_recall1=True)          # 6 full batches + one of 5
    assert cx.training                                                         # counterexamples.py:488
    x, q, aid, comp = oracle_inputs(ds["examples_list"], feats, q_table, ds["name_to_index"])
    p = O.to_torch(params)
    scores = O.neural_forward(p, torch.as_tensor(x), torch.as_tensor(q), torch.as_tensor(aid), L)
    gt = torch.as_tensor(comp)
    assert abs(res["loss"] - float(O.ce_sum(scores, gt)) / 53) <= 1e-3
    assert res["recall"] == float(O.recall_at_k(scores, gt, 5).sum()) / 53
    assert res["recall_1"] == float(O.recall_at_k(scores, gt, 1).sum()) / 53
    # the reference's own loop shape (getDataFromBatch -> model -> recallAtK) gives the same hits
    from vqa_counterexamples_b200.counterexamples import batchify, getDataFromBatch, recallAtK

    cx.eval()
    hits = 0
    for batch in batchify(ds["examples_list"], 8):
        f, w, a, c = getDataFromBatch(batch, feats, ds["name_to_index"])
        hits += int(recallAtK(cx(f, w, a), c, k=5).sum())
    assert hits / 53 == res["recall"]


def test_distance_baseline_eval_is_closed_form():
    from vqa_counterexamples_b200.counterexamples import eval_model
    from vqa_counterexamples_b200.cx import DistanceBaseline

    _, feats, _, ds = make_world(3, 40)
    res = eval_model(DistanceBaseline(24), ds, feats, batch_size=16, return_recall1=True)
    comp = np.array([ex["comp"]["knn_index"] for ex in ds["examples_list"]])
    assert res["recall"] == float((comp < 5).mean()) and res["recall_1"] == float((comp < 1).mean())


def test_sliced_global_batch_gradients_sum_to_full_batch():
    """Data-parallel semantics on one GPU: two ranks' slices of a global batch, each scaled by 1/global_batch, sum to
    the single-process gradient (what the SUM all-reduce of DataParallelCX delivers)."""
    from vqa_counterexamples_b200.counterexamples import DeviceFeatures, ExampleIndex, shard_batch

    params, feats, q_table, ds = make_world(4, 11)
    dev_feats = DeviceFeatures(feats)
    index = ExampleIndex(ds["examples_list"], ds["name_to_index"])
    ids = index.ids_of(ds["examples_list"])
    table = torch.as_tensor(q_table).cuda()

    def grads_of(sel, global_batch):
        cx = build_model(params, q_table)
        cx.train()
        img, wids, aid, comp = index.take(sel)
        cx.train_step(dev_feats.tensor, img, table[wids[:, 0]], aid, comp, lr=0.0, global_batch=global_batch)
        return cx.last_grads.double(), float(cx.metrics[0])

    full, loss_full = grads_of(ids, 11)
    parts = [grads_of(shard_batch(ids, r, 2), 11) for r in range(2)]
    summed = parts[0][0] + parts[1][0]
    scale = full.abs().max().item()
    assert (summed - full).abs().max().item() <= 2e-3 * scale          # TF32 products, different tile groupings
    assert abs(parts[0][1] + parts[1][1] - loss_full) <= 1e-3


def test_checkpoint_roundtrip_and_resume_key(tmp_path):
    from vqa_counterexamples_b200.counterexamples import eval_model, load_cx_checkpoint, save_cx_checkpoint

    params, feats, q_table, ds = make_world(5, 20)
    cx = build_model(params, q_table)
    os.makedirs(tmp_path / "ckpt")
    os.makedirs(tmp_path / "best")
    info = [eval_model(cx, ds, feats, 8, shuffle=False)]
    save_cx_checkpoint(cx, info, str(tmp_path), is_best=True)
    sd = torch.load(tmp_path / "best" / "model.ckpt")
    for name in ("answer_embedding.weight", "linear_1.weight", "linear_1.bias", "linear_2.weight", "out.weight", "out.bias",
                 "vqa_model.fusion.linear_v.weight", "vqa_model.linear_classif.bias"):
        assert name in sd, name
    assert tuple(sd["linear_1.weight"].shape) == (DIM_H, DV * 3 + 2400 * 2 + DF * 2 + DQ + 24 + 1)
    other = build_model(O.make_params(99, dim_v=DV, dim_q=DQ, dim_h_fusion=DF, R=R, n_ans=NA, dim_h=DIM_H, n_layers=L), q_table)
    info2, next_epoch, best = load_cx_checkpoint(other, str(tmp_path), resume_best=True)
    assert next_epoch == 2 and best == info[0]["recall"] and info2 == info
    again = eval_model(other, ds, feats, 8, shuffle=False)
    assert abs(again["loss"] - info[0]["loss"]) <= 1e-5 and again["recall"] == info[0]["recall"]


def test_train_epoch_learns_and_matches_reference_loop_shape():
    """A few epochs of the fused loop on a tiny set must drive the training loss down (the labels are learnable from
    v_rank alone), and print hooks fire at print_freq with the reference's metric names."""
    from vqa_counterexamples_b200.counterexamples import eval_model, train_epoch

    params, feats, q_table, ds = make_world(6, 64)
    for ex in ds["examples_list"]:
        ex["comp"]["knn_index"] = ex["question_id"] % 3                           # learnable from the rank one-hot
    cx = build_model(params, q_table, drop_p=0.0)
    random.seed(42)
    before = eval_model(cx, ds, feats, 16)
    seen = []
    for _ in range(30):
        n = train_epoch(cx, ds, feats, 16, lr=1e-2, print_freq=2, on_print=lambda b, m: seen.append((b, sorted(m))))
        assert n == 4
    after = eval_model(cx, ds, feats, 16)
    assert after["loss"] < before["loss"] - 0.5, (before, after)
    assert after["recall"] >= before["recall"]
    assert seen[0] == (2, ["loss", "recall"]) and seen[1][0] == 4
