"""Example builder join. The fixtures tests/golden/builder_{val,train}.npz come from EXECUTING the
reference notebooks' own cells on seeded synthetic inputs (oracle/gen_golden_builder.py); the CPU tests pin the oracle
restatement against them (and on a hand-checked case), the GPU tests compare vqacx_cx_build_examples with both."""
import os

import numpy as np
import pytest

from oracle.builder_oracle import build_examples, to_arrays

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _golden_case(tag):
    g = np.load(os.path.join(GOLDEN, f"builder_{tag}.npz"))
    coco = g["coco"]
    row_of = {int(c): i for i, c in enumerate(coco)}
    examples = [{"question_id": int(q), "image": int(coco[i])} for q, i in zip(g["qids"], g["q_img"])]
    knns = {}
    for i, row in enumerate(g["knn"].tolist()):       # the notebook's 24-neighbour lists, derived as stage 1's consumer does
        if i in row:
            row.remove(i)
        else:
            row = row[:-1]
        knns[int(coco[i])] = [int(coco[j]) for j in row]
    pairs = [tuple(p) for p in g["pairs"].tolist()]
    return g, row_of, examples, pairs, knns


@pytest.mark.parametrize("tag", ["val", "train"])
def test_oracle_matches_notebook_cells(tag):
    """The restatement against what the notebook's own cell code kept (order, comp.knn_index, neighbour lists)."""
    g, _, examples, pairs, knns = _golden_case(tag)
    out = build_examples(examples, pairs, knns)
    assert [e["question_id"] for e in out] == g["kept_qid"].tolist()
    assert [e["comp_question_id"] for e in out] == g["kept_comp_qid"].tolist()
    assert [e["comp_knn_index"] for e in out] == g["kept_knn_index"].tolist()
    assert [e["image"] for e in out] == g["kept_image"].tolist()
    assert [e["knns"] for e in out] == g["kept_knns"].tolist()


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["val", "train"])
def test_device_join_matches_notebook_cells(tag):
    """vqacx_cx_build_examples against the notebook-generated fixture directly (no restatement in between)."""
    import torch
    from vqa_counterexamples_b200.counterexamples import build_examples_from_knn

    g, row_of, examples, pairs, _ = _golden_case(tag)
    q_img, q_comp = to_arrays(examples, pairs)                       # id bookkeeping (host dictionaries)
    q_img = np.asarray([row_of[int(c)] for c in q_img], dtype=np.int32)
    kept, img_idx, comp_idx = build_examples_from_knn(torch.from_numpy(g["knn"]).cuda(), q_img, q_comp)
    coco = g["coco"]
    assert g["qids"][kept.cpu().numpy()].tolist() == g["kept_qid"].tolist()
    assert comp_idx.cpu().tolist() == g["kept_knn_index"].tolist()
    got = coco[img_idx.cpu().numpy()]
    assert got[:, 0].tolist() == g["kept_image"].tolist()
    assert got[:, 1:].tolist() == g["kept_knns"].tolist()


def test_oracle_hand_case():
    knns = {0: [1, 2, 3], 1: [0, 2, 3], 2: [3, 0, 1], 3: [2, 1, 0]}
    examples = [{"question_id": 10, "image": 0}, {"question_id": 11, "image": 2}, {"question_id": 12, "image": 1},
                {"question_id": 13, "image": 3}, {"question_id": 14, "image": 0}]
    pairs = [(10, 11), (12, 99), (13, 14)]                     # 99 is not an example -> 12 is dropped
    out = build_examples(examples, pairs, knns)
    assert [(e["question_id"], e["comp_knn_index"]) for e in out] == [(10, 1), (11, 1), (13, 2), (14, 2)]
    q_img, q_comp = to_arrays(examples, pairs)
    assert q_img.tolist() == [0, 2, 1, 3, 0] and q_comp.tolist() == [1, 0, -1, 4, 3]


def _synthetic(seed, n_img, n_q, k1):
    rng = np.random.default_rng(seed)
    # neighbour table with the image itself in column 0, except a few rows where "duplicates" displaced it
    knn = np.empty((n_img, k1), dtype=np.int64)
    for i in range(n_img):
        others = rng.permutation(np.delete(np.arange(n_img), i))[:k1 - 1]
        knn[i] = np.concatenate([[i], others])
    displaced = rng.choice(n_img, size=max(1, n_img // 50), replace=False)
    for i in displaced:
        knn[i] = rng.permutation(np.delete(np.arange(n_img), i))[:k1]      # self absent: the last entry is dropped
    swap = rng.choice(np.setdiff1d(np.arange(n_img), displaced), size=max(1, n_img // 50), replace=False)
    knn[swap, 0], knn[swap, 1] = knn[swap, 1].copy(), knn[swap, 0].copy()   # self in column 1 (an exact duplicate first)
    examples = [{"question_id": 1000 + q, "image": int(rng.integers(n_img))} for q in range(n_q)]
    qids = rng.permutation(n_q)
    pairs = [(1000 + int(a), 1000 + int(b)) for a, b in zip(qids[0:n_q // 2:2], qids[1:n_q // 2:2])]
    pairs += [(1000 + int(qids[-1]), 999_999)]                 # complement outside the example set
    # bias half of the pairs so that the complement's image IS a neighbour (otherwise almost nothing survives)
    for a, b in pairs[:len(pairs) // 2]:
        if b < 999_999:
            ia = examples[a - 1000]["image"]
            examples[b - 1000]["image"] = int(knn[ia, 1 + rng.integers(k1 - 1)])
    knns = {}
    for i in range(n_img):
        row = knn[i].tolist()
        if i in row:
            row.remove(i)
        else:
            row = row[:-1]
        knns[i] = row
    return knn, examples, pairs, knns


@pytest.mark.gpu
@pytest.mark.parametrize("seed,n_img,n_q", [(0, 300, 2000), (1, 60, 500)])
def test_device_join_matches_notebook_restatement(seed, n_img, n_q):
    import torch
    from vqa_counterexamples_b200.counterexamples import build_examples_from_knn

    k1 = 25
    knn, examples, pairs, knns = _synthetic(seed, n_img, n_q, k1)
    ref = build_examples(examples, pairs, knns)
    q_img, q_comp = to_arrays(examples, pairs)
    kept, img_idx, comp_idx = build_examples_from_knn(torch.from_numpy(knn).cuda(), q_img, q_comp)
    pos = {ex["question_id"]: i for i, ex in enumerate(examples)}
    assert len(ref) > 50
    assert kept.cpu().tolist() == [pos[e["question_id"]] for e in ref]
    assert comp_idx.cpu().tolist() == [e["comp_knn_index"] for e in ref]
    assert img_idx.cpu().tolist() == [[e["image"]] + e["knns"] for e in ref]
