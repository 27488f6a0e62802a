"""ORACLE (test infrastructure, not product code) for the other CX heads that share vqa_forward.

Only tests/, __graft_entry__.smoke() and bench.py's cpu legs may import this.

What it restates (numpy / torch on the CPU, plain ops in the reference's order), on top of ncx_oracle.vqa_forward:
  BlackBox.forward                          vqa/models/cx.py:120-136
  LinearContext.forward (+ autograd)        vqa/models/cx.py:149-156
  SemanticBaseline.set_answer_embedding     vqa/models/cx.py:172-175  (sklearn.metrics.pairwise.cosine_similarity:
                                            rows L2-normalised, zero rows left at zero, then X_n X_n^T)
  SemanticBaseline.forward                  vqa/models/cx.py:182-205  (numpy float loops; its softmax helper has no
                                            max subtraction, cx.py:177-180)

Pinning status: the reference has no tests or golden vectors for these heads. This restatement is
pinned against the reference's OWN classes imported from /root/reference in the build container: fixtures
tests/golden/heads_*.npz are produced by oracle/gen_golden_heads.py and tests/test_oracle_heads.py checks this file
against them.
"""
import numpy as np
import torch

from . import ncx_oracle as O


def cosine_similarity(emb):
    e = np.asarray(emb, dtype=np.float64)
    nrm = np.sqrt((e * e).sum(1, keepdims=True))
    nrm[nrm == 0.0] = 1.0
    en = e / nrm
    return en @ en.T


def _vqa(p, image_features, q_emb, dtype):
    pt = O.to_torch(p, dtype)
    x = torch.as_tensor(image_features, dtype=dtype)
    q = torch.as_tensor(q_emb, dtype=dtype)
    return pt, O.vqa_forward(pt, x, q)


def blackbox_scores(p, image_features, q_emb, answer_aids, dtype=torch.float64):
    _, (_, _, a_knns, _) = _vqa(p, image_features, q_emb, dtype)
    probs = torch.softmax(a_knns, dim=-1)
    aid = torch.as_tensor(answer_aids).view(-1, 1, 1).expand(-1, O.KNN, 1)
    return (-probs.gather(2, aid).squeeze(2)).numpy()


def semantic_scores(p, image_features, q_emb, answer_aids, emb_pairs, lam, dtype=torch.float64):
    _, (_, _, a_knns, _) = _vqa(p, image_features, q_emb, dtype)
    a_knns = a_knns.numpy()
    out = []
    for knn, aid in zip(a_knns, np.asarray(answer_aids)):
        scores = []
        for nb in knn:
            e = np.exp(nb)
            nb = e / e.sum()
            weighted = emb_pairs[aid, :].dot(nb) - nb[aid]
            logp = np.log(nb[aid] + 1e-8)
            scores.append(lam * weighted - (1.0 - lam) * logp)
        e = np.exp(np.array(scores))
        out.append(e / e.sum())
    return np.array(out)


def linear_context(p, image_features, q_emb, weight, bias, comp_idxs=None, dtype=torch.float64):
    """scores [B, 24]; with comp_idxs also (loss = sum-CE / B, dW, db) as counterexamples.py:325-337 computes them."""
    _, (_, _, _, z_knns) = _vqa(p, image_features, q_emb, dtype)
    B = z_knns.shape[0]
    w = torch.as_tensor(weight, dtype=dtype).clone().requires_grad_(True)
    b = torch.as_tensor(bias, dtype=dtype).clone().requires_grad_(True)
    scores = z_knns.reshape(B, -1) @ w.t() + b
    if comp_idxs is None:
        return scores.detach().numpy()
    loss = torch.nn.functional.cross_entropy(scores, torch.as_tensor(comp_idxs), reduction="sum") / B
    loss.backward()
    return scores.detach().numpy(), float(loss.detach()), w.grad.numpy(), b.grad.numpy()


def make_linear(seed, knn, dim_z):
    rng = np.random.Generator(np.random.Philox(seed))
    bound = 1.0 / np.sqrt(knn * dim_z)
    return (rng.uniform(-bound, bound, size=(knn, knn * dim_z)).astype(np.float32),
            rng.uniform(-bound, bound, size=(knn,)).astype(np.float32))
