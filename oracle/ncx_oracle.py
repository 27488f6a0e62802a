"""ORACLE (test infrastructure, not product code) for stage 2 — NeuralCX candidate scoring.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.

What it restates (torch on the CPU, float32 or float64, plain unfused ops in the reference's order):
  MutanFusion.forward            vqa/models/fusion.py:78-121   (eval mode: every F.dropout is the identity)
  AbstractNoAtt._classif         vqa/models/noatt.py:24-29
  CXModelBase.vqa_forward        vqa/models/cx.py:64-104       (frozen VQA model, q duplicated x25; trainable_vqa:
                                                               train-mode head with explicit dropout masks)
  NeuralModel.forward            vqa/models/cx.py:261-333      (24-way python loop, 10-way concat, MLP)
  CrossEntropyLoss(size_average=False) / len(batch)   counterexamples.py:310,334
  recallAtK                      counterexamples.py:501-506
  torch.optim.Adam(lr)           counterexamples.py:275-276,336-339

Parameters are a flat dict of float arrays under the reference's own state_dict names
(`answer_embedding.weight`, `linear_1.weight`, ..., `vqa_model.fusion.linear_v.weight`, ...); the question
encoder (seq2vec) is outside the hot path, so q_emb [B, dim_q] is an input here.

Pinning status: the reference has no tests or golden vectors for this path. The restatement is
pinned against the reference's OWN code imported from /root/reference in the build container: fixtures under
tests/golden/ncx_*.npz are produced by oracle/gen_golden_ncx.py (reference NeuralModel + MutanNoAtt, two shims
documented there) and tests/test_oracle_ncx.py checks this file against them.
"""
import math

import numpy as np
import torch

KNN = 24
DIM_A = 2400


def fusion_dims(p):
    wv, wq = p["vqa_model.fusion.linear_v.weight"], p["vqa_model.fusion.linear_q.weight"]
    R = 0
    while f"vqa_model.fusion.list_linear_hv.{R}.weight" in p:
        R += 1
    return {"dim_v": wv.shape[1], "dim_hv": wv.shape[0], "dim_q": wq.shape[1], "dim_hq": wq.shape[0],
            "dim_mm": p["vqa_model.fusion.list_linear_hv.0.weight"].shape[0], "R": R,
            "n_ans": p["vqa_model.linear_classif.weight"].shape[0]}


def mutan_fusion(p, v, q, mask_v=None, mask_q=None):
    """fusion.py:78-121 with activation_v = activation_q = tanh and no hv/hq/mm activation (the mutan_noatt yaml).
    mask_v / mask_q: train-mode F.dropout of the two inputs (fusion.py:86,94) as explicit masks already scaled by
    1/(1-p); None = eval mode. dropout_hv = dropout_hq = 0 in every reference option file."""
    if mask_v is not None:
        v = v * mask_v
    if mask_q is not None:
        q = q * mask_q
    x_v = torch.tanh(v @ p["vqa_model.fusion.linear_v.weight"].t() + p["vqa_model.fusion.linear_v.bias"])
    x_q = torch.tanh(q @ p["vqa_model.fusion.linear_q.weight"].t() + p["vqa_model.fusion.linear_q.bias"])
    x_mm = []
    r = 0
    while f"vqa_model.fusion.list_linear_hv.{r}.weight" in p:
        x_hv = x_v @ p[f"vqa_model.fusion.list_linear_hv.{r}.weight"].t() + p[f"vqa_model.fusion.list_linear_hv.{r}.bias"]
        x_hq = x_q @ p[f"vqa_model.fusion.list_linear_hq.{r}.weight"].t() + p[f"vqa_model.fusion.list_linear_hq.{r}.bias"]
        x_mm.append(x_hq * x_hv)
        r += 1
    return torch.stack(x_mm, dim=1).sum(1)


def classif(p, z, mask_c=None):
    """noatt.py:24-29 (no classif activation in the yaml; dropout is the identity in eval, an explicit mask in train)."""
    if mask_c is not None:
        z = z * mask_c
    return z @ p["vqa_model.linear_classif.weight"].t() + p["vqa_model.linear_classif.bias"]


def vqa_forward(p, image_features, q_emb, trainable=False, head_masks=None):
    """cx.py:64-104: all 25 images of every question through the VQA head. Frozen (the default): eval mode, outputs
    detached. trainable (cx.py:73-74): nothing is detached, and head_masks = (mask_v [B*25, dim_v], mask_q [B*25, dim_q],
    mask_c [B*25, dim_mm]) are the train-mode dropouts of the head in the reference's row order b*25 + i."""
    B = image_features.shape[0]
    mv, mq, mc = head_masks if head_masks is not None else (None, None, None)
    with torch.enable_grad() if trainable else torch.no_grad():
        v = image_features.reshape(B * (KNN + 1), -1)
        q_dup = q_emb.reshape(B, 1, -1).expand(B, KNN + 1, -1).reshape(B * (KNN + 1), -1)
        z = mutan_fusion(p, v, q_dup, mv, mq)
        a = classif(p, z, mc)
        a = a.reshape(B, KNN + 1, -1)
        z = z.reshape(B, KNN + 1, -1)
    return a[:, 0], z[:, 0], a[:, 1:], z[:, 1:]


def neural_forward(p, image_features, q_emb, answer_aids, n_layers, drop_masks=None, eps=1e-6, gates=None,
                   lesion=(), v_rank_rand=None, pre_out=None, trainable_vqa=False, head_masks=None):
    """cx.py:261-333. lesion: model_spec flags that are FALSE, out of {"v_mult", "v_dist", "v_rank"} (zeros for the
    first two, cx.py:297,301; v_rank_rand [24, B, 24] = the torch.rand([B, 24]) draw of each loop iteration, cx.py:306;
    the v_emb lesion replaces image_features before this call, cx.py:265-266). drop_masks: optional list (one per hidden layer) of
    [B, 24, dim_h] tensors already scaled by 1/(1-p) (nn.Dropout in training mode); None = eval / p = 0.
    gates: optional list (one per hidden layer) of [B, 24, dim_h] 0/1 tensors that REPLACE the relu decision
    (h = pre * gate instead of relu(pre)): a checker passes the decisions of the implementation under test so that a
    pre-activation within rounding error of zero does not turn into an O(1) gradient difference."""
    B = image_features.shape[0]
    v_orig = image_features[:, 0]
    v_knns = image_features[:, 1:]
    a_orig, z_orig, a_knns, z_knns = vqa_forward(p, image_features, q_emb, trainable_vqa, head_masks)
    E = p["answer_embedding.weight"]
    a_emb_gt = E[answer_aids]
    a_emb_knns = torch.softmax(a_knns, dim=-1) @ E          # bmm with E broadcast, cx.py:281-282
    scores = []
    for i in range(KNN):
        v_other = v_knns[:, i]
        v_mult = v_orig * v_other if "v_mult" not in lesion else torch.zeros_like(v_orig)
        v_dist = torch.norm(v_orig - v_other + eps, p=2, dim=1, keepdim=True)   # F.pairwise_distance, modern eps
        if "v_dist" in lesion:
            v_dist = torch.zeros_like(v_dist)
        if "v_rank" in lesion:
            v_rank = v_rank_rand[i].to(v_orig.dtype)
        else:
            v_rank = torch.zeros((B, KNN), dtype=v_orig.dtype, device=v_orig.device)
            v_rank[:, i] = 1
        inp = torch.cat((v_orig, v_other, v_mult, v_dist, v_rank, q_emb, z_orig, z_knns[:, i], a_emb_gt,
                         a_emb_knns[:, i]), dim=1)
        h = inp @ p["linear_1.weight"].t() + p["linear_1.bias"]
        if pre_out is not None:      # pre_out: dict layer -> list of 24 [B, dim_h] pre-activation tensors (inspection)
            pre_out.setdefault(1, []).append(h.detach())
        h = torch.relu(h) if gates is None else h * gates[0][:, i]
        if drop_masks is not None:
            h = h * drop_masks[0][:, i]
        for layer in range(2, n_layers + 1):
            h = h @ p[f"linear_{layer}.weight"].t() + p[f"linear_{layer}.bias"]
            if pre_out is not None:
                pre_out.setdefault(layer, []).append(h.detach())
            h = torch.relu(h) if gates is None else h * gates[layer - 1][:, i]
            if drop_masks is not None:
                h = h * drop_masks[layer - 1][:, i]
        scores.append(h @ p["out.weight"].t() + p["out.bias"])
    return torch.cat(scores, dim=1)


def ce_sum(scores, comp_idxs):
    """nn.CrossEntropyLoss(size_average=False): summed over the batch (counterexamples.py:310)."""
    return torch.nn.functional.cross_entropy(scores, comp_idxs, reduction="sum")


def recall_at_k(scores, comp_idxs, k):
    """counterexamples.py:501-506 as a rank test: hit iff fewer than k candidates beat the ground truth
    (ties: lower index first — torch.topk's order among exact ties is unspecified in the reference)."""
    s = scores.detach()
    gt = s.gather(1, comp_idxs.view(-1, 1))
    idx = torch.arange(s.shape[1], device=s.device).view(1, -1)
    ahead = (s > gt) | ((s == gt) & (idx < comp_idxs.view(-1, 1)))
    return (ahead.sum(1) < k).to(torch.int64)


TRAINABLE = ("answer_embedding.weight", "linear_1.weight", "linear_1.bias", "linear_2.weight", "linear_2.bias",
             "linear_3.weight", "linear_3.bias", "out.weight", "out.bias")


def to_torch(params, dtype=torch.float32):
    return {k: torch.as_tensor(np.asarray(v)).to(dtype) for k, v in params.items()}


def loss_and_grads(params, image_features, q_emb, answer_aids, comp_idxs, n_layers, drop_masks=None,
                   dtype=torch.float32, gates=None, lesion=(), v_rank_rand=None, trainable_vqa=False, head_masks=None):
    """One training forward + backward as the reference does it (loss = CE_sum / B, counterexamples.py:334-337).
    Returns (scores, loss, {name: grad}) as torch tensors. trainable_vqa: the head's parameters (`vqa_model.*`) get
    gradients too and grads["q_emb"] = dLoss/dq_emb (what the reference hands to seq2vec's backward)."""
    p = to_torch(params, dtype)
    names = list(TRAINABLE) + ([k for k in p if k.startswith("vqa_model.")] if trainable_vqa else [])
    for k in names:
        if k in p:
            p[k].requires_grad_(True)
    x = torch.as_tensor(np.asarray(image_features)).to(dtype)
    q = torch.as_tensor(np.asarray(q_emb)).to(dtype)
    if trainable_vqa:
        q.requires_grad_(True)
        if head_masks is not None:
            head_masks = tuple(None if m is None else torch.as_tensor(np.asarray(m)).to(dtype) for m in head_masks)
    aid = torch.as_tensor(np.asarray(answer_aids)).long()
    gt = torch.as_tensor(np.asarray(comp_idxs)).long()
    masks = None if drop_masks is None else [torch.as_tensor(np.asarray(m)).to(dtype) for m in drop_masks]
    if gates is not None:
        gates = [torch.as_tensor(np.asarray(g)).to(dtype) for g in gates]
    if v_rank_rand is not None:
        v_rank_rand = torch.as_tensor(np.asarray(v_rank_rand)).to(dtype)
    scores = neural_forward(p, x, q, aid, n_layers, masks, gates=gates, lesion=lesion, v_rank_rand=v_rank_rand,
                            trainable_vqa=trainable_vqa, head_masks=head_masks)
    loss = ce_sum(scores, gt) / x.shape[0]
    loss.backward()
    grads = {k: p[k].grad for k in names if k in p}
    if trainable_vqa:
        grads["q_emb"] = q.grad
    return scores.detach(), loss.detach(), grads


def adam_step(param, grad, m, v, step, lr=1e-4, b1=0.9, b2=0.999, eps=1e-8):
    """torch.optim.Adam defaults (counterexamples.py:275-276): returns updated (param, m, v) as float64 numpy."""
    param, grad, m, v = (np.asarray(a, dtype=np.float64) for a in (param, grad, m, v))
    m = b1 * m + (1 - b1) * grad
    v = b2 * v + (1 - b2) * grad * grad
    mhat = m / (1 - b1 ** step)
    vhat = v / (1 - b2 ** step)
    return param - lr * mhat / (np.sqrt(vhat) + eps), m, v


# ------------------------------------------------------------------------------------------------
# seeded synthetic parameters / batches (shared by the golden generator and the tests)
# ------------------------------------------------------------------------------------------------
def make_params(seed, dim_v=2048, dim_q=2400, dim_h_fusion=360, R=10, n_ans=2000, dim_h=300, n_layers=2,
                zero_emb_frac=0.02):
    """Reference-named parameter dict with torch-default-like uniform init (nn.Linear: U(-1/sqrt(fan_in), ..))."""
    rng = np.random.Generator(np.random.Philox(seed))

    def lin(out_f, in_f):
        b = 1.0 / math.sqrt(in_f)
        return (rng.uniform(-b, b, size=(out_f, in_f)).astype(np.float32),
                rng.uniform(-b, b, size=(out_f,)).astype(np.float32))

    p = {}
    p["vqa_model.fusion.linear_v.weight"], p["vqa_model.fusion.linear_v.bias"] = lin(dim_h_fusion, dim_v)
    p["vqa_model.fusion.linear_q.weight"], p["vqa_model.fusion.linear_q.bias"] = lin(dim_h_fusion, dim_q)
    for r in range(R):
        p[f"vqa_model.fusion.list_linear_hv.{r}.weight"], p[f"vqa_model.fusion.list_linear_hv.{r}.bias"] = lin(dim_h_fusion, dim_h_fusion)
        p[f"vqa_model.fusion.list_linear_hq.{r}.weight"], p[f"vqa_model.fusion.list_linear_hq.{r}.bias"] = lin(dim_h_fusion, dim_h_fusion)
    p["vqa_model.linear_classif.weight"], p["vqa_model.linear_classif.bias"] = lin(n_ans, dim_h_fusion)
    emb = rng.normal(0.0, 0.35, size=(n_ans, DIM_A)).astype(np.float32)
    nz = max(0, int(n_ans * zero_emb_frac))
    if nz:
        emb[rng.choice(n_ans, size=nz, replace=False)] = 0.0   # answers with OOV words (answer_embedding.ipynb cell 29)
    p["answer_embedding.weight"] = emb
    in1 = dim_v * 3 + DIM_A * 2 + dim_h_fusion * 2 + dim_q + KNN + 1
    p["linear_1.weight"], p["linear_1.bias"] = lin(dim_h, in1)
    for layer in range(2, n_layers + 1):
        p[f"linear_{layer}.weight"], p[f"linear_{layer}.bias"] = lin(dim_h, dim_h)
    p["out.weight"], p["out.bias"] = lin(1, dim_h)
    return p


def make_batch(seed, B, dim_v=2048, dim_q=2400, n_ans=2000):
    """Seeded [B,25,dim_v] non-negative features (neighbours = noisy copies of the original, sorted by distance
    like real kNN lists), q_emb, answer ids, ground-truth candidate indices."""
    rng = np.random.Generator(np.random.Philox(seed))
    base = np.maximum(rng.normal(0.35, 0.5, size=(B, 1, dim_v)), 0.0)
    scale = np.linspace(0.05, 0.4, KNN).reshape(1, KNN, 1)
    knn = np.maximum(base + scale * rng.normal(size=(B, KNN, dim_v)), 0.0)
    feats = np.concatenate([base, knn], axis=1).astype(np.float32)
    q_emb = np.tanh(rng.normal(0.0, 0.5, size=(B, dim_q))).astype(np.float32)
    aid = rng.integers(0, n_ans, size=B).astype(np.int64)
    comp = rng.integers(0, KNN, size=B).astype(np.int64)
    return feats, q_emb, aid, comp
