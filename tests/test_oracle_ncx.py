"""CPU: pin the NeuralCX oracle (oracle/ncx_oracle.py) against golden vectors produced by the reference's own
NeuralModel / MutanNoAtt code (oracle/gen_golden_ncx.py), and check the closed-form anchors of SURVEY.md §4."""
import glob
import math
import os

import numpy as np
import pytest
import torch

from oracle import ncx_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def load_case(path):
    z = np.load(path)
    dv, dq, df, R, na, dim_h, n_layers, B, seed = [int(v) for v in z["meta"]]
    params = O.make_params(seed, dim_v=dv, dim_q=dq, dim_h_fusion=df, R=R, n_ans=na, dim_h=dim_h, n_layers=n_layers)
    batch = O.make_batch(seed + 1000, B, dim_v=dv, dim_q=dq, n_ans=na)
    return z, params, batch, n_layers


def check_grads(z, grads, rtol, what):
    for key in z.files:
        if not key.startswith("gnorm:"):
            continue
        name = key[6:]
        g = np.asarray(grads[name], dtype=np.float64)
        gn = float(z[key])
        assert abs(np.sqrt((g ** 2).sum()) - gn) <= rtol * gn + 2e-6, (what, name, "norm")
        if "grad:" + name in z.files:
            ref = z["grad:" + name].astype(np.float64)
            assert np.abs(g - ref).max() <= rtol * np.abs(ref).max() + 2e-6, (what, name, "full")
        else:
            idx = z["gidx:" + name]
            ref = z["gval:" + name].astype(np.float64)
            assert np.abs(g.reshape(-1)[idx] - ref).max() <= rtol * np.abs(ref).max() + 2e-6, (what, name, "sampled")


FROZEN_CASES = sorted(p for p in glob.glob(os.path.join(GOLD, "ncx_*.npz")) if "ncx_tv_" not in p)
TV_CASES = sorted(glob.glob(os.path.join(GOLD, "ncx_tv_*.npz")))


@pytest.mark.parametrize("path", FROZEN_CASES)
def test_restatement_matches_reference_golden(path):
    z, params, (feats, q_emb, aid, comp), n_layers = load_case(path)
    # lesion fixtures: the flags that were false in the reference run; a lone q_emb lesion changes nothing there
    lesion = tuple(str(x) for x in z["lesion"]) if "lesion" in z.files else ()
    vr = z["v_rank_train"] if "v_rank" in lesion else None
    scores, loss, grads = O.loss_and_grads(params, feats, q_emb, aid, comp, n_layers, lesion=lesion, v_rank_rand=vr)
    # same fp32 arithmetic, different op order only inside BLAS: tight tolerances
    assert np.abs(scores.numpy() - z["scores"]).max() <= 2e-5
    if "v_rank" in lesion:      # the eval pass drew fresh random rank rows
        s_eval, _, _ = O.loss_and_grads(params, feats, q_emb, aid, comp, n_layers, lesion=lesion,
                                        v_rank_rand=z["v_rank_eval"])
        assert np.abs(s_eval.numpy() - z["scores_eval"]).max() <= 2e-5
    else:
        assert np.abs(scores.numpy() - z["scores_eval"]).max() <= 2e-5      # drop_p = 0: train == eval
    assert abs(float(loss) - float(z["loss"])) <= 1e-5
    check_grads(z, {k: v.numpy() for k, v in grads.items()}, 2e-3, os.path.basename(path))
    gt = torch.as_tensor(comp)
    assert np.array_equal(O.recall_at_k(torch.as_tensor(z["scores_eval"]), gt, 5).numpy(), z["recall5"])
    assert np.array_equal(O.recall_at_k(torch.as_tensor(z["scores_eval"]), gt, 1).numpy(), z["recall1"])


def tv_head_masks(z):
    """(mask_v, mask_q, mask_c) of a trainable_vqa fixture in the reference's row order b*25 + i, or None (p = 0)."""
    return (z["mask_v"], z["mask_q"], z["mask_c"]) if "mask_v" in z.files else None


@pytest.mark.parametrize("path", TV_CASES)
def test_trainable_vqa_restatement_matches_reference_golden(path):
    """cx.py:73-74: the reference run with trainable_vqa=True, module in train mode, its three head dropouts recorded;
    gradients of every NeuralCX AND every VQA-head parameter, and dLoss/dq_emb."""
    assert len(TV_CASES) >= 3
    z, params, (feats, q_emb, aid, comp), n_layers = load_case(path)
    scores, loss, grads = O.loss_and_grads(params, feats, q_emb, aid, comp, n_layers, trainable_vqa=True,
                                           head_masks=tv_head_masks(z))
    assert np.abs(scores.numpy() - z["scores"]).max() <= 2e-5
    assert abs(float(loss) - float(z["loss"])) <= 1e-5
    names = [k[5:] for k in z.files if k.startswith("grad:")]
    assert any(n.startswith("vqa_model.fusion.list_linear_hq") for n in names) and "q_emb" in names
    for name in names:
        ref = z["grad:" + name].astype(np.float64)
        g = grads[name].numpy().astype(np.float64)
        assert g.shape == ref.shape, name
        assert np.abs(g - ref).max() <= 2e-3 * np.abs(ref).max() + 2e-7, (name, np.abs(g - ref).max(), np.abs(ref).max())


def test_float64_restatement_close_to_float32():
    z, params, (feats, q_emb, aid, comp), n_layers = load_case(os.path.join(GOLD, "ncx_small_l2.npz"))
    s64, l64, _ = O.loss_and_grads(params, feats, q_emb, aid, comp, n_layers, dtype=torch.float64)
    assert np.abs(s64.numpy() - z["scores"]).max() <= 2e-5
    assert abs(float(l64) - float(z["loss"])) <= 1e-5


def test_closed_form_anchors():
    # CE over 24 equal scores is ln 24; DistanceBaseline scores [23..0] (cx.py:40-44) give Recall@k = P(idx < k)
    s = torch.zeros((7, 24))
    gt = torch.arange(7)
    assert abs(float(O.ce_sum(s, gt)) / 7 - math.log(24)) < 1e-6
    dist_scores = torch.arange(23, -1, -1, dtype=torch.float32).repeat(24, 1)
    gt = torch.arange(24)
    assert O.recall_at_k(dist_scores, gt, 5).tolist() == [1] * 5 + [0] * 19
    assert O.recall_at_k(dist_scores, gt, 1).tolist() == [1] + [0] * 23
    # ties: all-equal scores -> only ground truth 0..k-1 count as hits (lower index first)
    assert O.recall_at_k(torch.zeros((24, 24)), gt, 5).sum().item() == 5


def test_adam_matches_torch():
    rng = np.random.default_rng(0)
    w = rng.normal(size=(5, 7)); g1 = rng.normal(size=(5, 7)); g2 = rng.normal(size=(5, 7))
    p = torch.nn.Parameter(torch.tensor(w))
    opt = torch.optim.Adam([p], lr=1e-4)
    m = np.zeros_like(w); v = np.zeros_like(w); cur = w
    for step, g in enumerate((g1, g2), start=1):
        p.grad = torch.tensor(g)
        opt.step()
        cur, m, v = O.adam_step(cur, g, m, v, step)
        assert np.abs(cur - p.detach().numpy()).max() < 1e-12
