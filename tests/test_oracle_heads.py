"""CPU: pin the oracle of the other CX heads (oracle/heads_oracle.py) against golden vectors produced by the
reference's own BlackBox / LinearContext / SemanticBaseline classes (oracle/gen_golden_heads.py)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import heads_oracle as H
from oracle import ncx_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def load_case(path):
    z = np.load(path)
    dv, dq, df, R, na, B, seed = [int(v) for v in z["meta"]]
    params = O.make_params(seed, dim_v=dv, dim_q=dq, dim_h_fusion=df, R=R, n_ans=na, dim_h=12, n_layers=1)
    batch = O.make_batch(seed + 1000, B, dim_v=dv, dim_q=dq, n_ans=na)
    w, b = H.make_linear(seed + 9, 24, df)
    return z, params, batch, (w, b), float(z["lam"])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "heads_*.npz"))))
def test_restatement_matches_reference_golden(path):
    z, params, (feats, q_emb, aid, comp), (w, b), lam = load_case(path)
    # the reference ran in float32; the float64 restatement differs by float32 rounding only
    bb = H.blackbox_scores(params, feats, q_emb, aid)
    assert np.abs(bb - z["blackbox"]).max() <= 1e-6 + 1e-5 * np.abs(z["blackbox"]).max()
    sim = H.cosine_similarity(params["answer_embedding.weight"])
    ij = z["cos_ij"]
    assert np.abs(sim[ij[:, 0], ij[:, 1]] - z["cos_val"]).max() <= 1e-6
    sb = H.semantic_scores(params, feats, q_emb, aid, sim, lam)
    assert np.abs(sb - z["semantic"]).max() <= 2e-7
    assert np.abs(sb.sum(1) - 1.0).max() <= 1e-12
    scores, loss, gw, gb = H.linear_context(params, feats, q_emb, w, b, comp)
    assert np.abs(scores - z["linear_scores"]).max() <= 2e-5
    assert abs(loss - float(z["linear_loss"])) <= 1e-5
    assert np.abs(gw - z["linear_gw"]).max() <= 1e-5 * max(1.0, np.abs(z["linear_gw"]).max()) + 1e-6
    assert np.abs(gb - z["linear_gb"]).max() <= 1e-6


def test_closed_forms():
    # identical neighbours -> SemanticBaseline returns the uniform distribution; BlackBox rows are constant
    params = O.make_params(7, dim_v=64, dim_q=48, dim_h_fusion=40, R=3, n_ans=37, dim_h=12, n_layers=1)
    feats, q_emb, aid, _ = O.make_batch(8, 3, dim_v=64, dim_q=48, n_ans=37)
    feats[:, 1:] = feats[:, 1:2]
    sim = H.cosine_similarity(params["answer_embedding.weight"])
    assert np.allclose(np.diag(sim)[np.abs(params["answer_embedding.weight"]).sum(1) > 0], 1.0)
    sb = H.semantic_scores(params, feats, q_emb, aid, sim, 0.5)
    assert np.abs(sb - 1.0 / 24).max() <= 1e-12
    bb = H.blackbox_scores(params, feats, q_emb, aid)
    assert np.abs(bb - bb[:, :1]).max() <= 1e-12 and (bb < 0).all()
