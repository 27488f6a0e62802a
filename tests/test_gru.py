"""Question encoder: the oracle restatement against the golden vector produced by the
reference's own BayesianUniSkip (CPU), and the CUDA encoder against both (GPU)."""
import os

import numpy as np
import pytest

from oracle import gru_oracle as G

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gru_uniskip.npz")


def _case():
    z = np.load(GOLD)
    seed, B, T = [int(v) for v in z["meta"]]
    params = G.make_gru_params(seed)
    dico, vocab = G.make_vocab()
    return z, G.state_dict_from_params(params, dico, vocab), vocab


def test_oracle_matches_reference_golden():
    z, sd, _ = _case()
    got = G.encode(sd, z["wids"])
    assert np.abs(got - z["q_emb"]).max() <= 2e-6            # fp64 restatement vs the reference's fp32 run
    lengths = (z["wids"] != 0).sum(1)
    assert lengths.min() == 1 and lengths.max() == z["wids"].shape[1]


@pytest.mark.gpu
def test_cuda_encoder_matches_reference_golden():
    import torch
    from vqa_counterexamples_b200.vqa_models import BayesianUniSkip

    z, sd, vocab = _case()
    enc = BayesianUniSkip(len(vocab))
    enc.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    wids = torch.from_numpy(z["wids"]).cuda()
    got = enc(wids).cpu().numpy()
    # TF32 products (10-bit operands) through <= 26 recurrent steps; outputs lie in (-1, 1)
    assert np.abs(got - z["q_emb"]).max() <= 5e-3, np.abs(got - z["q_emb"]).max()
    assert np.abs(got - z["q_emb"]).mean() <= 5e-4
    # chunking and batch composition must not matter: every row depends on its own words only
    enc.chunk = 5
    got2 = enc(wids).cpu().numpy()
    assert np.abs(got2 - got).max() <= 5e-3
    # an all-padding row takes the reference's index -1 path (state after every step on zero inputs)
    w0 = torch.zeros((3, 26), dtype=torch.long, device="cuda")
    out0 = enc(w0).cpu().numpy()
    assert np.isfinite(out0).all() and np.abs(out0[0] - out0[1]).max() == 0
