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


GOLD_TRAIN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gru_uniskip_train.npz")
GRU_PARAMS = ["embedding.weight"] + ["rnn.gru_cell.weight_%s.%s" % (g, k) for g, ks in
                                     (("ir", ("weight", "bias")), ("ii", ("weight", "bias")), ("in", ("weight", "bias")),
                                      ("hr", ("weight",)), ("hi", ("weight",)), ("hn", ("weight",))) for k in ks]


def _train_case():
    z = np.load(GOLD_TRAIN)
    seed, B, T = [int(v) for v in z["meta"]]
    params = G.make_gru_params(seed)
    dico, vocab = G.make_vocab()
    masks = {k: z["mask_" + k] for k in ("ir", "ii", "in", "hr", "hi", "hn")}
    return z, G.state_dict_from_params(params, dico, vocab), vocab, masks


def check_gru_grads(z, grads, rtol):
    for name in GRU_PARAMS:
        g = np.asarray(grads[name], dtype=np.float64)
        gn = float(z["gnorm:" + name])
        assert abs(np.sqrt((g ** 2).sum()) - gn) <= rtol * gn + 1e-6, (name, "norm", np.sqrt((g ** 2).sum()), gn)
        if "grad:" + name in z.files:
            ref, got = z["grad:" + name].astype(np.float64), g
        else:
            ref, got = z["gval:" + name].astype(np.float64), g.reshape(-1)[z["gidx:" + name]]
        scale = max(np.abs(ref).max(), gn / np.sqrt(g.size))
        assert np.abs(got - ref).max() <= rtol * scale + 1e-6, (name, np.abs(got - ref).max(), scale)


def test_train_mode_oracle_matches_reference_golden():
    """BayesianGRU in TRAIN mode (six per-sequence dropout masks, read back from the reference's modules) + backward of
    sum(q_emb * gq) through all 26 steps: the restatement vs the reference's own autograd."""
    z, sd, _, masks = _train_case()
    q, grads = G.encode_train(sd, z["wids"], masks, z["gq"])
    assert np.abs(q - z["q_emb"]).max() <= 2e-6
    assert set(GRU_PARAMS) <= set(grads)
    assert np.abs(grads["embedding.weight"][0]).max() == 0           # padding_idx = 0 (skipthoughts.py:79-82)
    check_gru_grads(z, grads, 1e-4)
    assert float(z["mask_hr"].max()) == pytest.approx(1 / 0.75) and float(z["mask_hr"].min()) == 0.0


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


def _named_grads(enc, flat_grad):
    emb, w_i, b_i, w_h = enc._flat_views(flat_grad)
    H = enc.dim_h
    out = {"embedding.weight": emb}
    for g, name in enumerate(("ir", "ii", "in")):
        out[f"rnn.gru_cell.weight_{name}.weight"] = w_i[g * H:(g + 1) * H]
        out[f"rnn.gru_cell.weight_{name}.bias"] = b_i[g * H:(g + 1) * H]
    for g, name in enumerate(("hr", "hi", "hn")):
        out[f"rnn.gru_cell.weight_{name}.weight"] = w_h[g * H:(g + 1) * H]
    return {k: v.detach().cpu().numpy() for k, v in out.items()}


@pytest.mark.gpu
def test_cuda_train_mode_encoder_and_bptt_match_reference_golden():
    """vqacx_gru_forward_train / vqacx_gru_backward through the autograd bridge, with the reference's six dropout masks
    replayed: q_emb and the gradient of every encoder tensor vs the reference's own autograd (TF32 products through 26
    recurrent steps forward and back: 3e-2 of each tensor's scale, norms within 2e-2)."""
    import torch
    from vqa_counterexamples_b200.vqa_models import BayesianUniSkip

    z, sd, vocab, masks = _train_case()
    enc = BayesianUniSkip(len(vocab))
    enc.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    enc.make_trainable(float(z["p_drop"]))
    enc.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})     # reference names still load (into the views)
    assert "flat" not in enc.state_dict() and set(enc.state_dict()) == set(sd)
    enc.train()
    enc.train_masks = (torch.from_numpy(np.stack([masks["ir"], masks["ii"], masks["in"]])),
                       torch.from_numpy(np.stack([masks["hr"], masks["hi"], masks["hn"]])))
    wids = torch.from_numpy(z["wids"]).cuda()
    q = enc(wids)
    assert q.requires_grad
    assert np.abs(q.detach().cpu().numpy() - z["q_emb"]).max() <= 5e-3
    (q * torch.from_numpy(z["gq"]).cuda()).sum().backward()
    grads = _named_grads(enc, enc.flat.grad)
    assert np.abs(grads["embedding.weight"][0]).max() == 0
    check_gru_grads(z, grads, 3e-2)
    # built-in masks: two train-mode passes differ, eval mode is the deterministic eval encoder
    enc.train_masks = None
    with torch.no_grad():
        a, b = enc(wids).clone(), enc(wids).clone()
    assert float((a - b).abs().max()) > 1e-3
    enc.eval()
    e1 = enc(wids)
    assert not e1.requires_grad and torch.equal(e1, enc(wids))
    assert np.abs(e1.cpu().numpy() - G.encode(sd, z["wids"])).max() <= 5e-3
