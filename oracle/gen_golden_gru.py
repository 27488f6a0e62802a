"""Generates tests/golden/gru_uniskip.npz by running the REFERENCE's own BayesianUniSkip
(vqa/external/skip-thoughts.torch/pytorch/skipthoughts.py, imported read-only from /root/reference) in eval mode on
seeded synthetic word ids. Its three loader hooks (dictionary.txt / utable.npy / uni_skip.npz downloads,
skipthoughts.py:44-70) are overridden to return oracle.gru_oracle's seeded arrays; nothing else is touched.
Run in the build container:  python oracle/gen_golden_gru.py        (weights are regenerated from the seed, not stored)"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/vqa/external/skip-thoughts.torch/pytorch"
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)
sys.dont_write_bytecode = True

from oracle import gru_oracle as G  # noqa: E402


def main():
    import skipthoughts as st

    seed, B, T = 21, 12, 26
    params = G.make_gru_params(seed)
    dico, vocab = G.make_vocab()

    class Synth(st.BayesianUniSkip):
        def _load_dictionary(self):
            return dico

        def _load_emb_params(self):
            return params["utable"]

        def _load_rnn_params(self):
            return params

    model = Synth("/nonexistent", vocab, dropout=0.25)
    model.eval()
    wids = G.make_wids(seed + 1, B, T, len(vocab))
    wids[3] = 0
    wids[3, 0] = 5                                               # a one-word question
    with torch.no_grad():
        ref = model(torch.from_numpy(wids)).numpy()
    sd = G.state_dict_from_params(params, dico, vocab)
    for k, v in model.state_dict().items():                      # the restated state-dict construction is the reference's
        assert np.array_equal(sd[k], v.numpy()), k
    mine = G.encode(sd, wids)
    print("restatement vs reference max abs err:", np.abs(mine - ref).max())
    out = os.path.join(ROOT, "tests", "golden", "gru_uniskip.npz")
    np.savez_compressed(out, meta=np.array([seed, B, T]), wids=wids, q_emb=ref.astype(np.float32))
    print("wrote", out, os.path.getsize(out), "bytes")


def main_train():
    """TRAIN mode (trainable_vqa: the reference back-propagates the NeuralCX loss into seq2vec): the same class with its
    six SequentialDropout masks active (gru.py:49-84; one Bernoulli(1-p)/(1-p) mask per gate input and per sequence,
    dropout.py:44-66), forward + backward of loss = sum(q_emb * G). The masks are read back from the modules after the
    forward; gradients of the big tensors are stored as 2048 sampled entries + norm."""
    import skipthoughts as st

    seed, B, T = 31, 6, 26
    params = G.make_gru_params(seed)
    dico, vocab = G.make_vocab()

    class Synth(st.BayesianUniSkip):
        def _load_dictionary(self):
            return dico

        def _load_emb_params(self):
            return params["utable"]

        def _load_rnn_params(self):
            return params

    model = Synth("/nonexistent", vocab, dropout=0.25)
    model.train()
    wids = G.make_wids(seed + 1, B, T, len(vocab))
    wids[2] = 0
    wids[2, 0] = 7                                               # a one-word question
    torch.manual_seed(seed)
    q = model(torch.from_numpy(wids))
    cell = model.rnn.gru_cell
    masks = {k: getattr(cell, "drop_" + k).noise.detach().numpy().astype(np.float32).copy()
             for k in ("ir", "ii", "in", "hr", "hi", "hn")}
    gq = np.random.default_rng(seed + 2).standard_normal(q.shape).astype(np.float32)
    model.zero_grad()
    (q * torch.from_numpy(gq)).sum().backward()
    save = {"meta": np.array([seed, B, T]), "wids": wids, "q_emb": q.detach().numpy().astype(np.float32), "gq": gq,
            "p_drop": np.float64(0.25)}
    for k, m in masks.items():
        save["mask_" + k] = m
    rng = np.random.Generator(np.random.Philox(seed + 7))
    for k, prm in model.named_parameters():
        g = prm.grad.detach().numpy()
        save["gnorm:" + k] = np.float64(np.sqrt((g.astype(np.float64) ** 2).sum()))
        if g.size <= 8192:
            save["grad:" + k] = g
        else:
            idx = rng.choice(g.size, size=2048, replace=False)
            save["gidx:" + k] = idx.astype(np.int64)
            save["gval:" + k] = g.reshape(-1)[idx]
    sd = G.state_dict_from_params(params, dico, vocab)
    q_mine, g_mine = G.encode_train(sd, wids, masks, gq)
    print("train restatement vs reference: q_emb max abs err", np.abs(q_mine - save["q_emb"]).max(), "grad norms",
          {k[6:]: round(float(v), 4) for k, v in save.items() if k.startswith("gnorm:")})
    out = os.path.join(ROOT, "tests", "golden", "gru_uniskip_train.npz")
    np.savez_compressed(out, **save)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    if "train" in sys.argv[1:]:
        main_train()
    else:
        main()
