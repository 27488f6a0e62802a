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


if __name__ == "__main__":
    main()
