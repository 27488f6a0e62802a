"""The frozen VQA model NeuralCX scores with: parameter container mirroring vqa/models/noatt.py MutanNoAtt
(noatt.py:49-58) + fusion.MutanFusion (fusion.py:53-76), under the reference's own parameter names so that a
reference checkpoint (train.py:290-330 `best_model.pth.tar`) loads with load_state_dict.

The model is frozen on the CX path (cx.py:76-80), so it is never run through torch here: NeuralModel hands its
weights to the CUDA library, which evaluates `_fusion` and `_classif` for all 25 images of a question
(csrc/abi_ncx.cu). The question encoder (`seq2vec`, skip-thoughts GRU) is outside the hot path: with the model
frozen and in eval mode its output is a pure function of the question, consumed here as a precomputed table.
"""
import torch
import torch.nn as nn


class QuestionTable(nn.Module):
    """seq2vec stand-in: q_emb = table[question id]. The id travels in column 0 of `question_wids` (the remaining
    25 token slots are ignored), or call with a 1-D tensor of ids."""

    def __init__(self, table):
        super().__init__()
        self.register_buffer("table", torch.as_tensor(table, dtype=torch.float32), persistent=False)

    def forward(self, question_wids):
        ids = question_wids if question_wids.dim() == 1 else question_wids[:, 0]
        return self.table[ids.to(self.table.device)]


class MutanFusion(nn.Module):
    def __init__(self, opt):
        super().__init__()
        self.opt = opt
        self.linear_v = nn.Linear(opt["dim_v"], opt["dim_hv"])
        self.linear_q = nn.Linear(opt["dim_q"], opt["dim_hq"])
        self.list_linear_hv = nn.ModuleList([nn.Linear(opt["dim_hv"], opt["dim_mm"]) for _ in range(opt["R"])])
        self.list_linear_hq = nn.ModuleList([nn.Linear(opt["dim_hq"], opt["dim_mm"]) for _ in range(opt["R"])])


class MutanNoAtt(nn.Module):
    """`models.factory(options['model'], vocab_words, vocab_answers, ...)` result for arch MutanNoAtt
    (vqa/models/utils.py:14-30), as a weight container. `seq2vec` may be any module mapping question_wids
    [B, 26] -> [B, dim_q]."""

    def __init__(self, opt, vocab_words=(), vocab_answers=(), seq2vec=None):
        super().__init__()
        opt = dict(opt)
        opt["fusion"] = dict(opt["fusion"])
        opt["fusion"]["dim_h"] = opt["fusion"]["dim_mm"]                       # noatt.py:52
        for key, want in (("activation_v", "tanh"), ("activation_q", "tanh")):
            if opt["fusion"].get(key, want) != want:
                raise NotImplementedError(f"fusion.{key}={opt['fusion'][key]!r}: the CUDA head implements tanh")
        for key in ("activation_hv", "activation_hq", "activation_mm"):
            if key in opt["fusion"]:
                raise NotImplementedError(f"fusion.{key} is not used by the mutan_noatt options and is not implemented")
        if "activation" in opt.get("classif", {}):
            raise NotImplementedError("classif.activation is not used by the mutan_noatt options")
        self.opt = opt
        self.vocab_words = list(vocab_words)
        self.vocab_answers = list(vocab_answers)
        self.num_classes = len(self.vocab_answers)
        self.seq2vec = seq2vec if seq2vec is not None else nn.Identity()
        self.linear_classif = nn.Linear(opt["fusion"]["dim_h"], self.num_classes)
        self.fusion = MutanFusion(opt["fusion"])


def factory(opt, vocab_words, vocab_answers, cuda=True, data_parallel=True, seq2vec=None):
    """vqa/models/utils.py:14-30 for the one architecture on the CX path."""
    if opt["arch"] != "MutanNoAtt":
        raise NotImplementedError(f"arch {opt['arch']!r}: only MutanNoAtt is on the NeuralCX hot path")
    model = MutanNoAtt(opt, vocab_words, vocab_answers, seq2vec=seq2vec)
    return model.cuda() if cuda and torch.cuda.is_available() else model
