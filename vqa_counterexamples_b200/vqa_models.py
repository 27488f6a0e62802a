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


class BayesianUniSkip(nn.Module):
    """seq2vec = skip-thoughts `BayesianUniSkip` in eval mode (skipthoughts.py:265-300) as a weight container under the
    reference's parameter names (`embedding.weight`, `rnn.gru_cell.weight_{ir,ii,in,hr,hi,hn}.*`: a reference checkpoint's
    `seq2vec.*` entries load with load_state_dict); forward(question_wids [B, 26]) -> q_emb [B, 2400] runs the CUDA
    encoder (csrc/gru.cu: one input GEMM for all steps, one recurrent GEMM + gate kernel per step). Frozen and in eval
    mode on the CX path (cx.py:76-80): dropout is the identity and there is no backward."""

    def __init__(self, vocab_size, dim_emb=620, dim_h=2400, chunk=1024):
        super().__init__()
        self.dim_emb, self.dim_h, self.chunk = dim_emb, dim_h, chunk
        self.embedding = nn.Embedding(vocab_size + 1, dim_emb, padding_idx=0)     # skipthoughts.py:79-82
        cell = nn.Module()
        for name in ("ir", "ii", "in"):
            setattr(cell, "weight_" + name, nn.Linear(dim_emb, dim_h, bias=True))   # gru.py:18-20
        for name in ("hr", "hi", "hn"):
            setattr(cell, "weight_" + name, nn.Linear(dim_h, dim_h, bias=False))    # gru.py:21-23
        self.rnn = nn.Module()
        self.rnn.gru_cell = cell
        for p in self.parameters():
            p.requires_grad = False
        self._packed = None

    def load_state_dict(self, state_dict, strict=True):
        out = super().load_state_dict(state_dict, strict=strict)
        self._packed = None
        return out

    def _pack(self, dev):
        c = self.rnn.gru_cell
        f = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()
        self._packed = (f(self.embedding.weight),
                        f(torch.cat([c.weight_ir.weight, c.weight_ii.weight, c.weight_in.weight], 0)),
                        f(torch.cat([c.weight_ir.bias, c.weight_ii.bias, c.weight_in.bias], 0)),
                        f(torch.cat([c.weight_hr.weight, c.weight_hi.weight, c.weight_hn.weight], 0)))
        return self._packed

    @torch.no_grad()
    def forward(self, question_wids, lengths=None):
        import ctypes  # noqa: F401
        from ._lib import lib, check, ptr, current_stream

        if not torch.cuda.is_available():
            raise RuntimeError("BayesianUniSkip needs a CUDA device: there is no CPU fallback for this path")
        dev = question_wids.device if question_wids.is_cuda else torch.device("cuda", torch.cuda.current_device())
        emb, w_i, b_i, w_h = self._packed if self._packed is not None and self._packed[0].device == dev else self._pack(dev)
        wids = question_wids.to(dev, torch.int64).contiguous()
        B, T = wids.shape
        out = torch.empty((B, self.dim_h), dtype=torch.float32, device=dev)
        for b0 in range(0, B, self.chunk):
            w = wids[b0:b0 + self.chunk].contiguous()
            n = w.shape[0]
            need = lib.vqacx_gru_workspace_bytes(n, T, self.dim_emb, self.dim_h)
            ws = torch.empty(need, dtype=torch.uint8, device=dev)
            o = out[b0:b0 + n]
            check(lib.vqacx_gru_encode(ptr(emb), emb.shape[0], self.dim_emb, ptr(w), n, T, ptr(w_i), ptr(b_i), ptr(w_h),
                                       self.dim_h, 0, ptr(o), ptr(ws), ws.numel(), current_stream()), "vqacx_gru_encode")
        return out


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
