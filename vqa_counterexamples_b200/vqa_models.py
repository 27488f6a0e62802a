"""The VQA model NeuralCX scores with (frozen by default; trained along with NeuralCX under trainable_vqa=True, see
cx._TrainableHead and BayesianUniSkip.make_trainable): parameter container mirroring vqa/models/noatt.py MutanNoAtt
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


class _GruTrainFunction(torch.autograd.Function):
    """Autograd bridge of the train-mode encoder: forward = vqacx_gru_forward_train, backward = vqacx_gru_backward (BPTT)
    into the gradient of the module's flat parameter."""

    @staticmethod
    def forward(ctx, flat, module, wids):
        out, state = module._forward_train_raw(wids)
        ctx.module, ctx.state = module, state
        return out

    @staticmethod
    def backward(ctx, dq):
        return ctx.module._backward_raw(ctx.state, dq.contiguous()), None, None


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
        self.trainable = False       # make_trainable(): train-mode forward (Bayesian dropout) + BPTT on the device
        self.dropout = 0.0
        self.train_masks = None      # tests: (masks_x [3, B, 620], masks_h [3, B, 2400]) of 0 or 1/(1-p), gates r, i, n
        self.drop_seed = 0x6A09E667
        self._train_calls = 0

    def load_state_dict(self, state_dict, strict=True):
        out = super().load_state_dict(state_dict, strict=strict)
        self._packed = None
        return out

    # the flat training buffer is an implementation detail: checkpoints keep the reference's parameter names only
    def state_dict(self, *args, destination=None, prefix="", keep_vars=False):
        sd = super().state_dict(*args, destination=destination, prefix=prefix, keep_vars=keep_vars)
        sd.pop(prefix + "flat", None)
        return sd

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        if self.trainable and prefix + "flat" not in state_dict:
            state_dict[prefix + "flat"] = self.flat.data       # copied onto itself; the named views are loaded after it
        self._packed = None
        return super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)

    def make_trainable(self, dropout=0.25, device=None):
        """trainable_vqa (cx.py:73-74): the encoder trains with the rest of the VQA model. Moves the weights into ONE flat
        float32 parameter [embedding | W_ir;W_ii;W_in | b_ir;b_ii;b_in | W_hr;W_hi;W_hn] (the stacked layouts the CUDA
        encoder reads) of which the reference-named parameters become views; in train mode forward() then applies the six
        per-sequence dropout masks of BayesianGRUCell (gru.py:49-84) and back-propagates through time."""
        dev = device or torch.device("cuda", torch.cuda.current_device())
        c, E, H = self.rnn.gru_cell, self.dim_emb, self.dim_h
        V = self.embedding.weight.shape[0]
        r32 = lambda n: (n + 31) // 32 * 32
        self._off = {"emb": 0}
        self._off["w_i"] = r32(V * E)
        self._off["b_i"] = self._off["w_i"] + r32(3 * H * E)
        self._off["w_h"] = self._off["b_i"] + r32(3 * H)
        total = self._off["w_h"] + r32(3 * H * H)
        flat = torch.zeros(total, dtype=torch.float32, device=dev)

        def bind(param, off, shape):
            n = 1
            for d in shape:
                n *= d
            view = flat[off:off + n].view(*shape)
            view.copy_(param.detach().to(device=dev, dtype=torch.float32))
            param.data = view
            param.requires_grad = False          # trained through `flat`

        bind(self.embedding.weight, 0, (V, E))
        for g, name in enumerate(("ir", "ii", "in")):
            lin = getattr(c, "weight_" + name)
            bind(lin.weight, self._off["w_i"] + g * H * E, (H, E))
            bind(lin.bias, self._off["b_i"] + g * H, (H,))
        for g, name in enumerate(("hr", "hi", "hn")):
            bind(getattr(c, "weight_" + name).weight, self._off["w_h"] + g * H * H, (H, H))
        self.flat = nn.Parameter(flat)
        self.trainable, self.dropout, self._packed = True, float(dropout), None
        return self

    def _flat_views(self, flat=None):
        flat = self.flat.data if flat is None else flat
        V, E, H, o = self.embedding.weight.shape[0], self.dim_emb, self.dim_h, self._off
        return (flat[:V * E].view(V, E), flat[o["w_i"]:o["w_i"] + 3 * H * E].view(3 * H, E),
                flat[o["b_i"]:o["b_i"] + 3 * H], flat[o["w_h"]:o["w_h"] + 3 * H * H].view(3 * H, H))

    def _forward_train_raw(self, question_wids):
        from ._lib import lib, check, ptr, current_stream

        dev = self.flat.device
        emb, w_i, b_i, w_h = self._flat_views()
        wids = question_wids.to(dev, torch.int64).contiguous()
        B, T = wids.shape
        ws = torch.empty(lib.vqacx_gru_train_workspace_bytes(B, T, self.dim_emb, self.dim_h), dtype=torch.uint8, device=dev)
        out = torch.empty((B, self.dim_h), dtype=torch.float32, device=dev)
        mx, mh = self.train_masks if self.train_masks is not None else (None, None)
        mx = None if mx is None else mx.to(dev, torch.float32).contiguous()
        mh = None if mh is None else mh.to(dev, torch.float32).contiguous()
        seed = (self.drop_seed + 8 * self._train_calls) & 0xFFFFFFFFFFFFFFFF
        self._train_calls += 1
        check(lib.vqacx_gru_forward_train(ptr(emb), emb.shape[0], self.dim_emb, ptr(wids), B, T, ptr(w_i), ptr(b_i), ptr(w_h),
                                          self.dim_h, float(self.dropout), seed, ptr(mx), ptr(mh), ptr(out), ptr(ws),
                                          ws.numel(), current_stream()), "vqacx_gru_forward_train")
        return out, (wids, ws, mx, mh)

    def _backward_raw(self, state, dq):
        """dLoss/dq_emb [B, 2400] -> gradient of the flat parameter (same layout)."""
        from ._lib import lib, check, ptr, current_stream

        wids, ws = state[0], state[1]
        emb, w_i, b_i, w_h = self._flat_views()
        grads = torch.zeros_like(self.flat.data)
        g_emb, g_wi, g_bi, g_wh = self._flat_views(grads)
        B, T = wids.shape
        dq = dq.to(self.flat.device, torch.float32)
        check(lib.vqacx_gru_backward(emb.shape[0], self.dim_emb, ptr(wids), B, T, ptr(w_i), ptr(w_h), self.dim_h, ptr(dq),
                                     dq.stride(0), ptr(ws), ws.numel(), ptr(g_emb), ptr(g_wi), ptr(g_bi), ptr(g_wh),
                                     current_stream()), "vqacx_gru_backward")
        return grads

    def _pack(self, dev):
        if self.trainable:       # the stacked layouts ARE the storage: no copies, always current
            self._packed = self._flat_views()
            return self._packed
        c = self.rnn.gru_cell
        f = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()
        self._packed = (f(self.embedding.weight),
                        f(torch.cat([c.weight_ir.weight, c.weight_ii.weight, c.weight_in.weight], 0)),
                        f(torch.cat([c.weight_ir.bias, c.weight_ii.bias, c.weight_in.bias], 0)),
                        f(torch.cat([c.weight_hr.weight, c.weight_hi.weight, c.weight_hn.weight], 0)))
        return self._packed

    def forward(self, question_wids, lengths=None):
        if not torch.cuda.is_available():
            raise RuntimeError("BayesianUniSkip needs a CUDA device: there is no CPU fallback for this path")
        if self.trainable and self.training:
            if torch.is_grad_enabled():
                return _GruTrainFunction.apply(self.flat, self, question_wids)
            return self._forward_train_raw(question_wids)[0]
        with torch.no_grad():
            return self._forward_eval(question_wids)

    def _forward_eval(self, question_wids):
        import ctypes  # noqa: F401
        from ._lib import lib, check, ptr, current_stream

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
