"""The other CX models that share CXModelBase.vqa_forward (mirror of vqa/models/cx.py:20-210):

    RandomBaseline(knn_size)                                   cx.py:20-30
    BlackBox(vqa_model, knn_size=, trainable_vqa=)             cx.py:114-136   ("Hard Negative Mining" in the README table)
    LinearContext(vqa_model, knn_size=, trainable_vqa=)        cx.py:139-156   (trainable nn.Linear on the flattened z_knns)
    SemanticBaseline(vqa_model, knn_size=, trainable_vqa=)     cx.py:159-205   (+ set_lambda, set_answer_embedding)

Same call signature as the reference, `scores[B, knn] = model(image_features[B, knn+1, dim_v], question_wids, answer_aids)`,
and the same device-resident entry points as cx.NeuralModel (forward_indexed / eval_batch). The arithmetic is in
libvqacx.so: the frozen MUTAN head + classifier + softmax (the front end NeuralCX also runs) and one head kernel
(csrc/cx_heads.cu); LinearContext's Linear runs on the TF32 tensor-core GEMM, forward and backward. No CPU fallback.
"""
import ctypes

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from ._lib import check, current_stream, lib, ptr
from .cx import _Batch, _Frozen, ncx_dims


class RandomBaseline(nn.Module):
    """cx.py:20-30: uniform random scores (the README's 4.20 / 20.85 row)."""

    def __init__(self, knn_size):
        super().__init__()
        self.knn_size = knn_size
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.metrics = torch.zeros(4, dtype=torch.float32, device=self.device)
        self.hits = torch.zeros(2, dtype=torch.int32, device=self.device)
        self.vqa_model = None

    def reset_metrics(self):
        self.metrics.zero_()
        self.hits.zero_()

    def forward(self, image_features, question_wids, answer_aids):
        return torch.rand((image_features.size(0), self.knn_size), device=self.device)

    def eval_batch(self, features, img_idx, q_emb, answer_aids, comp_idxs):
        scores = torch.rand((int(img_idx.shape[0]), self.knn_size), device=self.device)
        check(lib.vqacx_ce_recall(ptr(scores), ptr(comp_idxs), scores.shape[0], self.knn_size, ptr(self.metrics), None, 1.0,
                                  ptr(self.hits), current_stream()), "vqacx_ce_recall")
        return scores


class _VqaHead(nn.Module):
    """CXModelBase (cx.py:51-104) for the heads that only read the frozen VQA model's outputs."""

    def __init__(self, vqa_model, knn_size, trainable_vqa=False):
        super().__init__()
        if trainable_vqa:
            raise NotImplementedError("trainable_vqa=True is outside the B200 hot path (the VQA head is frozen)")
        if not torch.cuda.is_available():
            raise RuntimeError(f"{type(self).__name__} needs a CUDA device: there is no CPU fallback for this path")
        self.vqa_model = vqa_model
        self.trainable_vqa = False
        self.vqa_model.eval()                                                    # cx.py:59-60
        for p in self.vqa_model.parameters():
            p.requires_grad = False                                              # cx.py:79-80
        self.knn_size = knn_size
        self.device = torch.device("cuda", torch.cuda.current_device())
        fo = vqa_model.opt["fusion"]
        self.dim_z = fo["dim_mm"]
        self.ans_size = len(vqa_model.vocab_answers)
        # dim_h / n_layers only size NeuralCX's part of the shared workspace: smallest legal values
        self.dims = ncx_dims(fo, self.ans_size, 32, 1, knn=knn_size)
        self._frozen = _Frozen(vqa_model, self.device)
        self._ws = None
        self.metrics = torch.zeros(4, dtype=torch.float32, device=self.device)   # [0] loss sum
        self.hits = torch.zeros(2, dtype=torch.int32, device=self.device)        # Recall@1, Recall@5 counts

    def cuda(self, device=None):
        return self

    def refresh_frozen(self):
        self._frozen = _Frozen(self.vqa_model, self.device)

    def load_state_dict(self, state_dict, strict=True):
        out = super().load_state_dict(state_dict, strict=strict)
        self.refresh_frozen()          # the C ABI reads packed copies of the VQA head's weights
        return out

    def reset_metrics(self):
        self.metrics.zero_()
        self.hits.zero_()

    def _make_batch(self, features, img_idx, q_emb, answer_aids):
        B = int(img_idx.shape[0])
        need = lib.vqacx_ncx_workspace_bytes(ctypes.byref(self.dims), B, 0)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        return _Batch(features, img_idx, q_emb.detach().to(self.device, torch.float32).contiguous(),
                      answer_aids.to(self.device).contiguous(), 0.0, 0, None, self._ws)

    def vqa_forward_indexed(self, features, img_idx, q_emb, answer_aids=None, want=("z_knns", "z_orig", "p_knns")):
        """vqa_forward (cx.py:64-104) on device-resident rows -> dict of z_knns [B, knn, dim_z], z_orig [B, dim_z],
        p_knns [B, knn, n_ans] = softmax(a_knns)."""
        B = int(img_idx.shape[0])
        aid = answer_aids if answer_aids is not None else torch.zeros(B, dtype=torch.int64, device=self.device)
        batch = self._make_batch(features, img_idx, q_emb, aid)
        out = {}
        if "z_knns" in want:
            out["z_knns"] = torch.empty((B, self.knn_size, self.dim_z), dtype=torch.float32, device=self.device)
        if "z_orig" in want:
            out["z_orig"] = torch.empty((B, self.dim_z), dtype=torch.float32, device=self.device)
        if "p_knns" in want:
            out["p_knns"] = torch.empty((B, self.knn_size, self.ans_size), dtype=torch.float32, device=self.device)
        check(lib.vqacx_cx_vqa_forward(ctypes.byref(self.dims), ctypes.byref(self._frozen.struct), ctypes.byref(batch.struct),
                                       ptr(batch.ws), batch.ws.numel(), ptr(out.get("z_knns")), ptr(out.get("z_orig")),
                                       ptr(out.get("p_knns")), current_stream()), "vqacx_cx_vqa_forward")
        return out

    def forward(self, image_features, question_wids, answer_aids):
        batch_size = image_features.size(0)
        assert image_features.size(1) == self.knn_size + 1                     # cx.py:65
        feats = image_features.to(self.device, torch.float32).contiguous().view(batch_size * (self.knn_size + 1), -1)
        img_idx = torch.arange(batch_size * (self.knn_size + 1), dtype=torch.int32, device=self.device).view(batch_size, -1)
        q_emb = self.vqa_model.seq2vec(question_wids)                           # cx.py:71
        return self.forward_indexed(feats, img_idx, q_emb, answer_aids.to(self.device))

    def eval_batch(self, features, img_idx, q_emb, answer_aids, comp_idxs):
        """One iteration of eval_model's loop (counterexamples.py:461-467), metrics accumulated on the device."""
        with torch.no_grad():
            scores = self.forward_indexed(features, img_idx, q_emb, answer_aids)
        check(lib.vqacx_ce_recall(ptr(scores), ptr(comp_idxs), scores.shape[0], self.knn_size, ptr(self.metrics), None, 1.0,
                                  ptr(self.hits), current_stream()), "vqacx_ce_recall")
        return scores

    def _head_scores(self, features, img_idx, q_emb, answer_aids, head, emb_pairs=None, lam=0.0):
        batch = self._make_batch(features, img_idx, q_emb, answer_aids)
        scores = torch.empty((batch.n, self.knn_size), dtype=torch.float32, device=self.device)
        check(lib.vqacx_cx_head_scores(ctypes.byref(self.dims), ctypes.byref(self._frozen.struct), ctypes.byref(batch.struct),
                                       ptr(batch.ws), batch.ws.numel(), head, ptr(emb_pairs), float(lam), ptr(scores),
                                       current_stream()), "vqacx_cx_head_scores")
        return scores


class BlackBox(_VqaHead):
    """cx.py:114-136: the VQA model's own probability of the original answer on each neighbour, sign flipped."""

    def __init__(self, vqa_model, knn_size=24, trainable_vqa=False):
        super().__init__(vqa_model, knn_size, trainable_vqa)

    def forward_indexed(self, features, img_idx, q_emb, answer_aids):
        return self._head_scores(features, img_idx, q_emb, answer_aids, _lib.VQACX_HEAD_BLACKBOX)


class SemanticBaseline(_VqaHead):
    """cx.py:159-205: lam * (answer-similarity-weighted probability mass, minus the answer's own) - (1 - lam) * log p,
    soft-maxed over the 24 neighbours. `set_answer_embedding` builds the cosine table on the device."""

    def __init__(self, vqa_model, knn_size=24, trainable_vqa=False):
        super().__init__(vqa_model, knn_size, trainable_vqa)
        self.lam = 0.5                                                           # cx.py:168
        self.emb = None
        self.emb_pairs = torch.zeros((self.ans_size, self.ans_size), dtype=torch.float32, device=self.device)

    def set_lambda(self, lam):
        self.lam = lam

    def set_answer_embedding(self, emb):
        e = torch.as_tensor(np.asarray(emb), dtype=torch.float32).to(self.device).contiguous()
        assert e.shape[0] == self.ans_size
        self.emb = e
        check(lib.vqacx_cosine_similarity(ptr(e), e.shape[0], e.shape[1], ptr(self.emb_pairs), current_stream()),
              "vqacx_cosine_similarity")

    def forward_indexed(self, features, img_idx, q_emb, answer_aids):
        return self._head_scores(features, img_idx, q_emb, answer_aids, _lib.VQACX_HEAD_SEMANTIC, self.emb_pairs, self.lam)


class _LinearContextFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, weight, bias, model, batch):
        B, K = batch.n, model.knn_size * model.dim_z
        z_flat = torch.empty((B, K), dtype=torch.float32, device=model.device)
        scores = torch.empty((B, model.knn_size), dtype=torch.float32, device=model.device)
        check(lib.vqacx_cx_linear_context_forward(ctypes.byref(model.dims), ctypes.byref(model._frozen.struct),
                                                  ctypes.byref(batch.struct), ptr(batch.ws), batch.ws.numel(), ptr(weight),
                                                  ptr(bias), ptr(z_flat), ptr(scores), current_stream()),
              "vqacx_cx_linear_context_forward")
        ctx.model, ctx.z_flat, ctx.ws = model, z_flat, batch.ws
        return scores

    @staticmethod
    def backward(ctx, dscores):
        m = ctx.model
        gw = torch.empty((m.knn_size, m.knn_size * m.dim_z), dtype=torch.float32, device=m.device)
        gb = torch.empty((m.knn_size,), dtype=torch.float32, device=m.device)
        ds = dscores.contiguous()
        check(lib.vqacx_cx_linear_context_backward(ctypes.byref(m.dims), ds.shape[0], ptr(ctx.ws), ctx.ws.numel(),
                                                   ptr(ctx.z_flat), ptr(ds), ptr(gw), ptr(gb), current_stream()),
              "vqacx_cx_linear_context_backward")
        return gw, gb, None, None


class LinearContext(_VqaHead):
    """cx.py:139-156: one nn.Linear(knn * dim_z, knn) over the 24 neighbours' fused embeddings, trained with the same
    loop as NeuralCX (counterexamples.py:322-339). State-dict names: linear.weight, linear.bias (+ vqa_model.*)."""

    def __init__(self, vqa_model, knn_size=24, trainable_vqa=False):
        super().__init__(vqa_model, knn_size, trainable_vqa)
        self.linear = nn.Linear(self.knn_size * self.dim_z, self.knn_size).to(self.device)
        self._step = 0
        self._adam = None

    def forward_indexed(self, features, img_idx, q_emb, answer_aids):
        batch = self._make_batch(features, img_idx, q_emb, answer_aids)
        return _LinearContextFn.apply(self.linear.weight, self.linear.bias, self, batch)

    def train_step(self, features, img_idx, q_emb, answer_aids, comp_idxs, lr=1e-4, grad_hook=None, global_batch=None):
        """One iteration of counterexamples.py:322-339 on the device: forward, CE / len(batch), Recall, backward, Adam
        (vqacx_adam_step, the same kernel NeuralCX uses)."""
        B = int(img_idx.shape[0])
        params = [self.linear.weight, self.linear.bias]
        if B == 0:   # an empty data-parallel slice of the last partial batch: zero gradients, still join the all-reduce
            scores = torch.empty((0, self.knn_size), dtype=torch.float32, device=self.linear.weight.device)
            for p in params:
                p.grad = torch.zeros_like(p.data)
        else:
            with torch.enable_grad():
                scores = self.forward_indexed(features, img_idx, q_emb, answer_aids)
            dscores = torch.empty_like(scores)
            check(lib.vqacx_ce_recall(ptr(scores), ptr(comp_idxs), B, self.knn_size, ptr(self.metrics), ptr(dscores),
                                      1.0 / float(global_batch or B), ptr(self.hits), current_stream()), "vqacx_ce_recall")
            scores.backward(dscores)
        if grad_hook is not None:
            for p in params:
                grad_hook(p.grad)
        if self._adam is None:
            self._adam = [(torch.zeros_like(p.data), torch.zeros_like(p.data)) for p in params]
        self._step += 1
        for p, (m, v) in zip(params, self._adam):
            check(lib.vqacx_adam_step(ptr(p.data), ptr(p.grad), ptr(m), ptr(v), p.numel(), float(lr), 0.9, 0.999, 1e-8,
                                      self._step, 1.0, current_stream()), "vqacx_adam_step")
            p.grad = None
        return scores.detach()
