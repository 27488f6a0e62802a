"""Stage 2 of the VQA-CX hot path: NeuralCX candidate scoring (mirror of vqa/models/cx.py).

`NeuralModel` keeps the reference's constructor and call signature
    NeuralModel(model_spec, dim_h, n_layers, emb, drop_p, vqa_model=, knn_size=24, trainable_vqa=False)   cx.py:220-259
    scores[B,24] = cx_model(image_features[B,25,dim_v], question_wids[B,26], answer_aids[B])               cx.py:261-333
and its state_dict names (answer_embedding.weight, linear_1.*, linear_2.*, linear_3.*, out.*), but all arithmetic
runs in libvqacx.so (include/vqacx.h): feature construction, the frozen MUTAN head, the MLP, softmax-CE, Recall@k,
backward and Adam. This file is host plumbing: parameter packing, device buffers, the autograd bridge.
There is no CPU or PyTorch fallback: without the CUDA library the import fails.
"""
import ctypes
import math

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from ._lib import NcxBatch, NcxDims, NcxFrozen, NcxHeadLayout, NcxHeadTrain, NcxLayout, check, current_stream, lib, ptr

KNN = 24
DIM_A = 2400
SPEC_KEYS = ("v_emb", "v_mult", "v_dist", "v_rank", "q_emb", "a_emb", "z_emb")


def recallAtK(scores, ground_truth, k=5):
    """counterexamples.py:501-506 drop-in (returns a per-sample 0/1 numpy array). Computed on the device by the
    fused CE/recall kernel's rank rule; only k in {1, 5} exist there, other k fall back to the same rank rule in torch
    ops on the device (plumbing: no CPU compute)."""
    assert scores.shape[0] == ground_truth.shape[0]
    s = scores.detach()
    gt = ground_truth.view(-1, 1)
    sg = s.gather(1, gt)
    idx = torch.arange(s.shape[1], device=s.device).view(1, -1)
    ahead = (s > sg) | ((s == sg) & (idx < gt))
    return (ahead.sum(1) < k).to(torch.int64).cpu().numpy()


class _Frozen:
    """Contiguous device copies of the frozen MutanNoAtt head in the layouts the C ABI takes (VqacxNcxFrozen)."""

    def __init__(self, vqa_model, device):
        f = vqa_model.fusion

        def dev(t):
            return t.detach().to(device=device, dtype=torch.float32).contiguous()

        self.wv, self.bv = dev(f.linear_v.weight), dev(f.linear_v.bias)
        self.wq, self.bq = dev(f.linear_q.weight), dev(f.linear_q.bias)
        self.whv = dev(torch.cat([l.weight for l in f.list_linear_hv], 0))
        self.bhv = dev(torch.cat([l.bias for l in f.list_linear_hv], 0))
        self.whq = dev(torch.cat([l.weight for l in f.list_linear_hq], 0))
        self.bhq = dev(torch.cat([l.bias for l in f.list_linear_hq], 0))
        self.wc, self.bc = dev(vqa_model.linear_classif.weight), dev(vqa_model.linear_classif.bias)
        # fp16 twin of the stacked W_hv for the fused MUTAN kernel (frozen: made once): [whv | bhv | 0...] with
        # roundup8(dim_hv + 1) columns -- the bias is one more k column against the constant-one column of x_v16
        k16 = (self.whv.shape[1] + 1 + 7) // 8 * 8
        w16 = torch.zeros((self.whv.shape[0], k16), dtype=torch.float16, device=device)
        w16[:, :self.whv.shape[1]] = self.whv.to(torch.float16)
        w16[:, self.whv.shape[1]] = self.bhv.to(torch.float16)
        self.whv_f16 = w16 if bool(torch.isfinite(w16).all()) else None
        self.struct = NcxFrozen(*[t.data_ptr() for t in (self.wv, self.bv, self.wq, self.bq, self.whv, self.bhv,
                                                         self.whq, self.bhq, self.wc, self.bc)],
                                None if self.whv_f16 is None else self.whv_f16.data_ptr())


def unpack_head(flat, dims, hl):
    """Flat head buffer (VqacxNcxHeadLayout: weights, gradients or Adam moments) -> dict under the reference's
    `vqa_model.*` parameter names (copies)."""
    dm, R = dims.dim_mm, dims.R
    out = {}

    def take(name, off, shape):
        out["vqa_model." + name] = flat[off:off + int(np.prod(shape))].view(*shape).clone()

    take("fusion.linear_v.weight", hl.off_wv, (dims.dim_hv, dims.dim_v))
    take("fusion.linear_v.bias", hl.off_bv, (dims.dim_hv,))
    take("fusion.linear_q.weight", hl.off_wq, (dims.dim_hq, dims.dim_q))
    take("fusion.linear_q.bias", hl.off_bq, (dims.dim_hq,))
    for r in range(R):
        take(f"fusion.list_linear_hv.{r}.weight", hl.off_whv + r * dm * dims.dim_hv, (dm, dims.dim_hv))
        take(f"fusion.list_linear_hv.{r}.bias", hl.off_bhv + r * dm, (dm,))
        take(f"fusion.list_linear_hq.{r}.weight", hl.off_whq + r * dm * dims.dim_hq, (dm, dims.dim_hq))
        take(f"fusion.list_linear_hq.{r}.bias", hl.off_bhq + r * dm, (dm,))
    take("linear_classif.weight", hl.off_wc, (dims.n_ans, dm))
    take("linear_classif.bias", hl.off_bc, (dims.n_ans,))
    return out


class _TrainableHead:
    """trainable_vqa=True (cx.py:73-74): the head's ten tensors in ONE flat float32 buffer (VqacxNcxHeadLayout, the layout
    the backward and Adam work on); the vqa_model's own parameters become views of it, so `vqa_model.state_dict()` and a
    checkpoint written from it always show the trained weights. `.struct` serves the evaluation-mode forward."""

    def __init__(self, vqa_model, dims, device):
        hl = NcxHeadLayout()
        check(lib.vqacx_ncx_head_layout(ctypes.byref(dims), ctypes.byref(hl)), "vqacx_ncx_head_layout")
        self.layout = hl
        self.flat = torch.zeros(int(hl.total), dtype=torch.float32, device=device)
        f = vqa_model.fusion
        dm = dims.dim_mm

        def bind(param, off, shape):
            n = int(np.prod(shape))
            view = self.flat[off:off + n].view(*shape)
            view.copy_(param.detach().to(device=device, dtype=torch.float32))
            param.data = view
            param.requires_grad = False      # trained through the flat buffer, not through torch autograd

        bind(f.linear_v.weight, hl.off_wv, (dims.dim_hv, dims.dim_v))
        bind(f.linear_v.bias, hl.off_bv, (dims.dim_hv,))
        bind(f.linear_q.weight, hl.off_wq, (dims.dim_hq, dims.dim_q))
        bind(f.linear_q.bias, hl.off_bq, (dims.dim_hq,))
        for r in range(dims.R):
            bind(f.list_linear_hv[r].weight, hl.off_whv + r * dm * dims.dim_hv, (dm, dims.dim_hv))
            bind(f.list_linear_hv[r].bias, hl.off_bhv + r * dm, (dm,))
            bind(f.list_linear_hq[r].weight, hl.off_whq + r * dm * dims.dim_hq, (dm, dims.dim_hq))
            bind(f.list_linear_hq[r].bias, hl.off_bhq + r * dm, (dm,))
        bind(vqa_model.linear_classif.weight, hl.off_wc, (dims.n_ans, dm))
        bind(vqa_model.linear_classif.bias, hl.off_bc, (dims.n_ans,))
        self.rebind(self.flat)

    def rebind(self, flat):
        base, hl = flat.data_ptr(), self.layout
        offs = (hl.off_wv, hl.off_bv, hl.off_wq, hl.off_bq, hl.off_whv, hl.off_bhv, hl.off_whq, hl.off_bhq, hl.off_wc,
                hl.off_bc)
        self.struct = NcxFrozen(*[base + 4 * int(o) for o in offs], None)
        self.whv_f16 = None


class _NcxFunction(torch.autograd.Function):
    """Autograd bridge so the reference's loop (`loss = criterion(scores, gt) / B; loss.backward()`,
    counterexamples.py:325-337) drives the CUDA forward / backward unchanged."""

    @staticmethod
    def forward(ctx, flat, head_flat, q_emb, model, batch):
        scores = model._forward_raw(batch, training=model.training)
        ctx.model, ctx.batch = model, batch
        return scores

    @staticmethod
    def backward(ctx, dscores):
        grads = ctx.model._backward_raw(ctx.batch, dscores.contiguous())
        if ctx.model.trainable_vqa:   # dLoss/dq_emb flows on into the question encoder when it takes part in autograd
            return grads[0], grads[1], (ctx.model.last_dq_emb if ctx.needs_input_grad[2] else None), None, None
        return grads, None, None, None, None


class _Batch:
    """Device tensors of one batch + the C struct pointing at them (kept alive together)."""

    def __init__(self, features, img_idx, q_emb, answer_aids, drop_p, drop_seed, drop_mask, ws, drop_seed_dev=None,
                 lesion=0, v_rank=None, hv_table=None):
        self.features, self.img_idx, self.q_emb, self.answer_aids = features, img_idx, q_emb, answer_aids
        self.drop_mask, self.ws, self.drop_seed_dev, self.v_rank = drop_mask, ws, drop_seed_dev, v_rank
        self.n = int(img_idx.shape[0])
        self.struct = NcxBatch(self.n, features.data_ptr(), features.stride(0), img_idx.data_ptr(), q_emb.data_ptr(),
                               q_emb.stride(0), answer_aids.data_ptr(), float(drop_p), int(drop_seed),
                               None if drop_mask is None else drop_mask.data_ptr(),
                               None if drop_seed_dev is None else drop_seed_dev.data_ptr(), int(lesion),
                               None if v_rank is None else v_rank.data_ptr(),
                               None if hv_table is None else hv_table.data_ptr())
        self.hv_table = hv_table


def ncx_dims(fusion_opt, n_ans, dim_h, n_layers, knn=KNN):
    return NcxDims(int(fusion_opt["dim_v"]), int(fusion_opt["dim_q"]), int(fusion_opt["dim_hv"]), int(fusion_opt["dim_hq"]),
                   int(fusion_opt["dim_mm"]), int(fusion_opt["R"]), int(n_ans), DIM_A, int(dim_h), int(n_layers), int(knn))


def ncx_layout(dims):
    lay = NcxLayout()
    check(lib.vqacx_ncx_layout(ctypes.byref(dims), ctypes.byref(lay)), "vqacx_ncx_layout")
    return lay


def _w1_groups(dims, lay):
    """(reference first column, packed first column, width) of each linear_1 input group, cx.py:309-320 order."""
    dv, dq, dz, da, k = dims.dim_v, dims.dim_q, dims.dim_mm, dims.dim_a, dims.knn
    groups, ref = [], 0
    for width, col in ((dv, lay.col_vorig), (dv, lay.col_vother), (dv, lay.col_vmult), (1, lay.col_dist),
                       (k, lay.col_rank), (dq, lay.col_q), (dz, lay.col_zorig), (dz, lay.col_zother),
                       (da, lay.col_agt), (da, lay.col_aother)):
        groups.append((ref, int(col), width))
        ref += width
    return groups, ref


def pack_state(state, dims, lay, out=None):
    """Reference-named tensors -> flat packed float32 buffer (VqacxNcxLayout). Works on any device."""
    H, hp = dims.dim_h, int(lay.ld_h)
    dev = state["linear_1.weight"].device
    flat = out if out is not None else torch.zeros(int(lay.total), dtype=torch.float32, device=dev)
    flat.zero_()
    flat[lay.off_emb:lay.off_emb + dims.n_ans * dims.dim_a] = state["answer_embedding.weight"].reshape(-1).float()
    w1 = flat[lay.off_w1:lay.off_w1 + H * lay.ld_w1].view(H, int(lay.ld_w1))
    groups, n_in = _w1_groups(dims, lay)
    src = state["linear_1.weight"].float()
    assert tuple(src.shape) == (H, n_in), (tuple(src.shape), (H, n_in))
    for ref, col, width in groups:
        w1[:, col:col + width] = src[:, ref:ref + width]
    flat[lay.off_b1:lay.off_b1 + H] = state["linear_1.bias"].float()
    for l, (ow, ob) in ((2, (lay.off_w2, lay.off_b2)), (3, (lay.off_w3, lay.off_b3))):
        if dims.n_layers >= l:
            flat[ow:ow + H * hp].view(H, hp)[:, :H] = state[f"linear_{l}.weight"].float()
            flat[ob:ob + H] = state[f"linear_{l}.bias"].float()
    flat[lay.off_wout:lay.off_wout + H] = state["out.weight"].reshape(-1).float()
    flat[lay.off_bout:lay.off_bout + 1] = state["out.bias"].reshape(-1).float()
    return flat


def unpack_state(flat, dims, lay):
    """Flat packed buffer -> dict of reference-named tensors (copies)."""
    H, hp = dims.dim_h, int(lay.ld_h)
    out = {"answer_embedding.weight": flat[lay.off_emb:lay.off_emb + dims.n_ans * dims.dim_a].view(dims.n_ans, dims.dim_a).clone()}
    w1 = flat[lay.off_w1:lay.off_w1 + H * lay.ld_w1].view(H, int(lay.ld_w1))
    groups, n_in = _w1_groups(dims, lay)
    ref_w1 = torch.empty((H, n_in), dtype=flat.dtype, device=flat.device)
    for ref, col, width in groups:
        ref_w1[:, ref:ref + width] = w1[:, col:col + width]
    out["linear_1.weight"] = ref_w1
    out["linear_1.bias"] = flat[lay.off_b1:lay.off_b1 + H].clone()
    for l, (ow, ob) in ((2, (lay.off_w2, lay.off_b2)), (3, (lay.off_w3, lay.off_b3))):
        if dims.n_layers >= l:
            out[f"linear_{l}.weight"] = flat[ow:ow + H * hp].view(H, hp)[:, :H].clone()
            out[f"linear_{l}.bias"] = flat[ob:ob + H].clone()
    out["out.weight"] = flat[lay.off_wout:lay.off_wout + H].view(1, H).clone()
    out["out.bias"] = flat[lay.off_bout:lay.off_bout + 1].clone()
    return out


class NeuralModel(nn.Module):
    """Drop-in for vqa.models.cx.NeuralModel (cx.py:218-333), including the lesion flags the options/cx/
    neuralcx_lesion_*.yaml files use: v_mult / v_dist false -> zeros (cx.py:297,301, in the feature kernel); v_rank false
    -> a torch.rand([B, 24]) row per neighbour (cx.py:306; drawn on the device here, fed to layer 1 as one more K
    segment); v_emb false -> torch.rand image features for everything downstream (cx.py:265-266); q_emb or z_emb false
    alone -> nothing changes, as in the reference (cx.py:270 still runs vqa_forward and uses its outputs). Both false
    crashes the reference (`a_knns` undefined at cx.py:281) and a_emb false is used by no config: those raise."""

    def __init__(self, model_spec, dim_h, n_layers, emb, drop_p, **kwargs):
        super().__init__()
        vqa_model = kwargs["vqa_model"]
        self.knn_size = kwargs["knn_size"]
        self.trainable_vqa = kwargs.get("trainable_vqa", False)
        assert self.knn_size == KNN                                              # cx.py:226
        off = {k for k in SPEC_KEYS if not model_spec.get(k, False)}
        if "a_emb" in off or {"q_emb", "z_emb"} <= off:
            raise NotImplementedError(f"lesioned model_spec flags {sorted(off)}: a_emb=false is used by no reference "
                                      "config and q_emb=z_emb=false crashes the reference (cx.py:270-281)")
        self.lesion = ((_lib.VQACX_LESION_V_MULT if "v_mult" in off else 0) | (_lib.VQACX_LESION_V_DIST if "v_dist" in off else 0) |
                       (_lib.VQACX_LESION_V_RANK if "v_rank" in off else 0))
        self.lesion_v_emb = "v_emb" in off
        self.v_rank_override = None   # tests: explicit [B*24, 24] stand-in rows instead of torch.rand
        if not torch.cuda.is_available():
            raise RuntimeError("NeuralModel needs a CUDA device: there is no CPU fallback for this path")
        self.model_spec = model_spec
        self.vqa_model = vqa_model
        if not self.trainable_vqa:
            self.vqa_model.eval()                                                # cx.py:60-61
        for p in self.vqa_model.parameters():
            p.requires_grad = False                                              # cx.py:79-80 (trainable: see _TrainableHead)
        self.device = torch.device("cuda", torch.cuda.current_device())
        fo = vqa_model.opt["fusion"]
        self.dim_v, self.dim_q, self.dim_z = fo["dim_v"], fo["dim_q"], fo["dim_mm"]
        self.ans_size = len(vqa_model.vocab_answers)
        self.dim_a, self.dim_h, self.n_layers, self.drop_p = DIM_A, dim_h, n_layers, float(drop_p)
        self.dims = ncx_dims(fo, self.ans_size, dim_h, n_layers)
        self.layout = ncx_layout(self.dims)
        # same module construction order as the reference -> same initial weights under the same torch seed
        init = {}
        answer_embedding = nn.Embedding(self.ans_size, self.dim_a)
        if emb is not None:
            assert emb.shape[1] == self.dim_a                                    # cx.py:241-243
            answer_embedding.weight.data = torch.as_tensor(np.asarray(emb), dtype=torch.float32)
        init["answer_embedding.weight"] = answer_embedding.weight.data
        input_size = self.dim_v * 3 + self.dim_a * 2 + self.dim_z * 2 + self.dim_q + self.knn_size + 1
        names = ["linear_1"] + [f"linear_{l}" for l in (2, 3) if n_layers >= l] + ["out"]
        for name in names:
            lin = nn.Linear(input_size if name == "linear_1" else dim_h, 1 if name == "out" else dim_h)
            init[name + ".weight"], init[name + ".bias"] = lin.weight.data, lin.bias.data
        self.flat = nn.Parameter(pack_state(init, self.dims, self.layout).to(self.device))
        if self.trainable_vqa:
            # cx.py:73-74: gradients flow into the MUTAN head (linear_v / linear_q / rank projections / classifier), which
            # runs in train mode (its dropouts active) whenever this module does. The question encoder stays frozen and
            # in eval mode: its backward (BPTT through the Bayesian GRU) is not built; `last_dq_emb` holds dLoss/dq_emb.
            fo_all = vqa_model.opt["fusion"]
            if fo_all.get("dropout_hv", 0) or fo_all.get("dropout_hq", 0):
                raise NotImplementedError("fusion.dropout_hv / dropout_hq > 0 is used by no reference option file")
            self._head = _TrainableHead(vqa_model, self.dims, self.device)
            self.head_flat = nn.Parameter(self._head.flat)
            self._frozen = self._head
            self._head_p = (float(fo_all.get("dropout_v", 0)), float(fo_all.get("dropout_q", 0)),
                            float(vqa_model.opt.get("classif", {}).get("dropout", 0)))
            self.head_masks = None     # tests: (mask_v, mask_q, mask_c) in library row order, 0 or 1/(1-p)
            self._head_exp_avg = self._head_exp_avg_sq = None
            self.last_dq_emb = None
            # the question encoder trains too when it is the device GRU (BayesianUniSkip.make_trainable: train-mode
            # Bayesian dropout + BPTT); any other seq2vec module (a precomputed table) is a constant of this model
            s2v = vqa_model.seq2vec
            self.trains_seq2vec = hasattr(s2v, "make_trainable")
            if self.trains_seq2vec:
                if not getattr(s2v, "trainable", False):
                    s2v.make_trainable(float(vqa_model.opt.get("seq2vec", {}).get("dropout", 0.25)), self.device)
                self._s2v_exp_avg = self._s2v_exp_avg_sq = None
            elif any(True for _ in getattr(s2v, "parameters", lambda: [])()):
                import warnings

                warnings.warn("trainable_vqa=True: seq2vec is not the device encoder (BayesianUniSkip) and stays frozen",
                              RuntimeWarning, stacklevel=2)
        else:
            self._frozen = _Frozen(vqa_model, self.device)
        self._ws_eval = None
        self._step = 0
        self._exp_avg = self._exp_avg_sq = None
        self._drop_calls = 0
        self.drop_seed = 0x5EED
        self.drop_mask = None      # tests: explicit [n_layers, B*24, dim_h] mask of 0 or 1/(1-p)
        self.metrics = torch.zeros(4, dtype=torch.float32, device=self.device)   # [0] loss sum
        self.hits = torch.zeros(2, dtype=torch.int32, device=self.device)        # Recall@1, Recall@5 counts
        # evaluation precision: "fp16" = the fast path (vqacx_ncx_fast_forward: per-image tables of the frozen head's
        # visual side, fp16 operands, fp32 accumulation) whenever a batch indexes a resident feature matrix;
        # "tf32" = always the general path (vqacx_ncx_forward), which also serves training and the reference-shaped call
        self.eval_precision = "fp16"
        self._own_version = 0          # bumped by every kernel that writes self.flat behind torch's back
        self._frozen_version = 0
        self._tables = {}              # (features ptr, rows, ld, frozen version) -> fp16 per-image table
        self._twins = None             # (version key, buffer)
        self._p1 = None                # (key, weakref(features), buffer): per-image v_other part of layer 1
        self.use_p1_table = True
        self.use_image_tables = True   # False: every path evaluates the frozen head's visual side per candidate
        if self.trainable_vqa:         # per-image tables / fp16 twins describe a FROZEN head
            self.eval_precision, self.use_image_tables = "tf32", False
        self._ws_fast = None
        self.last_eval_path = None

    # ---- nn.Module surface the reference relies on ---------------------------------------------
    def cuda(self, device=None):
        return self

    def train(self, mode=True):
        super().train(mode)
        if not self.trainable_vqa:
            self.vqa_model.eval()                                                # cx.py:76-78
        return self

    def refresh_frozen(self):
        """Re-read the VQA head's weights (after vqa_model.load_state_dict)."""
        if not self.trainable_vqa:       # trainable: the vqa_model's parameters ARE views of the flat head buffer
            self._frozen = _Frozen(self.vqa_model, self.device)
        self._frozen_version += 1
        self._tables.clear()
        self._twins = None

    def invalidate_image_tables(self):
        """Forget the per-image tables of the fast evaluation path (call after changing a feature matrix in place; the
        benchmarks call it so that every timed pass pays for its own table)."""
        self._tables.clear()
        self._p1 = None

    @property
    def compute_dtype(self):
        return ("eval: fp16 operands (per-image tables of the frozen visual side), fp32 accumulate; train: tf32 products, "
                "fp32 accumulate") if self.eval_precision == "fp16" else "tf32 tensor-core products, fp32 accumulate"

    # ---- fp16 fast path of the evaluation forward ------------------------------------------------
    def _tables_ok(self, features):
        """The per-image tables of the frozen head's visual side can serve this feature matrix (training and general
        path included: vqacx_ncx_forward takes them through VqacxNcxBatch.hv_table)."""
        return (self.use_image_tables and not self.lesion_v_emb and features.dim() == 2 and features.is_cuda
                and features.dtype == torch.float32 and features.stride(1) == 1 and features.stride(0) % 4 == 0
                and lib.vqacx_ncx_fast_twins_bytes(ctypes.byref(self.dims)) > 0)

    def _fast_ok(self, features):
        return (self.eval_precision == "fp16" and not (self.lesion & _lib.VQACX_LESION_V_RANK)
                and self._tables_ok(features))

    def _fast_ws(self, B):
        need = lib.vqacx_ncx_fast_workspace_bytes(ctypes.byref(self.dims), B)
        if self._ws_fast is None or self._ws_fast.numel() < need:
            self._ws_fast = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._ws_fast

    def _fast_tables(self, features, B):
        # keyed by the tensor OBJECT (weak reference) as well as its storage: a new matrix that the allocator placed at a
        # freed matrix's address must not inherit its table. In-place edits bump torch's version counter.
        key = (features.data_ptr(), int(features.shape[0]), int(features.stride(0)), features._version, self._frozen_version)
        hit = self._tables.get(key)
        tab = hit[1] if hit is not None and hit[0]() is features else None
        if tab is None:
            if len(self._tables) >= 2:
                self._tables.clear()
            n_img = int(features.shape[0])
            tab = torch.empty(lib.vqacx_ncx_fast_tables_bytes(ctypes.byref(self.dims), n_img), dtype=torch.uint8,
                              device=self.device)
            ws = self._fast_ws(B)
            check(lib.vqacx_ncx_fast_build_tables(ctypes.byref(self.dims), ctypes.byref(self._frozen.struct), ptr(features),
                                                  features.stride(0), n_img, ptr(tab), tab.numel(), ptr(ws), ws.numel(),
                                                  current_stream()), "vqacx_ncx_fast_build_tables")
            import weakref

            self._tables[key] = (weakref.ref(features), tab)
        return tab

    def _fast_twins(self, B):
        key = (self.flat._version, self._own_version, self._frozen_version, self.flat.data_ptr())
        if self._twins is None or self._twins[0] != key:
            buf = self._twins[1] if self._twins is not None else torch.empty(
                lib.vqacx_ncx_fast_twins_bytes(ctypes.byref(self.dims)), dtype=torch.uint8, device=self.device)
            ws = self._fast_ws(B)
            check(lib.vqacx_ncx_fast_build_twins(ctypes.byref(self.dims), ctypes.byref(self._frozen.struct), ptr(self.flat),
                                                 ptr(buf), buf.numel(), ptr(ws), ws.numel(), current_stream()),
                  "vqacx_ncx_fast_build_twins")
            self._twins = (key, buf)
        return self._twins[1]

    def _fast_p1(self, features, B):
        """P1[img] = features[img] . W1[:, v_other]^T: depends on the parameters AND the feature matrix; rebuilt (one
        product over all images) whenever either changed since the last evaluation batch."""
        if not self.use_p1_table:
            return None
        key = (features.data_ptr(), int(features.shape[0]), int(features.stride(0)), features._version,
               self.flat._version, self._own_version, self.flat.data_ptr())
        hit = self._p1
        if hit is not None and hit[0] == key and hit[1]() is features:
            return hit[2]
        n_img = int(features.shape[0])
        need = lib.vqacx_ncx_fast_p1_bytes(ctypes.byref(self.dims), n_img)
        buf = hit[2] if hit is not None and hit[2].numel() >= need else torch.empty(need, dtype=torch.uint8, device=self.device)
        ws = self._fast_ws(B)
        check(lib.vqacx_ncx_fast_build_p1(ctypes.byref(self.dims), ptr(self.flat), ptr(features), features.stride(0), n_img,
                                          ptr(buf), buf.numel(), ptr(ws), ws.numel(), current_stream()),
              "vqacx_ncx_fast_build_p1")
        import weakref

        self._p1 = (key, weakref.ref(features), buf)
        return buf

    def _forward_fast(self, features, img_idx, q_emb, answer_aids):
        B = int(img_idx.shape[0])
        scores = torch.empty((B, self.knn_size), dtype=torch.float32, device=self.device)
        if B == 0:
            return scores
        tab = self._fast_tables(features, B)
        twins = self._fast_twins(B)
        p1 = self._fast_p1(features, B)
        ws = self._fast_ws(B)
        batch = _Batch(features, img_idx.contiguous(), q_emb.detach().to(self.device, torch.float32).contiguous(),
                       answer_aids.to(self.device).contiguous(), 0.0, 0, None, ws, lesion=self.lesion)
        check(lib.vqacx_ncx_fast_forward(ctypes.byref(self.dims), ctypes.byref(self._frozen.struct), ptr(self.flat),
                                         ctypes.byref(batch.struct), ptr(tab), int(features.shape[0]), ptr(twins), ptr(p1),
                                         ptr(ws), ws.numel(), ptr(scores), current_stream()), "vqacx_ncx_fast_forward")
        self.last_eval_path = "fp16"
        return scores

    def state_dict(self, *args, **kwargs):
        sd = unpack_state(self.flat.detach(), self.dims, self.layout)
        for k, v in self.vqa_model.state_dict().items():
            sd["vqa_model." + k] = v
        return sd

    def load_state_dict(self, state_dict, strict=True):
        own = {k: torch.as_tensor(v).to(self.device) for k, v in state_dict.items() if not k.startswith("vqa_model.")}
        self._own_version += 1
        with torch.no_grad():
            pack_state(own, self.dims, self.layout, out=self.flat.data)
        vqa = {k[len("vqa_model."):]: v for k, v in state_dict.items() if k.startswith("vqa_model.")}
        if vqa:
            self.vqa_model.load_state_dict(vqa, strict=strict)
            self.refresh_frozen()
        return self

    # ---- forward ------------------------------------------------------------------------------
    def _make_batch(self, features, img_idx, q_emb, answer_aids, training, resident=False):
        B = int(img_idx.shape[0])
        hv_table = self._fast_tables(features, B) if resident and B > 0 and self._tables_ok(features) else None
        need = lib.vqacx_ncx_workspace_bytes(ctypes.byref(self.dims), B, 1 if training else 0)
        if training:
            ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        else:
            if self._ws_eval is None or self._ws_eval.numel() < need:
                self._ws_eval = torch.empty(need, dtype=torch.uint8, device=self.device)
            ws = self._ws_eval
        drop_p = self.drop_p if training else 0.0
        seed = self.drop_seed + 16 * self._drop_calls
        if training:
            self._drop_calls += 1
        v_rank = None
        if self.lesion_v_emb:                                                    # cx.py:265-266
            features = torch.rand((B * (self.knn_size + 1), self.dim_v), dtype=torch.float32, device=self.device)
            img_idx = torch.arange(B * (self.knn_size + 1), dtype=torch.int32, device=self.device).view(B, -1)
        if self.lesion & _lib.VQACX_LESION_V_RANK:                               # cx.py:306
            v_rank = torch.zeros((B * self.knn_size, 32), dtype=torch.float32, device=self.device)
            v_rank[:, :self.knn_size] = (torch.rand((B * self.knn_size, self.knn_size), device=self.device)
                                         if self.v_rank_override is None else self.v_rank_override)
        batch = _Batch(features, img_idx, q_emb.detach().to(self.device, torch.float32).contiguous(),
                       answer_aids.to(self.device).contiguous(), drop_p, seed,
                       self.drop_mask if training else None, ws, lesion=self.lesion, v_rank=v_rank, hv_table=hv_table)
        if self.trainable_vqa and training:
            # the head's train-mode state of this step: dropout probabilities, mask seed, activation workspace
            masks = self.head_masks or (None, None, None)
            batch.head_masks = tuple(None if m is None else m.to(self.device, torch.float32).contiguous() for m in masks)
            batch.head_train = NcxHeadTrain(self._head_p[0], self._head_p[1], self._head_p[2],
                                            (int(seed) * 0x9E3779B97F4A7C15 + 0x7F4A7C15) & 0xFFFFFFFFFFFFFFFF,
                                            *[None if m is None else m.data_ptr() for m in batch.head_masks])
            batch.tv_ws = torch.empty(lib.vqacx_ncx_tv_workspace_bytes(ctypes.byref(self.dims), B), dtype=torch.uint8,
                                      device=self.device)
        return batch

    def _forward_raw(self, batch, training):
        scores = torch.empty((batch.n, self.knn_size), dtype=torch.float32, device=self.device)
        if self.trainable_vqa and training:
            check(lib.vqacx_ncx_forward_tv(ctypes.byref(self.dims), ptr(self.head_flat), ptr(self.flat),
                                           ctypes.byref(batch.struct), ctypes.byref(batch.head_train), ptr(batch.ws),
                                           batch.ws.numel(), ptr(batch.tv_ws), batch.tv_ws.numel(), ptr(scores),
                                           current_stream()), "vqacx_ncx_forward_tv")
            return scores
        check(lib.vqacx_ncx_forward(ctypes.byref(self.dims), ctypes.byref(self._frozen.struct), ptr(self.flat),
                                    ctypes.byref(batch.struct), ptr(batch.ws), batch.ws.numel(), 1 if training else 0,
                                    ptr(scores), current_stream()), "vqacx_ncx_forward")
        return scores

    def _backward_raw(self, batch, dscores):
        grads = torch.empty_like(self.flat.data)
        if self.trainable_vqa:
            hgrads = torch.empty_like(self.head_flat.data)
            dq = torch.empty((batch.n, self.dim_q), dtype=torch.float32, device=self.device)
            check(lib.vqacx_ncx_backward_tv(ctypes.byref(self.dims), ptr(self.head_flat), ptr(self.flat),
                                            ctypes.byref(batch.struct), ctypes.byref(batch.head_train), ptr(batch.ws),
                                            batch.ws.numel(), ptr(batch.tv_ws), batch.tv_ws.numel(), ptr(dscores),
                                            ptr(grads), ptr(hgrads), ptr(dq), dq.stride(0), current_stream()),
                  "vqacx_ncx_backward_tv")
            self.last_dq_emb = dq
            return grads, hgrads
        check(lib.vqacx_ncx_backward(ctypes.byref(self.dims), ptr(self.flat), ctypes.byref(batch.struct), ptr(batch.ws),
                                     batch.ws.numel(), ptr(dscores), ptr(grads), current_stream()), "vqacx_ncx_backward")
        return grads

    def forward(self, image_features, question_wids, answer_aids):
        """cx.py:261-333. image_features [B, 25, dim_v] (already gathered, as getDataFromBatch returns them)."""
        batch_size = image_features.size(0)
        assert image_features.size(1) == self.knn_size + 1                     # cx.py:263
        feats = image_features.to(self.device, torch.float32).contiguous().view(batch_size * (self.knn_size + 1), -1)
        img_idx = torch.arange(batch_size * (self.knn_size + 1), dtype=torch.int32, device=self.device).view(batch_size, -1)
        q_emb = self.vqa_model.seq2vec(question_wids)                           # cx.py:71 (boundary: encoder output)
        return self.forward_indexed(feats, img_idx, q_emb, answer_aids.to(self.device), resident=False)

    def forward_indexed(self, features, img_idx, q_emb, answer_aids, resident=True):
        """Same scores from the device-resident feature matrix: img_idx[b] = [image row, 24 neighbour rows] (int32).
        resident=False: `features` is a transient gathered batch (the reference-shaped call), not worth a per-image
        table: general path."""
        if not self.training and resident and self._fast_ok(features):
            return self._forward_fast(features, img_idx, q_emb, answer_aids)
        batch = self._make_batch(features, img_idx, q_emb, answer_aids, self.training, resident=resident)
        if self.training and torch.is_grad_enabled():
            q_in = q_emb if (self.trainable_vqa and torch.is_tensor(q_emb) and q_emb.requires_grad) else None
            return _NcxFunction.apply(self.flat, self.head_flat if self.trainable_vqa else None, q_in, self, batch)
        if not self.training:
            self.last_eval_path = "tf32"
        return self._forward_raw(batch, self.training)

    # ---- fused steps: everything on the device, no host synchronisation -------------------------
    def ce_recall(self, scores, comp_idxs, dscores=None, grad_scale=1.0):
        """Accumulates sum-CE into self.metrics[0] and Recall@1/@5 hit counts into self.hits (device)."""
        check(lib.vqacx_ce_recall(ptr(scores), ptr(comp_idxs), scores.shape[0], self.knn_size, ptr(self.metrics),
                                  ptr(dscores), float(grad_scale), ptr(self.hits), current_stream()), "vqacx_ce_recall")

    def reset_metrics(self):
        self.metrics.zero_()
        self.hits.zero_()

    def adam_step(self, grads, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, grad_scale=1.0):
        """torch.optim.Adam(cx_model.parameters(), lr) + optimizer.step() (counterexamples.py:275-276,339)."""
        if self._exp_avg is None:
            self._exp_avg = torch.zeros_like(self.flat.data)
            self._exp_avg_sq = torch.zeros_like(self.flat.data)
        self._step += 1
        self._own_version += 1
        check(lib.vqacx_adam_step(ptr(self.flat), ptr(grads), ptr(self._exp_avg), ptr(self._exp_avg_sq), self.flat.numel(),
                                  float(lr), float(betas[0]), float(betas[1]), float(eps), self._step, float(grad_scale),
                                  current_stream()), "vqacx_adam_step")

    def adam_step_head(self, hgrads, lr=1e-4, betas=(0.9, 0.999), eps=1e-8):
        """The same Adam step over the trainable head's flat buffer (call after adam_step: shares its step count)."""
        if self._head_exp_avg is None:
            self._head_exp_avg = torch.zeros_like(self.head_flat.data)
            self._head_exp_avg_sq = torch.zeros_like(self.head_flat.data)
        self._frozen_version += 1
        check(lib.vqacx_adam_step(ptr(self.head_flat), ptr(hgrads), ptr(self._head_exp_avg), ptr(self._head_exp_avg_sq),
                                  self.head_flat.numel(), float(lr), float(betas[0]), float(betas[1]), float(eps),
                                  max(self._step, 1), 1.0, current_stream()), "vqacx_adam_step")

    def train_step(self, features, img_idx, q_emb, answer_aids, comp_idxs, lr=1e-4, grad_hook=None, global_batch=None,
                   question_wids=None):
        """One iteration of the loop at counterexamples.py:322-339 (forward, CE/B, Recall, backward, Adam) as
        CUDA kernels only. Loss sum and hit counts accumulate on the device (self.metrics / self.hits).
        grad_hook(grads) runs between backward and Adam (data-parallel all-reduce); global_batch is the `len(batch)`
        the loss is divided by (counterexamples.py:334) when this rank holds only a slice of it. Returns the scores."""
        s2v_state = None
        if self.trainable_vqa and getattr(self, "trains_seq2vec", False) and question_wids is not None:
            # the encoder runs in train mode for this step (its q_emb replaces the caller's) and is trained below
            q_emb, s2v_state = self.vqa_model.seq2vec._forward_train_raw(question_wids)
        batch = self._make_batch(features, img_idx, q_emb, answer_aids, True, resident=True)
        if batch.n == 0:   # a rank whose slice of the last partial batch is empty still joins the all-reduce
            grads = torch.zeros_like(self.flat.data)
            if grad_hook is not None:
                grads = grad_hook(grads)
            self.adam_step(grads, lr=lr)
            if self.trainable_vqa:
                hgrads = torch.zeros_like(self.head_flat.data)
                self.adam_step_head(grad_hook(hgrads) if grad_hook is not None else hgrads, lr=lr)
            return torch.empty((0, self.knn_size), dtype=torch.float32, device=self.device)
        scores = self._forward_raw(batch, True)
        dscores = torch.empty_like(scores)
        self.ce_recall(scores, comp_idxs, dscores, 1.0 / (global_batch or batch.n))
        grads = self._backward_raw(batch, dscores)
        hgrads = None
        if self.trainable_vqa:
            grads, hgrads = grads
        if grad_hook is not None:
            grads = grad_hook(grads)
            if hgrads is not None:
                hgrads = grad_hook(hgrads)
        self.adam_step(grads, lr=lr)
        if hgrads is not None:
            self.adam_step_head(hgrads, lr=lr)
            self.last_head_grads = hgrads
        if s2v_state is not None:
            s2v = self.vqa_model.seq2vec
            sgrads = s2v._backward_raw(s2v_state, self.last_dq_emb)
            if grad_hook is not None:
                sgrads = grad_hook(sgrads)
            if self._s2v_exp_avg is None:
                self._s2v_exp_avg = torch.zeros_like(s2v.flat.data)
                self._s2v_exp_avg_sq = torch.zeros_like(s2v.flat.data)
            check(lib.vqacx_adam_step(ptr(s2v.flat), ptr(sgrads), ptr(self._s2v_exp_avg), ptr(self._s2v_exp_avg_sq),
                                      s2v.flat.numel(), float(lr), 0.9, 0.999, 1e-8, max(self._step, 1), 1.0,
                                      current_stream()), "vqacx_adam_step")
            self.last_seq2vec_grads = sgrads
        self.last_grads = grads
        self._last_batch = batch
        return scores

    # ---- the same step replayed as CUDA graphs -------------------------------------------------------
    def _capture_train_graph(self, features, B, lr, global_batch, split):
        """Captures one train step (static input buffers, device-side Adam step / dropout-seed counters) as CUDA graphs:
        `fb` = gather + forward + CE/Recall + backward, `opt` = Adam + counter tick (one graph when no gradient hook
        runs in between). ~35 kernel launches + 3 memsets per step become one or two graph launches."""
        from types import SimpleNamespace

        dev = self.device
        st = SimpleNamespace()
        st.img = torch.zeros((B, self.knn_size + 1), dtype=torch.int32, device=dev)
        st.q = torch.zeros((B, self.dim_q), dtype=torch.float32, device=dev)
        st.aid = torch.zeros(B, dtype=torch.int64, device=dev)
        st.comp = torch.zeros(B, dtype=torch.int64, device=dev)
        st.ws = torch.empty(lib.vqacx_ncx_workspace_bytes(ctypes.byref(self.dims), B, 1), dtype=torch.uint8, device=dev)
        st.scores = torch.zeros((B, self.knn_size), dtype=torch.float32, device=dev)
        st.dscores = torch.zeros_like(st.scores)
        st.grads = torch.zeros_like(self.flat.data)
        st.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        st.seed_dev = torch.zeros(1, dtype=torch.int64, device=dev)      # uint64 counter (same bits)
        st.host_step = st.host_calls = -1
        if self._exp_avg is None:
            self._exp_avg = torch.zeros_like(self.flat.data)
            self._exp_avg_sq = torch.zeros_like(self.flat.data)
        st.tab = self._fast_tables(features, B) if self._tables_ok(features) else None   # kept alive with the graph
        st.batch = _Batch(features, st.img, st.q, st.aid, self.drop_p, self.drop_seed, self.drop_mask, st.ws, st.seed_dev,
                          lesion=self.lesion, hv_table=st.tab)   # only the deterministic lesions reach the graph path
        scale = 1.0 / (global_batch or B)

        def fb():
            check(lib.vqacx_ncx_forward(ctypes.byref(self.dims), ctypes.byref(self._frozen.struct), ptr(self.flat),
                                        ctypes.byref(st.batch.struct), ptr(st.ws), st.ws.numel(), 1, ptr(st.scores),
                                        current_stream()), "vqacx_ncx_forward")
            check(lib.vqacx_ce_recall(ptr(st.scores), ptr(st.comp), B, self.knn_size, ptr(self.metrics), ptr(st.dscores),
                                      float(scale), ptr(self.hits), current_stream()), "vqacx_ce_recall")
            check(lib.vqacx_ncx_backward(ctypes.byref(self.dims), ptr(self.flat), ctypes.byref(st.batch.struct), ptr(st.ws),
                                         st.ws.numel(), ptr(st.dscores), ptr(st.grads), current_stream()),
                  "vqacx_ncx_backward")

        def opt(p, g, m, v, n, step_dev, seed_dev):
            check(lib.vqacx_adam_step_dev(ptr(p), ptr(g), ptr(m), ptr(v), n, float(lr), 0.9, 0.999, 1e-8, ptr(step_dev),
                                          1.0, current_stream()), "vqacx_adam_step_dev")
            check(lib.vqacx_step_tick(ptr(step_dev), ptr(seed_dev), 16, current_stream()), "vqacx_step_tick")

        def opt_real():
            opt(self.flat, st.grads, self._exp_avg, self._exp_avg_sq, self.flat.numel(), st.step_dev, st.seed_dev)

        # warm-up outside capture (first-launch work: function attributes, module loading) without touching any state:
        # forward / backward on the zeroed static inputs with the metrics saved, Adam on a scratch buffer
        keep_m, keep_h = self.metrics.clone(), self.hits.clone()
        fb()
        d4 = [torch.zeros(8, dtype=torch.float32, device=dev) for _ in range(4)]
        opt(d4[0], d4[1], d4[2], d4[3], 8, torch.ones(1, dtype=torch.int32, device=dev),
            torch.zeros(1, dtype=torch.int64, device=dev))
        torch.cuda.synchronize()
        self.metrics.copy_(keep_m)
        self.hits.copy_(keep_h)

        def capture(fn):
            g = torch.cuda.CUDAGraph()
            n0 = lib.vqacx_launch_count()
            with torch.cuda.graph(g):
                fn()
            return g, int(lib.vqacx_launch_count() - n0)

        if split:
            st.g_fb, st.n_fb = capture(fb)
            st.g_opt, st.n_opt = capture(opt_real)
        else:
            st.g_fb, st.n_fb = capture(lambda: (fb(), opt_real()))
            st.g_opt, st.n_opt = None, 0
        return st

    def train_step_graph(self, features, img_idx, q_emb, answer_aids, comp_idxs, lr=1e-4, grad_hook=None,
                         global_batch=None, question_wids=None):
        """train_step with the kernels replayed from captured CUDA graphs (same arithmetic, same dropout stream, same
        Adam state). Inputs are copied into static device buffers; the returned scores tensor is overwritten by the
        next call. Falls back to train_step for an empty slice, when explicit dropout masks are injected, or when a
        lesion draws fresh random inputs every step (v_emb, v_rank)."""
        B = int(img_idx.shape[0])
        if (B == 0 or self.drop_mask is not None or self.lesion_v_emb or (self.lesion & _lib.VQACX_LESION_V_RANK)
                or self.trainable_vqa):
            return self.train_step(features, img_idx, q_emb, answer_aids, comp_idxs, lr=lr, grad_hook=grad_hook,
                                   global_batch=global_batch, question_wids=question_wids)
        if self._exp_avg is None:
            self._exp_avg = torch.zeros_like(self.flat.data)
            self._exp_avg_sq = torch.zeros_like(self.flat.data)
        key = (features.data_ptr(), features.stride(0), B, float(lr), global_batch or B, grad_hook is not None,
               self.drop_seed, self.drop_p, self._frozen.struct.wv, self.flat.data_ptr(),
               self._exp_avg.data_ptr(), features._version, self._frozen_version, self.use_image_tables)
        if not hasattr(self, "_graphs"):
            self._graphs = {}
        st = self._graphs.get(key)
        if st is None:
            if len(self._graphs) >= 4:
                self._graphs.clear()
            st = self._graphs[key] = self._capture_train_graph(features, B, lr, global_batch, grad_hook is not None)
        st.img.copy_(img_idx)
        st.q.copy_(q_emb.detach())
        st.aid.copy_(answer_aids)
        st.comp.copy_(comp_idxs)
        if st.host_step != self._step or st.host_calls != self._drop_calls:   # eager steps ran in between: re-sync
            st.step_dev.fill_(self._step + 1)
            st.seed_dev.fill_(16 * self._drop_calls)
        self._own_version += 1
        st.g_fb.replay()
        lib.vqacx_launch_count_add(st.n_fb)
        if st.g_opt is not None:
            grad_hook(st.grads)
            st.g_opt.replay()
            lib.vqacx_launch_count_add(st.n_opt)
        self._step += 1
        self._drop_calls += 1
        st.host_step, st.host_calls = self._step, self._drop_calls
        self.last_grads = st.grads
        self._last_batch = st.batch
        return st.scores

    def hidden_of_last_train_step(self, layer):
        """[B, 24, dim_h] post-relu/dropout activations of hidden layer `layer` (1-based) of the last train_step
        (inspection hook for the parity tests: which side of the relu each unit took)."""
        b = self._last_batch
        out = torch.empty((b.n, self.knn_size, self.dim_h), dtype=torch.float32, device=self.device)
        check(lib.vqacx_ncx_hidden(ctypes.byref(self.dims), b.n, 1, ptr(b.ws), b.ws.numel(), int(layer), ptr(out),
                                   current_stream()), "vqacx_ncx_hidden")
        return out

    def eval_batch(self, features, img_idx, q_emb, answer_aids, comp_idxs):
        """One iteration of eval_model's loop (counterexamples.py:461-467), metrics accumulated on the device."""
        if self._fast_ok(features):
            scores = self._forward_fast(features, img_idx, q_emb, answer_aids)
        else:
            batch = self._make_batch(features, img_idx, q_emb, answer_aids, False, resident=True)
            scores = self._forward_raw(batch, False)
            self.last_eval_path = "tf32"
        self.ce_recall(scores, comp_idxs)
        return scores


class DistanceBaseline(nn.Module):
    """cx.py:33-44: scores [23, 22, ..., 0] for every question (sanity anchor for the Recall kernel)."""

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

    def eval_batch(self, features, img_idx, q_emb, answer_aids, comp_idxs):
        n = int(img_idx.shape[0])
        scores = torch.arange(self.knn_size - 1, -1, -1, dtype=torch.float32, device=self.device).view(1, -1).expand(
            n, self.knn_size).contiguous()
        check(lib.vqacx_ce_recall(ptr(scores), ptr(comp_idxs), n, self.knn_size, ptr(self.metrics), None, 1.0,
                                  ptr(self.hits), current_stream()), "vqacx_ce_recall")
        return scores

    def forward(self, image_features, question_wids, answer_aids):
        batch_size = image_features.size(0)
        return torch.arange(self.knn_size - 1, -1, -1, dtype=torch.float32, device=image_features.device).view(1, -1).expand(
            batch_size, self.knn_size).contiguous()
