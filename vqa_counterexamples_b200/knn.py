"""Stage 1 of the VQA-CX hot path: exact k-nearest-neighbour retrieval over pooled image features.

Mirrors the reference entry point `knn.py` (CLI flags knn.py:11-21, flow knn.py:23-58) and the
sklearn-shaped pair it relies on: `NearestNeighbors(n_neighbors=k).fit(X)` / `.kneighbors(Q)`
(knn.py:41-42,49-51). All arithmetic runs in libvqacx.so (tcgen05 GEMM + fused top-k' + fp64
re-rank, see csrc/knn_gemm.cu); this file is host plumbing: device buffers, H2D/D2H, sharding.
"""
import argparse
import ctypes
import os

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, ptr, current_stream


class _Prepped:
    """Device-resident float32 rows plus the side arrays the tensor path needs."""

    def __init__(self, x_f32, max_bits=None, parts=None):
        assert x_f32.is_cuda and x_f32.dtype == torch.float32 and x_f32.dim() == 2 and x_f32.is_contiguous()
        n, d = x_f32.shape
        self.f32 = x_f32
        self.n, self.d = n, d
        dev = x_f32.device
        self.bad_checked = False
        if parts is not None:      # side arrays already produced by an upstream kernel (extract.PooledIndex)
            self.lp, self.norm_sq, self.max_bits, self.bad = parts
            return
        self.lp = torch.empty((n, d), dtype=torch.float16, device=dev)
        self.norm_sq = torch.empty((max(n, 1),), dtype=torch.float32, device=dev)
        self.max_bits = max_bits if max_bits is not None else torch.zeros(1, dtype=torch.int32, device=dev)
        self.bad = torch.zeros(1, dtype=torch.int32, device=dev)
        self.bad_checked = False
        if n > 0:
            check(
                lib.vqacx_knn_prep(ptr(x_f32), n, d, ptr(self.lp), ptr(self.norm_sq), ptr(self.max_bits),
                                   ptr(self.bad), current_stream()),
                "vqacx_knn_prep",
            )


def _any_unrepresentable(*prepped):
    """True if any row of these matrices is not representable in fp16 (|x| > 65504 or non-finite), in which case the
    tensor path would rank garbage. Read AFTER a search has synchronised the stream anyway (one 4-byte copy), so `fit`
    itself never blocks the host; each matrix is read once."""
    found = False
    for p in prepped:
        if p is not None and not p.bad_checked:
            p.bad_checked = True
            p.is_bad = int(p.bad.item()) != 0
        found = found or (p is not None and getattr(p, "is_bad", False))
    return found


def _as_device_f32(x, device):
    if isinstance(x, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
        return t.to(device, non_blocking=False)
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=torch.float32).contiguous()
    raise TypeError(f"expected numpy array or torch tensor, got {type(x)}")


def _to_host(*tensors):
    """Device results -> numpy through pinned staging blocks (torch's caching host allocator: a block goes back to the
    cache when the returned array is dropped, so a steady-state caller pays DMA time only, ~1 ms for [123287, 25] x 2
    instead of ~7 ms of pageable copies). Both copies are queued before the one synchronisation."""
    host = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in tensors]
    for h, t in zip(host, tensors):
        h.copy_(t, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return tuple(h.numpy() for h in host)


class NearestNeighbors:
    """Drop-in for `sklearn.neighbors.NearestNeighbors` as used by the reference (knn.py:41-51).

    Only what that call site exercises is supported: brute-force Euclidean search,
    `fit(X)` then `kneighbors(Q) -> (distances float64 [n,k], indices int64 [n,k])`, rows sorted by
    increasing distance (ties by index). numpy in -> numpy out; CUDA tensors in -> CUDA tensors out.
    """

    def __init__(self, n_neighbors=5, *, device=None, force_exact=False, row_offset=0, symmetric=True):
        self.n_neighbors = int(n_neighbors)
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.force_exact = bool(force_exact)
        # self-joins (queries are the fitted rows) use the symmetric sweep; False forces the plain one (A/B, tests)
        self.symmetric = bool(symmetric)
        self.row_offset = int(row_offset)  # global index of local row 0 (row-sharded databases)
        self._db = None
        self._fit_src = None
        self._ws = None
        self.last_stats = None
        self.profile = False  # True: record CUDA events around each phase -> last_phase_ms
        self.last_phase_ms = None

    # -- sklearn API -------------------------------------------------------------------------
    def fit(self, X, y=None):
        with torch.cuda.device(self.device):
            self._db = _Prepped(_as_device_f32(X, self.device))
            self._fit_src = X if isinstance(X, np.ndarray) else None  # knn.py queries slices of the fitted array
        return self   # no host sync here: the fp16 range check is read after the first search (_check_representable)

    def fit_prepared(self, f32, lp, norm_sq, max_bits, bad):
        """fit() from rows whose fp16 copy / norms were already produced (vqacx_pool_features): no pass over X."""
        with torch.cuda.device(self.device):
            self._db = _Prepped(f32, parts=(lp, norm_sq, max_bits, bad))
            self._fit_src = None
        return self

    def fit_kneighbors(self, X, n_neighbors=None, n_chunks=None, return_distance=True):
        """`fit(X)` followed by `kneighbors(X)` -- the whole knn.py workload (knn.py:41-56) -- as ONE pipelined call for a
        HOST matrix: the upload is cut into `n_chunks` pieces and the search of the rows that have landed overlaps the
        copy of the rest (vqacx_knn_selfjoin_host). Same results as the two calls; the instance is fitted afterwards.
        Device tensors (nothing to upload), force_exact and feature counts the tensor path does not take fall back
        to the two calls."""
        k = self.n_neighbors if n_neighbors is None else int(n_neighbors)
        if (not isinstance(X, np.ndarray)) or self.force_exact or X.ndim != 2 or X.shape[1] % 64 != 0 or X.shape[0] == 0 \
                or self.row_offset != 0:
            return self.fit(X).kneighbors(X, n_neighbors=k, return_distance=return_distance)
        if k > _lib.VQACX_KNN_MAX_K:
            raise ValueError(f"n_neighbors={k} exceeds the supported maximum {_lib.VQACX_KNN_MAX_K}")
        if k > X.shape[0]:
            raise ValueError(f"Expected n_neighbors <= n_samples_fit, but n_neighbors = {k}, n_samples_fit = {X.shape[0]}")
        xh = torch.from_numpy(np.ascontiguousarray(X, dtype=np.float32))
        n, d = xh.shape
        chunks = int(n_chunks) if n_chunks is not None else int(os.environ.get("VQACX_KNN_UPLOAD_CHUNKS", "2"))
        with torch.cuda.device(self.device):
            dev = self.device
            f32 = torch.empty((n, d), dtype=torch.float32, device=dev)
            lp = torch.empty((n, d), dtype=torch.float16, device=dev)
            norm_sq = torch.empty((n,), dtype=torch.float32, device=dev)
            max_bits = torch.zeros(1, dtype=torch.int32, device=dev)
            bad = torch.zeros(1, dtype=torch.int32, device=dev)
            dist = torch.empty((n, k), dtype=torch.float64, device=dev)
            idx = torch.empty((n, k), dtype=torch.int64, device=dev)
            need = lib.vqacx_knn_selfjoin_workspace_bytes(n, d, chunks)
            ws = torch.empty(need, dtype=torch.uint8, device=dev)
            stats = (ctypes.c_int32 * _lib.VQACX_KNN_NSTATS)()
            st = lib.vqacx_knn_selfjoin_host(ctypes.c_void_p(xh.data_ptr()), n, d, k, chunks, ptr(f32), ptr(lp), ptr(norm_sq),
                                             ptr(max_bits), ptr(bad), ptr(dist), ptr(idx), ptr(ws), ws.numel(), stats,
                                             current_stream())
            if st == _lib.VQACX_ERANGE:     # rows outside the fp16 range: the two calls, which fall back to the exact kernel
                del f32, lp, dist, idx, ws
                return self.fit(X).kneighbors(X, n_neighbors=k, return_distance=return_distance)
            check(st, "vqacx_knn_selfjoin_host")
            self._db = _Prepped(f32, parts=(lp, norm_sq, max_bits, bad))
            self._fit_src = X
            self.last_stats = {"flagged_rows": stats[0], "unresolved_rows": stats[1], "bound_violations": stats[2],
                               "tensor_path": bool(stats[3]),
                               "rerank_evaluated_avg": round(stats[4] / max(1, (n + 15) // 16), 2)}
            if stats[5]:   # symmetric sweep: column-side (mirror) keys per query, 1-in-16 sample
                self.last_stats["mirror_keys_avg"] = round(stats[5] / max(1, (n + 15) // 16), 2)
            _warn_inexact(self.last_stats)
            dist, idx = _to_host(dist, idx)
        return (dist, idx) if return_distance else idx

    def kneighbors(self, X=None, n_neighbors=None, return_distance=True):
        if self._db is None:
            raise RuntimeError("This NearestNeighbors instance is not fitted yet.")
        k = self.n_neighbors if n_neighbors is None else int(n_neighbors)
        want_numpy = X is None or isinstance(X, np.ndarray)
        with torch.cuda.device(self.device):
            if X is None:
                # sklearn semantics: the query set is the training set and each point is not its own neighbour
                dist, idx = self._search(self._db, k + 1)
                dist, idx = _drop_self(dist, idx, self.row_offset)
            else:
                if X is self._fit_src:
                    q = self._db.f32  # same host array that was fitted: already resident, no second upload
                else:
                    q = _as_device_f32(X, self.device)
                if q.dim() != 2 or q.shape[1] != self._db.d:
                    raise ValueError(f"query has {q.shape[1]} features, index has {self._db.d}")
                same = q.data_ptr() == self._db.f32.data_ptr() and q.shape == self._db.f32.shape
                qp = self._db if same else _Prepped(q, max_bits=self._db.max_bits.clone())
                dist, idx = self._search(qp, k)
        if want_numpy:
            dist, idx = _to_host(dist, idx)
        return (dist, idx) if return_distance else idx

    # -- device-side search -------------------------------------------------------------------
    def _search(self, qp, k, phase=0, hi=None):
        """phase 0: the whole search. phase 1 (row-sharded): candidates only, returns the [n_q, k] float32 upper bounds
        of this shard's k best candidates. phase 2: re-rank pruned by `hi` [n_q] (agreed global k-th bound)."""
        db = self._db
        if k > _lib.VQACX_KNN_MAX_K:
            raise ValueError(f"n_neighbors={k} exceeds the supported maximum {_lib.VQACX_KNN_MAX_K}")
        if k > db.n:
            raise ValueError(f"Expected n_neighbors <= n_samples_fit, but n_neighbors = {k}, n_samples_fit = {db.n}")
        dev = db.f32.device
        out_dist = out_idx = None
        if phase != 1:
            out_dist = torch.empty((qp.n, k), dtype=torch.float64, device=dev)
            out_idx = torch.empty((qp.n, k), dtype=torch.int64, device=dev)
        else:
            hi = torch.empty((qp.n, k), dtype=torch.float32, device=dev)
        need = lib.vqacx_knn_workspace_bytes(qp.n, db.n, db.d)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        stats = (ctypes.c_int32 * _lib.VQACX_KNN_NSTATS)()
        # the max norm must cover database AND query rows (error bound uses |x| + |y|max with y in the db only,
        # |x| is recomputed exactly per query) -> database max is what matters; qp.max_bits is a superset.
        flags = _lib.VQACX_KNN_FORCE_EXACT if self.force_exact else 0
        if not self.symmetric:
            flags |= _lib.VQACX_KNN_NO_SYMMETRIC
        phases = (ctypes.c_float * _lib.VQACX_KNN_NPHASES)() if self.profile else None
        args = (ptr(db.f32), ptr(db.lp), ptr(db.norm_sq), ptr(db.max_bits), db.n, self.row_offset,
                ptr(qp.f32), ptr(qp.lp), ptr(qp.norm_sq), qp.n, db.d, k, ptr(out_dist), ptr(out_idx),
                ptr(self._ws), self._ws.numel(), flags, stats, phases, current_stream())
        if phase == 0:
            check(lib.vqacx_knn_search(*args), "vqacx_knn_search")
            if not (flags & _lib.VQACX_KNN_FORCE_EXACT) and _any_unrepresentable(db, qp):
                # sklearn takes any finite float32; rows outside the fp16 range (|x| > 65504) cannot use the tensor-core
                # candidate pass, so the whole search is redone by the exact fp32 kernel (never a silent wrong answer)
                import warnings

                warnings.warn("vqacx kNN: features not representable in fp16 (|x| > 65504 or non-finite); "
                              "searching with the exact fp32 kernel instead of the tensor-core path", RuntimeWarning,
                              stacklevel=3)
                args = args[:16] + (flags | _lib.VQACX_KNN_FORCE_EXACT,) + args[17:]
                check(lib.vqacx_knn_search(*args), "vqacx_knn_search")
        else:
            check(lib.vqacx_knn_search_phase(*args, phase, ptr(hi)), "vqacx_knn_search_phase")
            if phase == 2 and not self.force_exact and _any_unrepresentable(db, qp):
                raise _lib.VqacxError("features are not representable in fp16 (|x| > 65504 or non-finite); run the "
                                      "row-sharded search with force_exact=True for such data")
        if phase == 1:
            self._phase1_arr = phases      # filled by phase 2 (phase 1 does not synchronise); kept alive here
            return hi
        self.last_stats = {
            "flagged_rows": stats[0], "unresolved_rows": stats[1], "bound_violations": stats[2],
            "tensor_path": bool(stats[3]),
            # candidates (of 64) whose float64 distance was evaluated, average over a 1-in-16 sample of the queries
            "rerank_evaluated_avg": round(stats[4] / max(1, (int(qp.n) + 15) // 16), 2),
        }
        if stats[5]:   # symmetric self-join: column-side (mirror) keys appended per query, same 1-in-16 sample
            self.last_stats["mirror_keys_avg"] = round(stats[5] / max(1, (int(qp.n) + 15) // 16), 2)
        if phases is not None:
            p1 = getattr(self, "_phase1_arr", None) if phase == 2 else None
            c0, f0 = (p1[0], p1[1]) if p1 is not None else (0.0, 0.0)
            self.last_phase_ms = {"candidates": phases[0] + c0, "finalize": phases[1] + f0, "rerank": phases[2],
                                  "fallback": phases[3]}
        _warn_inexact(self.last_stats)
        return out_dist, out_idx

    def _prep_queries(self, q):
        q = _as_device_f32(q, self.device)
        same = q.data_ptr() == self._db.f32.data_ptr() and q.shape == self._db.f32.shape
        return self._db if same else _Prepped(q, max_bits=self._db.max_bits.clone())


def _warn_inexact(stats):
    """The exactness proof failed for some rows and the exact fallback could not settle them either (more than k' - k
    exact ties at the k-th distance: membership / order among the tied rows is then arbitrary, as it is in sklearn), or
    the error model was violated: never silent."""
    if stats and (stats.get("unresolved_rows", 0) > 0 or stats.get("bound_violations", 0) > 0):
        import warnings

        warnings.warn("vqacx kNN: {} row(s) with unresolved exact ties at the k-th distance, {} error-model violation(s); "
                      "those rows were recomputed by the exact fp32 path".format(stats.get("unresolved_rows", 0),
                                                                                 stats.get("bound_violations", 0)),
                      RuntimeWarning, stacklevel=3)


def _drop_self(dist, idx, row_offset):
    """Remove each row's own index (or, if duplicates displaced it, the last column) -> k columns."""
    n, k1 = idx.shape
    own = torch.arange(n, device=idx.device, dtype=idx.dtype).unsqueeze(1) + row_offset
    is_self = idx == own
    has_self = is_self.any(dim=1, keepdim=True)
    drop = torch.where(has_self, is_self, torch.zeros_like(is_self))
    drop[:, -1] |= ~has_self.squeeze(1)
    keep = ~drop
    return dist[keep].view(n, k1 - 1), idx[keep].view(n, k1 - 1)


def merge_topk(dist_g, idx_g):
    """[G, n, k] per-shard results -> global [n, k] by (distance, index). Device tensors."""
    G, n, k = idx_g.shape
    out_d = torch.empty((n, k), dtype=torch.float64, device=idx_g.device)
    out_i = torch.empty((n, k), dtype=torch.int64, device=idx_g.device)
    check(lib.vqacx_topk_merge(ptr(dist_g.contiguous()), ptr(idx_g.contiguous()), G, n, k, ptr(out_d), ptr(out_i),
                               current_stream()), "vqacx_topk_merge")
    return out_d, out_i


def merge_topk_checked(dist_g, idx_g, prepped, k, row_ids=None):
    """Merge of the block-partitioned self-join's per-rank lists (dist_g [G, n, k + 1] with the exclusion bound in the
    last column, idx_g [G, n, k]) + completion of the exactness proof (vqacx_topk_merge_checked). Rows whose merged k-th
    distance does not lie below every rank's exclusion bound are recomputed by the exact fp32 search over the whole
    matrix `prepped` (their global query row = row_ids[r], default r). Returns (dist [n, k], idx [n, k], n_redone)."""
    G, n, k1 = dist_g.shape
    assert k1 == k + 1 and tuple(idx_g.shape) == (G, n, k)
    dev = idx_g.device
    out_d = torch.empty((n, k), dtype=torch.float64, device=dev)
    out_i = torch.empty((n, k), dtype=torch.int64, device=dev)
    fail_count = torch.zeros(1, dtype=torch.int32, device=dev)
    fail_rows = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    check(lib.vqacx_topk_merge_checked(ptr(dist_g.contiguous()), ptr(idx_g.contiguous()), G, n, k, ptr(out_d), ptr(out_i),
                                       ptr(fail_count), ptr(fail_rows), current_stream()), "vqacx_topk_merge_checked")
    n_fail = int(fail_count.item())
    if n_fail:
        rows = fail_rows[:n_fail].long()
        q_rows = rows if row_ids is None else row_ids.index_select(0, rows)
        ex = NearestNeighbors(n_neighbors=k, device=dev, force_exact=True)
        ex._db = prepped
        d_x, i_x = ex._search(_Prepped(prepped.f32.index_select(0, q_rows)), k)
        out_d.index_copy_(0, rows, d_x)
        out_i.index_copy_(0, rows, i_x)
    return out_d, out_i, n_fail


def shard_rows(n, rank, world):
    """Contiguous database rows [row0, row1) owned by `rank` (row0 = rank * ceil(n / world))."""
    per = (n + world - 1) // world
    return min(n, rank * per), min(n, (rank + 1) * per)


def _local_topk_cuda(local_rows, row0, queries, k, force_exact=False, profile=False):
    nn_ = NearestNeighbors(n_neighbors=k, device=local_rows.device, force_exact=force_exact, row_offset=row0)
    nn_.profile = profile
    if local_rows.shape[0] == 0:  # more ranks than rows: this shard contributes only padding
        nq = queries.shape[0]
        return (torch.empty((nq, 0), dtype=torch.float64, device=local_rows.device),
                torch.empty((nq, 0), dtype=torch.int64, device=local_rows.device), nn_)
    nn_.fit(local_rows)
    kk = min(k, nn_._db.n)
    d_loc, i_loc = nn_.kneighbors(queries, n_neighbors=kk)
    return d_loc, i_loc, nn_


def kth_bound(hi_all):
    """[G, n, k] float32 phase-1 bounds of all shards -> [n] upper bound on the global k-th distance (device)."""
    G, n, k = hi_all.shape
    out = torch.empty(n, dtype=torch.float32, device=hi_all.device)
    check(lib.vqacx_knn_kth_bound(ptr(hi_all.contiguous()), G, n, k, ptr(out), current_stream()), "vqacx_knn_kth_bound")
    return out


def _all_gather(t, world, group):
    import torch.distributed as dist

    out = torch.empty((world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
    if t.is_cuda:
        dist.all_gather_into_tensor(out, t.contiguous(), group=group)
    else:  # gloo has no all_gather_into_tensor
        dist.all_gather(list(out.unbind(0)), t.contiguous(), group=group)
    return out


def _pad_cols(t, k, value):
    if t.shape[1] >= k:
        return t
    pad = torch.full((t.shape[0], k - t.shape[1]), value, dtype=t.dtype, device=t.device)
    return torch.cat([t, pad], 1)


def _pad_rows(t, n, value):
    if t.shape[0] >= n:
        return t
    pad = torch.full((n - t.shape[0],) + tuple(t.shape[1:]), value, dtype=t.dtype, device=t.device)
    return torch.cat([t, pad], 0)


def _exchange_slices(t, world, group, fill, n_rows=None):
    """all_to_all of a per-rank [n_q, k] result: the queries are cut into `world` contiguous slices of
    per = ceil(n_q / world) rows (shard_rows), rank g receives everybody's rows of slice g -> [world, per, k] on every
    rank (rows past n_q are `fill`). 1/world of the bytes an all_gather of the full results moves."""
    import torch.distributed as dist

    per = ((t.shape[0] if n_rows is None else n_rows) + world - 1) // world
    send = _pad_rows(t, per * world, fill).contiguous()      # no copy when the producer allocated the padding rows
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)
    return recv.view(world, per, *t.shape[1:])


class _CudaShard:
    """The two halves of the row-sharded search on one rank (C ABI vqacx_knn_search_phase)."""

    def __init__(self, local_rows, row0, queries, k, force_exact=False, profile=False):
        self.nn = NearestNeighbors(n_neighbors=k, device=local_rows.device, force_exact=force_exact, row_offset=row0)
        self.nn.profile = profile
        self.nq, self.k, self.dev = queries.shape[0], k, local_rows.device
        self.empty = local_rows.shape[0] == 0       # more ranks than rows: this shard contributes only padding
        if not self.empty:
            self.nn.fit(local_rows)
            self.kk = min(k, self.nn._db.n)
            self.qp = self.nn._prep_queries(queries)

    def phase1(self):
        if self.empty:
            return torch.full((self.nq, self.k), float("inf"), dtype=torch.float32, device=self.dev)
        return _pad_cols(self.nn._search(self.qp, self.kk, phase=1), self.k, float("inf"))

    def phase2(self, bound):
        if self.empty:
            return (torch.full((self.nq, self.k), float("inf"), dtype=torch.float64, device=self.dev),
                    torch.full((self.nq, self.k), -1, dtype=torch.int64, device=self.dev))
        d, i = self.nn._search(self.qp, self.kk, phase=2, hi=bound)
        return _pad_cols(d, self.k, float("inf")), _pad_cols(i, self.k, -1)


class _Timeline:
    """Optional per-step timeline of the sharded search (profile=True): a CUDA event + a host timestamp at every stage
    boundary. `ms()` -> {stage: (device ms since the previous mark, host ms since the previous mark)}; reading it
    synchronises, so it is called after the step."""

    def __init__(self, on):
        self.on, self.marks = on, []
        if on:
            self.mark("start")

    def mark(self, name):
        if self.on:
            import time

            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self.marks.append((name, ev, time.perf_counter()))

    def ms(self):
        if not self.on:
            return None
        torch.cuda.synchronize()
        out = {}
        for (_, e0, t0), (name, e1, t1) in zip(self.marks, self.marks[1:]):
            out[name] = (round(e0.elapsed_time(e1), 4), round((t1 - t0) * 1e3, 4))
        return out


class _SymShard:
    """One rank's share of the block-partitioned SYMMETRIC self-join (C ABI vqacx_knn_selfjoin_shard): the three phases
    around the threshold all_reduce(MIN) and the bound all_gather."""

    def __init__(self, prepped, k, rank, world, profile=False):
        self.p, self.k, self.rank, self.world, self.profile = prepped, int(k), int(rank), int(world), profile
        dev = prepped.f32.device
        need = lib.vqacx_knn_selfjoin_shard_workspace_bytes(prepped.n, prepped.d, world)
        self.ws = torch.empty(need, dtype=torch.uint8, device=dev)
        self.thr = torch.empty(prepped.n, dtype=torch.float32, device=dev)
        self.hi = torch.empty((prepped.n + 1, self.k), dtype=torch.float32, device=dev)
        self.stats = (ctypes.c_int32 * _lib.VQACX_KNN_NSTATS)()
        self.ms = [0.0, 0.0, 0.0, 0.0]
        self.last_stats = self.last_phase_ms = None

    def _call(self, phase, hi, out_d=None, out_i=None):
        p = self.p
        # phases 1 and 2 do not synchronise: their event times are written into these arrays by phase 3 (kept alive here)
        phases = (ctypes.c_float * _lib.VQACX_KNN_NPHASES)() if self.profile else None
        self._phase_arrays[phase] = phases
        check(lib.vqacx_knn_selfjoin_shard(ptr(p.f32), ptr(p.lp), ptr(p.norm_sq), ptr(p.max_bits), p.n, p.d, self.k,
                                           self.rank, self.world, phase, ptr(self.thr), ptr(hi), ptr(out_d), ptr(out_i),
                                           ptr(self.ws), self.ws.numel(), self.stats, phases, current_stream()),
              "vqacx_knn_selfjoin_shard")
        if phase == 3 and self.profile:
            self.ms = [sum(a[j] for a in self._phase_arrays.values() if a is not None) for j in range(4)]

    def phase1(self):
        self.ms = [0.0, 0.0, 0.0, 0.0]
        self._phase_arrays = {}
        self._call(1, None)
        return self.thr

    def phase2(self, thr, want_hi=True):
        if thr.data_ptr() != self.thr.data_ptr():
            self.thr.copy_(thr)
        self._call(2, self.hi if want_hi else None)
        return self.hi if want_hi else None

    def pack(self, rows_alloc):
        """This rank's per-query candidate lists as exchange rows [rows_alloc, 130] int32 (after phase2(want_hi=False))."""
        p = self.p
        out = torch.empty((rows_alloc, _lib.VQACX_KNN_PACK_INTS), dtype=torch.int32, device=p.f32.device)
        check(lib.vqacx_knn_selfjoin_shard_pack(p.n, p.d, self.world, rows_alloc, ptr(out), ptr(self.ws), self.ws.numel(),
                                                current_stream()), "vqacx_knn_selfjoin_shard_pack")
        return out

    def owner(self, recv, per, q0, n_slice):
        """recv [world, per, 130]: every rank's lists for this rank's query slice -> final (dist, idx) [n_slice, k]."""
        p = self.p
        dev = p.f32.device
        out_d = torch.empty((n_slice, self.k), dtype=torch.float64, device=dev)
        out_i = torch.empty((n_slice, self.k), dtype=torch.int64, device=dev)
        phases = (ctypes.c_float * _lib.VQACX_KNN_NPHASES)() if self.profile else None
        self._phase_arrays[4] = phases
        check(lib.vqacx_knn_selfjoin_shard_owner(ptr(p.f32), ptr(p.norm_sq), ptr(p.max_bits), p.n, p.d, self.k, self.world,
                                                 ptr(recv), per, q0, n_slice, ptr(out_d), ptr(out_i), ptr(self.ws),
                                                 self.ws.numel(), self.stats, phases, current_stream()),
              "vqacx_knn_selfjoin_shard_owner")
        st = self.stats
        self.last_stats = {"flagged_rows": st[0], "unresolved_rows": st[1], "bound_violations": st[2], "tensor_path": True,
                           "rerank_evaluated_avg": round(st[4] / max(1, (n_slice + 15) // 16), 2),
                           "mirror_keys_avg": -1.0, "log_overflow": bool(st[6]),
                           "sweep": "symmetric, block-partitioned, candidate exchange"}
        if self.profile:
            ms = [sum(a[j] for a in self._phase_arrays.values() if a is not None) for j in range(4)]
            self.last_phase_ms = {"candidates": ms[0], "finalize": ms[1], "rerank": ms[2], "fallback": ms[3]}
        _warn_inexact(self.last_stats)
        return out_d, out_i

    def phase3(self, bound, rows_alloc=None):
        """rows_alloc >= n: allocate that many output rows (the tail is exchange padding: +inf / -1) so that the
        all_to_all needs no padding copy."""
        dev = self.p.f32.device
        rows = max(self.p.n, rows_alloc or 0)
        # column k of a distance row = this rank's exclusion bound (the proof is completed after the merge)
        out_d = torch.empty((rows, self.k + 1), dtype=torch.float64, device=dev)
        out_i = torch.empty((rows, self.k), dtype=torch.int64, device=dev)
        if rows > self.p.n:
            out_d[self.p.n:].fill_(float("inf"))
            out_i[self.p.n:].fill_(-1)
        self._call(3, bound, out_d, out_i)
        st, n = self.stats, self.p.n
        self.last_stats = {"flagged_rows": st[0], "unresolved_rows": st[1], "bound_violations": st[2], "tensor_path": True,
                           "rerank_evaluated_avg": round(st[4] / max(1, (n + 15) // 16), 2),
                           "mirror_keys_avg": round(st[5] / max(1, (n + 15) // 16), 2), "log_overflow": bool(st[6]),
                           "sweep": "symmetric, block-partitioned"}
        if self.profile:
            self.last_phase_ms = {"candidates": self.ms[0], "finalize": self.ms[1], "rerank": self.ms[2],
                                  "fallback": self.ms[3]}
        _warn_inexact(self.last_stats)
        return out_d, out_i


def _is_own_slice(local_rows, row0, queries, rank, world):
    """True when `local_rows` is exactly this rank's row slice of `queries` (same storage): a self-join."""
    if not (local_rows.is_cuda and queries.is_cuda and queries.dim() == 2 and queries.is_contiguous()):
        return False
    r0, r1 = shard_rows(queries.shape[0], rank, world)
    return (row0 == r0 and local_rows.shape[0] == r1 - r0 and local_rows.shape[1:] == queries.shape[1:] and
            local_rows.is_contiguous() and local_rows.dtype == queries.dtype == torch.float32 and
            local_rows.data_ptr() == queries.data_ptr() + r0 * queries.stride(0) * 4)


def sharded_kneighbors(local_rows, row0, queries, k, group=None, force_exact=False, profile=False,
                       return_index=False, local_topk=None, merge=None, shard_factory=None, kth=None,
                       result="replicated"):
    """Row-sharded exact kNN: every rank holds `local_rows` = global rows [row0, row0+n_local), all ranks
    hold the same `queries`. Protocol:
      1. each shard generates its candidates and publishes upper bounds hi[n, k] on the true distances of its k best
         (all_gather, float32); the k-th smallest of the G*k gathered bounds is an upper bound on the GLOBAL k-th
         distance of every query
      2. each shard re-ranks (float64) only the candidates that can beat that bound (~k/G rows per query instead of ~k:
         the re-rank, a per-rank cost that did not shrink with G, now scales), entries that cannot are +inf / -1
      3. all_to_all of the [n, k] per-shard results: rank g receives every shard's rows of query slice g
         (shard_rows(n, g, world)) and merges only those by (distance, index) -- 1/G of the NVLink traffic and of the
         merge work of "all_gather everything, merge everything everywhere".
    result = "sliced": returns the merged rows of THIS rank's query slice only ([q1 - q0, k], q0, q1 = shard_rows(n_q, rank,
    world)): the final answer stays distributed by query, which is how N processes each write their part of knn.py's
    output. result = "replicated" (default): one more all_gather so that every rank holds all [n_q, k] rows.
    `shard_factory(local_rows, row0, queries, k)` -> object with phase1() / phase2(bound) and `kth` / `merge` default to
    the CUDA paths; the world_size-2 gloo tests substitute CPU stand-ins to exercise this host protocol (padding of short
    shards, exchange layout, offsets) without a GPU. `local_topk` (single-exchange protocol: per-shard exact top-k, no
    bound exchange) is kept for the same purpose."""
    import torch.distributed as dist

    merge = merge or merge_topk
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    tl = _Timeline(profile and local_rows.is_cuda)
    if local_topk is not None or lib.vqacx_get_option(_lib.OPT_KNN_SINGLE_EXCHANGE) == 1 or world == 1:
        local_topk = local_topk or (lambda rows, r0, q, kk: _local_topk_cuda(rows, r0, q, kk, force_exact, profile))
        d_loc, i_loc, nn_ = local_topk(local_rows, row0, queries, k)
        d_loc, i_loc = _pad_cols(d_loc, k, float("inf")), _pad_cols(i_loc, k, -1)   # tiny (or empty) shard
    else:
        d_loc = None
        if (shard_factory is None and not force_exact and _is_own_slice(local_rows, row0, queries, rank, world) and
                k <= queries.shape[0] and
                lib.vqacx_knn_selfjoin_shard_supported(queries.shape[0], queries.shape[1], world)):
            # self-join (the knn.py workload): every rank holds the whole matrix anyway, so the ranks split the blocks
            # of the SYMMETRIC sweep instead of the database rows -- (groups + 1) / (2 groups) of the tensor work
            # CANDIDATE exchange: the per-rank candidate lists travel to the rank that owns the query (one all_to_all),
            # which merges them and runs the ordinary re-rank + exactness proof on its n / world queries
            sh = nn_ = _SymShard(_Prepped(queries), k, rank, world, profile)
            tl.mark("prep")
            thr = sh.phase1()
            tl.mark("phase1_diagonal")
            dist.all_reduce(thr, op=dist.ReduceOp.MIN, group=group)
            tl.mark("allreduce_thr")
            sh.phase2(thr, want_hi=False)
            n_q = queries.shape[0]
            per_q = (n_q + world - 1) // world
            send = sh.pack(per_q * world)
            tl.mark("phase2_mirrored_finalize_pack")
            recv = torch.empty_like(send)
            dist.all_to_all_single(recv, send, group=group)
            tl.mark("all_to_all_candidates")
            q0, q1 = shard_rows(n_q, rank, world)
            d, i = sh.owner(recv, per_q, q0, q1 - q0)                          # synchronises the stream
            tl.mark("owner_merge_rerank")
            if _any_unrepresentable(sh.p):
                raise _lib.VqacxError("features are not representable in fp16 (|x| > 65504 or non-finite); run the "
                                      "row-sharded search with force_exact=True for such data")
            if tl.on:
                nn_.last_timeline_ms = tl.ms()
            if result == "sliced":
                return d, i, ((lambda: nn_) if return_index else (lambda: nn_.last_stats))()
            d = _all_gather(_pad_rows(d, per_q, float("inf")), world, group).view(-1, k)[:n_q]
            i = _all_gather(_pad_rows(i, per_q, -1), world, group).view(-1, k)[:n_q]
            return d, i, ((lambda: nn_) if return_index else (lambda: nn_.last_stats))()
        if d_loc is None:
            factory = shard_factory or (lambda rows, r0, q, kk: _CudaShard(rows, r0, q, kk, force_exact, profile))
            shard = factory(local_rows, row0, queries, k)
            tl.mark("prep")
            hi = shard.phase1()
            tl.mark("phase1_candidates")
            hi_all = _all_gather(hi, world, group)
            bound = (kth or kth_bound)(hi_all)
            tl.mark("allgather_hi_kth")
            d_loc, i_loc = shard.phase2(bound)
            tl.mark("phase2_rerank")
            nn_ = getattr(shard, "nn", None)
    last = (lambda: nn_) if return_index else (lambda: getattr(nn_, "last_stats", None))
    if world == 1:
        return d_loc, i_loc, last()
    n_q = queries.shape[0]
    d_g = _exchange_slices(d_loc, world, group, float("inf"), n_rows=n_q)
    i_g = _exchange_slices(i_loc, world, group, -1, n_rows=n_q)
    tl.mark("all_to_all")
    if d_loc.shape[1] == k + 1:      # block-partitioned self-join: lists carry their exclusion bound -> checked merge
        q0, q1 = shard_rows(n_q, rank, world)
        per = d_g.shape[1]
        ids = torch.arange(q0, q0 + per, device=d_g.device).clamp_(max=n_q - 1)
        d, i, n_redone = merge_topk_checked(d_g, i_g, nn_.p, k, row_ids=ids)
        nn_.last_stats["rows_redone_after_merge"] = n_redone
    else:
        d, i = merge(d_g, i_g)
    tl.mark("merge")
    if tl.on and nn_ is not None:
        nn_.last_timeline_ms = tl.ms()
    if result == "sliced":
        q0, q1 = shard_rows(n_q, rank, world)
        return d[:q1 - q0], i[:q1 - q0], last()
    d = _all_gather(d, world, group).view(-1, k)[:n_q]
    i = _all_gather(i, world, group).view(-1, k)[:n_q]
    return d, i, last()


def sharded_selfjoin_host(x_local, row0, n_total, k, group=None, device=None, return_index=False):
    """knn.py's workload (knn.py:35-56: load, fit(features), kneighbors over the same rows) on N processes, one per GPU:
    every rank passes ITS rows [row0, row0 + len(x_local)) = shard_rows(n_total, rank, world) of the feature matrix as a
    HOST array and gets back, as host arrays, the neighbours of those same rows among all n_total rows.
    Each row crosses PCIe once, on its owner's link (n_total / world rows per rank instead of every rank uploading the
    whole matrix); the fp32 rows the other ranks' candidate sweeps and re-ranks need travel by ONE NCCL all_gather over
    NVLink; the search is `sharded_kneighbors` with the result left distributed by query; each rank downloads only its
    slice. Returns (distances float64 [n_local, k], indices int64 [n_local, k]) [, the fitted per-rank index]."""
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    r0, r1 = shard_rows(n_total, rank, world)
    if row0 != r0 or x_local.shape[0] != r1 - r0:
        raise ValueError(f"rank {rank} must hold rows [{r0}, {r1}) of the {n_total}-row matrix (shard_rows), "
                         f"got [{row0}, {row0 + x_local.shape[0]})")
    xh = torch.from_numpy(np.ascontiguousarray(x_local, dtype=np.float32)) if isinstance(x_local, np.ndarray) else x_local
    d_feat = int(xh.shape[1])
    per = (n_total + world - 1) // world
    with torch.cuda.device(dev):
        mine = torch.zeros((per, d_feat), dtype=torch.float32, device=dev) if r1 - r0 < per else \
            torch.empty((per, d_feat), dtype=torch.float32, device=dev)
        mine[:r1 - r0].copy_(xh, non_blocking=True)
        if world > 1:
            full = torch.empty((per * world, d_feat), dtype=torch.float32, device=dev)
            dist.all_gather_into_tensor(full, mine, group=group)
        else:
            full = mine
        x_all = full[:n_total]
        d_, i_, info = sharded_kneighbors(x_all[r0:r1], r0, x_all, k, group=group, result="sliced",
                                          return_index=return_index)
        d_h, i_h = _to_host(d_, i_)
    return (d_h, i_h, info) if return_index else (d_h, i_h)


# ------------------------------------------------------------------------------------------------
# CLI: same flags and output file as the reference's knn.py
# ------------------------------------------------------------------------------------------------
parser = argparse.ArgumentParser(description="Exact kNN over pooled image features (B200-native knn.py)")
parser.add_argument("base_dir", type=str, help="path to dir containing extracted hdf5 features")
parser.add_argument("--hdf5_file", type=str, default="trainset.hdf5")
parser.add_argument("--save_dir", type=str, default="data/coco/extract")
parser.add_argument("--save_file", type=str, default="knn_results_trainset.npy")
parser.add_argument("-k", "--k", default=25, type=int)
parser.add_argument("-b", "--batch_size", default=10, type=int,
                    help="kept for CLI compatibility; the GPU path searches all rows in one call")


def load_features(path):
    """`noatt` dataset of an extract.py HDF5 file (extract.py:105-109), or a .npy array of the same shape."""
    if path.endswith(".npy"):
        return np.load(path)
    try:
        import h5py
    except ImportError as e:  # pragma: no cover - h5py is not in this image
        raise ImportError("reading .hdf5 feature files needs h5py; pass a .npy file instead") from e
    with h5py.File(path, "r") as f:
        return np.array(f.get("noatt"))


def main(argv=None):
    args = parser.parse_args(argv)
    load_path = os.path.join(args.base_dir, args.hdf5_file)
    assert os.path.exists(load_path)
    assert os.path.isdir(args.save_dir)
    save_path = os.path.join(args.save_dir, args.save_file)
    if os.path.exists(save_path):
        print("Warning: {} already exists and will be overwritten.".format(save_path))
    print("Saving results to {}".format(save_path))

    features = load_features(load_path)
    print("Loaded features as array of size {}".format(features.shape))

    print("Fitting KNN model for k={}".format(args.k))
    nbrs = NearestNeighbors(n_neighbors=args.k)

    print("Starting KNN computations...")
    # fit(features) + kneighbors(features) in one pipelined call (the upload overlaps the search)
    all_distances, all_indices = nbrs.fit_kneighbors(features)
    print("kNN stats: {}".format(nbrs.last_stats))

    np.save(save_path, {"indices": all_indices, "distances": all_distances})
    return all_distances, all_indices


if __name__ == "__main__":
    main()
