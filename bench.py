#!/usr/bin/env python
"""bench.py — headline benchmark of the VQA-CX hot path on B200 (contract: task statement ④).

Primary metric (BASELINE.json): exact kNN queries/s, k = 25 (self + top-24), database = queries =
synthetic [123 287, 2048] float32 of the COCO feature shape (configs[1]).

  python bench.py --gpus N --steps K --warmup W          our CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference ...                   the reference's own CPU path (knn.py -> scikit-learn)

A step = one full pass: fit (fp16 copy + norms) + kneighbors of every row against the database.
  value : queries/s with the fp32 features already resident in HBM (device-timed, CUDA events)
  e2e   : the same through the public NearestNeighbors API with HOST (pinned) input and HOST output,
          H2D + D2H inside the timed region
N > 1 : database rows sharded over ranks (strong scaling: total work fixed), queries replicated, per-shard
        top-k -> NCCL all_gather -> on-device merge, all inside the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_DB = 123287
DIM = 2048
K = 25
SEED = 1234
METRIC = "knn_queries_per_sec"
UNIT = "queries/s"
WORKLOAD = "exact top-24 kNN (k=25 incl. self), self-join over synthetic 123287x2048 f32 FBResNet-152-shaped features"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--n-db", type=int, default=N_DB, help="override the database size (debugging only)")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the cpu_baseline leg (debugging only)")
    ap.add_argument("--no-ncx", action="store_true", help="skip the NeuralCX secondary numbers")
    ap.add_argument("--ref-queries", type=int, default=0, help="queries per reference step (0 = auto)")
    ap.add_argument("--no-config4", action="store_true", help="skip the BASELINE config 4 block (debugging only)")
    ap.add_argument("--no-config5", action="store_true", help="skip the BASELINE config 5 block (debugging only)")
    ap.add_argument("--cfg4-n-db", type=int, default=1_000_000, help="config 4 database rows (debugging only)")
    ap.add_argument("--cfg4-n-q", type=int, default=131072, help="config 4 queries (debugging only)")
    ap.add_argument("--cfg5-n", type=int, default=214354, help="config 5 questions (debugging only)")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def _sha16(path):
    import hashlib

    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()[:16]


def ncu_traffic_bytes():
    """dram__bytes_read.sum + dram__bytes_write.sum of the candidate phase of one search (the
    knn_gemm_topk_pair_kernel launches; for the symmetric self-join the diagonal + mirrored launches and the scatter
    pass, summed over the file's kernel blocks), from the committed `ncu --set full` summary of this same workload
    named by profiles/traffic_source.json. That file also records the sha256 of csrc/knn_gemm_pair.cu at capture time:
    when the kernel source has changed since, the number is stale and (None, reason) is returned instead."""
    side = os.path.join(ROOT, "profiles", "traffic_source.json")
    if not os.path.exists(side):
        return None, "no profiles/traffic_source.json"
    with open(side) as f:
        meta = json.load(f)
    kernel = os.path.join(ROOT, "vqa_counterexamples_b200", "csrc", "knn_gemm_pair.cu")
    if _sha16(kernel) != meta.get("knn_gemm_pair_cu_sha16"):
        return None, f"stale: {meta.get('csv')} was captured for another revision of knn_gemm_pair.cu"
    path = os.path.join(ROOT, "profiles", meta["csv"])
    vals = {}
    with open(path) as f:
        for ln in f:
            parts = ln.strip().split(",")
            if len(parts) == 3 and parts[0].startswith("dram__bytes_"):
                scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(parts[1])
                if scale:
                    vals[parts[0]] = vals.get(parts[0], 0.0) + float(parts[2]) * scale
    if "dram__bytes_read.sum" in vals and "dram__bytes_write.sum" in vals:
        return vals["dram__bytes_read.sum"] + vals["dram__bytes_write.sum"], meta["csv"]
    return None, f"{meta['csv']} holds no dram__bytes metrics"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def wait_ready(self, timeout=3.0):
        t_end = time.perf_counter() + timeout
        while self.proc is not None and not self.lines and time.perf_counter() < t_end:
            time.sleep(0.01)

    def window(self, t0, t1):
        """Only samples that arrived inside [t0, t1] (host clock) count: the sampler is started before the warm-up
        because nvidia-smi's own start-up (NVML init over every GPU of the box) stalls the devices for ~100 ms."""
        self.t0, self.t1 = t0, t1

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        t0, t1 = getattr(self, "t0", None), getattr(self, "t1", None)
        for ts, ln in self.lines:
            if t0 is not None and not (t0 <= ts <= t1 + 0.25):
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm), "power_w_max": max(power) if power else None}


# ------------------------------------------------------------------------------------------------
# CPU legs (oracle/ is test + baseline infrastructure: only these two functions touch it)
# ------------------------------------------------------------------------------------------------
def host_features(n_db):
    """Synthetic features on the host. With a GPU: generated on device (fast) and copied back, so both arms and
    the e2e leg see the same matrix; without: the numpy generator."""
    import torch
    from vqa_counterexamples_b200 import synth

    if torch.cuda.is_available():
        x = synth.clustered_features_cuda(n_db, DIM, seed=SEED)
        h = torch.empty((n_db, DIM), dtype=torch.float32, pin_memory=True)
        h.copy_(x)
        torch.cuda.synchronize()
        return h, x
    return torch.from_numpy(synth.clustered_features_np(n_db, DIM, seed=SEED)), None


def cpu_threads():
    try:
        from threadpoolctl import threadpool_info

        n = max([p.get("num_threads", 1) for p in threadpool_info()] + [1])
    except Exception:
        n = 1
    return max(n, os.cpu_count() or 1)


def time_reference_stock(x_np, n_queries):
    """knn.py:41-56 with its default batch_size=10 on the first n_queries rows; returns (queries/s, seconds)."""
    from sklearn.neighbors import NearestNeighbors

    nbrs = NearestNeighbors(n_neighbors=K).fit(x_np)
    t0 = time.perf_counter()
    for i in range(0, n_queries, 10):
        nbrs.kneighbors(x_np[i:i + 10, :])
    dt = time.perf_counter() - t0
    return n_queries / dt, dt


def time_reference_onecall(x_np, n_queries):
    from oracle.knn_oracle import knn_sklearn_onecall

    t0 = time.perf_counter()
    knn_sklearn_onecall(x_np, x_np[:n_queries], k=K)
    dt = time.perf_counter() - t0
    return n_queries / dt, dt


def cpu_baseline_leg(x_np):
    import sklearn

    time_reference_onecall(x_np, 64)  # warm the thread pools
    one_qps, one_s = time_reference_onecall(x_np, 1024)
    stock_qps, stock_s = time_reference_stock(x_np, 100)
    return {
        "value": stock_qps, "unit": UNIT, "cores": cpu_threads(), "kind": "reference",
        "sample": (f"knn.py:41-56 call sequence on scikit-learn {sklearn.__version__} (the library the reference "
                   f"calls), default batch_size=10, first 100 queries vs the full {x_np.shape[0]}x{DIM} database "
                   f"({stock_s:.1f} s)"),
        "onecall_value": one_qps,
        "onecall_sample": f"single kneighbors call, 1024 queries vs the full database ({one_s:.1f} s)",
    }


def use_all_host_threads():
    """The reference arm runs with every host thread it can use. torchrun exports OMP_NUM_THREADS=1 to its workers
    (seen at N = 2: 21 instead of 100 queries/s), so the limit is lifted explicitly: in the environment before the
    OpenMP / BLAS runtimes load, and through threadpoolctl / torch for the ones already loaded."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[var] = str(n)
    try:
        import numpy  # noqa: F401  (loads the BLAS / OpenMP runtimes threadpoolctl then finds)
        import sklearn.neighbors  # noqa: F401
        from threadpoolctl import threadpool_limits

        threadpool_limits(limits=n)
    except Exception:
        pass
    try:
        import torch

        torch.set_num_threads(n)
    except Exception:
        pass
    return n


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    use_all_host_threads()
    import numpy as np
    import sklearn

    h, _ = host_features(args.n_db)
    x_np = h.numpy()
    nq = args.ref_queries or 500
    for _ in range(args.warmup):
        time_reference_stock(x_np, min(nq, 20))
    t0 = time.perf_counter()
    for _ in range(args.steps):
        time_reference_stock(x_np, nq)
    dt = time.perf_counter() - t0
    qps = args.steps * nq / dt
    one_qps, _ = time_reference_onecall(x_np, 512)
    line = {
        "impl": "reference", "metric": METRIC, "value": qps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "n_db": args.n_db, "dim": DIM, "k": K,
                   "step": f"{nq} queries through knn.py's batch_size=10 loop against the full database"},
        "cpu_baseline": {"value": qps, "unit": UNIT, "cores": cpu_threads(), "kind": "reference",
                         "sample": f"knn.py:41-56 on scikit-learn {sklearn.__version__}, batch_size=10, {nq} queries/step",
                         "onecall_value": one_qps},
        "e2e": {"value": qps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ------------------------------------------------------------------------------------------------
# NeuralCX secondary numbers (BASELINE configs 1/3/5 shapes: MUTAN no-att dims, H = 300, L = 2, B = 64 x 24)
# ------------------------------------------------------------------------------------------------
NCX = {"dim_v": 2048, "dim_q": 2400, "dim_f": 360, "R": 10, "n_ans": 2000, "dim_h": 300, "n_layers": 2, "batch": 64,
       "eval_batch": 4096, "n_img": 82783, "n_questions": 8192, "drop_p": 0.25}
NCX_TRAIN_FLOP_PER_SAMPLE = 483.2e6     # de-duplicated fwd + bwd
NCX_EVAL_FLOP_PER_SAMPLE = 242.2e6
NCX_TRAIN_BYTES_PER_STEP = 0.40e9       # Adam w/g/m/v + gradient write + weights + gather
REFERENCE_DIR = "/root/reference"       # exists in the build container only, never on the GPU box


def _reference_neural_model():
    """The reference's own NeuralModel + MutanNoAtt (vqa/models/cx.py, noatt.py), imported read-only from
    /root/reference with the two shims oracle/gen_golden_ncx.py documents, at the bench dimensions. Returns None where
    the reference tree does not exist (the GPU box: a Python reference cannot travel, task statement (3))."""
    if not os.path.isdir(os.path.join(REFERENCE_DIR, "vqa", "models")):
        return None
    import numpy as np
    import torch

    from oracle import gen_golden_ncx as G
    from oracle import ncx_oracle as O

    import torch.nn.functional as F

    c = NCX
    cwd = os.getcwd()
    saved = (torch.Tensor.cuda, torch.nn.Module.cuda, F.pairwise_distance)

    def restore():      # the import shims (Tensor.cuda -> identity, ...) must not outlive the CPU leg
        torch.Tensor.cuda, torch.nn.Module.cuda, F.pairwise_distance = saved

    try:
        s2v, NeuralModel, MutanNoAtt = G.import_reference()
    finally:
        os.chdir(cwd)
    params = O.make_params(1, dim_v=c["dim_v"], dim_q=c["dim_q"], dim_h_fusion=c["dim_f"], R=c["R"], n_ans=c["n_ans"],
                           dim_h=c["dim_h"], n_layers=c["n_layers"])
    feats, q_emb, aid, comp = O.make_batch(2, c["batch"], dim_v=c["dim_v"], dim_q=c["dim_q"], n_ans=c["n_ans"])
    dims = {"dim_v": c["dim_v"], "dim_q": c["dim_q"], "dim_hv": c["dim_f"], "dim_hq": c["dim_f"], "dim_mm": c["dim_f"],
            "R": c["R"], "n_ans": c["n_ans"]}
    cx = G.build_reference(s2v, NeuralModel, MutanNoAtt, params, dims, q_emb, c["dim_h"], c["n_layers"], c["drop_p"])
    wids = torch.zeros((c["batch"], 26), dtype=torch.long)
    wids[:, 0] = torch.arange(c["batch"])
    return cx, torch.as_tensor(feats), wids, torch.as_tensor(aid), torch.as_tensor(np.asarray(comp)), restore


def ncx_cpu_baseline():
    """NeuralCX train step / eval forward on the host cores (BASELINE config 1; bounded: 3 steps of batch 64 x 24).
    kind 'reference' = the reference's own vqa.models.cx.NeuralModel imported from /root/reference (build container);
    kind 'port' = oracle/ncx_oracle.py, the restatement pinned against that class by tests/golden/ncx_*.npz, where the
    reference tree is absent (the GPU box)."""
    import torch

    c = NCX
    n = 3
    ref = None
    try:
        ref = _reference_neural_model()
    except Exception as e:      # an unimportable reference must not sink the bench line
        print(f"[bench] reference NeuralModel unavailable: {type(e).__name__}: {e}", file=sys.stderr)
    if ref is not None:
        cx, x, wids, aid, comp, restore = ref
        crit = torch.nn.CrossEntropyLoss(reduction="sum")
        opt = torch.optim.Adam([p for p in cx.parameters()], lr=1e-4)

        def step():
            cx.train()
            loss = crit(cx(x, wids, aid), comp) / c["batch"]
            opt.zero_grad()
            loss.backward()
            opt.step()

        step()
        t0 = time.perf_counter()
        for _ in range(n):
            step()
        dt = (time.perf_counter() - t0) / n
        cx.eval()
        cx(x, wids, aid)
        t0 = time.perf_counter()
        for _ in range(n):
            cx(x, wids, aid)       # eval_model builds the autograd graph too (counterexamples.py:450-490: no no_grad)
        dte = (time.perf_counter() - t0) / n
        restore()
        kind = "reference"
        sample = (f"vqa.models.cx.NeuralModel + MutanNoAtt imported from /root/reference, CrossEntropyLoss(sum)/B, "
                  f"backward, Adam: {n} steps of batch {c['batch']} x 24")
    else:
        from oracle import ncx_oracle as O

        params = O.make_params(1, dim_v=c["dim_v"], dim_q=c["dim_q"], dim_h_fusion=c["dim_f"], R=c["R"], n_ans=c["n_ans"],
                               dim_h=c["dim_h"], n_layers=c["n_layers"])
        feats, q_emb, aid, comp = O.make_batch(2, c["batch"], dim_v=c["dim_v"], dim_q=c["dim_q"], n_ans=c["n_ans"])
        O.loss_and_grads(params, feats, q_emb, aid, comp, c["n_layers"])      # warm-up
        t0 = time.perf_counter()
        for _ in range(n):
            O.loss_and_grads(params, feats, q_emb, aid, comp, c["n_layers"])
        dt = (time.perf_counter() - t0) / n
        p = O.to_torch(params)
        x, q, a = torch.as_tensor(feats), torch.as_tensor(q_emb), torch.as_tensor(aid)
        t0 = time.perf_counter()
        with torch.no_grad():
            for _ in range(n):
                O.neural_forward(p, x, q, a, c["n_layers"])
        dte = (time.perf_counter() - t0) / n
        kind = "port"
        sample = (f"oracle/ncx_oracle.py (restatement of cx.py:261-333 + backward, no Adam; pinned against the reference's "
                  f"own class by tests/golden/ncx_*.npz -- the Python reference cannot travel to this box), {n} steps of "
                  f"batch {c['batch']} x 24")
    return {"train_samples_per_sec": c["batch"] / dt, "eval_samples_per_sec": c["batch"] / dte, "unit": "samples/s",
            "cores": torch.get_num_threads(), "kind": kind, "sample": sample}


def build_ncx(dev):
    """NeuralModel at the bench dimensions (random-init weights of the real shapes) + its question table."""
    import torch
    from vqa_counterexamples_b200.cx import NeuralModel
    from vqa_counterexamples_b200.vqa_models import MutanNoAtt, QuestionTable

    c = NCX
    torch.manual_seed(42)
    opt = {"arch": "MutanNoAtt", "fusion": {"dim_v": c["dim_v"], "dim_q": c["dim_q"], "dim_hv": c["dim_f"], "dim_hq": c["dim_f"],
                                            "dim_mm": c["dim_f"], "R": c["R"]}, "classif": {"dropout": 0.5}}
    table = torch.tanh(torch.randn(c["n_questions"], c["dim_q"]) * 0.5).to(dev)
    vqa = MutanNoAtt(opt, ["w"], ["a"] * c["n_ans"], seq2vec=QuestionTable(table))
    spec = {k: True for k in ("v_emb", "v_mult", "v_dist", "v_rank", "q_emb", "a_emb", "z_emb")}
    emb = (torch.randn(c["n_ans"], 2400) * 0.35).numpy()
    cx = NeuralModel(spec, c["dim_h"], c["n_layers"], emb, c["drop_p"], vqa_model=vqa, knn_size=24, trainable_vqa=False)
    return cx, table


def gemm_peaks(dev):
    """Measured library GEMM rates that state the 'relevant roofline' of the NeuralCX products on this box
    (cuBLAS through torch.matmul; best of 5, CUDA events): an 8192^3 product and the layer-1 shape of an eval batch of
    4096 ([102400, 4480] x [4480, 320]), in TF32 (fp32 storage) and fp16."""
    import torch

    out = {}
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        for name, (m, n, k) in (("8192", (8192, 8192, 8192)), ("layer1_b4096", (102400, 320, 4480))):
            for dt, tag in ((torch.float32, "tf32"), (torch.float16, "fp16")):
                a = torch.randn((m, k), device=dev, dtype=dt)
                b = torch.randn((k, n), device=dev, dtype=dt)
                best = 1e30
                for _ in range(2):
                    a @ b
                for _ in range(5):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    a @ b
                    e1.record()
                    torch.cuda.synchronize()
                    best = min(best, e0.elapsed_time(e1))
                out[f"{tag}_tflops_{name}"] = 2.0 * m * n * k / (best * 1e-3) / 1e12
                del a, b
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
    return out


def ncx_leg(dev, world, rank, peaks, steps, with_cpu):
    import torch
    import torch.distributed as dist
    from vqa_counterexamples_b200 import synth
    from vqa_counterexamples_b200._lib import lib
    from vqa_counterexamples_b200.counterexamples import DataParallelCX

    c = NCX
    cx, table = build_ncx(dev)
    feats = synth.clustered_features_cuda(c["n_img"], c["dim_v"], seed=SEED, device=dev)
    dp = DataParallelCX()
    g = torch.Generator(device=dev).manual_seed(100 + rank)

    def make(B):
        return (torch.randint(0, c["n_img"], (B, 25), generator=g, device=dev, dtype=torch.int32),
                table[torch.randint(0, c["n_questions"], (B,), generator=g, device=dev)],
                torch.randint(0, c["n_ans"], (B,), generator=g, device=dev),
                torch.randint(0, 24, (B,), generator=g, device=dev))

    def timed(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        l0 = lib.vqacx_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / n], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), (lib.vqacx_launch_count() - l0) / n

    # weak scaling: 64 questions per GPU per step, gradients all-reduced (global batch 64 * world)
    batches = [make(c["batch"]) for _ in range(8)]
    it = [0]

    def train():
        b = batches[it[0] % len(batches)]
        it[0] += 1
        cx.train_step_graph(feats, *b, lr=1e-4, grad_hook=dp.allreduce_grads if world > 1 else None,
                            global_batch=c["batch"] * world)

    cx.train()
    ms_t, l_t = timed(train, steps)
    cx.eval()
    eb = make(c["eval_batch"])
    # One timed repetition = what an evaluation pass does: build the per-image tables of the frozen head's visual side
    # for this feature matrix (they are dropped first, so the build is INSIDE the timed region), then a val-pass worth
    # of batches. ms_per_batch carries its share of the table build.
    nb_eval = 52            # ceil(214 354 / 4096) batches: BASELINE config 5's pass length

    def eval_pass():
        cx.invalidate_image_tables()
        for _ in range(nb_eval):
            cx.eval_batch(feats, *eb)

    ms_pass, l_pass = timed(eval_pass, 3)
    ms_e, l_e = ms_pass / nb_eval, l_pass / nb_eval
    cx.invalidate_image_tables()
    ms_tab, _ = timed(lambda: (cx.invalidate_image_tables(), cx._fast_tables(feats, c["eval_batch"])), 3) \
        if cx.eval_precision == "fp16" else (0.0, 0)
    # e2e through the reference-shaped call: host feature batch [B,25,dim_v] (pinned) -> H2D -> forward -> loss to host
    img = batches[0][0].long()
    host_x = torch.empty((c["batch"], 25, c["dim_v"]), dtype=torch.float32, pin_memory=True)
    host_x.copy_(feats[img.view(-1)].view(c["batch"], 25, -1))
    wids = torch.zeros((c["batch"], 26), dtype=torch.long, device=dev)
    wids[:, 0] = torch.arange(c["batch"], device=dev)
    aid, comp = batches[0][2], batches[0][3]
    crit = torch.nn.CrossEntropyLoss(reduction="sum")
    cx.train()

    def ref_style_step():
        x = host_x.to(dev, non_blocking=True)
        scores = cx(x, wids, aid)
        loss = crit(scores, comp) / c["batch"]
        cx.zero_grad()
        loss.backward()
        cx.adam_step(cx.flat.grad)
        return float(loss)

    ms_r, _ = timed(ref_style_step, max(5, steps // 10))
    gp = gemm_peaks(dev) if rank == 0 else {}
    ev_tflops = NCX_EVAL_FLOP_PER_SAMPLE * c["eval_batch"] / (ms_e * 1e-3) / 1e12
    out = {
        "metric": "ncx_train_samples_per_sec", "value": c["batch"] * world / (ms_t * 1e-3), "unit": "samples/s",
        "scaling": "weak", "ms_per_step": ms_t, "launches_per_step": l_t,
        "config": {"workload": "NeuralCX train step (gather, MUTAN no-att head, MLP, CE/Recall, backward, Adam)",
                   "batch_per_gpu": c["batch"], "candidates": 24, "dim_h": c["dim_h"], "n_layers": c["n_layers"],
                   "drop_p": c["drop_p"], "n_images": c["n_img"],
                   "dtype": getattr(cx, "compute_dtype", "tf32 tensor-core products, fp32 accumulate")},
        "roofline": {"bound": "hbm", "achieved": NCX_TRAIN_BYTES_PER_STEP / (ms_t * 1e-3) / 1e9, "peak": peaks["hbm_gbs"],
                     "unit": "GB/s", "frac": NCX_TRAIN_BYTES_PER_STEP / (ms_t * 1e-3) / 1e9 / peaks["hbm_gbs"],
                     "algorithmic_bytes_per_step": NCX_TRAIN_BYTES_PER_STEP,
                     "tensor_tflops_algorithmic": NCX_TRAIN_FLOP_PER_SAMPLE * c["batch"] / (ms_t * 1e-3) / 1e12},
        "e2e": {"value": c["batch"] * world / (ms_r * 1e-3), "unit": "samples/s", "ms_per_step": ms_r,
                "h2d_bytes_per_step": c["batch"] * 25 * c["dim_v"] * 4, "d2h_bytes_per_step": 4,
                "api": "cx_model(image_features, wids, aids); CrossEntropyLoss(sum)/B; loss.backward(); Adam; float(loss)"},
        "eval": {"value": c["eval_batch"] * world / (ms_e * 1e-3), "unit": "samples/s", "batch_per_gpu": c["eval_batch"],
                 "ms_per_batch": ms_e, "launches_per_batch": l_e, "tensor_tflops_algorithmic": ev_tflops,
                 "timed_unit": f"one pass = per-image table build ({ms_tab:.3f} ms for {c['n_img']} images, inside the timed "
                               f"region) + {nb_eval} batches; ms_per_batch = pass / {nb_eval}",
                 "eval_path": cx.last_eval_path,
                 "roofline": {"bound": "tensor", "achieved": ev_tflops, "unit": "TFLOP/s",
                              "peak": peaks["bf16_tflops_sustained"],
                              "frac": ev_tflops / peaks["bf16_tflops_sustained"],
                              "peak_source": "cuBLAS bf16 sustained (MEASURED_PEAKS.json); the products run in the dtype "
                                             "named in config.dtype -- library TF32 / fp16 rates measured in this run "
                                             "are under gemm_peaks_measured"}},
        "gemm_peaks_measured": gp,
    }
    if with_cpu:
        out["cpu_baseline"] = ncx_cpu_baseline()
        cb = out["cpu_baseline"]
        out["vs_cpu_baseline"] = {"train_ratio": out["value"] / cb["train_samples_per_sec"],
                                  "train_e2e_ratio": out["e2e"]["value"] / cb["train_samples_per_sec"],
                                  "eval_ratio": out["eval"]["value"] / cb["eval_samples_per_sec"], "kind": cb["kind"]}
    del feats
    return out


# ------------------------------------------------------------------------------------------------
# parity checks that run inside the bench (checker = this library's exact fp32 path on ONE GPU; no oracle import)
# ------------------------------------------------------------------------------------------------
def knn_parity(d, i, ref_d, ref_i, rtol=1e-6):
    """Compares [m, k] results (device tensors). Index mismatches are 'tie_explained' when the two distances at that
    position agree within rtol (a float64 tie window: the order among exact ties is by index in both paths, near-ties
    can swap) AND the row's distance vector agrees everywhere; anything else is a real mismatch."""
    import torch

    neq = i != ref_i
    close = (d - ref_d).abs() <= rtol * torch.maximum(ref_d.abs(), torch.full_like(ref_d, 1e-12))
    row_ok = close.all(dim=1, keepdim=True)
    explained = neq & close & row_ok
    bad = (neq & ~explained) | ~close
    return {"rows": int(i.shape[0]), "k": int(i.shape[1]), "index_mismatch": int(neq.sum().item()),
            "tie_explained": int(explained.sum().item()), "unexplained": int(bad.sum().item()),
            "max_rel_dist_err": float(((d - ref_d).abs() / torch.clamp_min(ref_d.abs(), 1e-12)).max().item())}


def exact_reference(db, queries, k, dev):
    """Single-GPU exact search (force_exact: fp32 FMA kernel + float64 re-rank, no tensor-core candidate pass)."""
    from vqa_counterexamples_b200 import knn as knn_mod

    nn_ = knn_mod.NearestNeighbors(n_neighbors=k, device=dev, force_exact=True)
    nn_.fit(db)
    return nn_.kneighbors(queries)


def sample_rows(q0, q1, m, seed):
    import torch

    g = torch.Generator().manual_seed(seed)
    m = min(m, q1 - q0)
    return (torch.randperm(q1 - q0, generator=g)[:m].sort().values + q0)


# ------------------------------------------------------------------------------------------------
# BASELINE config 4: row-sharded top-24 kNN over a 1M x 2048 database, 131 072 replicated queries
# ------------------------------------------------------------------------------------------------
def config4_leg(args, dev, world, rank, peaks):
    import torch
    import torch.distributed as dist
    from vqa_counterexamples_b200 import knn as knn_mod, synth
    from vqa_counterexamples_b200._lib import lib

    n_db, n_q, k, passes = args.cfg4_n_db, args.cfg4_n_q, 24, 3
    n_centres = max(1, n_db // 60)
    centres = synth.cluster_centres_cuda(n_centres, DIM, seed=77, device=dev)
    r0, r1 = knn_mod.shard_rows(n_db, rank, world)
    shard = synth.clustered_rows_cuda(r0, r1, centres, seed=77, device=dev)          # this rank's rows only
    queries = synth.clustered_rows_cuda(0, n_q, centres, seed=78, device=dev)        # replicated
    phase_acc = {"candidates": 0.0, "finalize": 0.0, "rerank": 0.0, "fallback": 0.0}
    stats = {}
    timeline = {}

    def one(profile):
        d_, i_, nn_ = knn_mod.sharded_kneighbors(shard, r0, queries, k, profile=profile, return_index=True,
                                                 result="sliced")
        if profile and nn_ is not None and nn_.last_phase_ms:
            for k_, v in nn_.last_phase_ms.items():
                phase_acc[k_] += v
        if nn_ is not None:
            stats.update(nn_.last_stats or {})
            if profile and getattr(nn_, "last_timeline_ms", None):
                timeline.clear()
                timeline.update(nn_.last_timeline_ms)
        return d_, i_

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    one(False)
    barrier()
    l0 = lib.vqacx_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(passes):
        d_s, i_s = one(True)
    e1.record()
    barrier()
    launches = lib.vqacx_launch_count() - l0
    ms = torch.tensor([e0.elapsed_time(e1) / passes], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    # ---- parity: rank 0's slice of the sharded result against a single-GPU exact search over the WHOLE database
    parity = None
    if rank == 0:
        q0, q1 = knn_mod.shard_rows(n_q, 0, world)
        rows = sample_rows(q0, q1, 512, seed=4).to(dev)
        full = shard if world == 1 else torch.cat(
            [shard] + [synth.clustered_rows_cuda(*knn_mod.shard_rows(n_db, r, world), centres, seed=77, device=dev)
                       for r in range(1, world)], 0)
        ref_d, ref_i = exact_reference(full, queries.index_select(0, rows), k, dev)
        parity = knn_parity(d_s.index_select(0, rows - q0), i_s.index_select(0, rows - q0), ref_d, ref_i)
        parity["checker"] = "single-GPU exact fp32 search (force_exact) of the sampled queries over all database rows"
        del full, ref_d, ref_i
    cand_ms = phase_acc["candidates"] / passes
    flops = 2.0 * (r1 - r0) * n_q * DIM
    out = None
    if rank == 0:
        achieved = flops / (cand_ms * 1e-3) / 1e12 if cand_ms > 0 else 0.0
        peak = peaks["bf16_tflops_sustained"]
        out = {"metric": METRIC, "value": n_q / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "steps": passes, "warmup": 1,
               "scaling": "strong", "n_gpus": world,
               "config": {"workload": "BASELINE configs[3]: exact top-24 kNN, database rows sharded over the ranks (each "
                                      "rank generates only its shard), queries replicated, phase-1 bound all_gather + "
                                      "all_to_all of per-shard top-k + on-device merge inside the timed region",
                          "n_db": n_db, "n_queries": n_q, "dim": DIM, "k": k, "rows_per_rank": r1 - r0,
                          "result": "distributed by query slice (rank g holds queries shard_rows(n_q, g, world))"},
               "phases_ms": {k_: v / passes for k_, v in phase_acc.items()},
               "exchange_ms": ms - sum(phase_acc.values()) / passes,
               "timeline_ms_device_host": timeline or None,
               "roofline": {"bound": "tensor", "kernel": "knn_gemm_topk_pair_kernel", "achieved": achieved, "peak": peak,
                            "unit": "TFLOP/s", "frac": achieved / peak if peak else None, "kernel_ms": cand_ms,
                            "algorithmic_flop_per_launch": flops,
                            "algorithmic_flop_per_query": 2.0 * n_db * DIM},
               "gpu_launches": int(launches), "knn_stats": dict(stats), "parity_check": parity}
    del shard, queries, centres, d_s, i_s
    torch.cuda.empty_cache()
    return out


def config4_cpu_baseline(args):
    """sklearn one-call on a query subset against the (host) 1M database. The database is
    generated on the GPU and copied back; bounded to ~15 s of CPU work."""
    import numpy as np
    import torch
    from sklearn.neighbors import NearestNeighbors
    from vqa_counterexamples_b200 import synth

    n_db, k, nq = args.cfg4_n_db, 24, 128
    dev = torch.device("cuda", torch.cuda.current_device())
    centres = synth.cluster_centres_cuda(max(1, n_db // 60), DIM, seed=77, device=dev)
    x = synth.clustered_rows_cuda(0, n_db, centres, seed=77, device=dev).cpu().numpy()
    q = synth.clustered_rows_cuda(0, nq, centres, seed=78, device=dev).cpu().numpy()
    nbrs = NearestNeighbors(n_neighbors=k).fit(x)
    nbrs.kneighbors(q[:16])
    t0 = time.perf_counter()
    nbrs.kneighbors(q)
    dt = time.perf_counter() - t0
    return {"value": nq / dt, "unit": UNIT, "cores": cpu_threads(), "kind": "reference",
            "sample": f"one sklearn kneighbors call, {nq} queries vs the {n_db}x{DIM} database ({dt:.1f} s)"}


# ------------------------------------------------------------------------------------------------
# BASELINE config 5: full val-sized NeuralCX eval pass, data-parallel over the ranks
# ------------------------------------------------------------------------------------------------
def config5_leg(args, dev, world, rank, peaks, with_cpu):
    import numpy as np
    import torch
    import torch.distributed as dist
    from vqa_counterexamples_b200 import knn as knn_mod, synth
    from vqa_counterexamples_b200._lib import lib
    from vqa_counterexamples_b200.counterexamples import DataParallelCX, ExampleIndex, eval_indexed, shard_range

    c = NCX
    n_ex, n_img, bs, passes = args.cfg5_n, 40504, c["eval_batch"], 3
    cx, table = build_ncx(dev)                                   # identical on every rank (seeded)
    feats = synth.clustered_features_cuda(n_img, c["dim_v"], seed=SEED + 1, device=dev)
    # the candidates of a question are its image's true 24 nearest neighbours (stage 1 feeding stage 2)
    _, knn_idx = knn_mod.NearestNeighbors(n_neighbors=25, device=dev).fit(feats).kneighbors(feats)
    ex = synth.cx_examples_np(n_ex, n_img, knn_idx.cpu().numpy(), c["n_questions"], seed=4321, n_answers=c["n_ans"])
    index = ExampleIndex.from_arrays(ex["img_idx"], ex["q_idx"][:, None], ex["answer_aid"], ex["comp_idx"], device=dev)
    dp = DataParallelCX()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    res = eval_indexed(cx, index, feats, table, bs, dp=dp)       # warm-up pass (also the result that is checked)
    barrier()
    l0 = lib.vqacx_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(passes):
        cx.invalidate_image_tables()     # every timed pass builds its own per-image tables (inside the timed region)
        res_t = eval_indexed(cx, index, feats, table, bs, dp=dp)     # includes the 4-element all-reduce + its D2H read
    e1.record()
    barrier()
    launches = lib.vqacx_launch_count() - l0
    ms = torch.tensor([e0.elapsed_time(e1) / passes], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    out = None
    if rank == 0:
        # single-rank replay of every rank's batches: the (hits@1, hits@5, n) the N-rank pass must reproduce exactly
        solo = DataParallelCX()
        solo.world, solo.rank = 1, 0
        ref = eval_indexed(cx, index, feats, table, bs, dp=solo, spans=[shard_range(n_ex, r, world) for r in range(world)])
        comp = ex["comp_idx"]
        parity = {"n": res["n"], "hits_1": res["hits_1"], "hits_5": res["hits_5"], "single_rank_hits_1": ref["hits_1"],
                  "single_rank_hits_5": ref["hits_5"], "single_rank_n": ref["n"],
                  "recall_equal": bool(res["hits_1"] == ref["hits_1"] and res["hits_5"] == ref["hits_5"] and
                                       res["n"] == ref["n"] == n_ex and res_t["hits_5"] == res["hits_5"]),
                  "loss_abs_diff": abs(res["loss"] - ref["loss"]),
                  "checker": "one process replaying all ranks' batches (same kernels, no collective)"}
        tfl = NCX_EVAL_FLOP_PER_SAMPLE * n_ex / world / (ms * 1e-3) / 1e12
        out = {"metric": "ncx_eval_samples_per_sec", "value": n_ex / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms,
               "steps": passes, "warmup": 1, "scaling": "strong", "n_gpus": world,
               "config": {"workload": "BASELINE configs[4]: full VQA-CX val-sized eval pass (eval_model, "
                                      "counterexamples.py:450-490): examples sharded over the ranks, sum-CE / Recall@1 / "
                                      "Recall@5 accumulated on the device, ONE 4-element all-reduce per pass",
                          "n_questions": n_ex, "candidates": 24, "n_images": n_img, "batch": bs,
                          "candidates_from": "stage-1 kNN of the 40 504 synthetic features (this library)",
                          "dtype": getattr(cx, "compute_dtype", "tf32 tensor-core products, fp32 accumulate")},
               "results": {"loss": res["loss"], "recall_1": res["recall_1"], "recall_5": res["recall"],
                           "distance_baseline_recall_1": float((comp < 1).mean()),
                           "distance_baseline_recall_5": float((comp < 5).mean())},
               "roofline": {"bound": "tensor", "achieved": tfl, "unit": "TFLOP/s per GPU",
                            "peak": peaks["bf16_tflops_sustained"], "frac": tfl / peaks["bf16_tflops_sustained"],
                            "algorithmic_flop_per_sample": NCX_EVAL_FLOP_PER_SAMPLE},
               "gpu_launches": int(launches), "parity_check": parity}
        if with_cpu:
            out["cpu_baseline"] = config5_cpu_baseline()
    del feats, index, cx
    torch.cuda.empty_cache()
    return out


def config5_cpu_baseline():
    """NeuralCX eval forward on the host cores on a bounded subset (8 batches of 64)."""
    b = ncx_cpu_baseline()
    return {"value": b["eval_samples_per_sec"], "unit": "samples/s", "cores": b["cores"], "kind": b["kind"],
            "sample": "eval forward, " + b["sample"]}


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl b200) needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL's own log lines (e.g. "NCCL version ..." under NCCL_DEBUG=VERSION/INFO) go to stderr: stdout carries
        # the one JSON line only
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)

    from vqa_counterexamples_b200 import knn as knn_mod
    from vqa_counterexamples_b200._lib import lib

    n = args.n_db
    h, x = host_features(n)  # identical on every rank (same seed)
    x_np = h.numpy()
    peaks = load_peaks()
    with_cpu = rank == 0 and world == 1 and not args.no_cpu_baseline

    cpu_base = None
    if with_cpu:
        cpu_base = cpu_baseline_leg(x_np)

    # rows of the database owned by this rank (= the query slice whose final result this rank holds)
    r0, r1 = knn_mod.shard_rows(n, rank, world)
    phase_acc = {"candidates": 0.0, "finalize": 0.0, "rerank": 0.0, "fallback": 0.0}
    stats_last = {}
    timeline = {}     # last timed step at N > 1: stage -> (device ms, host ms) between consecutive marks

    def step_device(profile=False):
        if world == 1:
            nn_ = knn_mod.NearestNeighbors(n_neighbors=K, device=dev)
            nn_.profile = profile
            nn_.fit(x)
            d_, i_ = nn_.kneighbors(x)
        else:
            d_, i_, nn_ = knn_mod.sharded_kneighbors(x[r0:r1], r0, x, K, profile=profile, return_index=True,
                                                     result="sliced")
        if profile and nn_.last_phase_ms:
            for k_, v in nn_.last_phase_ms.items():
                phase_acc[k_] += v
        if profile and getattr(nn_, "last_timeline_ms", None):
            timeline.clear()
            timeline.update(nn_.last_timeline_ms)
        stats_last.update(nn_.last_stats or {})
        return d_, i_

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()          # before the warm-up: nvidia-smi's start-up must not land in the timed region
        sampler.wait_ready()
    for _ in range(args.warmup):
        step_device()
    barrier()
    l0 = lib.vqacx_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_host0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        d_dev, i_dev = step_device(profile=True)
    e1.record()
    barrier()
    sampler.window(t_host0, time.perf_counter())
    launches = lib.vqacx_launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step = float(ms.item()) / args.steps
    timeline_ranks = None
    if world > 1 and timeline:
        # the last timed step's stage timeline of EVERY rank: the spread between ranks is what the collectives wait for
        boxes = [None] * world
        dist.all_gather_object(boxes, dict(timeline))
        timeline_ranks = {st: {"device_ms_min": min(b[st][0] for b in boxes if st in b),
                               "device_ms_max": max(b[st][0] for b in boxes if st in b)} for st in timeline}

    # ---- e2e: host in -> host out through the public API ------------------------------------
    def step_e2e(src=None):
        if world == 1:
            # knn.py's whole workload on a host matrix: fit(features) + kneighbors(features), one pipelined call
            return knn_mod.NearestNeighbors(n_neighbors=K, device=dev).fit_kneighbors(x_np if src is None else src)
        # one process per GPU: every rank uploads ITS rows only, the fp32 rows travel by one NVLink all_gather, each
        # rank downloads the neighbours of its own rows (knn.sharded_selfjoin_host)
        return knn_mod.sharded_selfjoin_host(x_np[r0:r1], r0, n, K, device=dev)

    def time_e2e(steps, src=None):
        # warm-up holds its results exactly like the timed loop (two result sets alive across a step boundary), so
        # both sets of pinned staging blocks exist in the host allocator's cache before the timed region
        for _ in range(3):
            out = step_e2e(src)
        barrier()
        e0.record()
        for _ in range(steps):
            out = step_e2e(src)
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps, out

    ms_e2e, (d_host, i_host) = time_e2e(args.steps)
    ms_e2e_pageable = None
    if world == 1:
        # knn.py:35-37 loads the features with np.array(...): pageable memory. Same call, same region, 3 steps.
        x_pageable = np.array(x_np)
        ms_e2e_pageable, _ = time_e2e(3, x_pageable)
        del x_pageable

    # sanity: self is the nearest neighbour of every row, distances ascend (cheap size-independent property)
    own = torch.arange(r0, r1, device=dev) if world > 1 else torch.arange(n, device=dev)
    assert bool((i_dev[:, 0] == own).all()), "self is not the first neighbour"
    assert bool((d_dev[:, 1:] >= d_dev[:, :-1]).all()), "distances not sorted"
    # parity inside the run: sampled rows of this run's result (rank 0's query slice) and of the e2e result against a
    # single-GPU exact search; an unexplained mismatch fails the run
    parity = None
    if rank == 0:
        rows = sample_rows(r0 if world > 1 else 0, r1 if world > 1 else n, 512, seed=3).to(dev)
        ref_d, ref_i = exact_reference(x, x.index_select(0, rows), K, dev)
        off = r0 if world > 1 else 0
        parity = knn_parity(d_dev.index_select(0, rows - off), i_dev.index_select(0, rows - off), ref_d, ref_i)
        pe = knn_parity(torch.from_numpy(d_host).to(dev).index_select(0, rows - off),
                        torch.from_numpy(i_host).to(dev).index_select(0, rows - off), ref_d, ref_i)
        parity["e2e_unexplained"] = pe["unexplained"]
        parity["checker"] = "single-GPU exact fp32 search (force_exact) of the sampled rows over all database rows"
        del ref_d, ref_i

    if rank == 0:
        n_local = r1 - r0
        cand_ms = phase_acc["candidates"] / args.steps
        flops = 2.0 * n * n_local * DIM
        achieved = flops / (cand_ms * 1e-3) / 1e12 if cand_ms > 0 else 0.0
        peak = peaks["bf16_tflops_sustained"]
        # The symmetric self-join sweep (abi_knn.cu make_sym_plan) computes every off-diagonal block of the distance
        # matrix once: it ISSUES (groups + 1) / (2 groups) of the algorithmic 2 N^2 d FLOP. `achieved`/`frac` keep
        # the algorithmic numerator, so they can exceed 1; `executed_*` is the tensor-pipe view.
        executed_frac = 1.0
        if "mirror_keys_avg" in stats_last:
            g = max(3, min(lib.vqacx_get_option(2), 15)) | 1      # VQACX_OPT_KNN_SYM_GROUPS
            executed_frac = (g + 1) / (2.0 * g)
        traffic, traffic_src = ncu_traffic_bytes() if world == 1 else (None, "single-GPU capture only")
        line = {
            "metric": METRIC, "value": n / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f16 tensor-core candidates (fp32 accumulate) + f64 re-rank", "data": "synthetic",
            "config": {"workload": WORKLOAD, "n_db": n, "n_queries": n, "dim": DIM, "k": K,
                       "parallelism": "single GPU" if world == 1 else
                       f"database rows sharded x{world}, bound all_gather + all_to_all of per-shard top-k + merge of the "
                       "rank's query slice",
                       "l2": "inputs (1.0 GB f32 + 0.5 GB f16) exceed the 126 MB L2; no explicit flush",
                       "seed": SEED},
            "e2e": {"value": n / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(n) * DIM * 4, "d2h_bytes_per_step": int(n) * K * 16,
                    "host_memory": "pinned",
                    "api": "NearestNeighbors(k).fit_kneighbors(X_host)" if world == 1 else
                           "knn.sharded_selfjoin_host(X_host[r0:r1], r0, n, k): each rank uploads its rows, one NCCL "
                           "all_gather of the fp32 rows, each rank downloads its query slice"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "tensor", "kernel": "knn_gemm_topk_pair_kernel", "achieved": achieved, "peak": peak,
                         "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": f"{peaks['source']} cuBLAS bf16 sustained (MEASURED_PEAKS.json)",
                         "algorithmic_flop_per_launch": flops, "kernel_ms": cand_ms,
                         "sweep": "symmetric (diagonal + mirrored launches)" if executed_frac < 1.0 else "plain",
                         "executed_flop_per_step": flops * executed_frac,
                         "executed_tflops": achieved * executed_frac,
                         "frac_executed": achieved * executed_frac / peak if peak else None},
            "phases_ms": {k_: v / args.steps for k_, v in phase_acc.items()},
            "knn_stats": stats_last,
            "timeline_ms_device_host": timeline or None,
            "timeline_over_ranks": timeline_ranks,
            "clocks": clocks,
            "parity_check": parity,
        }
        if ms_e2e_pageable is not None:
            line["e2e"]["pageable_value"] = n / (ms_e2e_pageable * 1e-3)
            line["e2e"]["pageable_ms_per_step"] = ms_e2e_pageable
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
    del d_dev, i_dev, d_host, i_host
    torch.cuda.empty_cache()
    ncx = cfg4 = cfg5 = None
    if not args.no_ncx:
        ncx = ncx_leg(dev, world, rank, peaks, max(20, 10 * args.steps), with_cpu)
        torch.cuda.empty_cache()
    del x
    torch.cuda.empty_cache()
    if not args.no_config5:
        cfg5 = config5_leg(args, dev, world, rank, peaks, with_cpu)
    if not args.no_config4:
        cfg4 = config4_leg(args, dev, world, rank, peaks)
        if with_cpu and cfg4 is not None:
            cfg4["cpu_baseline"] = config4_cpu_baseline(args)
    failed = []
    if rank == 0:
        if ncx is not None:
            line["ncx"] = ncx
        if cfg4 is not None:
            line["config4"] = cfg4
        if cfg5 is not None:
            line["config5"] = cfg5
        emit(line)
        if parity and (parity["unexplained"] or parity["e2e_unexplained"]):
            failed.append(f"headline kNN parity: {parity}")
        if cfg4 is not None and cfg4["parity_check"]["unexplained"]:
            failed.append(f"config4 parity: {cfg4['parity_check']}")
        if cfg5 is not None and not cfg5["parity_check"]["recall_equal"]:
            failed.append(f"config5 parity: {cfg5['parity_check']}")
        for f in failed:
            print("[bench] PARITY FAILURE " + f, file=sys.stderr)
    if world > 1:
        dist.destroy_process_group()
    return 1 if failed else 0


_JSON_FD = None


def emit(line):
    """The ONE JSON line, written to the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD
    args = parse_args()
    # stdout carries the JSON line only: whatever a library prints on fd 1 (NCCL's "NCCL version ..." line still
    # landed there at N = 2 despite NCCL_DEBUG_FILE) is sent to stderr for the rest of the run
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
