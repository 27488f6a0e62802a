"""Host side of stage 2: the reference's `counterexamples.py` train / eval driver for `--cx_model NeuralModel`,
re-hosted over libvqacx.so. Same CLI flags (counterexamples.py:39-87), same YAML option keys merged the same way
(vqa/lib/utils.py:6-12), same free functions with the same argument meaning:

    batchify(example_list, batch_size, shuffle=True)                       counterexamples.py:509-516
    getDataFromBatch(batch, features, name_to_index, pairwise=False)       counterexamples.py:519-547
    recallAtK(scores, ground_truth, k=5)                                   counterexamples.py:501-506
    eval_model(cx_model, valset, features_val, batch_size)                 counterexamples.py:450-490
    save_cx_checkpoint / load_cx_checkpoint                                counterexamples.py:550-580

What changes is where the work happens. The reference gathers 25 feature rows per question on the host with numpy
fancy indexing and ships 13 MB over PCIe every step (counterexamples.py:540-541), then synchronises every step for
recallAtK. Here the feature matrix is uploaded once (`DeviceFeatures`), an example list is resolved once into index
tensors on the device (`ExampleIndex`: what the name_to_index loops of getDataFromBatch compute), and a step is CUDA
kernels only: gather + NeuralCX forward + CE/Recall + backward + Adam, with loss and hit counts accumulated on the
device and read once per `print_freq` / epoch. Multi-GPU is one process per GPU: examples sharded by rank, gradients
averaged with one all-reduce over the flat parameter buffer (`DataParallelCX`), eval reduced with one 4-element
all-reduce. All arithmetic stays in libvqacx.so; there is no CPU fallback.
"""
import argparse
import json
import os
import pickle
import random
import shutil
from datetime import datetime

import numpy as np
import torch

from .cx import DistanceBaseline, NeuralModel, recallAtK  # noqa: F401  (recallAtK re-exported: reference name)
from .cx_heads import BlackBox, LinearContext, RandomBaseline, SemanticBaseline
from . import vqa_models as models

KNN = 24

parser = argparse.ArgumentParser(description="NeuralCX training / evaluation (B200-native counterexamples.py)")
parser.add_argument("--path_opt", default="options/cx/counterexamples_default.yaml", type=str,
                    help="path to a yaml options file")
parser.add_argument("-cx", "--cx_model", required=True, type=str, help="Counterexample model type")
parser.add_argument("-lr", "--learning_rate", type=float, help="initial learning rate")
parser.add_argument("-lb", "--sb_lambda", type=float, help="semantic baseline lambda (unused by NeuralModel)")
parser.add_argument("-b", "--batch_size", type=int, help="mini-batch size")
parser.add_argument("--epochs", type=int, help="number of total epochs to run")
parser.add_argument("--project_dir", default="/datadrive/vqa.pytorch/", type=str,
                    help="path to project root whose data to use")
parser.add_argument("--resume", default="", type=str, help="run name to resume")
parser.add_argument("--best", action="store_true", help="whether to resume best checkpoint")
parser.add_argument("-c", "--comment", type=str, default="")
parser.add_argument("-p", "--print_freq", default=100, type=int, help="print frequency")
parser.add_argument("-v", "--eval_freq", default=-1, type=int, help="eval frequency")
parser.add_argument("-t", "--test", action="store_true", help="Run eval on full testset after training")
parser.add_argument("--viz", action="store_true", help="accepted for CLI compatibility; plotting is out of scope")
parser.add_argument("--pairwise", action="store_true", help="Pairwise training (not part of the NeuralModel path)")
_group = parser.add_mutually_exclusive_group(required=False)
_group.add_argument("--pretrained_vqa", dest="pretrained_vqa", action="store_true")
_group.add_argument("--untrained_vqa", dest="pretrained_vqa", action="store_false")
parser.set_defaults(pretrained_vqa=True)
parser.add_argument("--trainable_vqa", action="store_true", help="If true, backprop through VQA model")
parser.add_argument("-dev", "--dev_mode", action="store_true")
parser.add_argument("-y", "--yes", action="store_true", help="overwrite an existing save directory without asking")


# ------------------------------------------------------------------------------------------------
# options (vqa/lib/utils.py:6-12)
# ------------------------------------------------------------------------------------------------
def update_values(dict_from, dict_to):
    """Non-None leaves of `dict_from` (the CLI) overwrite `dict_to` (the YAML); returns dict_to. The reference
    recurses with `dict_to[key]` and so raises KeyError when the YAML lacks a CLI section; kept."""
    for key, value in dict_from.items():
        if isinstance(value, dict):
            update_values(dict_from[key], dict_to[key])
        elif value is not None:
            dict_to[key] = dict_from[key]
    return dict_to


def load_options(args):
    """counterexamples.py:98-115: CLI dict merged over the YAML file (yaml.safe_load: PyYAML >= 6 has no bare load)."""
    import yaml

    options = {
        "optim": {"lr": args.learning_rate, "batch_size": args.batch_size, "epochs": args.epochs},
        "cx_model": {"pretrained_vqa": args.pretrained_vqa, "trainable_vqa": args.trainable_vqa},
    }
    with open(args.path_opt, "r") as handle:
        options_yaml = yaml.safe_load(handle)
    options = update_values(options, options_yaml)
    options["vgenome"] = None
    return options


# ------------------------------------------------------------------------------------------------
# batching (host; order matters for trajectory parity with the reference's python RNG)
# ------------------------------------------------------------------------------------------------
def batchify(example_list, batch_size, shuffle=True):
    """counterexamples.py:509-516: in-place random.shuffle (python MT, seeded by main) then contiguous slices."""
    if shuffle:
        random.shuffle(example_list)
    return [example_list[i:min(i + batch_size, len(example_list))] for i in range(0, len(example_list), batch_size)]


class DeviceFeatures:
    """The `noatt` feature matrix [N, dim_v] resident in HBM (uploaded once instead of 25 rows per question per
    step). Behaves like the numpy array for the one thing the reference does with it: `len`, `.shape`."""

    def __init__(self, features, device=None):
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if isinstance(features, DeviceFeatures):
            features = features.tensor
        if isinstance(features, np.ndarray):
            features = torch.from_numpy(np.ascontiguousarray(features, dtype=np.float32))
        self.tensor = features.to(device=dev, dtype=torch.float32, non_blocking=False).contiguous()
        self.shape = tuple(self.tensor.shape)

    def __len__(self):
        return self.shape[0]


class ExampleIndex:
    """An example list resolved once into device index tensors: exactly the values getDataFromBatch derives per
    batch with python loops (counterexamples.py:526-537): image row + 24 neighbour rows through `name_to_index`,
    `question_wids`, `answer_aid`, `comp.knn_index`. `order` follows the list, so a `batchify` shuffle of the list is
    mirrored by permuting `ids` (each example carries its position in `_cx_pos`)."""

    def __init__(self, examples_list, name_to_index, device=None):
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        n = len(examples_list)
        img = np.empty((n, KNN + 1), dtype=np.int32)
        wids = np.zeros((n, len(examples_list[0]["question_wids"]) if n else 26), dtype=np.int64)
        aid = np.empty((n,), dtype=np.int64)
        comp = np.empty((n,), dtype=np.int64)
        for i, ex in enumerate(examples_list):
            img[i, 0] = name_to_index[ex["image_name"]]
            knns = ex["knns"]
            assert len(knns) == KNN, "each example carries its 24 nearest-neighbour image names"
            img[i, 1:] = [name_to_index[name] for name in knns]
            wids[i] = ex["question_wids"]
            aid[i] = ex["answer_aid"]
            comp[i] = ex["comp"]["knn_index"]
            ex["_cx_pos"] = i
        self.n = n
        self.img_idx = torch.from_numpy(img).to(dev)
        self.question_wids = torch.from_numpy(wids).to(dev)
        self.answer_aids = torch.from_numpy(aid).to(dev)
        self.comp_idxs = torch.from_numpy(comp).to(dev)
        self.device = dev

    @classmethod
    def from_arrays(cls, img_idx, question_wids, answer_aids, comp_idxs, device=None):
        """Index-form examples (e.g. built on the device from stage-1 output by `examples_from_knn`)."""
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self = cls.__new__(cls)
        self.n = int(len(img_idx))
        self.img_idx = torch.as_tensor(img_idx).to(dev, torch.int32).contiguous()
        self.question_wids = torch.as_tensor(question_wids).to(dev, torch.int64).contiguous()
        self.answer_aids = torch.as_tensor(answer_aids).to(dev, torch.int64).contiguous()
        self.comp_idxs = torch.as_tensor(comp_idxs).to(dev, torch.int64).contiguous()
        self.device = dev
        return self

    def ids_of(self, batch):
        return torch.tensor([ex["_cx_pos"] for ex in batch], dtype=torch.int64, device=self.device)

    def take(self, ids):
        """(img_idx [b,25] i32, question_wids [b,26], answer_aids [b], comp_idxs [b]) for example ids (device)."""
        return (self.img_idx.index_select(0, ids), self.question_wids.index_select(0, ids),
                self.answer_aids.index_select(0, ids), self.comp_idxs.index_select(0, ids))


def getDataFromBatch(batch, features, name_to_index, pairwise=False):
    """counterexamples.py:519-547, same return value: (image_features [B,25,dim_v] cuda f32, question_wids [B,26],
    answer_aids [B], comp_idxs [B]) on the device. `features` may be the host numpy matrix (reference behaviour:
    host gather + H2D) or a `DeviceFeatures` (gather on the device, no PCIe traffic)."""
    if pairwise:
        raise NotImplementedError("pairwise batches belong to PairwiseModel, outside the NeuralModel hot path")
    image_idxs, question_wids, answer_aids, comp_idxs = [], [], [], []
    for ex in batch:
        image_idx = [name_to_index[ex["image_name"]]]
        knn_idxs = [name_to_index[name] for name in ex["knns"]]
        image_idxs.append(image_idx + knn_idxs)
        question_wids.append(ex["question_wids"])
        answer_aids.append(ex["answer_aid"])
        comp_idxs.append(ex["comp"]["knn_index"])
    if isinstance(features, DeviceFeatures):
        idx = torch.as_tensor(np.asarray(image_idxs, dtype=np.int64), device=features.tensor.device)
        image_features = features.tensor.index_select(0, idx.view(-1)).view(len(batch), KNN + 1, -1)
    else:
        image_features = torch.from_numpy(np.array([features[idxs] for idxs in image_idxs])).cuda()
    question_wids = torch.as_tensor(np.asarray(question_wids, dtype=np.int64)).cuda()
    answer_aids = torch.as_tensor(np.asarray(answer_aids, dtype=np.int64)).cuda()
    comp_idxs = torch.as_tensor(np.asarray(comp_idxs, dtype=np.int64)).cuda()
    return image_features, question_wids, answer_aids, comp_idxs


# ------------------------------------------------------------------------------------------------
# data-parallel plumbing (one process per GPU; NCCL on the B200 box, gloo in the CPU tests)
# ------------------------------------------------------------------------------------------------
def build_examples_from_knn(knn_idx, q_img, q_comp):
    """The example builder join on the device (vqacx_valset_builder.ipynb cell 13): stage-1 output `knn_idx` [n_img, 25]
    (int64, each image among its own neighbours) x complementary question pairs -> the index tensors stage 2 consumes,
    without the notebooks' Python dict loops or the JSON round trip. q_img [n_q] / q_comp [n_q] (int32; -1 = no
    complement in the set). Returns (kept [m] int64 question positions in input order, img_idx [m, 25] int32,
    comp_idx [m] int64) for the m examples the notebook keeps."""
    from ._lib import lib, check, ptr, current_stream

    dev = knn_idx.device
    knn_idx = knn_idx.to(torch.int64).contiguous()
    q_img = torch.as_tensor(q_img, dtype=torch.int32, device=dev).contiguous()
    q_comp = torch.as_tensor(q_comp, dtype=torch.int32, device=dev).contiguous()
    n_q, k1 = int(q_img.shape[0]), int(knn_idx.shape[1])
    img_idx = torch.empty((n_q, k1), dtype=torch.int32, device=dev)
    comp_idx = torch.empty(n_q, dtype=torch.int32, device=dev)
    check(lib.vqacx_cx_build_examples(ptr(knn_idx), knn_idx.shape[0], k1, ptr(q_img), ptr(q_comp), n_q, ptr(img_idx),
                                      ptr(comp_idx), current_stream()), "vqacx_cx_build_examples")
    kept = torch.nonzero(comp_idx >= 0).squeeze(1)               # index bookkeeping only
    return kept, img_idx.index_select(0, kept), comp_idx.index_select(0, kept).to(torch.int64)


def shard_range(n, rank, world):
    """Contiguous [begin, end) of n units owned by `rank` (ceil split: the last ranks may own less or nothing)."""
    per = (n + world - 1) // world
    return min(n, rank * per), min(n, (rank + 1) * per)


def shard_batch(batch_ids, rank, world):
    """The slice of one global batch a rank works on (global batch semantics: 64 examples split `world` ways)."""
    b, e = shard_range(len(batch_ids), rank, world)
    return batch_ids[b:e]


class DataParallelCX:
    """Gradient / metric reduction for one-process-per-GPU NeuralCX. Every rank holds the full
    parameters and Adam state; a global batch is split by `shard_batch`; each rank back-propagates the SUM-reduced
    loss of its slice divided by the GLOBAL batch size, so the all-reduce (sum) of the flat gradient buffer is exactly
    the single-process gradient and the replicated Adam steps stay bit-identical across ranks."""

    def __init__(self, group=None):
        import torch.distributed as dist

        self.dist = dist
        self.group = group
        self.on = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if self.on else 1
        self.rank = dist.get_rank(group) if self.on else 0

    def allreduce_grads(self, flat_grads):
        if self.world > 1:
            self.dist.all_reduce(flat_grads, op=self.dist.ReduceOp.SUM, group=self.group)
        return flat_grads

    def reduce_metrics(self, loss_sum, hits, count):
        """(sum loss, hits@1, hits@5, n) over all ranks -> python floats/ints. One 4-element all-reduce."""
        t = torch.stack([loss_sum.double().reshape(()), hits[0].double(), hits[1].double(),
                         torch.as_tensor(float(count), dtype=torch.float64, device=loss_sum.device)])
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        v = t.tolist()
        return v[0], int(round(v[1])), int(round(v[2])), int(round(v[3]))


# ------------------------------------------------------------------------------------------------
# evaluation
# ------------------------------------------------------------------------------------------------
def _q_emb(cx_model, question_wids):
    if getattr(cx_model, "vqa_model", None) is None:
        return None
    with torch.no_grad():
        return cx_model.vqa_model.seq2vec(question_wids)


def eval_model(cx_model, valset, features_val, batch_size, pairwise=False, dp=None, shuffle=True,
               return_recall1=False):
    """counterexamples.py:450-490: {'loss': sum CE / n, 'recall': Recall@5} over `valset['examples_list']` (shuffled in
    place like the reference unless shuffle=False), model left in train mode. Scores, CE and Recall@1/@5 are computed
    and accumulated on the device; the only D2H read is the 4-number result. With `dp` the batches are sharded across
    ranks and reduced once at the end."""
    if pairwise:
        raise NotImplementedError("pairwise evaluation belongs to PairwiseModel, outside the NeuralModel hot path")
    cx_model.eval()
    dp = dp or DataParallelCX()
    feats = features_val if isinstance(features_val, DeviceFeatures) else DeviceFeatures(features_val)
    index = valset.get("_cx_index")
    if index is None or index.n != len(valset["examples_list"]):
        index = valset["_cx_index"] = ExampleIndex(valset["examples_list"], valset["name_to_index"])
    cx_model.reset_metrics()
    n_local = 0
    for batch in batchify(valset["examples_list"], batch_size=batch_size, shuffle=shuffle):
        cx_model.knn_size = KNN
        ids = shard_batch(index.ids_of(batch), dp.rank, dp.world)
        if ids.numel() == 0:
            continue
        img_idx, wids, aid, comp = index.take(ids)
        cx_model.eval_batch(feats.tensor, img_idx, _q_emb(cx_model, wids), aid, comp)
        n_local += int(ids.numel())
    loss_sum, h1, h5, n = dp.reduce_metrics(cx_model.metrics[0], cx_model.hits, n_local)
    results = {"loss": loss_sum / max(n, 1), "recall": h5 / max(n, 1)}
    if return_recall1:
        results["recall_1"] = h1 / max(n, 1)
    cx_model.train()
    return results


def eval_indexed(cx_model, index, features, q_emb_table, batch_size, dp=None, spans=None):
    """Index-form evaluation pass (BASELINE config 5; eval_model's loop, counterexamples.py:450-490, over examples that
    are already index tensors): this rank's examples shard_range(index.n, rank, world) in batches of `batch_size`,
    question embeddings looked up in a device table by `index.question_wids[:, 0]`, sum-CE / Recall@1 / Recall@5
    accumulated on the device and reduced over the ranks by ONE 4-element all-reduce. `spans` overrides the example
    ranges this call walks (a list of [begin, end); used to replay every rank's batches in one process).
    Returns {'loss', 'recall', 'recall_1', 'n', 'hits_1', 'hits_5'}."""
    dp = dp or DataParallelCX()
    cx_model.eval()
    cx_model.reset_metrics()
    spans = [shard_range(index.n, dp.rank, dp.world)] if spans is None else list(spans)
    feats = features.tensor if isinstance(features, DeviceFeatures) else features
    n_local = 0
    for b, e in spans:
        for i in range(b, e, batch_size):
            j = min(e, i + batch_size)
            q = q_emb_table.index_select(0, index.question_wids[i:j, 0])
            cx_model.eval_batch(feats, index.img_idx[i:j], q, index.answer_aids[i:j], index.comp_idxs[i:j])
        n_local += max(0, e - b)
    loss_sum, h1, h5, n = dp.reduce_metrics(cx_model.metrics[0], cx_model.hits, n_local)
    cx_model.train()
    return {"loss": loss_sum / max(n, 1), "recall": h5 / max(n, 1), "recall_1": h1 / max(n, 1), "n": n,
            "hits_1": h1, "hits_5": h5}


def log_results(writer, mode, epoch, i, metrics):
    """counterexamples.py:493-498."""
    metrics_str = ""
    for k, v in metrics.items():
        if writer is not None:
            writer.add_scalar(k, v, i)
        metrics_str += "{}: {:.4f}, ".format(k, v)
    print("Epoch {} {}: {}".format(epoch, mode, metrics_str))


# ------------------------------------------------------------------------------------------------
# checkpoints (counterexamples.py:550-580; same files, same state_dict names)
# ------------------------------------------------------------------------------------------------
def save_cx_checkpoint(cx_model, info, save_dir, is_best=True):
    path_ckpt_model = os.path.join(save_dir, "ckpt", "model.ckpt")
    path_ckpt_info = os.path.join(save_dir, "ckpt", "info.ckpt")
    path_best_model = os.path.join(save_dir, "best", "model.ckpt")
    path_best_info = os.path.join(save_dir, "best", "info.ckpt")
    torch.save({k: v.detach().cpu() for k, v in cx_model.state_dict().items()}, path_ckpt_model)
    torch.save(info, path_ckpt_info)
    if is_best:
        shutil.copyfile(path_ckpt_model, path_best_model)
        shutil.copyfile(path_ckpt_info, path_best_info)
    print("{}Saved checkpoint to {}".format("* " if is_best else "", save_dir))


def load_cx_checkpoint(cx_model, save_dir, resume_best=True):
    sub = "best" if resume_best else "ckpt"
    path_ckpt_model = os.path.join(save_dir, sub, "model.ckpt")
    path_ckpt_info = os.path.join(save_dir, sub, "info.ckpt")
    cx_model.load_state_dict(torch.load(path_ckpt_model, map_location="cpu"))
    print("Loaded model from {}".format(path_ckpt_model))
    info = torch.load(path_ckpt_info, weights_only=False)
    assert len(info) > 0
    last_epoch = len(info)
    print("Epoch {}: {}".format(last_epoch, info[-1]))
    # the reference reads info[-1]['recall_5'] (counterexamples.py:580) although eval_model stores 'recall'
    # (KeyError on --resume); both spellings are accepted here
    last = info[-1]
    return info, last_epoch + 1, last["recall_5"] if "recall_5" in last else last["recall"]


# ------------------------------------------------------------------------------------------------
# training
# ------------------------------------------------------------------------------------------------
def train_epoch(cx_model, trainset, features_train, batch_size, lr, dp=None, print_freq=0, on_print=None,
                max_batches=None, eval_freq=-1, on_eval=None):
    """One pass of the loop at counterexamples.py:312-361 over `trainset['examples_list']` (shuffled in place by
    batchify). Per step: CUDA kernels only (gather, forward, CE / Recall@5, backward, [all-reduce], Adam). Returns
    the number of batches run. on_print(train_b, {'loss', 'recall'}) fires every print_freq batches with the metrics
    of the LAST batch, as the reference logs them. on_eval(train_b) fires every eval_freq batches (> 0) except after
    the last batch, whose evaluation the caller runs itself (counterexamples.py:357-361); the model is put back in train
    mode afterwards."""
    dp = dp or DataParallelCX()
    feats = features_train if isinstance(features_train, DeviceFeatures) else DeviceFeatures(features_train)
    index = trainset.get("_cx_index")
    if index is None or index.n != len(trainset["examples_list"]):
        index = trainset["_cx_index"] = ExampleIndex(trainset["examples_list"], trainset["name_to_index"])
    cx_model.train()
    if dp.world > 1:
        cx_model.drop_seed = 0x5EED + 0x9E3779B97F4A7C15 * dp.rank % (1 << 62)   # independent dropout streams per rank
    train_b = 0
    batches = batchify(trainset["examples_list"], batch_size=batch_size)
    for batch in batches:
        assert cx_model.training
        ids = shard_batch(index.ids_of(batch), dp.rank, dp.world)
        img_idx, wids, aid, comp = index.take(ids)
        log_now = print_freq > 0 and (train_b + 1) % print_freq == 0
        if log_now:
            cx_model.reset_metrics()
        # CUDA-graph replay of the step (VQACX_NO_GRAPH=1: launch the kernels one by one)
        graph_step = getattr(cx_model, "train_step_graph", None)    # NeuralModel only; LinearContext steps eagerly
        step = cx_model.train_step if graph_step is None or os.environ.get("VQACX_NO_GRAPH") == "1" else graph_step
        extra = {"question_wids": wids} if getattr(cx_model, "trains_seq2vec", False) else {}   # trainable_vqa: cx.py:73-74
        step(feats.tensor, img_idx, None if extra else _q_emb(cx_model, wids), aid, comp, lr=lr,
             grad_hook=dp.allreduce_grads if dp.world > 1 else None, global_batch=len(batch), **extra)
        train_b += 1
        if log_now and on_print is not None:
            loss_sum, _, h5, n = dp.reduce_metrics(cx_model.metrics[0], cx_model.hits, int(ids.numel()))
            on_print(train_b, {"loss": loss_sum / max(n, 1), "recall": h5 / max(n, 1)})
        if max_batches is not None and train_b >= max_batches:
            break
        if on_eval is not None and eval_freq > 0 and train_b % eval_freq == 0 and train_b != len(batches):
            on_eval(train_b)
            cx_model.train()
    return train_b


def _load_pickle(path):
    with open(path, "rb") as f:
        return pickle.load(f)


def _load_features(dirname, fname):
    from .knn import load_features

    for cand in (os.path.join(dirname, fname), os.path.join(dirname, os.path.splitext(fname)[0] + ".npy")):
        if os.path.exists(cand):
            return load_features(cand)
    raise FileNotFoundError(os.path.join(dirname, fname))


def load_vqa_checkpoint(vqa_model, own_encoder, path):
    """counterexamples.py:250-252 (`load_vqa_checkpoint` -> model.load_state_dict, strict): a missing file, a missing
    head tensor or an unexpected key raises instead of leaving a randomly initialised head behind. `own_encoder` is
    the CUDA BayesianUniSkip that takes the checkpoint's `seq2vec.*` entries (None: the caller supplied its own
    question encoder, e.g. a precomputed QuestionTable, and those entries are not consumed)."""
    if not os.path.exists(path):
        raise FileNotFoundError("pretrained_vqa=True but there is no VQA checkpoint at {} (pass --untrained_vqa to "
                                "run with the initialised head)".format(path))
    state = torch.load(path, map_location="cpu")
    if isinstance(state, dict) and "state_dict" in state and not any(k.startswith("fusion.") for k in state):
        state = state["state_dict"]
    state = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in state.items()}   # nn.DataParallel
    head = {k: v for k, v in state.items() if not k.startswith("seq2vec.")}
    enc = {k[len("seq2vec."):]: v for k, v in state.items() if k.startswith("seq2vec.")}
    want = {k for k in vqa_model.state_dict().keys() if not k.startswith("seq2vec.")}
    missing, unexpected = sorted(want - set(head)), sorted(set(head) - want)
    if missing or unexpected:
        raise RuntimeError("VQA checkpoint {} does not match the {} head: missing {}, unexpected {}".format(
            path, type(vqa_model).__name__, missing[:8], unexpected[:8]))
    vqa_model.load_state_dict(head, strict=False)            # strict on the head (checked above); seq2vec handled below
    if own_encoder is not None:
        if not enc:
            raise RuntimeError("VQA checkpoint {} carries no seq2vec.* entries for the question encoder".format(path))
        own_encoder.load_state_dict(enc, strict=True)         # the checkpoint's fine-tuned encoder weights


def build_cx_model(cx_model_name, options, trainset, pretrained_vqa=True, seq2vec=None, sb_lambda=None):
    """counterexamples.py:213-273 for the models on this path."""
    if cx_model_name == "RandomBaseline":
        return RandomBaseline(knn_size=KNN), None
    if cx_model_name == "DistanceBaseline":
        return DistanceBaseline(knn_size=KNN), None
    if cx_model_name not in ("NeuralModel", "BlackBox", "LinearContext", "SemanticBaseline"):
        if cx_model_name in ("PairwiseModel", "PairwiseLinearModel"):
            raise NotImplementedError(f"cx_model {cx_model_name} (pairwise training) is outside the B200 hot path")
        raise ValueError("Unrecognized cx_model {}".format(cx_model_name))
    if cx_model_name == "SemanticBaseline" and sb_lambda is None:
        raise ValueError("If semantic baseline is selected then --sb_lambda must also be provided.")   # :237-239
    own_encoder = seq2vec is None
    if own_encoder:
        # the reference builds seq2vec from options['model']['seq2vec'] (models/seq2vec.py:10-18); the CUDA encoder
        # implements the arch the mutan_noatt options name (skipthoughts BayesianUniSkip, 620 -> 2400) and nothing else
        s2v = options["model"].get("seq2vec", {})
        if s2v.get("arch", "skipthoughts") != "skipthoughts" or s2v.get("type", "BayesianUniSkip") != "BayesianUniSkip":
            raise NotImplementedError("seq2vec {}/{}: only skipthoughts/BayesianUniSkip is built on this path".format(
                s2v.get("arch"), s2v.get("type")))
        from .vqa_models import BayesianUniSkip
        seq2vec = BayesianUniSkip(len(trainset["vocab_words"]))
    vqa_model = models.factory(options["model"], trainset["vocab_words"], trainset["vocab_answers"], cuda=True,
                               data_parallel=True, seq2vec=seq2vec)
    if pretrained_vqa:
        load_vqa_checkpoint(vqa_model, seq2vec if own_encoder else None,
                            os.path.join(options["logs"]["dir_logs"], "best_model.pth.tar"))     # train.py:355-359
    if cx_model_name == "BlackBox":                                                  # counterexamples.py:230-232
        return BlackBox(vqa_model, knn_size=KNN), vqa_model
    if cx_model_name == "LinearContext":                                             # :233-235
        return LinearContext(vqa_model, knn_size=KNN), vqa_model
    if cx_model_name == "SemanticBaseline":                                          # :236-246
        cx_model = SemanticBaseline(vqa_model, knn_size=KNN)
        cx_model.set_lambda(sb_lambda)
        cx_model.set_answer_embedding(_load_pickle(os.path.join(options["vqa"]["path_trainset"],
                                                                "answer_embedding.pickle")))
        return cx_model, vqa_model
    model_spec = options["cx_model"]
    emb = None
    if model_spec.get("pretrained_emb"):
        emb_path = os.path.join(options["vqa"]["path_trainset"], "answer_embedding.pickle")
        emb = _load_pickle(emb_path)
    cx_model = NeuralModel(model_spec=model_spec, dim_h=model_spec["dim_h"], n_layers=model_spec["n_layers"], emb=emb,
                           drop_p=model_spec["drop_p"], vqa_model=vqa_model, knn_size=KNN,
                           trainable_vqa=model_spec.get("trainable_vqa", False))
    return cx_model, vqa_model


def init_distributed():
    """One process per GPU under torchrun: binds this process to cuda:LOCAL_RANK and joins the process group (NCCL;
    VQACX_DIST_BACKEND=gloo for the single-GPU 2-rank smoke test). No-op for a plain `python counterexamples.py` run or
    when the caller initialised torch.distributed itself. Returns True when this call created the group."""
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1 or "RANK" not in os.environ or not dist.is_available():
        return False
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local % max(1, torch.cuda.device_count()))
    if dist.is_initialized():
        return False
    backend = os.environ.get("VQACX_DIST_BACKEND", "nccl")
    if backend == "nccl":
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    else:
        dist.init_process_group(backend)
    return True


def _broadcast_obj(dp, obj):
    """rank 0's python object on every rank (run names, confirm answers): identical paths on all ranks."""
    if dp.world == 1:
        return obj
    box = [obj]
    dp.dist.broadcast_object_list(box, src=0, group=dp.group)
    return box[0]


def _barrier(dp):
    if dp.world > 1:
        dp.dist.barrier(group=dp.group)


def main(argv=None, seq2vec=None):
    """counterexamples.py:90-389 for `--cx_model NeuralModel` (and the other heads on this path). `seq2vec`: the
    question encoder module (question_wids -> [B, dim_q]). None = the CUDA `BayesianUniSkip` (vqa_models.py; weights
    from the VQA checkpoint's `seq2vec.*` entries); callers may also hand in a `QuestionTable` of precomputed encodings.
    Under torchrun every rank runs this function: rank 0 owns the run name, the directories, the writers and the
    checkpoints; the others join the collectives and wait at the barriers around them."""
    args = parser.parse_args(argv)
    if args.pairwise:
        raise NotImplementedError("--pairwise trains PairwiseModel, outside the NeuralModel hot path")
    options = load_options(args)
    owns_group = init_distributed()
    dp = DataParallelCX()

    random.seed(42)
    torch.manual_seed(42)
    torch.cuda.manual_seed_all(42)

    if args.cx_model == "NeuralModel" and not args.comment:
        args.comment = options["cx_model"]["name"]
    if args.resume:
        run_name = args.resume
        save_dir = os.path.join(args.project_dir, "logs", "cx", run_name)
        assert os.path.isdir(save_dir)
        log_dir = None
        if dp.rank == 0:
            i = 1
            log_dir = os.path.join(args.project_dir, "runs", run_name, "resume_{}".format(i))
            while os.path.isdir(log_dir):
                i += 1
                log_dir = os.path.join(args.project_dir, "runs", run_name, "resume_{}".format(i))
        log_dir = _broadcast_obj(dp, log_dir)
    else:
        run_name = None
        if dp.rank == 0:                     # one clock: ranks can straddle a second boundary
            run_name = datetime.now().strftime("%b%d-%H-%M-%S")
            if args.comment:
                run_name += "_" + args.comment
        run_name = _broadcast_obj(dp, run_name)
        save_dir = os.path.join(args.project_dir, "logs", "cx", run_name)
        go = True
        if dp.rank == 0:                     # rank 0 alone asks, erases and creates; the others wait for its answer
            if os.path.isdir(save_dir):
                if not args.yes:
                    import click

                    go = bool(click.confirm("Save directory already exists in {}. Erase?".format(save_dir)))
                if go:
                    shutil.rmtree(save_dir)
            if go:
                os.makedirs(os.path.join(save_dir, "ckpt"))
                os.makedirs(os.path.join(save_dir, "best"))
        go = _broadcast_obj(dp, go)
        if not go:
            return None
        log_dir = os.path.join(args.project_dir, "runs", run_name)
    _barrier(dp)

    train_writer = val_writer = None
    if dp.rank == 0:
        try:
            from torch.utils.tensorboard import SummaryWriter

            train_writer = SummaryWriter(log_dir=os.path.join(log_dir, "train"))
            val_writer = SummaryWriter(log_dir=os.path.join(log_dir, "val"))
        except Exception:  # tensorboard is optional plumbing
            pass
    print("Saving model to {}".format(save_dir))
    print("Logging results to {}".format(log_dir))

    print("=> Loading VQA dataset...")
    pdir = os.path.join(options["vqa"]["path_trainset"], "pickle_old")
    trainset = _load_pickle(os.path.join(pdir, "trainset_augmented_small.pickle" if args.dev_mode
                                         else "trainset_augmented.pickle"))
    valset = _load_pickle(os.path.join(pdir, "valset_augmented_small.pickle"))
    testset = _load_pickle(os.path.join(pdir, "valset_augmented.pickle")) if args.test else None

    print("=> Loading COCO image features...")
    features_train = DeviceFeatures(_load_features(options["coco"]["path_features"], "trainset.hdf5"))
    features_val = DeviceFeatures(_load_features(options["coco"]["path_features"], "valset.hdf5"))

    print("=> Building model...")
    cx_model, vqa_model = build_cx_model(args.cx_model, options, trainset, pretrained_vqa=args.pretrained_vqa,
                                         seq2vec=seq2vec, sb_lambda=args.sb_lambda)
    print("Built {}".format(args.cx_model))
    trainable = isinstance(cx_model, (NeuralModel, LinearContext))

    if args.resume:
        info, start_epoch, best_recall = load_cx_checkpoint(cx_model, save_dir, resume_best=args.best)
    else:
        info, start_epoch, best_recall = [], 1, 0
    cx_model.cuda()

    print("=> Starting training...")
    epoch = None
    bs, lr = options["optim"]["batch_size"], options["optim"]["lr"]
    for epoch in range(start_epoch, options["optim"]["epochs"] + 1):
        n_batches = (len(trainset["examples_list"]) + bs - 1) // bs
        if trainable:
            def on_print(train_b, metrics, epoch=epoch, n_batches=n_batches):
                if dp.rank == 0:
                    log_results(train_writer, mode="train", epoch=epoch, i=(epoch - 1) * n_batches + train_b,
                                metrics=metrics)

            def on_eval(train_b, epoch=epoch, n_batches=n_batches):          # counterexamples.py:357-361 (-v)
                res = eval_model(cx_model, valset, features_val, bs, dp=dp)
                if dp.rank == 0:
                    log_results(val_writer, mode="val", epoch=epoch, i=(epoch - 1) * n_batches + train_b, metrics=res)

            train_epoch(cx_model, trainset, features_train, bs, lr, dp=dp, print_freq=args.print_freq,
                        on_print=on_print, eval_freq=args.eval_freq, on_eval=on_eval)
        eval_results = eval_model(cx_model, valset, features_val, bs, dp=dp)
        if dp.rank == 0:
            log_results(val_writer, mode="val", epoch=epoch, i=epoch * n_batches, metrics=eval_results)
        info.append(eval_results)
        # a diverged model (NaN scores) must not become "best": its recall is meaningless
        is_best = bool(np.isfinite(info[-1]["loss"])) and info[-1]["recall"] > best_recall
        if is_best:
            best_recall = info[-1]["recall"]
        if dp.rank == 0 and trainable:
            save_cx_checkpoint(cx_model, info, save_dir, is_best=is_best)
        _barrier(dp)                          # nobody reads ckpt/ or best/ before rank 0 has written them

    test_results = None
    if args.test:
        best_epoch = 0
        if epoch is not None and trainable and os.path.exists(os.path.join(save_dir, "best", "model.ckpt")):
            _, best_epoch, _ = load_cx_checkpoint(cx_model, save_dir, resume_best=True)
        test_results = eval_model(cx_model, testset, features_val, bs, dp=dp)
        test_results["best_epoch"] = best_epoch
        if dp.rank == 0:
            with open(os.path.join(save_dir, "final_results.txt"), "w") as outfile:
                outfile.write(json.dumps(test_results))
            print("Saved results to {}".format(save_dir))
            print("FINAL RESULTS ON BEST EPOCH {}".format(best_epoch), test_results)
    _barrier(dp)
    if owns_group:
        dp.dist.destroy_process_group()
    return {"info": info, "test": test_results, "save_dir": save_dir}


if __name__ == "__main__":
    main()
