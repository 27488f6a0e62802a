"""The two command-line entry points of the path, end to end on the GPU:
`knn.main([...])` (knn.py:23-58: load -> fit -> kneighbors -> np.save({"indices", "distances"})) and
`counterexamples.main([...])` (counterexamples.py:90-389: options, pickles, features, train epochs, eval, checkpoints,
--test), single process and two ranks (one process per rank, torch.distributed)."""
import json
import os
import pickle
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import ncx_oracle as O
from oracle.knn_oracle import compare_knn, knn_restatement_f64
from vqa_counterexamples_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DV, DQ, DF, R, NA, DIM_H, L = 64, 48, 40, 3, 37, 16, 2
N_IMG, N_Q = 90, 40


def test_knn_main_writes_the_reference_output_dict(tmp_path, capsys):
    """knn.py CLI on a .npy feature file: the saved file is the pickled dict knn.py:58 writes --
    {"indices": int64 [N, k], "distances": float64 [N, k]} -- and equals the oracle."""
    from vqa_counterexamples_b200 import knn

    n, d, k = 1500, 128, 25
    x = synth.clustered_features_np(n, d, seed=31)
    x[17] = x[3]                                   # an exact duplicate: tie order by index
    base, save = tmp_path / "extract", tmp_path / "out"
    base.mkdir()
    save.mkdir()
    np.save(base / "trainset.npy", x)
    ret_d, ret_i = knn.main([str(base), "--hdf5_file", "trainset.npy", "--save_dir", str(save), "--save_file", "knn.npy",
                             "-k", str(k), "-b", "10"])
    out = capsys.readouterr().out
    assert "Loaded features as array of size (1500, 128)" in out and "Fitting KNN model for k=25" in out
    res = np.load(save / "knn.npy", allow_pickle=True).item()
    assert sorted(res.keys()) == ["distances", "indices"]
    assert res["indices"].dtype == np.int64 and res["indices"].shape == (n, k)
    assert res["distances"].dtype == np.float64 and res["distances"].shape == (n, k)
    assert np.array_equal(res["indices"], ret_i) and np.array_equal(res["distances"], ret_d)
    ref_d, ref_i = knn_restatement_f64(x, x, k=k, direct=True)
    rep = compare_knn(res["distances"], res["indices"], ref_d, ref_i)
    assert rep["ok"], rep
    # a second run overwrites, with the reference's warning
    knn.main([str(base), "--hdf5_file", "trainset.npy", "--save_dir", str(save), "--save_file", "knn.npy", "-k", "5"])
    assert "already exists and will be overwritten" in capsys.readouterr().out
    assert np.load(save / "knn.npy", allow_pickle=True).item()["indices"].shape == (n, 5)


def _write_project(root, seed=8):
    """A miniature of the reference's data layout: options yaml, pickle_old/{train,val}set_augmented*.pickle with the
    example schema of SURVEY 3.1, feature matrices, a VQA checkpoint under logs.dir_logs and an answer embedding."""
    rng = np.random.default_rng(seed)
    params = O.make_params(seed, dim_v=DV, dim_q=DQ, dim_h_fusion=DF, R=R, n_ans=NA, dim_h=DIM_H, n_layers=L)
    names = ["COCO_val2014_%012d.jpg" % i for i in range(N_IMG)]
    name_to_index = {nm: i for i, nm in enumerate(names)}

    def dataset(n_examples):
        ex = []
        for e in range(n_examples):
            img = int(rng.integers(N_IMG))
            knn = rng.choice([i for i in range(N_IMG) if i != img], size=24, replace=False)
            ex.append({"image_name": names[img], "knns": [names[j] for j in knn],
                       "question_wids": [int(rng.integers(N_Q))] + [0] * 25, "answer_aid": int(rng.integers(NA)),
                       "comp": {"knn_index": e % 3}, "question_id": e})
        return {"examples_list": ex, "name_to_index": name_to_index, "vocab_words": ["w"] * 7, "vocab_answers": ["a"] * NA}

    trainset = os.path.join(root, "trainset")
    os.makedirs(os.path.join(trainset, "pickle_old"))
    for fname, n in (("trainset_augmented.pickle", 96), ("trainset_augmented_small.pickle", 48),
                     ("valset_augmented_small.pickle", 40), ("valset_augmented.pickle", 56)):
        with open(os.path.join(trainset, "pickle_old", fname), "wb") as f:
            pickle.dump(dataset(n), f)
    with open(os.path.join(trainset, "answer_embedding.pickle"), "wb") as f:
        pickle.dump(params["answer_embedding.weight"], f)
    feats_dir = os.path.join(root, "extract")
    os.makedirs(feats_dir)
    feats = np.maximum(rng.normal(0.3, 0.5, size=(N_IMG, DV)), 0).astype(np.float32)
    np.save(os.path.join(feats_dir, "trainset.npy"), feats)
    np.save(os.path.join(feats_dir, "valset.npy"), feats)
    logs = os.path.join(root, "logs", "vqa")
    os.makedirs(logs)
    torch.save({k[len("vqa_model."):]: torch.as_tensor(v) for k, v in params.items() if k.startswith("vqa_model.")},
               os.path.join(logs, "best_model.pth.tar"))
    q_table = np.tanh(rng.normal(0, 0.5, size=(N_Q, DQ))).astype(np.float32)
    np.save(os.path.join(root, "q_table.npy"), q_table)
    opt = {"logs": {"dir_logs": logs}, "vqa": {"path_trainset": trainset}, "coco": {"path_features": feats_dir},
           "optim": {"lr": 1e-2, "batch_size": 16, "epochs": 2},
           "model": {"arch": "MutanNoAtt", "seq2vec": {"arch": "skipthoughts", "type": "BayesianUniSkip"},
                     "fusion": {"dim_v": DV, "dim_q": DQ, "dim_hv": DF, "dim_hq": DF, "dim_mm": DF, "R": R},
                     "classif": {"dropout": 0.5}},
           "cx_model": {"name": "neuralcx_test", "dim_h": DIM_H, "n_layers": L, "drop_p": 0.0, "pretrained_emb": True,
                        "trainable_vqa": False, **{k: True for k in ("v_emb", "v_mult", "v_dist", "v_rank", "q_emb",
                                                                     "a_emb", "z_emb")}}}
    import yaml

    path_opt = os.path.join(root, "opt.yaml")
    with open(path_opt, "w") as f:
        yaml.safe_dump(opt, f)
    return path_opt, q_table


def test_counterexamples_main_trains_evaluates_checkpoints_and_tests(tmp_path, capsys):
    from vqa_counterexamples_b200 import counterexamples as cxm
    from vqa_counterexamples_b200.vqa_models import QuestionTable

    root = str(tmp_path)
    path_opt, q_table = _write_project(root)
    argv = ["--cx_model", "NeuralModel", "--path_opt", path_opt, "--project_dir", root, "-p", "2", "-v", "4", "-t", "-y",
            "-c", "clitest"]
    out = cxm.main(argv, seq2vec=QuestionTable(q_table).cuda())
    text = capsys.readouterr().out
    assert "Built NeuralModel" in text and "=> Starting training..." in text and "FINAL RESULTS ON BEST EPOCH" in text
    assert text.count(" train: ") == 2 * 3 and text.count(" val: ") >= 2 + 2      # -p 2 over 6 batches, -v 4 + epoch ends
    save_dir = out["save_dir"]
    assert os.path.basename(save_dir).endswith("_clitest")
    for sub in ("ckpt", "best"):
        assert os.path.exists(os.path.join(save_dir, sub, "model.ckpt")) and os.path.exists(os.path.join(save_dir, sub, "info.ckpt"))
    assert len(out["info"]) == 2 and all(np.isfinite(e["loss"]) for e in out["info"])
    assert out["info"][1]["loss"] < out["info"][0]["loss"] + 0.05            # labels are learnable from the rank one-hot
    final = json.load(open(os.path.join(save_dir, "final_results.txt")))
    assert set(final) >= {"loss", "recall", "best_epoch"} and final["best_epoch"] in (2, 3)
    # --resume continues from the checkpoint (the reference's own KeyError on 'recall_5' is documented in SURVEY A.1)
    out2 = cxm.main(["--cx_model", "NeuralModel", "--path_opt", path_opt, "--project_dir", root, "--resume",
                     os.path.basename(save_dir), "--epochs", "3", "-y"], seq2vec=QuestionTable(q_table).cuda())
    assert len(out2["info"]) == 3
    # a missing pretrained checkpoint is an error, not a silently random head
    os.remove(os.path.join(root, "logs", "vqa", "best_model.pth.tar"))
    with pytest.raises(FileNotFoundError):
        cxm.main(argv, seq2vec=QuestionTable(q_table).cuda())


def test_counterexamples_main_trainable_vqa(tmp_path, capsys):
    """--trainable_vqa (cx.py:73-74, options/cx/neuralcx_trainable_vqa.yaml): the driver trains the MUTAN head along with
    NeuralCX; the checkpoint's vqa_model.* entries are the TRAINED head, and a resumed run starts from them."""
    from vqa_counterexamples_b200 import counterexamples as cxm
    from vqa_counterexamples_b200.vqa_models import QuestionTable

    root = str(tmp_path)
    path_opt, q_table = _write_project(root)
    pre = torch.load(os.path.join(root, "logs", "vqa", "best_model.pth.tar"))
    out = cxm.main(["--cx_model", "NeuralModel", "--path_opt", path_opt, "--project_dir", root, "--trainable_vqa", "-y",
                    "-c", "tv"], seq2vec=QuestionTable(q_table).cuda())
    assert len(out["info"]) == 2 and all(np.isfinite(e["loss"]) for e in out["info"])
    ckpt = torch.load(os.path.join(out["save_dir"], "ckpt", "model.ckpt"), map_location="cpu")
    sd = ckpt["state_dict"] if "state_dict" in ckpt else ckpt
    for name in ("fusion.linear_v.weight", "fusion.list_linear_hq.1.bias", "linear_classif.weight"):
        assert not torch.equal(sd["vqa_model." + name].float().cpu(), pre[name].float()), name
    out2 = cxm.main(["--cx_model", "NeuralModel", "--path_opt", path_opt, "--project_dir", root, "--trainable_vqa",
                     "--resume", os.path.basename(out["save_dir"]), "--epochs", "3", "-y"],
                    seq2vec=QuestionTable(q_table).cuda())
    assert len(out2["info"]) == 3 and np.isfinite(out2["info"][2]["loss"])


WORKER = r"""
import os, sys, json
sys.path.insert(0, {root!r})
import numpy as np, torch
from vqa_counterexamples_b200 import counterexamples as cxm
from vqa_counterexamples_b200.vqa_models import QuestionTable
q_table = np.load(os.path.join({proj!r}, "q_table.npy"))
out = cxm.main(["--cx_model", "NeuralModel", "--path_opt", {opt!r}, "--project_dir", {proj!r}, "-t", "-y", "-c", "two",
                "--dev_mode"], seq2vec=QuestionTable(q_table).cuda())
rank = int(os.environ["RANK"])
with open(os.path.join({proj!r}, "rank%d.json" % rank), "w") as f:
    json.dump({{"save_dir": out["save_dir"], "info": out["info"], "test": out["test"]}}, f)
"""


def test_counterexamples_main_two_ranks(tmp_path):
    """torchrun-style launch of main() with two ranks: both bind a device and join the group, rank 0 alone
    names / creates the run directory and writes checkpoints, every rank reads `best` only after the barrier, and both
    report the same metrics. Two GPUs -> NCCL; one GPU -> both ranks share it over gloo (the collectives of this path,
    all_reduce / broadcast_object_list / barrier, exist there)."""
    root = str(tmp_path)
    path_opt, _ = _write_project(root)
    script = os.path.join(root, "worker.py")
    with open(script, "w") as f:
        f.write(WORKER.format(root=ROOT, proj=root, opt=path_opt))
    env = dict(os.environ)
    if torch.cuda.device_count() < 2:
        env["VQACX_DIST_BACKEND"] = "gloo"
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29671", script], env=env, capture_output=True,
                         text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    a = json.load(open(os.path.join(root, "rank0.json")))
    b = json.load(open(os.path.join(root, "rank1.json")))
    assert a["save_dir"] == b["save_dir"] and a["info"] == b["info"] and a["test"] == b["test"]
    assert len(os.listdir(os.path.join(root, "logs", "cx"))) == 1               # one run directory, made by rank 0
    assert os.path.exists(os.path.join(a["save_dir"], "final_results.txt"))
