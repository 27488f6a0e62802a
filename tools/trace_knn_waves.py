import os, sys, torch
sys.path.insert(0, ".")
from vqa_counterexamples_b200 import synth
from vqa_counterexamples_b200.knn import NearestNeighbors
x = synth.clustered_features_cuda(123287, 2048, seed=1234)
nn_ = NearestNeighbors(n_neighbors=25).fit(x)
for _ in range(3): nn_.kneighbors(x)
os.environ["VQACX_KNN_WAVE_TRACE"] = "1"
os.environ["VQACX_KNN_OVERLAP"] = "1"
nn_.kneighbors(x)
torch.cuda.synchronize()
