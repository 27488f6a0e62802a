"""B200-native VQA-CX hot path: exact top-24 visual kNN (knn.py) feeding NeuralCX scoring
(counterexamples.py --cx_model NeuralModel). Host code is plumbing over libvqacx.so (include/vqacx.h).

Submodules import the CUDA library on first use and fail loudly if it has not been built; there is no
CPU / PyTorch fallback on the product path.
"""
__version__ = "0.1.0"
