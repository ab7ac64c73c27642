"""Timing of the ContactPCA permutation test on the shape SURVEY names: 1 000 permutations of a
20 x 4 300 table.  GPU: chacra_b200.pca.run_permuted_pca (CUDA events around the call's kernels, plus wall time with
the transfers); CPU: the reference's way, sklearn PCA().fit on a de-correlated copy per permutation, one process.
  python tools/pca_run.py [--perms 1000] [--cpu-perms 50]"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--perms", type=int, default=1000)
    ap.add_argument("--cpu-perms", type=int, default=50)
    ap.add_argument("--states", type=int, default=20)
    ap.add_argument("--contacts", type=int, default=4300)
    a = ap.parse_args()
    import torch
    from sklearn.decomposition import PCA
    from chacra_b200.pca import run_permuted_pca
    rng = np.random.default_rng(1)
    melt = np.linspace(1.0, 0.0, a.states)[:, None] ** rng.uniform(0.3, 3.0, a.contacts)[None, :]
    vals = np.clip(melt * rng.random(a.contacts)[None, :] + 0.02 * rng.standard_normal((a.states, a.contacts)), 0, 1)
    run_permuted_pca(vals, 8, seed=0)                   # library load, first launch
    torch.cuda.synchronize()
    best_wall, best_dev = 1e9, 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        evr = run_permuted_pca(vals, a.perms, seed=3)
        e1.record()
        torch.cuda.synchronize()
        best_wall = min(best_wall, time.perf_counter() - t0)
        best_dev = min(best_dev, e0.elapsed_time(e1) * 1e-3)
    t0 = time.perf_counter()
    pca = PCA()
    for _ in range(a.cpu_perms):
        idx = np.random.rand(*vals.shape).argsort(0)
        pca.fit(vals[idx, np.arange(vals.shape[1])]).explained_variance_ratio_
    cpu = (time.perf_counter() - t0) / a.cpu_perms
    print("permuted PCA %d x %d: GPU %d permutations in %.2f ms on the stream (%.2f ms wall with transfers) = %.0f permutations/s | "
          "sklearn on one core %.2f ms per permutation = %.0f permutations/s | first ratios %s" % (
              a.states, a.contacts, a.perms, best_dev * 1e3, best_wall * 1e3, a.perms / best_wall, cpu * 1e3, 1.0 / cpu,
              np.round(evr.mean(axis=0)[:3], 4)))


if __name__ == "__main__":
    main()
