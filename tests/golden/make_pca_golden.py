"""Golden vectors for the ContactPCA permutation test, made in the build container by running the
REFERENCE's own ``de_correlate`` / ``run_permuted_pca`` (chacra/ContactFrequencies.py:631-656).  The module cannot be
imported whole (Bio, MDAnalysis, pymol helpers are absent), so the two function definitions are taken from the file
with ``ast`` and executed unchanged with numpy and sklearn's PCA in scope.  Data: every 11th contact column of
chacra/examples/total_contacts.pd (20 states).  The row orders are recovered by replaying numpy's global generator
(``de_correlate`` draws exactly one ``rand(*a.shape)`` per call; the full-SVD PCA draws nothing).
    python tests/golden/make_pca_golden.py          # needs /root/reference; writes tests/golden/pca_golden.npz
"""
import ast
import os

import numpy as np
import pandas as pd
from sklearn.decomposition import PCA

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/chacra/ContactFrequencies.py"


def reference_functions():
    tree = ast.parse(open(REF).read())
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("de_correlate", "run_permuted_pca")]
    assert len(keep) == 2
    ns = {"np": np, "PCA": PCA}
    exec(compile(ast.Module(body=keep, type_ignores=[]), REF, "exec"), ns)
    return ns["de_correlate"], ns["run_permuted_pca"]


def main():
    de_correlate, run_permuted_pca = reference_functions()
    df = pd.read_pickle("/root/reference/chacra/examples/total_contacts.pd")
    vals = np.ascontiguousarray(df.values[:, ::11], dtype=np.float64)
    n_perm = 6
    np.random.seed(20240611)
    evr = np.array(run_permuted_pca(vals, n_perm))
    np.random.seed(20240611)
    idx = np.stack([np.random.rand(*vals.shape).argsort(0) for _ in range(n_perm)])
    for k in range(n_perm):                              # the replay really is what the reference permuted
        again = PCA().fit(vals[idx[k], np.arange(vals.shape[1])]).explained_variance_ratio_
        assert np.array_equal(again, evr[k])
    np.random.seed(3)
    assert np.array_equal(de_correlate(vals), vals[np.random.RandomState(3).rand(*vals.shape).argsort(0), np.arange(vals.shape[1])])
    real = PCA().fit(vals).explained_variance_ratio_     # ContactPCA.__init__, ContactFrequencies.py:700-703
    np.savez_compressed(os.path.join(HERE, "pca_golden.npz"), vals=vals, idx=idx.astype(np.uint8), evr=evr, real=real)
    print("vals", vals.shape, "evr", evr.shape, "first ratios", evr[0][:3], "real", real[:3])


if __name__ == "__main__":
    main()
