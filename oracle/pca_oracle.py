"""TEST INFRASTRUCTURE ONLY -- CPU restatement of ChACRA's PCA permutation test.  Nothing on the
product path imports this file.  Pinned: tests/test_pca_oracle.py checks it against vectors produced by the
reference's own functions (tests/golden/make_pca_golden.py).

Follows chacra/ContactFrequencies.py:
  de_correlate          :631-639   out[i, c] = a[idx[i, c], c], idx = rand(*a.shape).argsort(0)
  run_permuted_pca      :641-656   PCA().fit(de_correlate(vals)).explained_variance_ratio_ per permutation
  ContactPCA.permuted_pca :949-962 |diff| of the ratios, p = mean(permuted > real), deepest component with p <= 0.05
sklearn's PCA (full SVD for these shapes) is restated with numpy: centre the columns, singular values S,
explained variance S^2/(n-1), ratio = that over its sum.
"""
import numpy as np


def de_correlate_indices(shape, rng):
    return rng.random(shape).argsort(0)


def apply_indices(a, idx):
    return a[idx, np.arange(a.shape[1])]


def explained_variance_ratio(x):
    x = np.asarray(x, dtype=np.float64)
    xc = x - x.mean(axis=0)
    s = np.linalg.svd(xc, compute_uv=False)
    ev = s ** 2 / max(x.shape[0] - 1, 1)
    tot = ev.sum()
    return ev / tot if tot > 0 else np.zeros_like(ev)


def run_permuted_pca(vals, indices):
    return np.array([explained_variance_ratio(apply_indices(vals, idx)) for idx in indices])


def pvalues(real_ratio, permuted_ratio):
    real_diffs = np.abs(np.diff(real_ratio, prepend=0))
    perm_diffs = np.abs(np.diff(permuted_ratio, axis=1, prepend=0))
    pvals = np.mean(perm_diffs > real_diffs, axis=0)
    hits = np.where(pvals <= 0.05)[0]
    return pvals, (int(hits[-1]) + 1 if len(hits) else 0)
