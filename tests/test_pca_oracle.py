"""The PCA permutation-test restatement against vectors made by the reference's own de_correlate / run_permuted_pca
(tests/golden/make_pca_golden.py), and the host-side p-value step of the product."""
import os

import numpy as np

from conftest import GOLDEN
import pca_oracle as po
from chacra_b200.pca import permutation_pvalues

G = np.load(os.path.join(GOLDEN, "pca_golden.npz"))
TOL = 1e-12


def test_oracle_reproduces_the_reference_ratios():
    vals, idx = G["vals"], G["idx"].astype(np.int64)
    got = po.run_permuted_pca(vals, idx)
    assert got.shape == G["evr"].shape == (6, 20)
    assert np.abs(got - G["evr"]).max() < TOL
    assert np.abs(po.explained_variance_ratio(vals) - G["real"]).max() < TOL
    assert np.all(np.diff(got, axis=1) <= 1e-15) and np.allclose(got.sum(axis=1), 1.0)


def test_every_golden_index_column_is_a_row_permutation():
    idx = G["idx"]
    assert (np.sort(idx, axis=1) == np.arange(20)[None, :, None]).all()


def test_pvalue_step_matches_the_restatement_and_known_cases():
    real, perm = G["real"], G["evr"]
    p1, top = permutation_pvalues(real, perm)
    p2, deepest = po.pvalues(real, perm)
    assert np.array_equal(p1, p2) and top == list(range(1, deepest + 1))
    assert p1[0] == 0.0 and top[:1] == [1]              # the real first component towers over any permuted one
    p, top = permutation_pvalues(np.array([0.6, 0.3, 0.1]), np.array([[0.4, 0.35, 0.25]]))
    assert np.array_equal(p, [0.0, 0.0, 0.0]) and top == [1, 2, 3]
    p, top = permutation_pvalues(np.array([0.2, 0.7, 0.1]), np.array([[0.5, 0.3, 0.2]]))
    assert np.array_equal(p, [1.0, 0.0, 0.0]) and top == [1, 2, 3]      # the deepest significant index decides, as in the reference
    p, top = permutation_pvalues(np.array([0.5, 0.5]), np.array([[0.9, 0.1]]))
    assert np.array_equal(p, [1.0, 1.0]) and top == []
