"""ContactPCA permutation test on the device against the reference-made vectors and the numpy
restatement.  Floating point: every explained-variance ratio within 1e-10 (absolute) of sklearn's / numpy's SVD."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
import pca_oracle as po

pytestmark = pytest.mark.gpu
TOL = 1e-10
G = np.load(os.path.join(GOLDEN, "pca_golden.npz"))


def test_reference_permutations_give_reference_ratios():
    from chacra_b200.pca import explained_variance_ratio, run_permuted_pca
    vals, idx = G["vals"], G["idx"].astype(np.int64)
    got = run_permuted_pca(vals, len(idx), permutations=idx)
    assert got.shape == G["evr"].shape
    assert np.abs(got - G["evr"]).max() < TOL
    assert np.abs(explained_variance_ratio(vals) - G["real"]).max() < TOL


@pytest.mark.parametrize("n,p", [(2, 7), (5, 3), (20, 4300), (21, 257), (32, 1000), (33, 300), (64, 513), (1, 4), (20, 1)])
def test_generated_permutations_are_permutations_and_ratios_match_the_restatement(n, p):
    from chacra_b200.pca import run_permuted_pca
    rng = np.random.default_rng(n * 1000 + p)
    vals = rng.random((n, p)) * rng.random(p)[None, :] + np.linspace(0, 1, n)[:, None] * rng.standard_normal(p)[None, :]
    vals[:, p // 2] = 0.25                                                  # a constant column
    n_perm = 5
    evr, idx = run_permuted_pca(vals, n_perm, seed=11, return_permutations=True)
    assert idx.shape == (n_perm, n, p) and evr.shape == (n_perm, min(n, p))
    assert (np.sort(idx, axis=1) == np.arange(n)[None, :, None]).all()
    want = po.run_permuted_pca(vals, idx.astype(np.int64))[:, :min(n, p)]
    assert np.abs(evr - want).max() < TOL
    again = run_permuted_pca(vals, n_perm, seed=11)
    assert np.array_equal(again, evr)                                       # same seed, same answer
    other = run_permuted_pca(vals, n_perm, seed=12, return_permutations=True)[1]
    if n > 3 and p > 3:
        assert not np.array_equal(other, idx)
        assert len({idx[0, :, c].tobytes() for c in range(min(p, 50))}) > 1    # columns are shuffled independently


def test_generator_is_uniform_enough():
    from chacra_b200.pca import run_permuted_pca
    vals = np.random.default_rng(0).random((4, 6000))
    _, idx = run_permuted_pca(vals, 4, seed=99, return_permutations=True)
    orders = idx.transpose(0, 2, 1).reshape(-1, 4)                           # 24 000 draws over the 24 orders of 4 rows
    codes = orders @ np.array([64, 16, 4, 1])
    _, counts = np.unique(codes, return_counts=True)
    assert len(counts) == 24
    chi2 = ((counts - 1000.0) ** 2 / 1000.0).sum()
    assert chi2 < 60.0                                                       # 23 degrees of freedom: P(chi2 > 60) ~ 3e-5


def test_significance_result_on_the_example_table():
    from chacra_b200.pca import permuted_pca
    res = permuted_pca(G["vals"], 200, seed=5)
    assert res["permuted_explained_variance"].shape == (200, 20)
    assert np.abs(res["explained_variance_ratio"] - G["real"]).max() < TOL
    assert res["chacra_pvals"][0] == 0.0 and res["top_chacras"][:2] == [1, 2]
    ref_like, _ = po.pvalues(G["real"], res["permuted_explained_variance"])
    assert np.array_equal(ref_like, res["chacra_pvals"])


def test_bad_input():
    from chacra_b200.pca import run_permuted_pca
    with pytest.raises(ValueError):
        run_permuted_pca(np.zeros((65, 3)), 1, seed=0)
    with pytest.raises(ValueError):
        run_permuted_pca(np.zeros((3, 3)), 1, seed=0, permutations=np.full((1, 3, 3), 3))
    assert run_permuted_pca(np.zeros((3, 5)), 0).shape == (0, 3)
    zero = run_permuted_pca(np.ones((3, 5)), 2, seed=1)                      # no variance at all: ratios 0, not NaN
    assert np.array_equal(zero, np.zeros((2, 3)))
