"""ContactPCA permutation test on the GPU.

Mirrors ``run_permuted_pca`` / ``de_correlate`` (chacra/ContactFrequencies.py:631-656) and the p-value step of
``ContactPCA.permuted_pca`` (:910-964): every contact column of the [states x contacts] frequency table gets its rows
reordered independently, a PCA is fitted, and the explained-variance ratios of N such permutations are compared with
those of the real data.  The device kernel (``csrc/pca_perm.cuh`` behind ``chc_permuted_pca``) works on the Gram matrix
of the centred columns in float64; tolerance against sklearn's full-SVD PCA: 1e-10 absolute on every ratio
(tests/test_gpu_pca.py).  There is no CPU path here: without the CUDA library the call raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

MAX_SAMPLES = 64


def _run(vals, n_perm, perms, seed, want_perms, device):
    import torch
    if not torch.cuda.is_available():
        raise _lib.ChacraLibraryError("the permutation test runs on the GPU (sm_100a); there is no CPU fallback")
    vals = np.ascontiguousarray(vals, dtype=np.float64)
    if vals.ndim != 2 or vals.shape[0] < 1 or vals.shape[1] < 1:
        raise ValueError("vals must be a non-empty [samples][features] table")
    n, p = vals.shape
    if n > MAX_SAMPLES:
        raise ValueError("at most %d samples (thermodynamic states), got %d" % (MAX_SAMPLES, n))
    dev = torch.device("cuda", device)
    lib = _lib.load()
    d_vals = torch.from_numpy(vals).to(dev)
    d_work = torch.empty(n * p, dtype=torch.float64, device=dev)
    ncomp = min(n, p)
    d_evr = torch.zeros((n_perm, ncomp), dtype=torch.float64, device=dev)
    d_in = d_out = None
    if perms is not None:
        perms = np.asarray(perms)
        if perms.shape != (n_perm, n, p):
            raise ValueError("permutations must have shape [n_perm][samples][features] = %s" % ((n_perm, n, p),))
        if perms.min() < 0 or perms.max() >= n:
            raise ValueError("permutation entries must be row numbers 0..%d" % (n - 1))
        d_in = torch.from_numpy(np.ascontiguousarray(perms.transpose(0, 2, 1).astype(np.uint8))).to(dev)
    if want_perms:
        d_out = torch.empty((n_perm, p, n), dtype=torch.uint8, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    rc = lib.chc_permuted_pca(d_vals.data_ptr(), n, p, n_perm, d_in.data_ptr() if d_in is not None else None,
                              d_out.data_ptr() if d_out is not None else None, C.c_uint64(seed & (2 ** 64 - 1)), d_work.data_ptr(),
                              d_evr.data_ptr(), device, stream)
    if rc != 0:
        raise RuntimeError("chc_permuted_pca failed (%d)" % rc)
    evr = d_evr.cpu().numpy()
    if want_perms:
        return evr, d_out.cpu().numpy().transpose(0, 2, 1)
    return evr


def explained_variance_ratio(vals, device: int = 0) -> np.ndarray:
    """PCA().fit(vals).explained_variance_ratio_ of the data as they are (ContactPCA.__init__, ContactFrequencies.py:700-703)."""
    vals = np.asarray(vals)
    ident = np.broadcast_to(np.arange(vals.shape[0])[None, :, None], (1,) + vals.shape)
    return _run(vals, 1, ident, 0, False, device)[0]


def run_permuted_pca(vals, n_permutations: int, seed: int | None = None, permutations=None, return_permutations: bool = False, device: int = 0):
    """[n_permutations][n_components] explained-variance ratios of column-wise row-permuted data
    (``run_permuted_pca``, ContactFrequencies.py:641-656; the reference returns the same rows as a list).

    ``seed``: the device generator's key (None: drawn from numpy's global generator, like the reference's
    unseeded ``np.random.rand``).  ``permutations``: int array [n_permutations][samples][features] of source rows, the
    ``idx`` of ``de_correlate`` (:631-639), used instead of the generator."""
    if seed is None:
        seed = int(np.random.randint(0, 2 ** 63 - 1, dtype=np.int64))
    if n_permutations <= 0:
        vals = np.asarray(vals)
        return np.zeros((0, min(vals.shape)), np.float64)
    return _run(vals, int(n_permutations), permutations, int(seed), return_permutations, device)


def permutation_pvalues(real_ratio, permuted_ratio):
    """(chacra_pvals, top_chacras) exactly as ContactPCA.permuted_pca computes them (ContactFrequencies.py:949-962)."""
    real_diffs = np.abs(np.diff(np.asarray(real_ratio), prepend=0))
    perm_diffs = np.abs(np.diff(np.asarray(permuted_ratio), axis=1, prepend=0))
    pvals = np.mean(perm_diffs > real_diffs, axis=0)
    hits = np.where(pvals <= 0.05)[0]
    deepest = int(hits[-1]) + 1 if len(hits) else 0
    return pvals, list(range(1, deepest + 1))


def permuted_pca(freqs, N_permutations: int = 500, seed: int | None = None, device: int = 0) -> dict:
    """The significance test of ``ContactPCA.permuted_pca`` for a frequency table (DataFrame or array, states x contacts)."""
    vals = np.asarray(getattr(freqs, "values", freqs), dtype=np.float64)
    real = explained_variance_ratio(vals, device)
    variance = run_permuted_pca(vals, N_permutations, seed, device=device)
    pvals, top = permutation_pvalues(real, variance)
    return {"explained_variance_ratio": real, "permuted_explained_variance": variance, "chacra_pvals": pvals, "top_chacras": top}
