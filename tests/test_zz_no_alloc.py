"""chc_permuted_pca_static: the allocation-free variant of the permutation test (scratch in a static device array inside the
library).  No counterpart in the reference; the check is against the existing entry point chc_permuted_pca on the same inputs,
bit for bit (same kernels, same arithmetic), on two calls in a row with different shapes so that the second call finds the first
call's contents in the shared array.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ELEMS = 1 << 19          # CHC_PCA_STATIC_MAX_ELEMS in include/chacra_b200.h


def _both(lib, torch, vals, n_perm, seed):
    """(ratios, permutations) through chc_permuted_pca and through chc_permuted_pca_static"""
    dev = torch.device("cuda", 0)
    n, p = vals.shape
    d_vals = torch.from_numpy(vals).to(dev)
    d_work = torch.empty(n * p, dtype=torch.float64, device=dev)
    ncomp = min(n, p)
    evr = [torch.full((n_perm, ncomp), -1.0, dtype=torch.float64, device=dev) for _ in range(2)]
    perm = [torch.zeros((n_perm, p, n), dtype=torch.uint8, device=dev) for _ in range(2)]
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    rc = lib.chc_permuted_pca(d_vals.data_ptr(), n, p, n_perm, None, perm[0].data_ptr(), C.c_uint64(seed), d_work.data_ptr(),
                              evr[0].data_ptr(), 0, stream)
    assert rc == 0
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    rc = lib.chc_permuted_pca_static(d_vals.data_ptr(), n, p, n_perm, None, perm[1].data_ptr(), C.c_uint64(seed), evr[1].data_ptr(),
                                     0, stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] >= free_before          # nothing was taken from the device heap
    return [e.cpu().numpy() for e in evr], [q.cpu().numpy() for q in perm]


def test_static_scratch_entry_point_matches_the_existing_one_twice():
    import torch
    from chacra_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(2024)
    for n, p, n_perm, seed in ((20, 4300, 12, 7), (33, 257, 9, 8), (20, 4300, 12, 7)):     # large, small (stale tail), large again
        vals = rng.random((n, p))
        evr, perm = _both(lib, torch, vals, n_perm, seed)
        assert np.array_equal(perm[0], perm[1])
        assert np.array_equal(evr[0], evr[1])                     # bit-exact: identical kernels on identical data
        assert np.all(evr[1] >= 0.0) and np.allclose(evr[1].sum(axis=1), 1.0)


def test_static_scratch_entry_point_rejects_tables_beyond_its_array():
    import torch
    from chacra_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    n, p = 64, MAX_ELEMS // 64 + 1
    d_vals = torch.zeros((n, p), dtype=torch.float64, device=dev)
    d_evr = torch.zeros((2, n), dtype=torch.float64, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    assert lib.chc_permuted_pca_static(d_vals.data_ptr(), n, p, 2, None, None, C.c_uint64(1), d_evr.data_ptr(), 0, stream) == -3   # CHC_E_TOO_LARGE
    # the largest table that fits still runs
    p = MAX_ELEMS // 64
    d_vals = torch.rand((n, p), dtype=torch.float64, device=dev)
    assert lib.chc_permuted_pca_static(d_vals.data_ptr(), n, p, 2, None, None, C.c_uint64(1), d_evr.data_ptr(), 0, stream) == 0
    torch.cuda.synchronize()
    assert np.allclose(d_evr.cpu().numpy().sum(axis=1), 1.0)
