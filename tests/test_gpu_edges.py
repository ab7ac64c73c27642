"""Edge cases of the C ABI and size-independent properties on the bench shape (config 2 of BASELINE.json: 6 516 atoms,
all interaction types) -- the counts are integers, so every identity below is exact."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
ALL = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]


@pytest.fixture(scope="module")
def c2():
    import torch
    import bench
    from chacra_b200 import compile_topology
    from chacra_b200.synthetic import FrameGenerator
    path, atoms = bench.make_topology(4)
    topo = compile_topology(atoms)
    gen = FrameGenerator(atoms.xyz, 2001, amplitude=1.0, device="cuda")
    frames = torch.cat([gen.frames(500) for _ in range(6)])          # 3 000 frames = 30 fragments
    return topo, frames


def _counts(topo, submits, n_states=1, **kw):
    from chacra_b200.engine import ContactEngine
    eng = ContactEngine(topo, ALL, n_states=n_states, **kw)
    for xyz, state, first in submits:
        eng.submit(xyz, state, first)
    eng.synchronize()
    out = [eng.counts(s) for s in range(n_states)]
    eng.close()
    return out


def test_counts_are_additive_over_fragment_aligned_pieces_and_chunking_is_invisible(c2, monkeypatch):
    topo, x = c2
    whole = _counts(topo, [(x, 0, 0)])[0]
    assert whole.max() <= len(x) and whole.sum() > 0
    a = _counts(topo, [(x[:1300], 0, 0)])[0]
    b = _counts(topo, [(x[1300:], 0, 1300)])[0]
    assert np.array_equal(a + b, whole)                                     # linearity
    pieces = _counts(topo, [(x[s:s + 700], 0, s) for s in range(0, 3000, 700)])[0]
    assert np.array_equal(pieces, whole)                                    # continued fragments (700 is not a multiple of 100)
    monkeypatch.setenv("CHACRA_CHUNK_FRAMES", "333")                        # kernel chunks that cut fragments
    assert np.array_equal(_counts(topo, [(x, 0, 0)])[0], whole)
    monkeypatch.delenv("CHACRA_CHUNK_FRAMES")
    host = _counts(topo, [(x.cpu().numpy(), 0, 0)])[0]                       # host-buffer entry point
    assert np.array_equal(host, whole)
    twice = _counts(topo, [(x, 0, 0), (x, 0, 0)])[0]
    assert np.array_equal(twice, 2 * whole)


def test_states_do_not_mix_and_fragment_order_is_free(c2):
    topo, x = c2
    s0, s1, s2 = _counts(topo, [(x[:1000], 0, 0), (x[1000:2000], 2, 0)], n_states=3)
    assert s1.sum() == 0 and s0.sum() > 0 and s2.sum() > 0
    assert np.array_equal(s0, _counts(topo, [(x[:1000], 0, 0)])[0])
    # whole fragments may arrive in any order: the hydrophobic set belongs to the fragment, not to the history
    fwd = _counts(topo, [(x[:1000], 0, 0)])[0]
    rev = _counts(topo, [(x[k:k + 100], 0, k) for k in range(900, -1, -100)])[0]
    assert np.array_equal(fwd, rev)


def test_empty_single_and_invalid_submissions(c2):
    import torch
    from chacra_b200.engine import ContactEngine, ChacraError
    topo, x = c2
    eng = ContactEngine(topo, ALL, n_states=2)
    eng.submit(x[:0], 0, 0)                                                 # zero frames: a no-op
    eng.synchronize()
    assert eng.counts(0).sum() == 0
    eng.submit(x[:1], 1, 0)                                                 # one frame
    eng.synchronize()
    one = eng.counts(1)
    assert one.max() == 1 and one.sum() > 1000 and eng.counts(0).sum() == 0
    with pytest.raises(ChacraError):
        eng.submit(x[:10], 2, 0)                                            # no such state
    with pytest.raises(ChacraError):
        eng.submit(x[:10], 0, 150)                                          # starts inside a fragment it does not continue
    with pytest.raises(ChacraError):
        eng.submit(x[:10, :-1], 0, 0)                                       # wrong atom count
    with pytest.raises(ChacraError):
        eng.submit(x[:10].double(), 0, 0)                                   # wrong dtype
    eng.submit(x[:150], 0, 0)
    eng.submit(x[150:260], 0, 150)                                          # continuing the open fragment is fine
    eng.synchronize()
    assert np.array_equal(eng.counts(0), _counts(topo, [(x[:260], 0, 0)])[0])
    eng.close()


def test_degenerate_frames_do_not_disturb_their_neighbours():
    """Frames are independent: NaN, infinite, absurdly large or fully collapsed coordinates in some frames of one fragment
    leave the counts of the other fragments exactly as they are (and nothing hangs or faults)."""
    import torch
    import bench
    from chacra_b200 import compile_topology
    from chacra_b200.synthetic import FrameGenerator
    path, atoms = bench.make_topology(1)
    topo = compile_topology(atoms)
    x = FrameGenerator(atoms.xyz, 77, amplitude=1.0, device="cuda").frames(300)
    clean = _counts(topo, [(x[:100], 0, 0), (x[200:], 0, 200)])[0]
    bad = x.clone()
    bad[120, 5] = float("nan")
    bad[121, ::7] = float("inf")
    bad[122] *= 1e30
    bad[123] = bad[123, 0]                                    # every atom on one point
    bad[124, :, 2] = 0.0                                      # flattened into a plane
    for env in ({}, {"CHACRA_BIN_VARIANT": "0"}, {"CHACRA_BIN_VARIANT": "4"}, {"CHACRA_FORCE_LARGE": "1"}):
        os.environ.update(env)
        try:
            a = _counts(topo, [(bad[:100], 0, 0), (bad[200:], 0, 200)])[0]
            whole = _counts(topo, [(bad, 0, 0)])[0]
            middle = _counts(topo, [(bad[100:200], 0, 100)])[0]
        finally:
            for k in env:
                os.environ.pop(k, None)
        assert np.array_equal(a, clean)
        assert np.array_equal(whole, clean + middle)          # still additive over fragments
        assert middle.max() <= 100
