"""The XTC codec restatement (oracle/xtc_oracle.py) and the host-side frame index (chacra_b200/xtc.py) -- CPU only.

Parity is *unpinned* against real XTC files (none exist in the reference tree or in this image); what is pinned here is
that decoder and encoder, both written from the published xdrfile algorithm, agree on every stream feature."""
import os
import struct

import numpy as np
import pytest

import xtc_oracle as xo
from conftest import GOLDEN


def _walk(rng, n, wide=False):
    steps = rng.integers(-150, 150, (n, 3))
    jumps = (rng.random((n, 1)) < 0.1) * rng.integers(-5000, 5000, (n, 3))
    ints = np.cumsum(steps + jumps, 0) + rng.integers(-100000, 100000, 3)
    if wide:
        ints[:, 0] *= 300                  # one axis wider than 2^24: the three-separate-fields branch
    return ints


@pytest.mark.parametrize("seed", range(40))
def test_integer_streams_round_trip(seed):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(10, 500))
    ints = _walk(rng, n, wide=seed % 7 == 0)
    data, mn, mx, si = xo.encode_ints(ints, adapt=bool(seed % 2), max_run_atoms=int(rng.integers(1, 9)))
    assert np.array_equal(xo.decode_ints(data, n, mn, mx, si), ints)


def test_water_like_runs_and_radix_adaptation():
    """O H H triples: small differences inside a molecule, large jumps between molecules; the stream must contain runs
    (flag bit 1, run > 0), unchanged-run flags (0) and both directions of the small-radix index."""
    rng = np.random.default_rng(3)
    o = rng.integers(0, 30000, (200, 3))
    ints = np.stack([o, o + rng.integers(-90, 90, (200, 3)), o + rng.integers(-90, 90, (200, 3))], 1).reshape(-1, 3)
    data, mn, mx, si = xo.encode_ints(ints)
    assert np.array_equal(xo.decode_ints(data, len(ints), mn, mx, si), ints)
    assert len(data) < 0.6 * ints.size * 4


def test_sizeofints_is_the_bit_length_of_the_product():
    assert xo.sizeofints([1, 1, 1]) == 1 and xo.sizeofints([256, 1, 1]) == 9 and xo.sizeofints([30000, 30000, 30000]) == 45
    assert xo.sizeofint(0) == 0 and xo.sizeofint(1) == 1 and xo.sizeofint(255) == 8 and xo.sizeofint(256) == 9


def test_file_round_trip_and_frame_index(tmp_path):
    from chacra_b200.trajectory import read_dcd
    from chacra_b200.xtc import XTCFile
    frames = read_dcd(os.path.join(GOLDEN, "5xnd_trajectory.dcd"))[:4]
    path = str(tmp_path / "t.xtc")
    xo.write_xtc(path, frames)
    back = xo.read_xtc(path)
    want = xo.ints_to_angstrom(xo.quantize_nm(frames / np.float32(10.0)))
    assert np.array_equal(back, want)
    assert np.abs(back - frames).max() < 0.0051 + 1e-4          # 0.001 nm grid
    x = XTCFile(path)                                            # header walk only: no GPU needed
    assert (x.n_frames, x.n_atoms) == (4, frames.shape[1])
    assert x.compressed_bytes() < 0.45 * frames.nbytes
    assert all(int(o) % 4 == 0 for o in x._index["offset"])
    # a truncated last frame is dropped
    raw = open(path, "rb").read()
    open(path, "wb").write(raw[:-100])
    assert XTCFile(path).n_frames == 3
    open(path, "wb").write(struct.pack(">ii", 1994, 5) + raw[8:])
    with pytest.raises(ValueError):
        XTCFile(path)


def test_small_systems_are_stored_raw(tmp_path):
    xyz = np.arange(27, dtype=np.float32).reshape(1, 9, 3)
    path = str(tmp_path / "s.xtc")
    xo.write_xtc(path, xyz)
    assert os.path.getsize(path) == 52 + 4 + 108
    assert np.array_equal(xo.read_xtc(path), (xyz / np.float32(10.0)) * np.float32(10.0))
