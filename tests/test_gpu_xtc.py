"""XTC frames decoded by the CUDA kernel behind chc_xtc_decode vs the CPU codec restatement (oracle/xtc_oracle.py), and
the drop-in script on an XTC trajectory (what ChACRA's process-output passes: state_{i}.xtc)."""
import os
import struct
import sys
from collections import Counter

import numpy as np
import pytest

import xtc_oracle as xo
from conftest import GOLDEN, ROOT

pytestmark = pytest.mark.gpu
GC = os.path.join(ROOT, "chacra_b200", "getcontacts")
TOP = os.path.join(GOLDEN, "5xnd_topology.pdb")
TRJ = os.path.join(GOLDEN, "5xnd_trajectory.dcd")


def _write_int_frames(path, frames_ints, precision=1000.0, **kw):
    with open(path, "wb") as fh:
        for k, ints in enumerate(frames_ints):
            n = len(ints)
            data, mn, mx, si = xo.encode_ints(ints, **kw)
            fh.write(struct.pack(">iiif", 1995, n, k, float(k)) + b"\0" * 36 + struct.pack(">i", n))
            fh.write(struct.pack(">f3i3iii", precision, *mn, *mx, si, len(data)) + data + b"\0" * ((-len(data)) % 4))


@pytest.mark.parametrize("wide", [False, True])
def test_decoder_matches_oracle_on_every_stream_feature(tmp_path, wide, monkeypatch):
    """Random walks with jumps (runs of every length, both radix directions, flag-0 repeats); `wide`: one axis spans
    more than 2^24 grid steps, so the large integers are three separate bit fields instead of one packed number
    (also exercises packed numbers of more than 64 bits when narrow axes are still large)."""
    from chacra_b200.xtc import XTCFile
    rng = np.random.default_rng(11 + wide)
    n = 777
    frames = []
    for f in range(37):
        steps = rng.integers(-150, 150, (n, 3))
        jumps = (rng.random((n, 1)) < 0.1) * rng.integers(-5000, 5000, (n, 3))
        ints = np.cumsum(steps + jumps, 0) + rng.integers(-100000, 100000, 3)
        if wide:
            ints[:, 0] *= 300
        elif f % 3 == 0:
            ints = ints * np.array([60, 70, 80])       # ~2^22 per axis: a packed triple of about 67 bits
        elif f % 3 == 1:
            ints = ints // 16                          # compact box: packed triples of at most 52 bits (reciprocal division path)
        frames.append(ints)
    path = str(tmp_path / "r.xtc")
    _write_int_frames(path, frames, adapt=True)
    x = XTCFile(path)
    assert (x.n_frames, x.n_atoms) == (37, n)
    got = x.read()
    want = np.stack([xo.ints_to_angstrom(i) for i in frames])
    assert np.array_equal(got, want)
    assert np.array_equal(got, xo.read_xtc(path))
    # strided and partial reads gather the right frames
    assert np.array_equal(x.read(3, 30, 4), want[3:31:4])
    assert np.array_equal(x.read(36, None, 1), want[36:])
    # the pipelined read (file read of chunk k+1 under the decode of chunk k) is the same function
    x.CHUNK_BYTES = 20000
    assert np.array_equal(x.read(), want)
    assert np.array_equal(x.read(1, 35, 3), want[1:36:3])
    # the first decoder (one thread walks and decodes a whole frame) is kept behind a switch: same bits
    monkeypatch.setenv("CHACRA_XTC_ONEPASS", "1")
    assert np.array_equal(XTCFile(path).read(), want)


def test_small_raw_frames_and_corrupt_streams(tmp_path):
    from chacra_b200.xtc import XTCFile
    xyz = np.linspace(-3, 7, 2 * 7 * 3, dtype=np.float32).reshape(2, 7, 3)
    p = str(tmp_path / "s.xtc")
    xo.write_xtc(p, xyz)
    assert np.array_equal(XTCFile(p).read(), xo.read_xtc(p))
    # a bit stream that announces more atoms in a run than the frame has left
    rng = np.random.default_rng(0)
    ints = np.cumsum(rng.integers(-50, 50, (64, 3)), 0)
    q = str(tmp_path / "c.xtc")
    _write_int_frames(q, [ints])
    raw = bytearray(open(q, "rb").read())
    raw[4:8] = struct.pack(">i", 40)                 # both atom counts say 40: the last run overshoots
    raw[52:56] = struct.pack(">i", 40)
    open(q, "wb").write(bytes(raw))
    with pytest.raises(ValueError):
        XTCFile(q).read()


def test_dynamic_contacts_from_xtc(tmp_path):
    """The script on state.xtc = the script on a DCD holding the coordinates the XTC grid (0.001 nm) stores."""
    from chacra_b200.trajectory import read_dcd
    if GC not in sys.path:
        sys.path.insert(0, GC)
    import get_dynamic_contacts as dyn
    frames = read_dcd(TRJ)
    xtc = str(tmp_path / "state_0.xtc")
    xo.write_xtc(xtc, frames)
    out_x = tmp_path / "cont_xtc.tsv"
    dyn.main(["--topology", TOP, "--trajectory", xtc, "--output", str(out_x), "--itypes", "all", "--distout"])
    # the same quantised coordinates through the engine's host path, contact sets compared by atom index
    import torch
    import getcontacts_oracle as o
    grid = xo.ints_to_angstrom(xo.quantize_nm(frames / np.float32(10.0)))
    res = o.run(TOP, TOP, ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"], frames=grid)
    want = Counter((str(c[0]), c[1]) for c in res["contacts"])
    body = [ln.rstrip("\n").split("\t") for ln in open(out_x)][2:]
    got = Counter((t[0], t[1]) for t in body)
    assert got == want and len(body) == len(res["contacts"])
    assert {int(t[0]) for t in body} == set(range(20))


def test_random_corruption_never_faults(tmp_path):
    """Random byte flips inside the compressed streams (coding parameters and lengths untouched): the decoder either
    returns coordinates or reports an inconsistent stream -- no fault, no hang, and a later clean read still works."""
    from chacra_b200.xtc import XTCFile
    rng = np.random.default_rng(17)
    n = 300
    frames = [np.cumsum(rng.integers(-120, 120, (n, 3)), 0) + rng.integers(-3000, 3000, 3) for _ in range(6)]
    good = str(tmp_path / "good.xtc")
    _write_int_frames(good, frames, adapt=True)
    want = XTCFile(good).read()
    blob = bytearray(open(good, "rb").read())
    idx = XTCFile(good)._index
    outcomes = Counter()
    for trial in range(60):
        bad = bytearray(blob)
        for _ in range(int(rng.integers(1, 6))):
            f = int(rng.integers(0, len(idx)))
            at = int(idx["offset"][f]) + int(rng.integers(0, int(idx["nbytes"][f])))
            bad[at] ^= int(rng.integers(1, 256))
        path = str(tmp_path / "bad.xtc")
        open(path, "wb").write(bad)
        try:
            got = XTCFile(path).read()
            assert got.shape == want.shape
            outcomes["decoded"] += 1
        except ValueError:
            outcomes["rejected"] += 1
    assert outcomes["decoded"] + outcomes["rejected"] == 60 and outcomes["rejected"] > 0
    assert np.array_equal(XTCFile(good).read(), want)


def test_workspace_sizes_and_the_allocating_entry_point(tmp_path):
    """chc_xtc_decode_ws with a workspace for all frames, with one that holds three frames (several batches) and with one that is too
    small, and chc_xtc_decode (which allocates its own): same coordinates, bit for bit."""
    import ctypes as C
    import torch
    from chacra_b200 import _lib
    from chacra_b200.xtc import XTCFile, _SLACK
    rng = np.random.default_rng(5)
    n = 411
    frames = [np.cumsum(rng.integers(-140, 140, (n, 3)), 0) + rng.integers(-2000, 2000, 3) for _ in range(11)]
    path = str(tmp_path / "w.xtc")
    _write_int_frames(path, frames, adapt=True)
    x = XTCFile(path)
    want = np.stack([xo.ints_to_angstrom(i) for i in frames])
    assert np.array_equal(x.read(), want)
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    meta, used = x._stage(np.arange(x.n_frames), 0)
    d_blob = x._stage_bufs[0][:used + _SLACK].to(dev)
    d_meta = torch.from_numpy(meta.view(np.uint8).reshape(-1)).to(dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    full = int(lib.chc_xtc_work_bytes(x.n_frames, n))
    three = 64 + 3 * (8 * n + 8) + 16
    assert full >= 64 + x.n_frames * (8 * n + 8)
    for bytes_ in (full, three):
        out = torch.zeros((x.n_frames, n, 3), dtype=torch.float32, device=dev)
        work = torch.empty(bytes_, dtype=torch.uint8, device=dev)
        assert lib.chc_xtc_decode_ws(d_blob.data_ptr(), d_meta.data_ptr(), x.n_frames, n, out.data_ptr(), 0, work.data_ptr(), bytes_, 0, stream) == 0
        assert np.array_equal(out.cpu().numpy(), want)
    small = torch.empty(64 + 8 * n, dtype=torch.uint8, device=dev)           # not even one frame's share
    out = torch.zeros((x.n_frames, n, 3), dtype=torch.float32, device=dev)
    assert lib.chc_xtc_decode_ws(d_blob.data_ptr(), d_meta.data_ptr(), x.n_frames, n, out.data_ptr(), 0, small.data_ptr(), small.numel(), 0, stream) == -1
    assert lib.chc_xtc_decode(d_blob.data_ptr(), d_meta.data_ptr(), x.n_frames, n, out.data_ptr(), 0, 0, stream) == 0
    assert np.array_equal(out.cpu().numpy(), want)
