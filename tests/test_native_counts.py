"""The threaded contact-list reader (chc_count_tsv, csrc/tsv_counts.h) against the line-by-line readers of
chacra_b200/frequencies.py -- which tests/test_frequencies.py pins to the reference's own script."""
import gzip
import os

import numpy as np
import pytest

from conftest import GOLDEN
from chacra_b200 import frequencies as fq


def _both(path, itypes, shipped):
    lines = open(path).read().splitlines(True)
    want = (fq.shipped_counts if shipped else fq.intended_counts_tsv)(lines, itypes)
    got = fq.native_counts(path, itypes, shipped=shipped, threads=7)
    return got, want


@pytest.mark.parametrize("shipped", [True, False])
@pytest.mark.parametrize("itypes", [fq.ALL_FREQ_ITYPES, ["vdw"], ["hbss", "sb"], None])
def test_golden_contact_list(tmp_path, shipped, itypes):
    path = str(tmp_path / "c.tsv")
    open(path, "w").write(gzip.open(os.path.join(GOLDEN, "test_5xnd_expected.tsv.gz"), "rt").read())
    got, want = _both(path, itypes, shipped)
    assert got == want and got[0] == 20 and len(got[1]) > 10


def _synthetic(path, frames, rng, start=0, step=1, comments=True, crlf=False):
    res = ["%s:%s:%d" % (c, n, k) for c in "AB" for n in ("ALA", "ARG", "HOH") for k in range(1, 14)]
    names = ["vdw", "sb", "hbss", "hp", "wb", "wb2", "pc"]
    nl = "\r\n" if crlf else "\n"
    with open(path, "w", newline="") as fh:
        fh.write("# total_frames:%d beg:0 end:%d stride:1 interaction_types:all%s" % (frames, frames - 1, nl))
        fh.write("# Columns: frame, interaction_type, atom_1, atom_2[, atom_3[, atom_4]], distance_1-2" + nl)
        for f in range(frames):
            for _ in range(int(rng.integers(0, 60))):
                it = names[int(rng.integers(0, len(names)))]
                a, b = rng.choice(len(res), 2)
                cols = [str(start + f * step), it, res[a] + ":CA", res[b] + ":CB"]
                if it in ("wb", "wb2"):
                    cols.append(res[int(rng.integers(0, len(res)))] + ":O")
                if it == "wb2":
                    cols.append(res[int(rng.integers(0, len(res)))] + ":O")
                cols.append("%.3f" % (rng.random() * 4))
                fh.write("\t".join(cols) + nl)
            if comments and f % 37 == 5:
                fh.write("# a comment in the middle" + nl + nl)


@pytest.mark.parametrize("shipped", [True, False])
@pytest.mark.parametrize("kw", [dict(), dict(start=5, step=3), dict(crlf=True), dict(comments=False, start=0, step=1)])
def test_random_files_any_thread_count(tmp_path, shipped, kw):
    rng = np.random.default_rng(len(kw) * 7 + shipped)
    path = str(tmp_path / "r.tsv")
    _synthetic(path, 2500, rng, **kw)                       # about 2.5 MB: several pieces
    lines = open(path, newline="").read().splitlines(True)
    want = (fq.shipped_counts if shipped else fq.intended_counts_tsv)(lines, fq.ALL_FREQ_ITYPES)
    for threads in (1, 2, 5, 16):
        assert fq.native_counts(path, fq.ALL_FREQ_ITYPES, shipped=shipped, threads=threads) == want


def test_edge_files(tmp_path):
    p = str(tmp_path / "e.tsv")
    open(p, "w").write("")
    assert fq.native_counts(p) == (0, {})
    open(p, "w").write("# total_frames:3 beg:0\n")
    assert fq.native_counts(p) == (3, {})
    # one frame only: the shipped reader never flushes it
    open(p, "w").write("# total_frames:1 x\n0\tvdw\tA:ALA:1:CA\tA:ARG:2:CB\n0\tvdw\tA:ARG:2:CB\tA:ALA:1:CA\n")
    assert fq.native_counts(p, None, shipped=True) == (1, {})
    assert fq.native_counts(p, None, shipped=False) == (1, {("A:ALA:1", "A:ARG:2"): 1})
    # a file that does not start at frame 0 loses its first line; labels with fewer than three fields are kept whole
    open(p, "w").write("4\tsb\tX:1\tA:B:3:N\n4\tsb\tA:B:3:N\tX:1\n5\tsb\tq\tq\n6\tsb\tq\tq\n")
    lines = open(p).read().splitlines(True)
    for shipped in (True, False):
        assert fq.native_counts(p, None, shipped=shipped) == (fq.shipped_counts if shipped else fq.intended_counts_tsv)(lines, None)
    # frame numbers that go back: not a file in frame order
    open(p, "w").write("0\tsb\ta:b:1\ta:b:2\n1\tsb\ta:b:1\ta:b:2\n0\tsb\ta:b:1\ta:b:2\n")
    assert fq.native_counts(p) is None
    for bad, msg in (("x\tsb\ta\tb\n", "First column"), ("0\tsb\ta\n", "Invalid number of tokens"), ("0\tsb\ta\tb\tc\td\te\tf\n", "Invalid number")):
        open(p, "w").write(bad)
        with pytest.raises(fq.ParseError, match=msg):
            fq.native_counts(p)
        with pytest.raises(fq.ParseError, match=msg):
            fq.shipped_counts(open(p).read().splitlines(True))
    with pytest.raises(OSError):
        fq.native_counts(str(tmp_path / "missing.tsv"))


def test_frequency_script_uses_the_reader_and_matches_the_reference_table(tmp_path):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(GOLDEN), "..", "chacra_b200", "getcontacts"))
    from chacra_b200.getcontacts import get_contact_frequencies
    src = str(tmp_path / "c.tsv")
    open(src, "w").write(gzip.open(os.path.join(GOLDEN, "test_5xnd_expected.tsv.gz"), "rt").read())
    out = str(tmp_path / "f.tsv")
    get_contact_frequencies.main(argv=["--input_files", src, "--output_file", out])
    got = sorted(open(out).read().splitlines()[2:])
    want = sorted(open(os.path.join(GOLDEN, "freq_5xnd_reference.tsv")).read().splitlines()[2:])
    assert got == want and len(got) == 397
