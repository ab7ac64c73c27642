"""Contact-list TSV writer on hand-made records: no GPU needed."""
import io

import numpy as np

from chacra_b200.engine import CONTACT_DTYPE, ITYPE_NAMES
from chacra_b200.tsv import ContactWriter, header_lines

LABELS = ["A:ARG:76:NH2", "A:GLU:82:OE1", "A:ALA:5:CB", "A:LEU:9:CD1", "B:HOH:7:O", "A:HOH:3:O", "A:ALA:5:CB"]


def _rec(rows):
    r = np.zeros(len(rows), CONTACT_DTYPE)
    for k, (f, it, a, b, w1, w2, d) in enumerate(rows):
        r[k] = (f, ITYPE_NAMES.index(it), a, b, w1, w2, d, 0)
    return r


def test_header():
    assert header_lines(20, 0, 19, 1, ["hp", "sb"], True) == (
        "# total_frames:20 beg:0 end:19 stride:1 interaction_types:hp,sb\n"
        "# Columns: frame, interaction_type, atom_1, atom_2[, atom_3[, atom_4]], distance_1-2\n")
    assert header_lines(1, 0, 0, 1, ["sb"], False).endswith("atom_2[, atom_3[, atom_4]]\n")


def test_canonical_order_blocks_and_frames():
    fd = io.StringIO()
    w = ContactWriter(fd, LABELS, beg=10, stride=5, distout=True)
    w.write(_rec([(1, "vdw", 3, 2, -1, -1, 3.14159), (0, "hp", 2, 3, -1, -1, 3.5), (0, "sb", 1, 0, -1, -1, 2.8714),
                  (0, "wb2", 1, 0, 4, 5, 5.0), (0, "hbss", 0, 1, -1, -1, 2.9)]))
    lines = fd.getvalue().splitlines()
    assert lines == [
        "10\tsb\tA:ARG:76:NH2\tA:GLU:82:OE1\t2.871",          # atoms swapped into string order
        "10\thbss\tA:ARG:76:NH2\tA:GLU:82:OE1\t2.900",
        "10\twb2\tA:ARG:76:NH2\tA:GLU:82:OE1\tA:HOH:3:O\tB:HOH:7:O\t5.000",   # waters swapped along with the atoms
        "10\thp\tA:ALA:5:CB\tA:LEU:9:CD1\t3.500",
        "15\tvdw\tA:ALA:5:CB\tA:LEU:9:CD1\t3.142"]             # frame = beg + index * stride


def test_duplicate_lines_dropped_per_fragment():
    fd = io.StringIO()
    w = ContactWriter(fd, LABELS, beg=0, stride=1, distout=False)
    # atoms 2 and 6 carry the same label: without --distout their lines are identical and written once
    n = w.write(_rec([(0, "vdw", 2, 3, -1, -1, 3.0), (0, "vdw", 6, 3, -1, -1, 3.2), (100, "vdw", 2, 3, -1, -1, 3.0)]))
    assert n == 2 and fd.getvalue().splitlines() == ["0\tvdw\tA:ALA:5:CB\tA:LEU:9:CD1", "100\tvdw\tA:ALA:5:CB\tA:LEU:9:CD1"]
    fd = io.StringIO()
    w = ContactWriter(fd, LABELS, beg=0, stride=1, distout=True)
    assert w.write(_rec([(0, "vdw", 2, 3, -1, -1, 3.0), (0, "vdw", 6, 3, -1, -1, 3.2)])) == 2


def _random_records(rng, n_atoms, frames, per_frame, dup_share=0.05):
    n = frames * per_frame
    r = np.zeros(n, CONTACT_DTYPE)
    r["frame"] = rng.permutation(np.repeat(np.arange(frames) * 3 + 7, per_frame))      # any order, gaps between frames
    r["itype"] = rng.integers(0, len(ITYPE_NAMES), n)
    r["atom1"] = rng.integers(0, n_atoms, n)
    r["atom2"] = rng.integers(0, n_atoms, n)
    wb = np.isin(r["itype"], [ITYPE_NAMES.index(x) for x in ("wb", "lwb", "wb2", "lwb2")])
    wb2 = np.isin(r["itype"], [ITYPE_NAMES.index(x) for x in ("wb2", "lwb2")])
    r["water1"] = np.where(wb, rng.integers(0, n_atoms, n), -1)
    r["water2"] = np.where(wb2, rng.integers(0, n_atoms, n), -1)
    r["distance"] = (rng.integers(0, 1 << 16, n) / 4096.0).astype(np.float32)              # many ties at the third decimal
    k = int(n * dup_share)
    r[rng.integers(0, n, k)] = r[rng.integers(0, n, k)]                                  # exact duplicate records
    return r


def test_native_writer_equals_the_python_restatement(tmp_path):
    import tsv_oracle
    rng = np.random.default_rng(3)
    for labels, distout, threads in [([("A:RES:%d:X%d" % (k // 3, k % 3)) for k in range(300)], True, 5),
                                     ([("A:RES:%d:X%d" % (k // 3, k % 2)) for k in range(300)], False, 3),     # shared labels
                                     ([("A:RES:%d:X%d" % (k // 3, k % 2)) for k in range(300)], True, 1),
                                     (LABELS, True, 64)]:
        rec = _random_records(rng, len(labels), 23, 400)
        want = io.StringIO()
        n_want = tsv_oracle.ContactWriterPy(want, labels, 100, 10, distout).write(rec.copy())
        path = tmp_path / "out.tsv"
        with open(path, "w") as fd:
            fd.write("# header\n")
            w = ContactWriter(fd, labels, 100, 10, distout, threads=threads)
            n_got = w.write(rec)
            fd.write("# trailer\n")                                                    # Python-level writes stay in order
        text = open(path).read()
        assert text == "# header\n" + want.getvalue() + "# trailer\n"
        assert n_got == n_want == len(want.getvalue().splitlines()) and n_got < len(rec)


def test_distance_text_is_printf_exact():
    import ctypes as C
    from chacra_b200 import _lib
    lib = _lib.load()
    buf = C.create_string_buffer(64)
    rng = np.random.default_rng(9)
    vals = np.concatenate([rng.random(20000).astype(np.float32) * 12, (rng.integers(0, 1 << 18, 20000) / 16.0).astype(np.float32),
                           np.array([0, 0.0005, 0.0625, 0.1875, 0.3125, 1e-8, 1e-30, 3.9995, 123456.789, 1e9, 8388608.5, 1e-45], np.float32)])
    for v in vals.tolist():
        lib.chc_format_fixed3(C.c_float(v), buf)
        assert buf.value.decode() == "%.3f" % v


def test_bad_records_are_refused():
    import pytest
    fd = io.StringIO()
    w = ContactWriter(fd, LABELS, 0, 1, True)
    bad = _rec([(0, "vdw", 2, 99, -1, -1, 3.0)])
    with pytest.raises(RuntimeError):
        w.write(bad)
    assert w.write(_rec([])) == 0 and fd.getvalue() == ""
