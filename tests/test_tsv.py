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
