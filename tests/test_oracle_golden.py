"""The CPU oracle against the reference's own golden contact lists.

Pins: external/getcontacts/tests/test_5xnd_expected.tsv, test_5y0t_interface_expected.tsv and the
assertions of external/getcontacts/tests/test_get_dynamic_contacts.py:22-45,:65-76 and
test_get_static_contacts.py:20-43.
"""
import os
from collections import Counter

import getcontacts_oracle as o
from conftest import GOLDEN, read_golden_tsv

ALL = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]


def _lines(res):
    top = res["top"]
    return [(str(c[0]), c[1], top.label(c[2]), top.label(c[3])) + tuple(top.label(w) for w in c[4:]) for c in res["contacts"]]


def test_5xnd_all_itypes_match_golden(oracle_5xnd):
    gold = [tuple(x[:4]) for x in read_golden_tsv(os.path.join(GOLDEN, "test_5xnd_expected.tsv.gz"))]
    mine = _lines(oracle_5xnd)
    assert len(gold) == 22587                      # test_get_dynamic_contacts.py:22 (22589 lines incl. 2 comments)
    assert Counter(mine) == Counter(gold)           # same multiset: nothing missing, nothing extra, no duplicates
    assert {int(x[0]) for x in mine} == set(range(20))      # :31
    per = Counter(x[1] for x in mine)
    assert per == {"vdw": 18253, "hbbb": 1355, "hbsb": 292, "hbss": 81, "hp": 2441, "sb": 71, "pc": 51, "ps": 18, "ts": 25}


def test_5xnd_no_duplicates_even_when_swapped(oracle_5xnd):
    seen = set()
    for ln in _lines(oracle_5xnd):                   # test_get_dynamic_contacts.py:34-45
        a = (ln[0], ln[1], ln[2], ln[3])
        b = (ln[0], ln[1], ln[3], ln[2])
        assert a not in seen and b not in seen
        seen.add(a)


def test_5xnd_salt_bridges_asp_lys():
    res = o.run(os.path.join(GOLDEN, "5xnd_topology.pdb"), os.path.join(GOLDEN, "5xnd_trajectory.dcd"), ["sb"],
                sele1="resname ASP LYS", sele2="resname ASP LYS")
    assert len(res["contacts"]) == 20               # 22 lines incl. 2 header lines (:65)
    assert {c[0] for c in res["contacts"]} == {1, 2, 3, 4, 5, 8, 12, 13, 15, 16, 18, 19}     # :74


def test_5y0t_interface_matches_golden(pdb_5y0t):
    geom = dict(o.DEFAULT_GEOM)
    geom["HBOND_CUTOFF_ANGLE"] = 180               # get_static_contacts.py hbond angle default
    res = o.run(pdb_5y0t, pdb_5y0t, ALL, geom=geom, sele1="chain A", sele2="chain B")
    gold = read_golden_tsv(os.path.join(GOLDEN, "test_5y0t_interface_expected.tsv"))
    assert len(gold) == 269                         # test_get_static_contacts.py:20 (271 lines incl. comments)

    def norm(t):                                    # the golden predates the water swap (compute_contacts.py:346-348)
        return t[:4] + tuple(sorted(t[4:]))
    assert Counter(norm(x) for x in _lines(res)) == Counter(norm(tuple(x)) for x in gold)
    assert {x[1] for x in _lines(res)} == {"hbbb", "hbsb", "hbss", "hp", "lwb", "lwb2", "ps", "sb", "vdw", "wb", "wb2"}   # :29


def test_intended_counts_vs_reference_frequency_table(oracle_5xnd):
    """The as-shipped table (reference script output, committed fixture) is the intended count table
    minus the never-flushed last frame and the first-line-of-frame drops."""
    counts = o.intended_counts(oracle_5xnd)
    assert len(counts) == 398 and sum(counts.values()) == 5831
    rows = read_golden_tsv(os.path.join(GOLDEN, "freq_5xnd_reference.tsv"))
    assert len(rows) == 397 and max(float(r[2]) for r in rows) == 0.95
    for r1, r2, f in rows:
        shipped = round(float(f) * 20)
        assert 0 <= counts[(r1, r2)] - shipped <= 4
