"""Frequency reduction against the reference script's own output and its unit tests."""
import gzip
import io
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

from chacra_b200 import frequencies as fq

GC = os.path.join(ROOT, "chacra_b200", "getcontacts")


def _golden_lines():
    return gzip.open(os.path.join(GOLDEN, "test_5xnd_expected.tsv.gz"), "rt").read().splitlines(True)


def _read_table(path_or_text):
    text = open(path_or_text).read() if os.path.exists(path_or_text) else path_or_text
    rows = [ln.split("\t") for ln in text.splitlines() if ln and not ln.startswith("#")]
    return {(a, b): f for a, b, f in rows}


def test_shipped_semantics_match_reference_table():
    """freq_5xnd_reference.tsv is the output of the reference's unchanged script (tests/golden/make_golden.py)."""
    total, counts = fq.shipped_counts(_golden_lines(), fq.ALL_FREQ_ITYPES)
    assert total == 20 and len(counts) == 397 and max(counts.values()) == 19      # never 20: last frame is not flushed
    ref = _read_table(os.path.join(GOLDEN, "freq_5xnd_reference.tsv"))
    assert {k: "%.3f" % (v / total) for k, v in counts.items()} == ref


def test_intended_semantics():
    total, counts = fq.intended_counts_tsv(_golden_lines(), fq.ALL_FREQ_ITYPES)
    assert total == 20 and len(counts) == 398 and sum(counts.values()) == 5831
    assert all(a <= b for a, b in counts)
    # hp lines are not a counted type (get_contact_frequencies.py:107-109)
    assert "hp" not in fq.ALL_FREQ_ITYPES
    _, only_hp = fq.intended_counts_tsv(_golden_lines(), ["hp"])
    assert sum(only_hp.values()) > 0


def test_reader_quirks_small_case():
    lines = ["# total_frames:3 beg:0 end:2 stride:1 interaction_types:vdw\n", "# Columns: ...\n",
             "0\tvdw\tA:ALA:1:CB\tA:LEU:9:CD1\n", "0\tvdw\tA:LEU:9:CD2\tA:ALA:1:CB\n",
             "1\tvdw\tA:ALA:1:CB\tA:LEU:9:CD1\n",          # lost: first line of frame 1
             "1\tvdw\tA:GLY:2:CA\tA:LEU:9:CD1\n",
             "2\tvdw\tA:GLY:2:CA\tA:LEU:9:CD1\n"]          # last frame: never flushed
    total, shipped = fq.shipped_counts(lines, ["vdw"])
    assert total == 3 and shipped == {("A:ALA:1", "A:LEU:9"): 1, ("A:GLY:2", "A:LEU:9"): 1}
    _, intended = fq.intended_counts_tsv(lines, ["vdw"])
    assert intended == {("A:ALA:1", "A:LEU:9"): 2, ("A:GLY:2", "A:LEU:9"): 2}


def test_parse_errors_like_reference():
    # contact_calc/unittests/test_transformations.py: non-integer frame and wrong token count raise
    with pytest.raises(fq.ParseError):
        fq.shipped_counts(["x\tvdw\tA:ALA:1:CB\tA:LEU:9:CD1\n"])
    with pytest.raises(fq.ParseError):
        fq.shipped_counts(["0\tvdw\tA:ALA:1:CB\n"])


def test_counts_to_table_inverts_pair_index():
    labels = np.array(sorted(["A:ALA:1", "A:GLY:2", "A:LEU:9", "B:ALA:1"]), dtype=object)
    r = len(labels)
    counts = np.zeros(r * (r + 1) // 2, np.int32)

    def idx(lo, hi):
        return lo * r - lo * (lo - 1) // 2 + (hi - lo)
    counts[idx(0, 2)] = 7
    counts[idx(1, 3)] = 2
    counts[idx(3, 3)] = 1
    assert fq.counts_to_table(labels, counts) == {("A:ALA:1", "A:LEU:9"): 7, ("A:GLY:2", "B:ALA:1"): 2, ("B:ALA:1", "B:ALA:1"): 1}


def test_cli_drop_in_writes_reference_table(tmp_path):
    sys.path.insert(0, GC)
    import get_contact_frequencies as cli
    src = tmp_path / "cont.tsv"
    src.write_text("".join(_golden_lines()))
    out = tmp_path / "freqs.tsv"
    cli.main(["--input_files", str(src), "--output_file", str(out)])
    text = out.read_text().splitlines()
    assert text[0] == "#\ttotal_frames:20\tinteraction_types:" + ",".join(fq.ALL_FREQ_ITYPES)
    assert text[1] == "#\tColumns:\tresidue_1,\tresidue_2\tframe_count\tcontact_frequency"
    assert _read_table(str(out)) == _read_table(os.path.join(GOLDEN, "freq_5xnd_reference.tsv"))
    with pytest.raises(SystemExit) as e:            # usage + exit status 2, like the reference's parser
        cli.main(["--output_file", str(out)])
    assert e.value.code == 2


def test_dynamic_cli_flag_errors():
    sys.path.insert(0, GC)
    import get_dynamic_contacts as cli
    with pytest.raises(SystemExit) as e:
        cli.main(["--topology", "a.pdb"])
    assert e.value.code == 2
    with pytest.raises(SystemExit) as e:
        cli.main(["--topology", "a.pdb", "--trajectory", "a.dcd", "--output", "o.tsv", "--itypes", "nonsense"])
    assert e.value.code == 2


def test_count_sidecar_round_trip_and_exact_merge_of_runs(tmp_path):
    from chacra_b200 import frequencies as fq
    runs = [(700, {("A:ALA:1", "A:ARG:2"): 333, ("A:ALA:1", "A:GLY:9"): 1}),
            (1300, {("A:ALA:1", "A:ARG:2"): 1001, ("A:ARG:2", "A:GLY:9"): 650})]
    files = []
    for k, (frames, table) in enumerate(runs):
        p = str(tmp_path / ("counts_%d.tsv" % k))
        with open(p, "w") as fd:
            fq.write_count_table(fd, frames, fq.ALL_FREQ_ITYPES, table)
        assert fq.read_count_table(p) == (frames, table)
        files.append(p)
    out = str(tmp_path / "freqs_state_0.tsv")
    total, merged = fq.merge_runs_exact(files, out)
    assert total == 2000 and merged == {("A:ALA:1", "A:ARG:2"): 1334, ("A:ALA:1", "A:GLY:9"): 1, ("A:ARG:2", "A:GLY:9"): 650}
    rows = [ln.split("\t") for ln in open(out).read().splitlines()]
    assert rows[0][:2] == ["#", "total_frames:2000"]
    assert rows[2:] == [["A:ALA:1", "A:ARG:2", "0.667"], ["A:ALA:1", "A:GLY:9", "0.001"], ["A:ARG:2", "A:GLY:9", "0.325"]]
    # the reference's way (frame-weighted three-decimal frequencies) is off in the last digit here: 0.476*0.35 + 0.770*0.65 = 0.6671 vs 1334/2000 = 0.667
    weighted = round(333 / 700, 3) * 0.35 + round(1001 / 1300, 3) * 0.65
    assert abs(weighted - 1334 / 2000) > 5e-5
