"""The drop-in scripts end to end on the GPU: same assertions as the reference's own integration tests
(external/getcontacts/tests/test_get_dynamic_contacts.py:11-76) plus equality with its golden contact list."""
import os
import sys
from collections import Counter

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, read_golden_tsv

pytestmark = pytest.mark.gpu
GC = os.path.join(ROOT, "chacra_b200", "getcontacts")
TOP = os.path.join(GOLDEN, "5xnd_topology.pdb")
TRJ = os.path.join(GOLDEN, "5xnd_trajectory.dcd")


def _cli():
    if GC not in sys.path:
        sys.path.insert(0, GC)
    import get_dynamic_contacts
    import get_contact_frequencies
    return get_dynamic_contacts, get_contact_frequencies


def _lines(path):
    return [ln.rstrip("\n") for ln in open(path)]


def test_dynamic_contacts_all_itypes(tmp_path):
    dyn, freq = _cli()
    out = tmp_path / "cont.tsv"
    dyn.main(["--topology", TOP, "--trajectory", TRJ, "--output", str(out), "--itypes", "all"])
    lines = _lines(out)
    assert len(lines) == 22589                                   # test_get_dynamic_contacts.py:22
    assert lines[0] == "# total_frames:20 beg:0 end:19 stride:1 interaction_types:hp,sb,pc,ps,ts,vdw,hb"
    assert lines[1] == "# Columns: frame, interaction_type, atom_1, atom_2[, atom_3[, atom_4]]"
    body = [ln.split("\t") for ln in lines[2:]]
    assert {int(t[0]) for t in body} == set(range(20))           # :31
    seen = set()
    for t in body:                                               # :34-45 no duplicate, even with swapped atoms
        assert (t[0], t[1], t[2], t[3]) not in seen and (t[0], t[1], t[3], t[2]) not in seen
        seen.add((t[0], t[1], t[2], t[3]))
    gold = read_golden_tsv(os.path.join(GOLDEN, "test_5xnd_expected.tsv.gz"))
    assert Counter(tuple(t) for t in body) == Counter(tuple(t) for t in gold)
    assert [int(t[0]) for t in body] == sorted(int(t[0]) for t in body)     # frames ascending
    # the frequency step on OUR contact list, shipped semantics: the pair set and every count that does not depend on
    # which line happens to be first in a frame equal the reference table; the rest differ by the documented bound
    f_out = tmp_path / "freqs.tsv"
    freq.main(["--input_files", str(out), "--output_file", str(f_out)])
    mine = {(a, b): round(float(f) * 20) for a, b, f in (ln.split("\t") for ln in _lines(f_out)[2:])}
    ref = {(a, b): round(float(f) * 20) for a, b, f in read_golden_tsv(os.path.join(GOLDEN, "freq_5xnd_reference.tsv"))}
    assert set(mine) == set(ref)
    diff = {k: mine[k] - ref[k] for k in ref if mine[k] != ref[k]}
    assert sum(abs(v) for v in diff.values()) <= 2 * 19         # at most one first-line drop per frame on either side
    # intended semantics = the device count table
    f_int = tmp_path / "freqs_intended.tsv"
    freq.main(["--input_files", str(out), "--output_file", str(f_int), "--semantics", "intended"])
    inten = {(a, b): round(float(f) * 20) for a, b, f in (ln.split("\t") for ln in _lines(f_int)[2:])}
    assert len(inten) == 398 and sum(inten.values()) == 5831


def test_dynamic_contacts_distout_and_device_frequencies(tmp_path):
    dyn, _ = _cli()
    out, fout, cout = tmp_path / "cont.tsv", tmp_path / "freqs_dev.tsv", tmp_path / "counts_dev.tsv"
    dyn.main(["--topology", TOP, "--trajectory", TRJ, "--output", str(out), "--itypes", "all", "--distout",
              "--sele", "protein", "--sele2", "protein", "--cores", "4", "--frequency_output", str(fout),
              "--count_output", str(cout)])                                                                 # ChACRA's call + extensions
    lines = _lines(out)
    assert lines[1].endswith(", distance_1-2") and len(lines) == 22589
    import getcontacts_oracle as o
    xyz = o.read_dcd(TRJ)
    top = o.Topology(TOP)
    lab = {top.label(i): i for i in range(top.n)}
    for t in [ln.split("\t") for ln in lines[2:400]]:
        a, b, f = xyz[int(t[0])][lab[t[2]]], xyz[int(t[0])][lab[t[3]]], t[-1]
        assert f == "%.3f" % float(np.sqrt(o._d2(a, b), dtype=np.float32))
    rows = [ln.split("\t") for ln in _lines(fout)[2:]]
    assert len(rows) == 398 and sum(round(float(r[2]) * 20) for r in rows) == 5831
    # the integer sidecar: the counts behind that table, and the ones the threaded reader finds in the emitted text
    crow = [ln.split("\t") for ln in _lines(cout)[2:]]
    assert _lines(cout)[0].startswith("#\ttotal_frames:20\t") and [r[:2] for r in crow] == [r[:2] for r in rows]
    assert all("%.3f" % (int(c[2]) / 20) == r[2] for c, r in zip(crow, rows)) and sum(int(c[2]) for c in crow) == 5831
    from chacra_b200 import frequencies as fq
    total, from_text = fq.native_counts(str(out), fq.ALL_FREQ_ITYPES, shipped=False)
    assert total == 20 and from_text == {(c[0], c[1]): int(c[2]) for c in crow}


def test_dynamic_contacts_salt_bridges_asp_lys(tmp_path):
    dyn, _ = _cli()
    out = tmp_path / "sb.tsv"
    dyn.main(["--topology", TOP, "--trajectory", TRJ, "--output", str(out), "--itypes", "sb", "--sele", "resname ASP LYS"])
    lines = _lines(out)
    assert len(lines) == 22                                      # test_get_dynamic_contacts.py:65
    assert {int(ln.split("\t")[0]) for ln in lines[2:]} == {1, 2, 3, 4, 5, 8, 12, 13, 15, 16, 18, 19}    # :74


def test_beg_end_stride_and_empty_selection(tmp_path):
    dyn, _ = _cli()
    out = tmp_path / "sub.tsv"
    dyn.main(["--topology", TOP, "--trajectory", TRJ, "--output", str(out), "--itypes", "vdw", "--beg", "3", "--end", "15",
              "--stride", "4"])
    lines = _lines(out)
    assert lines[0].startswith("# total_frames:4 beg:3 end:15 stride:4 ")
    # the reference never clamps a fragment to --end (compute_contacts.py:253-254: end_frame = beg_frame + 100*stride - 1), so
    # the last fragment also writes frame 19 although the header says total_frames:4
    assert {int(ln.split("\t")[0]) for ln in lines[2:]} == {3, 7, 11, 15, 19}
    gold = [t for t in read_golden_tsv(os.path.join(GOLDEN, "test_5xnd_expected.tsv.gz")) if t[1] == "vdw" and int(t[0]) in (3, 7, 11, 15, 19)]
    assert Counter(tuple(ln.split("\t")) for ln in lines[2:]) == Counter(tuple(t) for t in gold)
    with pytest.raises(SystemExit) as e:                         # compute_contacts.py:240-245
        dyn.main(["--topology", TOP, "--trajectory", TRJ, "--output", str(out), "--itypes", "sb", "--sele", "resname XYZ"])
    assert e.value.code == 1


def test_counts_only_run_and_near_cutoff_file(tmp_path):
    """Extensions of the drop-in script: --no_contact_output (nothing but the count table leaves the GPU; --output may be
    omitted) gives the same frequency / count files as the run that also writes the contact list, and
    --near_cutoff_output lists the oracle's near-cutoff observations with atom labels."""
    import getcontacts_oracle as o
    dyn, _ = _cli()
    a = {k: str(tmp_path / k) for k in ("list.tsv", "f1.tsv", "c1.tsv", "n1.tsv", "f2.tsv", "c2.tsv", "n2.tsv")}
    dyn.main(["--topology", TOP, "--trajectory", TRJ, "--output", a["list.tsv"], "--itypes", "all", "--distout",
              "--frequency_output", a["f1.tsv"], "--count_output", a["c1.tsv"], "--near_cutoff_output", a["n1.tsv"]])
    dyn.main(["--topology", TOP, "--trajectory", TRJ, "--itypes", "all", "--no_contact_output",
              "--frequency_output", a["f2.tsv"], "--count_output", a["c2.tsv"], "--near_cutoff_output", a["n2.tsv"]])
    for x, y in (("f1.tsv", "f2.tsv"), ("c1.tsv", "c2.tsv"), ("n1.tsv", "n2.tsv")):
        assert open(a[x]).read() == open(a[y]).read()
    res = o.run(TOP, TRJ, ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"], collect_near=True)
    top = res["top"]
    want = sorted({(f, kind) + tuple(sorted((top.label(i), top.label(j)))) for f, kind, i, j, _ in res["near"]})
    rows = [ln.rstrip("\n").split("\t") for ln in open(a["n1.tsv"]) if not ln.startswith("#")]
    assert sorted({(int(r[0]), r[1]) + tuple(sorted((r[2], r[3]))) for r in rows}) == want and len(want) > 0
    assert all(abs(float(r[4])) < 1e-3 for r in rows)
    with pytest.raises(SystemExit) as e:                         # --output stays required when the list is written
        dyn.main(["--topology", TOP, "--trajectory", TRJ, "--itypes", "all"])
    assert e.value.code == 2


def test_all_states_launcher_single_rank(tmp_path):
    """get_state_frequencies.py (the multi-GPU launcher) run as one process: per-state tables equal to what
    get_dynamic_contacts --frequency_output writes for each state trajectory on its own; with --beg/--end/--stride too."""
    import numpy as np
    from test_gpu_dcd import write_dcd
    from chacra_b200.trajectory import DCDFile
    from chacra_b200.getcontacts import get_state_frequencies
    dyn, _ = _cli()
    x = DCDFile(TRJ).read()
    states = []
    for s in range(3):                                   # three "states": the fixture frames in different orders, 130 / 20 / 250 frames
        n = (130, 20, 250)[s]
        p = str(tmp_path / ("state_%d.dcd" % s))
        write_dcd(p, np.ascontiguousarray(x[(np.arange(n) * (2 * s + 1) + s) % 20]), "<", False)
        states.append(p)
    for extra in ([], ["--beg", "5", "--end", "117", "--stride", "3"]):
        out = tmp_path / ("all%d" % len(extra))
        get_state_frequencies.main(["--topology", TOP, "--trajectories"] + states[::-1] + ["--output_dir", str(out), "--counts",
                                    "--near_cutoff_output", str(out) + "_near.tsv"] + extra)
        for s in range(3):
            one = tmp_path / ("one%d_%d" % (s, len(extra)))
            dyn.main(["--topology", TOP, "--trajectory", states[s], "--itypes", "all", "--no_contact_output",
                      "--frequency_output", str(one) + "_f.tsv", "--count_output", str(one) + "_c.tsv"] + extra)
            assert open(out / ("freqs_state_%d.tsv" % s)).read() == open(str(one) + "_f.tsv").read()
            assert open(out / ("counts_state_%d.tsv" % s)).read() == open(str(one) + "_c.tsv").read()
        assert sum(1 for ln in open(str(out) + "_near.tsv") if not ln.startswith("#")) > 0
