"""The frequency tables this repo writes must load through ChACRA's OWN consumers unchanged (north star: "process-output,
ContactFrequencies PCA and the writers run unchanged").  ``load_contact_file`` / ``make_contact_dataframe``
(chacra/ContactFrequencies.py:38-102) and ``freq_frames`` (chacra/trajectories/process_hremd.py:590-608) are lifted out of
the reference sources with ``ast`` (the modules cannot be imported whole: Bio / MDAnalysis are absent) and executed as they
are, the way tests/golden/make_pca_golden.py does for the PCA functions.  Needs /root/reference (this container only)."""
import ast
import gzip
import os
import re

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

REF = "/root/reference/chacra"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="the reference checkout is not on this machine")


def _lift(path, names):
    import pandas as pd
    tree = ast.parse(open(path).read())
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert len(keep) == len(names)
    ns = {"pd": pd, "os": os, "re": re, "np": np}
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return [ns[n] for n in names]


def test_frequency_files_load_through_the_reference_consumers(tmp_path):
    from chacra_b200 import frequencies as fq
    from chacra_b200.getcontacts import get_contact_frequencies as gcf
    load_contact_file, make_contact_dataframe = _lift(os.path.join(REF, "ContactFrequencies.py"), ["load_contact_file", "make_contact_dataframe"])
    (freq_frames,) = _lift(os.path.join(REF, "trajectories", "process_hremd.py"), ["freq_frames"])
    cont = tmp_path / "cont_state_0.tsv"
    cont.write_bytes(gzip.open(os.path.join(GOLDEN, "test_5xnd_expected.tsv.gz"), "rb").read())
    d = tmp_path / "freqs"
    d.mkdir()
    # state 0: the drop-in frequency script on the reference's golden contact list (as-shipped semantics);
    # state 1: the table written straight from an integer count table (what --frequency_output / the multi-GPU launcher emit)
    gcf.main(["--input_files", str(cont), "--output_file", str(d / "freqs_state_0.tsv")])
    total, counts = fq.intended_counts_tsv(open(cont).read().splitlines(True), fq.ALL_FREQ_ITYPES)
    with open(d / "freqs_state_1.tsv", "w") as fd:
        fq.write_frequency_table(fd, total, fq.ALL_FREQ_ITYPES, {k: (c, c / total) for k, c in counts.items()})
    # 1. one file, against the table the reference script itself produced (tests/golden/make_golden.py)
    s0 = load_contact_file(str(d / "freqs_state_0.tsv"))
    ref = load_contact_file(os.path.join(GOLDEN, "freq_5xnd_reference.tsv"))
    assert len(s0) == 397 and s0.sort_index().equals(ref.sort_index())
    # 2. the folder form process-output uses (file names parsed by int(re.split(r"_|\.", x)[-2]), process_hremd_output.py:182)
    df = make_contact_dataframe(str(d))
    assert list(df.index) == [0, 1] and df.shape[1] == 398 and not df.isna().any().any()
    assert df.loc[1].max() == 1.0 and abs(df.loc[0].max() - 0.95) < 1e-12          # shipped reader never flushes the last frame
    s1 = load_contact_file(str(d / "freqs_state_1.tsv"))
    assert {k: round(v * total) for k, v in s1.items()} == {"%s-%s" % k: c for k, c in counts.items()}
    # 3. the frame count ChACRA re-weights merged runs with (process_hremd_output.py:205-222)
    assert freq_frames(str(d / "freqs_state_0.tsv")) == 20 and freq_frames(str(d / "freqs_state_1.tsv")) == 20
