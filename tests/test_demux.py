"""Replica -> state index logic against vectors produced by the reference's own
sort_replica_trajectories (tests/golden/make_demux_golden.py)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from chacra_b200.demux import sort_replica_trajectories, state_sources

G = np.load(os.path.join(GOLDEN, "demux_golden.npz"))
CASES = sorted({k.split("_")[0] for k in G.files})


@pytest.mark.parametrize("case", CASES)
def test_sort_replica_trajectories_matches_reference(case):
    table = G[case + "_table"]
    interval, traj_len = (int(v) for v in G[case + "_args"])
    got = sort_replica_trajectories(list(table), interval, traj_len)
    n = table.shape[1]
    assert sorted(got) == list(range(n))
    for r in range(n):
        assert len(got[r]) == n
        for s in range(n):
            assert np.array_equal(got[r][s], G["%s_r%d_s%d" % (case, r, s)])


@pytest.mark.parametrize("case", CASES)
def test_state_sources_is_the_inverse_view(case):
    table = G[case + "_table"]
    interval, traj_len = (int(v) for v in G[case + "_args"])
    src = state_sources(list(table), interval, traj_len)
    n = table.shape[1]
    saved = table[::interval][:traj_len]
    assert src.shape == (n, len(saved)) and src.dtype == np.int64
    for s in range(n):
        for r in range(n):
            frames = G["%s_r%d_s%d" % (case, r, s)]
            assert (src[s, frames] == r).all()                 # the frames the reference takes from replica r for state s
        assert (saved[np.arange(len(saved)), src[s]] == s).all()


def test_dataframe_like_input_and_bad_rows():
    import pandas as pd
    table = np.array([[0, 1, 2], [2, 0, 1], [1, 2, 0], [0, 2, 1]])
    df = pd.DataFrame({"replica_to_state_idx": list(table)})
    assert np.array_equal(state_sources(df), np.argsort(table, axis=1).T)
    assert state_sources(df, 2).shape == (3, 2) and state_sources(df, 1, 3).shape == (3, 3)
    with pytest.raises(ValueError, match="row 1"):
        state_sources([[0, 1, 2], [0, 0, 1]])
    with pytest.raises(ValueError):
        state_sources([[0, 1, 3]])
    assert state_sources([]).shape[1] == 0
