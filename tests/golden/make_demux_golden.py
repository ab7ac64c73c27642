"""Golden vectors for the replica -> state index logic, made in the build container by calling the
REFERENCE's own ``sort_replica_trajectories`` (chacra/trajectories/process_hremd.py:56-94).  MDAnalysis is not installed
here, so the module is imported with empty stand-ins for it (the function itself only uses numpy/pandas).
    python tests/golden/make_demux_golden.py          # needs /root/reference; writes tests/golden/demux_golden.npz
"""
import importlib.util
import os
import sys
import types

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))


def reference_function():
    for name in ("MDAnalysis", "MDAnalysis.analysis", "MDAnalysis.analysis.align", "chacra", "chacra.utils"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["MDAnalysis.analysis"].align = sys.modules["MDAnalysis.analysis.align"]
    sys.modules["chacra.utils"].get_resources = lambda *a, **k: None
    spec = importlib.util.spec_from_file_location("ref_process_hremd", "/root/reference/chacra/trajectories/process_hremd.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.sort_replica_trajectories


def main():
    fn = reference_function()
    out = {}
    for case, (n_states, cycles, interval, traj_len, seed) in enumerate([(4, 37, 1, 37, 1), (6, 101, 5, 20, 2), (20, 400, 4, 100, 3), (3, 10, 3, 3, 4)]):
        rng = np.random.default_rng(seed)
        table = np.stack([rng.permutation(n_states) for _ in range(cycles)])
        df = pd.DataFrame({"u_kn": [np.zeros((n_states, n_states))] * cycles, "replica_to_state_idx": list(table)})
        res = fn(df, interval, traj_len)
        out["case%d_table" % case] = table
        out["case%d_args" % case] = np.array([interval, traj_len])
        for r in range(n_states):
            for s in range(n_states):
                out["case%d_r%d_s%d" % (case, r, s)] = np.asarray(res[r][s], np.int64)
    np.savez_compressed(os.path.join(HERE, "demux_golden.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
