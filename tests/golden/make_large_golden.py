"""Golden vectors for the large-system path (BASELINE.json config 4 shape: 61 tiled copies of 5xnd, 99 369 atoms).

The numpy oracle needs minutes per frame at this size, so its output is committed: one synthetic frame
(float32 coordinates) plus the oracle's contact list for it (all interaction types) as integer records.
Run from the repository root:  python tests/golden/make_large_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

ITYPES = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]


def main(copies=61, seed=4000):
    import getcontacts_oracle as o
    from chacra_b200.engine import ITYPE_NAMES
    from chacra_b200.synthetic import FrameGenerator, tile_pdb
    from chacra_b200 import read_pdb
    pdb = os.path.join("/tmp", "c4_%d.pdb" % copies)
    tile_pdb(os.path.join(HERE, "5xnd_topology.pdb"), pdb, copies)
    atoms = read_pdb(pdb)
    frames = FrameGenerator(atoms.xyz, seed, device="cpu").frames(1).numpy()
    res = o.run(pdb, pdb, ITYPES, frames=frames)
    rec = np.array([(c[0], ITYPE_NAMES.index(c[1]), min(c[2], c[3]), max(c[2], c[3])) for c in res["contacts"]], np.int32)
    rec = rec[np.lexsort((rec[:, 3], rec[:, 2], rec[:, 1], rec[:, 0]))]
    counts = o.intended_counts(res)
    np.savez_compressed(os.path.join(HERE, "c4_tiled61_frame.npz"), copies=copies, frames=frames, contacts=rec,
                        n_pairs=len(counts), pair_frames=sum(counts.values()))
    print("atoms %d contacts %d residue pairs %d" % (atoms.n, len(rec), len(counts)))


if __name__ == "__main__":
    main()
