"""Oracle goldens at the full sizes of BASELINE.json configs 3 and 5 (the numpy oracle is too slow to run them inside the
GPU test session).  Run from the repository root:  python tests/golden/make_config_goldens.py [c3] [c5]

  c3_tiled10_101frames.npz  10 tiled copies of 5xnd (16 290 atoms, in-place pair path), 101 frames = one full 100-frame
                            fragment plus the first frame of the next one: per-frame / per-itype contact numbers, an
                            order-free 64-bit digest of every frame's contact records, the whole residue-pair count table.
  c5_solv200k_frame.npz     4 tiled copies in TIP3 water, 200 k atoms (second cell list for unselected solvent, water
                            bridges): one frame, every contact record incl. the mediating waters, the count table.
Frames are NOT stored: they are regenerated from the seed by chacra_b200.synthetic.FrameGenerator on the CPU (bit-exact
for a given torch build); a digest of the float32 bits guards that.
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

ITYPES = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]
MIX = np.uint64(0x9E3779B97F4A7C15)


def record_digest(frame_ids, itype, a, b, w1, w2, n_frames):
    """Order-free digest per frame: sum (mod 2^64) of a 64-bit mix of (itype, lo atom, hi atom, waters)."""
    with np.errstate(over="ignore"):
        lo, hi = np.minimum(a, b).astype(np.uint64), np.maximum(a, b).astype(np.uint64)
        swap = b < a
        x1 = np.where(swap, w2, w1).astype(np.int64).astype(np.uint64)       # the waters travel with their atoms
        x2 = np.where(swap, w1, w2).astype(np.int64).astype(np.uint64)
        h = (itype.astype(np.uint64) + np.uint64(1)) * MIX
        for v in (lo, hi, x1 + np.uint64(2), x2 + np.uint64(2)):
            h = (h ^ (v + np.uint64(0x7F4A7C15))) * MIX
            h ^= h >> np.uint64(29)
        out = np.zeros(n_frames, np.uint64)
        np.add.at(out, frame_ids, h)
    return out


def frames_digest(frames):
    return int(np.bitwise_xor.reduce(frames.view(np.uint32).ravel().astype(np.uint64) * MIX >> np.uint64(7)))


def oracle_arrays(res, n_frames):
    from chacra_b200.engine import ITYPE_NAMES
    c = res["contacts"]
    f = np.array([r[0] for r in c], np.int64)
    it = np.array([ITYPE_NAMES.index(r[1]) for r in c], np.int64)
    a = np.array([r[2] for r in c], np.int64)
    b = np.array([r[3] for r in c], np.int64)
    w1 = np.array([r[4] if len(r) > 4 else -1 for r in c], np.int64)
    w2 = np.array([r[5] if len(r) > 5 else -1 for r in c], np.int64)
    per = np.zeros((n_frames, len(ITYPE_NAMES)), np.int32)
    np.add.at(per, (f, it), 1)
    return f, it, a, b, w1, w2, per


def count_table(res, topo_labels):
    import getcontacts_oracle as o
    want = o.intended_counts(res)
    rank = {lab: k for k, lab in enumerate(topo_labels)}
    r = len(topo_labels)
    idx, val = [], []
    for (r1, r2), c in want.items():
        lo, hi = sorted((rank[r1], rank[r2]))
        idx.append(lo * r - (lo * (lo - 1)) // 2 + (hi - lo))
        val.append(c)
    order = np.argsort(idx)
    return np.array(idx, np.int64)[order], np.array(val, np.int32)[order]


def make_c3(copies=10, seed=3000, n_frames=101, amplitude=0.8):
    from parallel import run_parallel
    from chacra_b200 import read_pdb, compile_topology
    from chacra_b200.synthetic import FrameGenerator, tile_pdb
    pdb = os.path.join("/tmp", "c3_%d.pdb" % copies)
    tile_pdb(os.path.join(HERE, "5xnd_topology.pdb"), pdb, copies)
    atoms = read_pdb(pdb)
    frames = FrameGenerator(atoms.xyz, seed, amplitude=amplitude, device="cpu").frames(n_frames).numpy()
    t0 = time.time()
    res = run_parallel(pdb, frames, ITYPES, collect_near=True)
    f, it, a, b, w1, w2, per = oracle_arrays(res, n_frames)
    topo = compile_topology(atoms, frames[0])
    idx, val = count_table(res, topo.residue_label)
    from chacra_b200.engine import NEAR_KINDS
    near = np.array(sorted({(t[0], NEAR_KINDS.index(t[1]), min(t[2], t[3]), max(t[2], t[3])) for t in res["near"]}), np.int32).reshape(-1, 4)
    np.savez_compressed(os.path.join(HERE, "c3_tiled10_101frames.npz"), copies=copies, seed=seed, n_frames=n_frames, amplitude=amplitude,
                        frames_digest=np.uint64(frames_digest(frames)), per_frame_itype=per,
                        digest=record_digest(f, it, a, b, w1, w2, n_frames), count_idx=idx, count_val=val, near=near)
    print("c3: atoms %d, %d contacts, %d residue pairs, %d near, %.0f s" % (atoms.n, len(f), len(idx), len(near), time.time() - t0))


def make_c5(target=200000, seed=5000, amplitude=0.5):
    import getcontacts_oracle as o
    from chacra_b200 import read_pdb, compile_topology
    from chacra_b200.synthetic import FrameGenerator, solvate_pdb, tile_pdb
    prot, pdb = os.path.join("/tmp", "c5_prot.pdb"), os.path.join("/tmp", "c5_%d.pdb" % target)
    tile_pdb(os.path.join(HERE, "5xnd_topology.pdb"), prot, 4)
    solvate_pdb(prot, pdb, target)
    atoms = read_pdb(pdb)
    frames = FrameGenerator(atoms.xyz, seed, amplitude=amplitude, device="cpu").frames(1).numpy()
    t0 = time.time()
    res = o.run(pdb, pdb, ITYPES, frames=frames, collect_near=True)
    f, it, a, b, w1, w2, per = oracle_arrays(res, 1)
    topo = compile_topology(atoms, frames[0])
    idx, val = count_table(res, topo.residue_label)
    rec = np.stack([f, it, a, b, w1, w2], 1).astype(np.int32)
    from chacra_b200.engine import NEAR_KINDS
    near = np.array(sorted({(t[0], NEAR_KINDS.index(t[1]), min(t[2], t[3]), max(t[2], t[3])) for t in res["near"]}), np.int32).reshape(-1, 4)
    np.savez_compressed(os.path.join(HERE, "c5_solv200k_frame.npz"), target=target, seed=seed, amplitude=amplitude, n_atoms=atoms.n,
                        frames_digest=np.uint64(frames_digest(frames)), records=rec, per_frame_itype=per, count_idx=idx, count_val=val,
                        near=near)
    print("c5: atoms %d, %d contacts (%s), %d residue pairs, %d near, %.0f s" % (
        atoms.n, len(f), ", ".join("%s %d" % (n, c) for n, c in zip(__import__("chacra_b200.engine", fromlist=["x"]).ITYPE_NAMES, per[0]) if c),
        len(idx), len(near), time.time() - t0))


if __name__ == "__main__":
    which = sys.argv[1:] or ["c3", "c5"]
    if "c3" in which:
        make_c3()
    if "c5" in which:
        make_c5()
