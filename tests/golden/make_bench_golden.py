"""Committed oracle count table for bench.py's in-run parity check (outside the timed region): 100 frames (one fragment) of the
BASELINE config-2 shape (bench.make_topology(4), 6 516 atoms) at the amplitude of the hottest state, generated on the CPU by
chacra_b200.synthetic.FrameGenerator(seed 2100) and pushed through the numpy oracle.
Run from the repository root:  python tests/golden/make_bench_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, HERE)


def main(seed=2100, n_frames=100):
    import bench
    from make_config_goldens import count_table, frames_digest
    from parallel import run_parallel
    from chacra_b200 import compile_topology
    from chacra_b200.synthetic import FrameGenerator
    path, atoms = bench.make_topology(4)
    amp = bench.state_amplitude(19, 20)
    frames = FrameGenerator(atoms.xyz, seed, amplitude=amp, device="cpu").frames(n_frames).numpy()
    res = run_parallel(path, frames, bench.ITYPES)
    topo = compile_topology(atoms)
    idx, val = count_table(res, topo.residue_label)
    np.savez_compressed(os.path.join(HERE, "bench_c2_100frames.npz"), seed=seed, n_frames=n_frames, amplitude=amp,
                        frames_digest=np.uint64(frames_digest(frames)), count_idx=idx, count_val=val)
    print("pairs %d pair-frames %d" % (len(idx), int(val.sum())))


if __name__ == "__main__":
    main()
