"""A small XTC file of the BASELINE config-2 shape for bench.py's `e2e_xtc` leg (what ChACRA hands getcontacts is state_{i}.xtc):
24 frames of bench.make_topology(4) (6 516 atoms) from chacra_b200.synthetic.FrameGenerator(seed 2200), encoded with the
test-only encoder of oracle/xtc_oracle.py (precision 1000).  bench.py tiles the compressed frames in pinned host memory.
Run from the repository root:  python tests/golden/make_bench_xtc.py"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def main(seed=2200, n_frames=24):
    import bench
    import xtc_oracle as xo
    from chacra_b200.synthetic import FrameGenerator
    path, atoms = bench.make_topology(4)
    frames = FrameGenerator(atoms.xyz, seed, amplitude=bench.state_amplitude(10, 20), device="cpu").frames(n_frames).numpy()
    out = os.path.join(HERE, "bench_c2_24frames.xtc")
    with open(out, "wb") as fh:
        for k, f in enumerate(frames):
            fh.write(xo.encode_frame(f, step=k))
    print("%s: %d bytes, %.2f of float32" % (out, os.path.getsize(out), os.path.getsize(out) / (12.0 * frames.shape[1] * n_frames)))


if __name__ == "__main__":
    main()
