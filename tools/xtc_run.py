"""Decode throughput of the XTC path on a config-2-shaped trajectory: 200 distinct synthetic frames
are encoded once by the test-only oracle encoder, the file is repeated to --frames frames, then read through
chacra_b200.xtc.XTCFile (host header walk + H2D of the compressed bytes + device decode) and fed to the engine.
  python tools/xtc_run.py [--frames 10000]"""
import argparse
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=10000)
    a = ap.parse_args()
    import torch
    import xtc_oracle as xo
    from chacra_b200 import compile_topology
    from chacra_b200.engine import ContactEngine
    from chacra_b200.synthetic import FrameGenerator
    from chacra_b200.xtc import XTCFile
    path, atoms = bench.make_topology(4)
    gen = FrameGenerator(atoms.xyz, 2000)
    base = gen.frames(200).numpy()
    t0 = time.time()
    one = b"".join(xo.encode_frame(f, step=k) for k, f in enumerate(base))
    t_enc = time.time() - t0
    xtc = os.path.join(tempfile.gettempdir(), "xtc_run_%d.xtc" % os.getpid())
    reps = (a.frames + 199) // 200
    with open(xtc, "wb") as fh:
        for _ in range(reps):
            fh.write(one)
    t0 = time.time()
    x = XTCFile(xtc)
    t_index = time.time() - t0
    n = min(a.frames, x.n_frames)
    x.read_device(0, 199)                       # warm-up (library load, pinned allocation)
    torch.cuda.synchronize()
    t0 = time.time()
    dev = x.read_device(0, n - 1)
    torch.cuda.synchronize()
    t_read = time.time() - t0
    t_host, t_dev = x.last_timing
    want = xo.ints_to_angstrom(xo.quantize_nm(base[:3] / 10.0))
    assert (dev[:3].cpu().numpy() == want).all() and (dev[200:203].cpu().numpy() == want).all()
    eng = ContactEngine(compile_topology(atoms), bench.ITYPES, n_states=1)
    eng.submit(dev, 0, 0)
    eng.synchronize()
    t0 = time.time()
    eng.reset_counts()
    dev2 = x.read_device(0, n - 1)
    eng.submit(dev2, 0, 0)
    eng.synchronize()
    t_all = time.time() - t0
    comp = x.compressed_bytes() * n / x.n_frames
    print("xtc: %d frames x %d atoms, %.1f MB compressed (%.2f of float32) | encode(oracle, 200 frames) %.1f s | header walk %.2f s | "
          "read_device %.3f s (file->pinned %.3f, H2D+decode %.3f) = %.0f frames/s (%.2f GB/s of float32 out) | read + contacts %.3f s = %.0f frames/s" % (
              n, x.n_atoms, comp / 1e6, comp / (12.0 * x.n_atoms * n), t_enc, t_index, t_read, t_host, t_dev, n / t_read,
              12.0 * x.n_atoms * n / t_read / 1e9, t_all, n / t_all))
    os.unlink(xtc)


if __name__ == "__main__":
    main()
