"""DCD input throughput on the config-2 shape: a 10 000-frame DCD written once, then read (a) by the
host reader (numpy interleave, what round-1 sessions before this one shipped) and (b) through DCDFile.read_device
(file bytes over PCIe + chc_dcd_unpack), and fed to the contact engine.
  python tools/dcd_run.py [--frames 10000]"""
import argparse
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=10000)
    a = ap.parse_args()
    import numpy as np
    import torch
    from chacra_b200 import compile_topology
    from chacra_b200.engine import ContactEngine
    from chacra_b200.synthetic import FrameGenerator
    from chacra_b200.trajectory import DCDFile
    from test_gpu_dcd import write_dcd
    path, atoms = bench.make_topology(4)
    gen = FrameGenerator(atoms.xyz, 2000)
    base = gen.frames(500).numpy()
    dcd = os.path.join(tempfile.gettempdir(), "dcd_run_%d.dcd" % os.getpid())
    write_dcd(dcd, base, "<", True)
    blob = open(dcd, "rb").read()
    x0 = DCDFile(dcd)
    with open(dcd, "ab") as fh:                          # repeat the frame records up to --frames
        for _ in range((a.frames + 499) // 500 - 1):
            fh.write(blob[x0.data_offset:])
    x = DCDFile(dcd)
    n = min(a.frames, x.n_frames)
    eng = ContactEngine(compile_topology(atoms), bench.ITYPES, n_states=1)
    x.read_device(0, 499)
    eng.submit(x.read_device(0, 499), 0, 0)
    eng.synchronize()
    t0 = time.perf_counter()
    host = x.read(0, min(n, 2000) - 1)
    t_host = (time.perf_counter() - t0) * n / min(n, 2000)
    assert np.array_equal(host[:500], base)
    res = []
    for _ in range(3):
        eng.reset_counts()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dev = x.read_device(0, n - 1)
        torch.cuda.synchronize()
        t_read = time.perf_counter() - t0
        eng.submit(dev, 0, 0)
        eng.synchronize()
        res.append((time.perf_counter() - t0, t_read))
    t_all, t_read = min(res)
    assert np.array_equal(dev[:500].cpu().numpy(), base)
    gb = 12.0 * x.n_atoms * n / 1e9
    print("dcd: %d frames x %d atoms (%.2f GB) | host reader %.2f s = %.0f frames/s | read_device %.3f s = %.0f frames/s (%.1f GB/s) | "
          "read_device + contacts %.3f s = %.0f frames/s" % (n, x.n_atoms, gb, t_host, n / t_host, t_read, n / t_read, gb / t_read, t_all, n / t_all))
    os.unlink(dcd)


if __name__ == "__main__":
    main()
