"""Small driver for ncu: N launches of the contact kernel on config-2-shaped frames (no oracle, no host path)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=1480)
    ap.add_argument("--launches", type=int, default=3)
    ap.add_argument("--copies", type=int, default=4)
    ap.add_argument("--no-rotate", action="store_true")
    ap.add_argument("--itypes", default=",".join(bench.ITYPES))
    ap.add_argument("--emit", action="store_true", help="records mode: contact records are emitted (and dropped) -- the kernel the drop-in script runs")
    ap.add_argument("--near", action="store_true", help="near-cutoff channel on")
    a = ap.parse_args()
    import torch
    from chacra_b200 import compile_topology
    from chacra_b200.engine import ContactEngine
    from chacra_b200.synthetic import FrameGenerator
    path, atoms = bench.make_topology(a.copies)
    topo = compile_topology(atoms)
    eng = ContactEngine(topo, a.itypes.split(","), n_states=1, emit_contacts=a.emit, track_near=a.near, contact_capacity=1 << 20, near_capacity=1 << 16)
    gen = FrameGenerator(atoms.xyz, 2000, device="cuda", rotate=not a.no_rotate)
    x = torch.cat([gen.frames(min(500, a.frames - b)) for b in range(0, a.frames, 500)])
    eng.set_timing(True)
    for _ in range(a.launches):
        eng.submit(x, 0, 0)
    eng.synchronize()
    st = eng.stats()
    pair_ms, bin_ms = st["pair_ms_total"] / a.launches, st["bin_ms_total"] / a.launches
    ms = pair_ms + bin_ms
    print("frames %d atoms %d path %d: pair %.3f ms + bin %.3f ms per pass, %.0f frames/s, pair %.2f us/frame/SM, "
          "bin %.0f GB/s, smem %d, chunk %d, record %d B, pairs/frame %.1f" % (
              a.frames, atoms.n, st["path"], pair_ms, bin_ms, a.frames / ms * 1e3, pair_ms * 1e3 / (a.frames / 148.0),
              12.0 * atoms.n * a.frames / (bin_ms * 1e-3) / 1e9, st["smem_bytes"], st["chunk_frames"], st["record_bytes"],
              float(eng.counts(0).sum()) / a.frames / a.launches))


if __name__ == "__main__":
    main()
