"""Throughput of the BASELINE.json shapes that are not the bench workload (configs 3, 4, 5): one state, a few hundred
frames generated on the device, CUDA-event times of the binning and pair kernels.
  python tools/config_run.py --config 3|4|5 [--frames F] [--passes K]"""
import argparse
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=3)
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--passes", type=int, default=2)
    a = ap.parse_args()
    import torch
    from chacra_b200 import read_pdb, compile_topology
    from chacra_b200.engine import ContactEngine
    from chacra_b200.synthetic import FrameGenerator, solvate_pdb, tile_pdb
    src = os.path.join(bench.GOLDEN, "5xnd_topology.pdb")
    tmp = tempfile.gettempdir()
    t0 = time.time()
    if a.config == 5:           # 4 protein copies + TIP3 to 200k atoms
        prot, pdb = os.path.join(tmp, "c5_prot_%d.pdb" % os.getpid()), os.path.join(tmp, "c5_%d.pdb" % os.getpid())
        tile_pdb(src, prot, 4)
        solvate_pdb(prot, pdb, 200000)
        frames, amp = a.frames or 296, 0.5
    else:
        copies = {3: 10, 4: 61}[a.config]
        pdb = os.path.join(tmp, "c%d_%d.pdb" % (a.config, os.getpid()))
        tile_pdb(src, pdb, copies)
        frames, amp = a.frames or {3: 1480, 4: 296}[a.config], 0.8
    atoms = read_pdb(pdb)
    t1 = time.time()
    topo = compile_topology(atoms)
    t2 = time.time()
    eng = ContactEngine(topo, bench.ITYPES, n_states=1)
    gen = FrameGenerator(atoms.xyz, 1000 * a.config, amplitude=amp, device="cuda")
    step = max(1, min(frames, (256 << 20) // (12 * atoms.n)))
    x = torch.cat([gen.frames(min(step, frames - b)) for b in range(0, frames, step)])
    eng.submit(x, 0, 0)            # warm-up
    eng.synchronize()
    eng.reset_counts()
    eng.set_timing(True)
    for _ in range(a.passes):
        eng.submit(x, 0, 0)
    eng.synchronize()
    st = eng.stats()
    pair_ms, bin_ms = st["pair_ms_total"] / a.passes, st["bin_ms_total"] / a.passes
    print("config %d: atoms %d pair atoms %d (grid A %d, solvent %d) residues %d | read %.1f s compile %.1f s | path %d chunk %d "
          "record %d B | %d frames: pair %.2f ms + bin %.2f ms -> %.0f frames/s (bin %.0f GB/s) | residue pairs/frame %.1f" % (
              a.config, atoms.n, len(topo.pair_atom), topo.n_pair_a, topo.n_solvent_pair, topo.n_residues, t1 - t0, t2 - t1,
              st["path"], st["chunk_frames"], st["record_bytes"], frames, pair_ms, bin_ms, frames / (pair_ms + bin_ms) * 1e3,
              12.0 * atoms.n * frames / (bin_ms * 1e-3) / 1e9, float(eng.counts(0).sum()) / frames / a.passes))


if __name__ == "__main__":
    main()
