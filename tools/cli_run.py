"""The drop-in script end to end on the config-2 shape: DCD in, contact-list TSV out (what ChACRA's process-output runs
per state), plus the frequency table straight from the device counts.
  python tools/cli_run.py [--frames 2000]"""
import argparse
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=2000)
    a = ap.parse_args()
    import io
    import numpy as np
    from chacra_b200.synthetic import FrameGenerator
    from chacra_b200.trajectory import DCDFile
    from chacra_b200.getcontacts import get_dynamic_contacts
    from test_gpu_dcd import write_dcd
    top, atoms = bench.make_topology(4)
    tmp = tempfile.gettempdir()
    dcd, out = os.path.join(tmp, "cli_run_%d.dcd" % os.getpid()), os.path.join(tmp, "cli_run_%d.tsv" % os.getpid())
    gen = FrameGenerator(atoms.xyz, 2000)
    write_dcd(dcd, np.concatenate([gen.frames(500).numpy() for _ in range((a.frames + 499) // 500)])[:a.frames], "<", True)
    argv = ["--topology", top, "--trajectory", dcd, "--output", out, "--itypes", "all", "--distout", "--sele", "protein", "--sele2", "protein",
            "--cores", "8"]
    get_dynamic_contacts.main(argv=argv[:6] + ["--itypes", "sb"] + argv[8:] + ["--end", "99"])      # warm-up: library, CUDA context
    t0 = time.perf_counter()
    get_dynamic_contacts.main(argv=argv)
    dt = time.perf_counter() - t0
    size = os.path.getsize(out)
    lines = sum(1 for _ in open(out)) - 2
    # second script of the flow: the frequency table from that file (threaded reader), and the same table from the device counts
    from chacra_b200 import frequencies as fq
    t0 = time.perf_counter()
    total, shipped = fq.native_counts(out, fq.ALL_FREQ_ITYPES, shipped=True)
    t_freq = time.perf_counter() - t0
    _, intended = fq.native_counts(out, fq.ALL_FREQ_ITYPES, shipped=False)
    sample = []
    with open(out) as fh:
        for k, line in enumerate(fh):
            sample.append(line)
            if k >= 400000:
                break
    t0 = time.perf_counter()
    fq.shipped_counts(sample, fq.ALL_FREQ_ITYPES)
    t_py = (time.perf_counter() - t0) / len(sample)
    # the Python line loop this writer replaced, on the records of 20 frames
    import tsv_oracle
    from chacra_b200 import compile_topology
    from chacra_b200.engine import ContactEngine
    eng = ContactEngine(compile_topology(atoms), bench.ITYPES, n_states=1, emit_contacts=True, contact_capacity=1 << 20)
    eng.submit(DCDFile(dcd).read_device(0, 19), 0, 0)
    eng.synchronize()
    rec = eng.drain_contacts()
    # the count table the kernels accumulated for ALL frames == the table read back from the text they emitted
    full = ContactEngine(compile_topology(atoms), bench.ITYPES, n_states=1)
    full.submit(DCDFile(dcd).read_device(), 0, 0)
    full.synchronize()
    table = fq.counts_to_table(compile_topology(atoms).residue_label, full.counts(0))
    assert table == intended and total == a.frames, (len(table), len(intended), total)
    t0 = time.perf_counter()
    tsv_oracle.ContactWriterPy(io.StringIO(), atoms.atom_labels(), 0, 1, True).write(rec)
    py = (time.perf_counter() - t0) / 20
    print("cli: %d frames x %d atoms -> %d lines, %.1f MB in %.2f s = %.0f frames/s (%.0f lines/frame, %.2f GB/s of text) | "
          "Python line loop: %.1f ms per frame = %.0f frames/s | frequency table from the file: %.2f s = %.1f M lines/s (Python reader: "
          "%.2f M lines/s), %d pairs, equal to the device count table" % (
              a.frames, atoms.n, lines, size / 1e6, dt, a.frames / dt, lines / a.frames, size / dt / 1e9, py * 1e3, 1.0 / py,
              t_freq, lines / t_freq / 1e6, 1e-6 / t_py, len(shipped)))
    os.unlink(dcd); os.unlink(out)


if __name__ == "__main__":
    main()
