"""Multi-GPU product path check (run on a box with >= N GPUs): get_state_frequencies.py under torchrun with N ranks must write
exactly the files one rank writes.  python tools/launcher_check.py [N]"""
import filecmp
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    import bench
    from test_gpu_dcd import write_dcd
    from chacra_b200.synthetic import FrameGenerator
    tmp = tempfile.mkdtemp(prefix="chacra_launcher_")
    path, atoms = bench.make_topology(4)
    states = []
    for s, frames in enumerate((1300, 250, 730, 99, 1000)):
        x = FrameGenerator(atoms.xyz, 7000 + s, amplitude=bench.state_amplitude(s, 5), device="cpu").frames(frames).numpy()
        p = os.path.join(tmp, "state_%d.dcd" % s)
        write_dcd(p, x, "<", False)
        states.append(p)
    script = os.path.join(ROOT, "chacra_b200", "getcontacts", "get_state_frequencies.py")
    common = ["--topology", path, "--trajectories"] + states + ["--counts"]
    one, many = os.path.join(tmp, "one"), os.path.join(tmp, "many")
    subprocess.run([sys.executable, script] + common + ["--output_dir", one, "--near_cutoff_output", one + "_near.tsv"], check=True)
    subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
                    "--master-port", "29577", script] + common + ["--output_dir", many, "--near_cutoff_output", many + "_near.tsv"], check=True)
    names = sorted(os.listdir(one))
    assert names == sorted(os.listdir(many)) and len(names) == 10, names
    match, mismatch, errors = filecmp.cmpfiles(one, many, names, shallow=False)
    assert not mismatch and not errors, (mismatch, errors)
    assert filecmp.cmp(one + "_near.tsv", many + "_near.tsv", shallow=False)
    rows = sum(1 for _ in open(one + "_near.tsv"))
    print("launcher check ok: %d ranks, %d files identical to the single-rank run, %d near-cutoff lines" % (n, len(names), rows))


if __name__ == "__main__":
    main()
