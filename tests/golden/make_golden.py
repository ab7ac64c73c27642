"""Regenerates tests/golden/ from the read-only reference checkout (run in the build container only).

The files are DATA the reference ships for its own tests (external/getcontacts/tests/): the 5xnd
topology + 20-frame DCD with the expected all-itype contact list, and the hydrogenated 5Y0T
structure with its expected chain A / chain B interface contacts (incl. water bridges).  The large
ones are gzip-compressed.  `freq_5xnd_reference.tsv` is the OUTPUT of the reference's unchanged
get_contact_frequencies.py on the 5xnd golden contact list (pure Python, runs here).
"""
import gzip
import os
import shutil
import subprocess
import sys

REF = "/root/reference/external/getcontacts"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    t = os.path.join(REF, "tests")
    shutil.copy(os.path.join(t, "5xnd_topology.pdb"), os.path.join(HERE, "5xnd_topology.pdb"))
    shutil.copy(os.path.join(t, "5xnd_trajectory.dcd"), os.path.join(HERE, "5xnd_trajectory.dcd"))
    shutil.copy(os.path.join(t, "test_5y0t_interface_expected.tsv"), os.path.join(HERE, "test_5y0t_interface_expected.tsv"))
    for name in ("test_5xnd_expected.tsv", "5Y0T_h.pdb"):
        with open(os.path.join(t, name), "rb") as src, gzip.GzipFile(os.path.join(HERE, name + ".gz"), "wb", mtime=0) as dst:
            shutil.copyfileobj(src, dst)
    out = os.path.join(HERE, "freq_5xnd_reference.tsv")
    subprocess.run([sys.executable, "get_contact_frequencies.py", "--input_files", os.path.join(t, "test_5xnd_expected.tsv"),
                    "--output_file", out], cwd=REF, check=True, env=dict(os.environ, PYTHONHASHSEED="0"))
    # row order depends on the string-hash seed; store sorted so that the fixture is stable
    lines = open(out).read().splitlines()
    with open(out, "w") as fh:
        fh.write("\n".join(lines[:2] + sorted(lines[2:])) + "\n")


if __name__ == "__main__":
    main()
