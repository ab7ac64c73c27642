import gzip
import os
import shutil
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def pdb_5y0t(tmp_path_factory):
    """The hydrogenated 5Y0T structure is stored gzip-compressed; the readers want a plain file."""
    dst = tmp_path_factory.mktemp("golden") / "5Y0T_h.pdb"
    with gzip.open(os.path.join(GOLDEN, "5Y0T_h.pdb.gz"), "rb") as src, open(dst, "wb") as out:
        shutil.copyfileobj(src, out)
    return str(dst)


@pytest.fixture(scope="session")
def oracle_5xnd():
    """All-itype oracle run on the 5xnd fixture (20 frames, ~4 s), shared by the tests."""
    import getcontacts_oracle as o
    return o.run(os.path.join(GOLDEN, "5xnd_topology.pdb"), os.path.join(GOLDEN, "5xnd_trajectory.dcd"),
                 ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"], collect_near=True)


def read_golden_tsv(path):
    op = gzip.open if path.endswith(".gz") else open
    with op(path, "rt") as fh:
        return [ln.rstrip("\n").split("\t") for ln in fh if ln.strip() and not ln.startswith("#")]
