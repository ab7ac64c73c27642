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


def tune_pair_distance(xyz, i, j, target, above):
    """Move atom j along the line from atom i until the float32 ``sqrtf(d2)`` of the pair (the recipe of SURVEY A.4/A.7)
    lies within 5e-5 of ``target``, on the side asked for.  Returns a float32 copy of the frame."""
    import numpy as np
    x = np.array(xyz, dtype=np.float32, copy=True)
    u = (x[j].astype(np.float64) - x[i].astype(np.float64))
    u /= np.linalg.norm(u)

    def dist(t):
        y = x.copy()
        y[j] = (x[i].astype(np.float64) + u * t).astype(np.float32)
        d = (y[i] - y[j]).astype(np.float32)
        d2 = np.float32(np.float32(d[0] * d[0] + d[1] * d[1]) + d[2] * d[2])
        return float(np.float32(np.sqrt(d2))), y
    want = target + (3e-5 if above else -3e-5)
    lo, hi = target - 1.0, target + 1.0
    for _ in range(60):
        mid = 0.5 * (lo + hi)
        d, y = dist(mid)
        if d < want:
            lo = mid
        else:
            hi = mid
    for t in (mid, lo, hi):
        d, y = dist(t)
        if abs(d - target) < 9e-5 and (d > target) == above:
            return y
    raise AssertionError("could not tune the pair distance (%.6f)" % d)
