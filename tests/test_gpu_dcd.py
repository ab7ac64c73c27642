"""DCD frames interleaved on the device (chc_dcd_unpack) against the host reader and against files
written here with known contents (both byte orders, with and without the unit-cell record)."""
import os
import struct

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
TRJ = os.path.join(GOLDEN, "5xnd_trajectory.dcd")


def write_dcd(path, xyz, endian="<", cell=False, titles=2):
    e = endian
    f, n = xyz.shape[:2]
    icntrl = [0] * 20
    icntrl[0], icntrl[1], icntrl[2], icntrl[10], icntrl[19] = f, 1, 1, 1 if cell else 0, 24
    with open(path, "wb") as fh:
        fh.write(struct.pack(e + "i", 84) + b"CORD" + struct.pack(e + "20i", *icntrl) + struct.pack(e + "i", 84))
        tl = 4 + 80 * titles
        fh.write(struct.pack(e + "ii", tl, titles) + (b"REMARK".ljust(80)) * titles + struct.pack(e + "i", tl))
        fh.write(struct.pack(e + "iii", 4, n, 4))
        for k in range(f):
            if cell:
                fh.write(struct.pack(e + "i", 48) + struct.pack(e + "6d", 50, 90, 50, 90, 90, 50) + struct.pack(e + "i", 48))
            for d in range(3):
                fh.write(struct.pack(e + "i", 4 * n) + xyz[k, :, d].astype(e + "f4").tobytes() + struct.pack(e + "i", 4 * n))


def test_golden_dcd_device_equals_host():
    from chacra_b200.trajectory import DCDFile
    x = DCDFile(TRJ)
    host = x.read()
    assert host.shape == (20, 1629, 3)
    assert np.array_equal(x.read_device().cpu().numpy(), host)
    assert np.array_equal(x.read_device(3, 17, 2).cpu().numpy(), host[3:18:2])
    assert np.array_equal(x.read_device(1, 19, 5).cpu().numpy(), host[1:20:5])        # gathered frame by frame
    assert x.read_device(5, 4).shape[0] == 0
    x.CHUNK_BYTES = 3 * x.frame_bytes + 100                                           # pipelined: seven chunks
    assert np.array_equal(x.read_device().cpu().numpy(), host)
    assert np.array_equal(x.read_device(0, 18, 2).cpu().numpy(), host[0:19:2])


@pytest.mark.parametrize("endian,cell,n", [("<", False, 1), (">", False, 255), ("<", True, 256), (">", True, 1000), ("<", True, 70001)])
def test_written_files_round_trip(tmp_path, endian, cell, n):
    from chacra_b200.trajectory import DCDFile
    rng = np.random.default_rng(n)
    xyz = (rng.standard_normal((7, n, 3)) * 40).astype(np.float32)
    path = str(tmp_path / "t.dcd")
    write_dcd(path, xyz, endian, cell)
    x = DCDFile(path)
    assert (x.n_frames, x.n_atoms) == (7, n)
    assert np.array_equal(x.read(), xyz)
    assert np.array_equal(x.read_device().cpu().numpy(), xyz)
    assert np.array_equal(x.read_device(2, 6, 3).cpu().numpy(), xyz[2:7:3])


def test_contacts_from_dcd_use_the_device_reader(tmp_path):
    """compute_contacts on the 5xnd DCD: the frames reach the engine through read_device; the counts equal those of
    the host-read frames submitted directly."""
    from chacra_b200 import compile_topology, read_pdb
    from chacra_b200.engine import ContactEngine
    from chacra_b200.trajectory import DCDFile
    atoms = read_pdb(os.path.join(GOLDEN, "5xnd_topology.pdb"))
    topo = compile_topology(atoms)
    x = DCDFile(TRJ)
    a = ContactEngine(topo, ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"])
    a.submit(x.read_device(), 0, 0)
    a.synchronize()
    b = ContactEngine(topo, ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"])
    b.submit(x.read(), 0, 0)
    b.synchronize()
    assert a.counts(0).sum() > 0 and np.array_equal(a.counts(0), b.counts(0))
    a.close(); b.close()
