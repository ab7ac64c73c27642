"""Trajectory input: CHARMM/NAMD DCD -> float32 [F][N][3] Angstrom.

Replaces VMD's molfile plugins on the path (``contact_utils.py:89-116`` ``load_traj``,
``:119-131`` ``simulation_length``).  XTC decode is the next row.
"""
from __future__ import annotations

import struct

import numpy as np


class DCDFile:
    """Random access to the frames of a DCD file through a memory map."""

    def __init__(self, path: str):
        self.path = path
        head = np.fromfile(path, dtype=np.uint8, count=4096)
        raw = head.tobytes()
        if len(raw) < 100:
            raise ValueError("%s: not a DCD file" % path)
        if struct.unpack("<i", raw[:4])[0] == 84:
            self.endian = "<"
        elif struct.unpack(">i", raw[:4])[0] == 84:
            self.endian = ">"
        else:
            raise ValueError("%s: not a DCD file (bad first record)" % path)
        e = self.endian
        if raw[4:8] != b"CORD":
            raise ValueError("%s: not a coordinate DCD" % path)
        icntrl = struct.unpack(e + "20i", raw[8:88])
        self.has_cell = icntrl[10] == 1
        if icntrl[11] == 1:
            raise ValueError("%s: 4-dimensional DCD files are not supported" % path)
        pos = 92
        n = struct.unpack(e + "i", raw[pos:pos + 4])[0]
        pos += 8 + n
        if pos + 16 > len(raw):                     # long title block: re-read what is needed
            raw = open(path, "rb").read(pos + 16)
        n = struct.unpack(e + "i", raw[pos:pos + 4])[0]
        self.n_atoms = struct.unpack(e + "i", raw[pos + 4:pos + 8])[0]
        pos += 8 + n
        self.data_offset = pos
        self.cell_bytes = 56 if self.has_cell else 0
        self.frame_bytes = self.cell_bytes + 3 * (8 + 4 * self.n_atoms)
        import os
        self.n_frames = (os.path.getsize(path) - pos) // self.frame_bytes
        self._mm = np.memmap(path, dtype=np.uint8, mode="r") if self.n_frames else None

    def read(self, beg: int = 0, end: int | None = None, stride: int = 1) -> np.ndarray:
        """Frames beg, beg+stride, ... <= end (inclusive, like the reference's --beg/--end)."""
        end = self.n_frames - 1 if end is None else min(end, self.n_frames - 1)
        idx = range(beg, end + 1, max(1, stride))
        out = np.empty((len(idx), self.n_atoms, 3), np.float32)
        rec = 4 * self.n_atoms
        for k, f in enumerate(idx):
            p = self.data_offset + f * self.frame_bytes + self.cell_bytes
            for d in range(3):
                out[k, :, d] = np.frombuffer(self._mm, dtype=self.endian + "f4", count=self.n_atoms, offset=p + 4)
                p += 8 + rec
        return out


def read_dcd(path: str) -> np.ndarray:
    return DCDFile(path).read()
