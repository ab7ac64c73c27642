"""Trajectory input: CHARMM/NAMD DCD -> float32 [F][N][3] Angstrom.

Replaces VMD's molfile plugins on the path (``contact_utils.py:89-116`` ``load_traj``,
``:119-131`` ``simulation_length``).  XTC is decoded on the GPU (``xtc.py``).
"""
from __future__ import annotations

import ctypes as C
import struct

import numpy as np

from ._hostio import FileStager


class DCDFile(FileStager):
    """Random access to the frames of a DCD file.  ``read`` interleaves on the host (numpy, no GPU needed: topology
    set-up reads frame 0 this way); ``read_device`` ships the file bytes over PCIe as they are and interleaves them
    on the GPU (``chc_dcd_unpack``, csrc/dcd_unpack.cuh), which is what the contact path uses."""

    CHUNK_BYTES = 48 << 20           # raw bytes per pipeline stage (file read of chunk k+1 under the H2D copy of chunk k)

    def __init__(self, path: str, device: int = 0):
        self.path, self.device = path, int(device)
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

    def read_device(self, beg: int = 0, end: int | None = None, stride: int = 1):
        """The same frames as ``read`` as a CUDA float32 [F][N][3] tensor."""
        import torch
        from . import _lib
        if not torch.cuda.is_available():
            raise _lib.ChacraLibraryError("read_device needs the GPU (sm_100a); use read() for host arrays")
        end = self.n_frames - 1 if end is None else min(end, self.n_frames - 1)
        stride = max(1, stride)
        idx = np.arange(beg, end + 1, stride)
        dev = torch.device("cuda", self.device)
        out = torch.empty((len(idx), self.n_atoms, 3), dtype=torch.float32, device=dev)
        if len(idx) == 0:
            return out
        lib = _lib.load()
        span = stride <= 2                               # read the skipped frames too rather than gather frame by frame
        pitch = self.frame_bytes * (stride if span else 1)
        per_chunk = max(1, self.CHUNK_BYTES // pitch)
        cuts = list(range(0, len(idx), per_chunk)) + [len(idx)]
        swap = 1 if self.endian != ("<" if np.little_endian else ">") else 0
        d_raw = [None, None]
        copied = [torch.cuda.Event(), torch.cuda.Event()]
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

        def stage(c, which):
            frames = idx[cuts[c]:cuts[c + 1]]
            if span:
                lo = self.data_offset + int(frames[0]) * self.frame_bytes
                nbytes = (int(frames[-1]) - int(frames[0])) * self.frame_bytes + self.frame_bytes
                host = self._pinned(nbytes, which)
                self._pread(memoryview(host.numpy())[:nbytes], lo)
            else:
                nbytes = len(frames) * self.frame_bytes
                host = self._pinned(nbytes, which)
                view = host.numpy()
                for k, f in enumerate(frames.tolist()):
                    p = self.data_offset + f * self.frame_bytes
                    view[k * self.frame_bytes:(k + 1) * self.frame_bytes] = self._mm[p:p + self.frame_bytes]
            return nbytes

        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(1) as reader:
            pending = reader.submit(stage, 0, 0)
            for c in range(len(cuts) - 1):
                nbytes = pending.result()
                w = c & 1
                if d_raw[w] is None or d_raw[w].numel() < nbytes:
                    d_raw[w] = torch.empty(nbytes, dtype=torch.uint8, device=dev)
                d_raw[w][:nbytes].copy_(self._stage_bufs[w][:nbytes], non_blocking=True)
                copied[w].record()
                if c + 2 < len(cuts):
                    if c >= 1:
                        copied[(c + 1) & 1].synchronize()   # the copy out of that pinned buffer has finished
                    pending = reader.submit(stage, c + 1, (c + 1) & 1)
                n = cuts[c + 1] - cuts[c]
                rc = lib.chc_dcd_unpack(d_raw[w].data_ptr(), self.cell_bytes + 4, pitch, n, self.n_atoms, swap,
                                        out[cuts[c]:].data_ptr(), 0, self.device, stream)
                if rc != 0:
                    raise RuntimeError("chc_dcd_unpack failed (%d)" % rc)
            # the pinned staging buffers belong to the reader: the next read_device call refills them straight away, so the
            # copies out of them must have finished before this call returns (the unpack kernels may still be running)
            for k in range(min(2, len(cuts) - 1)):
                copied[k].synchronize()
        return out


def read_dcd(path: str) -> np.ndarray:
    return DCDFile(path).read()


class PDBFrames:
    """A PDB file used as a trajectory: one frame per MODEL (a file without MODEL records is one frame).
    VMD's pdb plugin behaves the same way (``contact_utils.py:106-114`` loads it with ``molecule.read``)."""

    def __init__(self, path: str):
        frames, cur = [], []
        with open(path) as fh:
            for line in fh:
                rec = line[:6]
                if rec in ("ATOM  ", "HETATM"):
                    cur.append((float(line[30:38]), float(line[38:46]), float(line[46:54])))
                elif rec.strip() == "ENDMDL":
                    if cur:
                        frames.append(cur)
                    cur = []
        if cur:
            frames.append(cur)
        if not frames:
            raise ValueError("%s: no ATOM/HETATM records" % path)
        n = len(frames[0])
        frames = [f for f in frames if len(f) == n]
        self.xyz = np.array(frames, np.float64).astype(np.float32)
        self.n_frames, self.n_atoms = self.xyz.shape[0], n

    def read(self, beg: int = 0, end: int | None = None, stride: int = 1) -> np.ndarray:
        end = self.n_frames - 1 if end is None else min(end, self.n_frames - 1)
        return np.ascontiguousarray(self.xyz[beg:end + 1:max(1, stride)])


def open_trajectory(path: str):
    """Trajectory reader by file suffix (the reference maps suffixes to VMD molfile plugins,
    ``contact_utils.py:40-66``).  Every reader offers ``n_frames``, ``n_atoms`` and
    ``read(beg, end, stride)`` -> float32 [F][N][3] Angstrom with ``end`` inclusive."""
    low = path.lower()
    if low.endswith(".dcd"):
        return DCDFile(path)
    if low.endswith(".pdb") or low.endswith(".ent"):
        return PDBFrames(path)
    if low.endswith(".xtc"):
        from .xtc import XTCFile
        return XTCFile(path)
    raise ValueError("unsupported trajectory format: %s (dcd, xtc and pdb are read)" % path)
