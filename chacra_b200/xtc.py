"""XTC trajectory input, decoded on the GPU.

ChACRA hands getcontacts ``state_{i}.xtc`` (``chacra/scripts/process_hremd_output.py:147``), which the reference reads
through VMD's gromacs molfile plugin (``contact_calc/contact_utils.py:89-116``).  Here the host only walks the frame
headers (52 bytes + 40 bytes of coding parameters per frame); the compressed bit streams go to the device as they are
and ``chc_xtc_decode`` (``csrc/xtc_decode.cuh``) turns them into float32 [F][N][3] Angstrom (group boundaries by one thread per frame, then one thread per group).
There is no CPU decoder on the product path.

Parity note: no XTC fixture, xdrfile source or XTC-capable package exists in the reference tree or in this image, so
the format is restated from the published xdrfile algorithm (``oracle/xtc_oracle.py``): *unpinned* against real files.
"""
from __future__ import annotations

import ctypes as C
import os
import struct
import time

import numpy as np

from . import _lib
from ._hostio import FileStager

XTC_MAGIC = 1995
_SLACK = 128                      # zero bytes after the last stream of a transfer: a corrupt run may read this far past its end
_HEADER = 52                      # magic, natoms, step, time, box[9]
META_DTYPE = np.dtype([("offset", "<i8"), ("nbytes", "<i4"), ("precision", "<f4"), ("minint", "<i4", 3),
                       ("maxint", "<i4", 3), ("smallidx", "<i4"), ("raw", "<i4"), ("bitsize", "<i4"),
                       ("bitsizeint", "<i4", 3)])
assert META_DTYPE.itemsize == 64


def _sizeofint(size: int) -> int:
    num, bits = 1, 0
    while size >= num and bits < 32:
        bits += 1
        num <<= 1
    return bits


def _sizeofints(sizes) -> int:
    prod = 1
    for s in sizes:
        prod *= int(s)
    return max(prod.bit_length(), 1)


class XTCFile(FileStager):
    """Frame index of an XTC file + device decode.  ``read`` returns host float32 [F][N][3] (like the other readers),
    ``read_device`` a CUDA tensor that ``ContactEngine.submit`` takes without another copy."""

    def __init__(self, path: str, device: int = 0):
        self.path, self.device = path, int(device)
        self._mm = np.memmap(path, dtype=np.uint8, mode="r")
        raw = self._mm
        size = raw.shape[0]
        pos, recs = 0, []
        self.n_atoms = None
        while pos + _HEADER + 4 <= size:
            head = bytes(raw[pos:pos + _HEADER + 40])
            magic, natoms = struct.unpack_from(">ii", head, 0)
            if magic != XTC_MAGIC:
                raise ValueError("%s: bad XTC magic %d at byte %d" % (path, magic, pos))
            if self.n_atoms is None:
                self.n_atoms = natoms
            elif natoms != self.n_atoms:
                raise ValueError("%s: frame %d has %d atoms, frame 0 has %d" % (path, len(recs), natoms, self.n_atoms))
            (lsize,) = struct.unpack_from(">i", head, _HEADER)
            if lsize != natoms:
                raise ValueError("%s: inconsistent atom counts in frame %d" % (path, len(recs)))
            rec = np.zeros((), META_DTYPE)
            if natoms <= 9:
                rec["raw"], rec["nbytes"], rec["precision"] = 1, 12 * natoms, 1.0
                data = pos + _HEADER + 4
                nxt = data + 12 * natoms
            else:
                if len(head) < _HEADER + 40:
                    raise ValueError("%s: truncated frame %d" % (path, len(recs)))
                precision, = struct.unpack_from(">f", head, _HEADER + 4)
                mn = struct.unpack_from(">3i", head, _HEADER + 8)
                mx = struct.unpack_from(">3i", head, _HEADER + 20)
                smallidx, nbytes = struct.unpack_from(">ii", head, _HEADER + 32)
                sizeint = [mx[k] - mn[k] + 1 for k in range(3)]
                if min(sizeint) <= 0 or nbytes < 0 or not precision > 0:
                    raise ValueError("%s: corrupt coding parameters in frame %d" % (path, len(recs)))
                rec["precision"], rec["minint"], rec["maxint"], rec["smallidx"], rec["nbytes"] = precision, mn, mx, smallidx, nbytes
                if (sizeint[0] | sizeint[1] | sizeint[2]) > 0xffffff:
                    rec["bitsize"], rec["bitsizeint"] = 0, [_sizeofint(s) for s in sizeint]
                else:
                    rec["bitsize"] = _sizeofints(sizeint)
                data = pos + _HEADER + 40
                nxt = data + ((nbytes + 3) & ~3)
            if nxt > size:
                break                                   # a truncated last frame is dropped, like VMD does
            rec["offset"] = data                        # file offset for now; rebased per read
            recs.append(rec)
            pos = nxt
        self._index = np.array(recs, META_DTYPE) if recs else np.zeros(0, META_DTYPE)
        self.n_frames = len(self._index)
        if self.n_atoms is None:
            self.n_atoms = 0

    # ------------------------------------------------------------------------------------------------
    CHUNK_BYTES = 48 << 20           # compressed bytes per pipeline stage (file read of chunk k+1 under the H2D copy of chunk k)

    def read_device(self, beg: int = 0, end: int | None = None, stride: int = 1):
        """Frames beg, beg+stride, ... <= end (inclusive) as a CUDA float32 [F][N][3] tensor, Angstrom."""
        import torch
        if not torch.cuda.is_available():
            raise _lib.ChacraLibraryError("XTC input is decoded on the GPU (sm_100a); there is no CPU fallback")
        end = self.n_frames - 1 if end is None else min(end, self.n_frames - 1)
        idx = np.arange(beg, end + 1, max(1, stride))
        dev = torch.device("cuda", self.device)
        out = torch.empty((len(idx), self.n_atoms, 3), dtype=torch.float32, device=dev)
        if len(idx) == 0:
            return out
        t0 = time.perf_counter()
        sizes = (self._index["nbytes"][idx].astype(np.int64) + 3) & ~3
        cuts = [0]
        acc = 0
        for k, n in enumerate(sizes.tolist()):         # chunk boundaries by compressed size
            if acc + n > self.CHUNK_BYTES and acc > 0:
                cuts.append(k)
                acc = 0
            acc += n
        cuts.append(len(idx))
        # the file is read in chunks by a helper thread while the previous chunk crosses PCIe; the decode is ONE launch
        # over all frames (a frame's bit stream is serial, so a launch takes as long as its slowest thread: more frames
        # per launch cost nothing until every SM sub-partition has a warp)
        d_blob = torch.empty(int(sizes.sum()) + 64 * len(idx) * 2 + 4096, dtype=torch.uint8, device=dev) if len(cuts) > 2 else None
        metas, off, t_host = [], 0, 0.0
        if len(cuts) == 2:
            meta, used = self._stage(idx, 0)
            d_blob = self._stage_bufs[0][:used + _SLACK].to(dev, non_blocking=True)
            metas.append(meta)
            t_host = time.perf_counter() - t0
        else:
            from concurrent.futures import ThreadPoolExecutor
            done = [torch.cuda.Event(), torch.cuda.Event()]
            with ThreadPoolExecutor(1) as reader:
                pending = reader.submit(self._stage, idx[cuts[0]:cuts[1]], 0)
                for c in range(len(cuts) - 1):
                    t1 = time.perf_counter()
                    meta, used = pending.result()
                    t_host += time.perf_counter() - t1          # time the device side had to wait for the file
                    if off + used + _SLACK > d_blob.numel():         # spans carry the frame headers too; grow if the estimate was short
                        bigger = torch.empty(off + used + _SLACK + (d_blob.numel() >> 2), dtype=torch.uint8, device=dev)
                        bigger[:off].copy_(d_blob[:off])
                        d_blob = bigger
                    d_blob[off:off + used + _SLACK].copy_(self._stage_bufs[c & 1][:used + _SLACK], non_blocking=True)
                    done[c & 1].record()
                    meta["offset"] += off
                    metas.append(meta)
                    off += used
                    if c + 2 < len(cuts):
                        if c >= 1:
                            done[(c + 1) & 1].synchronize()     # the copy out of that pinned buffer has finished
                        pending = reader.submit(self._stage, idx[cuts[c + 1]:cuts[c + 2]], (c + 1) & 1)
        t1 = time.perf_counter()
        meta = np.concatenate(metas)
        d_meta = torch.from_numpy(meta.view(np.uint8).reshape(-1)).to(dev)
        lib = _lib.load()
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        # workspace of the two-pass decoder (group records), kept between reads: all frames in one batch when 2 GB allow it
        need = min(int(lib.chc_xtc_work_bytes(len(meta), self.n_atoms)), 2 << 30)
        work = self.__dict__.get("_work")
        if work is None or work.numel() < need or work.device != dev:
            work = self._work = torch.empty(need, dtype=torch.uint8, device=dev)
        rc = lib.chc_xtc_decode_ws(d_blob.data_ptr(), d_meta.data_ptr(), len(meta), self.n_atoms, out.data_ptr(), 0, work.data_ptr(),
                                   work.numel(), self.device, stream)
        if rc != 0:
            raise ValueError("%s: XTC bit stream inconsistent with %d atoms (chc_xtc_decode returned %d)" % (self.path, self.n_atoms, rc))
        self.last_timing = (t_host, time.perf_counter() - t1)   # (waiting for file -> pinned, rest of H2D + decode) seconds
        return out

    def _stage(self, idx, which: int):
        """compressed streams of frames idx -> pinned buffer `which`; returns (frame table with buffer offsets, bytes used)"""
        meta = self._index[idx].copy()
        sizes = (meta["nbytes"].astype(np.int64) + 3) & ~3
        lo, hi = int(meta["offset"][0]), int(meta["offset"][-1] + sizes[-1])
        if hi - lo <= 1.25 * sizes.sum() + 4096:        # (nearly) consecutive frames: one span, headers and all
            host = self._pinned(hi - lo + _SLACK, which)
            self._pread(memoryview(host.numpy())[:hi - lo], lo)        # straight from the file into pinned memory
            meta["offset"] -= lo                        # file offsets are multiples of 4 (XDR), so these are too
            used = hi - lo
        else:
            starts = np.concatenate([[0], np.cumsum(sizes)[:-1]])
            used = int(sizes.sum())
            host = self._pinned(used + _SLACK, which)
            view = host.numpy()
            for src, n, dst in zip(meta["offset"], sizes, starts):
                view[dst:dst + n] = self._mm[src:src + n]
            meta["offset"] = starts
        host.numpy()[used:used + _SLACK] = 0
        return meta, used

    def read(self, beg: int = 0, end: int | None = None, stride: int = 1) -> np.ndarray:
        return self.read_device(beg, end, stride).cpu().numpy()

    def compressed_bytes(self) -> int:
        return int(self._index["nbytes"].sum())
