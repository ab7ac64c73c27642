"""Host-side file staging shared by the trajectory readers that decode on the GPU (``xtc.py``, ``trajectory.py``):
bytes go from the page cache straight into pinned memory with parallel ``preadv`` calls, into two buffers that are kept
between reads so that the read of chunk k+1 can run under the H2D copy of chunk k."""
from __future__ import annotations

import os


class FileStager:
    """Mixin: needs ``self.path``."""

    def _pread(self, dst: memoryview, offset: int, piece: int = 8 << 20, workers: int = 16) -> None:
        """fill dst from the file at offset; large spans as parallel preads (the page-cache copy of one thread, about
        6 GB/s, is otherwise the slowest stage of the XTC path)"""
        fd = os.open(self.path, os.O_RDONLY)
        try:
            def one(beg):
                end = min(beg + piece, len(dst))
                while beg < end:
                    got = os.preadv(fd, [dst[beg:end]], offset + beg)
                    if got <= 0:
                        raise ValueError("%s: short read" % self.path)
                    beg += got
            starts = range(0, len(dst), piece)
            if len(starts) <= 1:
                for b in starts:
                    one(b)
            else:
                from concurrent.futures import ThreadPoolExecutor
                with ThreadPoolExecutor(min(workers, len(starts), os.cpu_count() or 1)) as pool:
                    list(pool.map(one, starts))
        finally:
            os.close(fd)

    def _pinned(self, nbytes: int, which: int = 0):
        """pinned staging buffers kept between reads (allocating pinned memory costs more than the decode)"""
        import torch
        bufs = self.__dict__.setdefault("_stage_bufs", [None, None])
        buf = bufs[which]
        if buf is None or buf.numel() < nbytes:
            if buf is not None:
                torch.cuda.synchronize()                # an earlier asynchronous copy may still read the old buffer
            buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, pin_memory=True)
            bufs[which] = buf
        return buf
