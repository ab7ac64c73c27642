"""Contact-list TSV writer: the text format of ``contact_consumer``
(``external/getcontacts/contact_calc/compute_contacts.py:306-359``).

Input is the binary record array drained from the device (``engine.CONTACT_DTYPE``); output lines are
``frame <TAB> itype <TAB> atom_1 <TAB> atom_2 [<TAB> water_1 [<TAB> water_2]] [<TAB> distance]`` with
* atom labels ``chain:resname:resid+icode:name`` (``atom.py:40-48`` minus the VMD index, ``:338-341``),
* atom_1/atom_2 ordered as Python strings and the two waters of a ``wb2``/``lwb2`` line swapped along with them
  (``:343-348``),
* identical lines written once per 100-frame fragment (``:350-359``),
* frames ascending; inside a frame the interaction-type blocks follow the reference's compute order
  sb, pc, ps, ts, vdw, hb*, wb*, hp (``:85-101``) and, inside a block, ascending atom indices.  The
  reference's order inside a block is VMD's pair order times CPython's set order, i.e. undefined; this
  writer picks a deterministic one.
"""
from __future__ import annotations

import os

import numpy as np

from . import _lib
from .engine import CONTACT_DTYPE, ITYPE_NAMES

# rank of every record itype inside a frame (compute order of compute_contacts.py:85-101; hbonds.py emits the
# hydrogen bonds first and the water bridges after them)
_BLOCK = {"sb": 0, "sblp": 0, "sbpl": 0, "sbll": 0, "pc": 1, "ps": 2, "ts": 3, "vdw": 4,
          "hbbb": 5, "hbsb": 5, "hbss": 5, "hbll": 5, "hblb": 5, "hbls": 5, "wb": 6, "wb2": 6, "lwb": 6, "lwb2": 6,
          "hp": 7, "hplp": 7, "hpll": 7}
BLOCK_RANK = np.array([_BLOCK[n] for n in ITYPE_NAMES], np.int64)


def header_lines(total_frames: int, beg: int, end: int, stride: int, itypes, distout: bool) -> str:
    """compute_contacts.py:309-316."""
    cols = "# Columns: frame, interaction_type, atom_1, atom_2[, atom_3[, atom_4]]"
    return "# total_frames:%d beg:%d end:%d stride:%d interaction_types:%s\n%s%s\n" % (
        total_frames, beg, end, stride, ",".join(itypes), cols, ", distance_1-2" if distout else "")


class ContactWriter:
    """Orders, de-duplicates and formats drained contact records through the native writer of libchacra_b200
    (``chc_write_contacts_tsv``, csrc/tsv_writer.h: host threads, straight to the file descriptor)."""

    def __init__(self, fd, atom_labels, beg: int, stride: int, distout: bool, fragment_frames: int = 100, threads: int | None = None):
        self.fd = fd
        labels = [str(x) for x in atom_labels]
        # atoms with identical label strings share a rank, so that line identity is decided on integers
        uniq, inv = np.unique(np.asarray(labels, dtype=str), return_inverse=True)
        self.label_rank = np.ascontiguousarray(inv, dtype=np.int32)     # order of ranks == Python string order of the labels
        self.ranks_unique = int(len(uniq) == len(labels))
        enc = [x.encode() for x in labels]
        self.label_off = np.zeros(len(enc) + 1, np.int64)
        np.cumsum([len(x) for x in enc], out=self.label_off[1:])
        self.label_text = b"".join(enc)
        names = [x.encode() for x in ITYPE_NAMES]
        self.itype_off = np.zeros(len(names) + 1, np.int32)
        np.cumsum([len(x) for x in names], out=self.itype_off[1:])
        self.itype_text = b"".join(names)
        self.block_rank = np.ascontiguousarray(BLOCK_RANK, dtype=np.int32)
        self.beg, self.stride, self.distout, self.frag = int(beg), int(stride), bool(distout), int(fragment_frames)
        self.threads = int(threads or min(32, os.cpu_count() or 1))
        self.lines_written = 0
        self.lib = _lib.load()

    def write(self, rec: np.ndarray) -> int:
        """Write the records of a set of WHOLE frames (any order; the array is reordered in place).  Returns the number
        of lines."""
        if len(rec) == 0:
            return 0
        if not (isinstance(rec, np.ndarray) and rec.dtype == CONTACT_DTYPE and rec.flags.c_contiguous and rec.flags.writeable):
            rec = np.array(rec, dtype=CONTACT_DTYPE, order="C")
        try:
            fileno = self.fd.fileno()
        except (AttributeError, OSError):
            fileno = None
        if fileno is not None:
            self.fd.flush()
            n = self._native(fileno, rec)
        else:                                            # in-memory text streams: through a temporary file
            import tempfile
            with tempfile.TemporaryFile() as tmp:
                n = self._native(tmp.fileno(), rec)
                tmp.seek(0)
                self.fd.write(tmp.read().decode())
        self.lines_written += n
        return n

    def _native(self, fileno: int, rec: np.ndarray) -> int:
        n = self.lib.chc_write_contacts_tsv(fileno, rec.ctypes.data, len(rec), self.label_text, self.label_off.ctypes.data,
                                            self.label_rank.ctypes.data, len(self.label_rank), self.ranks_unique, self.itype_text,
                                            self.itype_off.ctypes.data, self.block_rank.ctypes.data, len(self.block_rank),
                                            self.beg, self.stride, int(self.distout), self.threads)
        if n < 0:
            raise RuntimeError("chc_write_contacts_tsv failed (%d)" % n)
        return int(n)
