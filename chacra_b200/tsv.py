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

import numpy as np

from .engine import ITYPE_NAMES

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
    def __init__(self, fd, atom_labels, beg: int, stride: int, distout: bool, fragment_frames: int = 100):
        self.fd = fd
        self.labels = np.asarray(atom_labels, dtype=object)
        # atoms with identical label strings share a rank, so that line identity is decided on integers
        uniq, inv = np.unique(self.labels.astype(str), return_inverse=True)
        self.label_rank = inv.astype(np.int64)          # order of ranks == Python string order of the labels
        self.beg, self.stride, self.distout, self.frag = int(beg), int(stride), bool(distout), int(fragment_frames)
        self.lines_written = 0

    def write(self, rec: np.ndarray) -> int:
        """Write the records of a set of WHOLE fragments (any order).  Returns the number of lines."""
        if len(rec) == 0:
            return 0
        a1, a2 = rec["atom1"].astype(np.int64), rec["atom2"].astype(np.int64)
        w1, w2 = rec["water1"].astype(np.int64), rec["water2"].astype(np.int64)
        swap = self.label_rank[a2] < self.label_rank[a1]                 # "if interaction[3] < interaction[2]"
        a1, a2 = np.where(swap, a2, a1), np.where(swap, a1, a2)
        two = w2 >= 0
        w1, w2 = np.where(swap & two, w2, w1), np.where(swap & two, w1, w2)
        frame = rec["frame"].astype(np.int64)
        it = rec["itype"].astype(np.int64)
        order = np.lexsort((w2, w1, a2, a1, BLOCK_RANK[it], frame))
        frame, it, a1, a2, w1, w2 = frame[order], it[order], a1[order], a2[order], w1[order], w2[order]
        dist = rec["distance"][order]
        dtxt = np.array(["%.3f" % d for d in dist.tolist()], dtype=object) if self.distout else None
        # per-fragment de-duplication of identical text lines (label ranks + the printed distance decide identity)
        lr1, lr2 = self.label_rank[a1], self.label_rank[a2]
        lw1 = np.where(w1 >= 0, self.label_rank[np.maximum(w1, 0)], -1)
        lw2 = np.where(w2 >= 0, self.label_rank[np.maximum(w2, 0)], -1)
        seen, out = set(), []
        lab = self.labels
        names = ITYPE_NAMES
        frag_of = frame // self.frag
        cur_frag = -1
        fr_l, it_l, a1_l, a2_l, w1_l, w2_l = frame.tolist(), it.tolist(), a1.tolist(), a2.tolist(), w1.tolist(), w2.tolist()
        k1, k2, k3, k4 = lr1.tolist(), lr2.tolist(), lw1.tolist(), lw2.tolist()
        fg = frag_of.tolist()
        for n in range(len(fr_l)):
            if fg[n] != cur_frag:
                cur_frag, seen = fg[n], set()
            key = (fr_l[n], it_l[n], k1[n], k2[n], k3[n], k4[n], dtxt[n] if dtxt is not None else None)
            if key in seen:
                continue
            seen.add(key)
            cols = [str(self.beg + fr_l[n] * self.stride), names[it_l[n]], lab[a1_l[n]], lab[a2_l[n]]]
            if w1_l[n] >= 0:
                cols.append(lab[w1_l[n]])
                if w2_l[n] >= 0:
                    cols.append(lab[w2_l[n]])
            if dtxt is not None:
                cols.append(dtxt[n])
            out.append("\t".join(cols))
        if out:
            self.fd.write("\n".join(out))
            self.fd.write("\n")
        self.lines_written += len(out)
        return len(out)
