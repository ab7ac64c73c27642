"""TEST INFRASTRUCTURE ONLY -- line-by-line Python restatement of the contact-list TSV writer (contact_consumer,
gc/contact_calc/compute_contacts.py:306-359), kept to check the native writer (csrc/tsv_writer.h).  Nothing on the
product path imports this file.  Order inside a frame: interaction block, atom indices, waters, itype, distance bits
(the reference's order inside a block is undefined)."""
import numpy as np

from chacra_b200.engine import ITYPE_NAMES
from chacra_b200.tsv import BLOCK_RANK


class ContactWriterPy:
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
        order = np.lexsort((rec["distance"].view(np.uint32), it, w2, w1, a2, a1, BLOCK_RANK[it], frame))
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
