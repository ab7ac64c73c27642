"""Residue-pair contact frequencies: the reduction of ``get_contact_frequencies.py``
(``external/getcontacts/get_contact_frequencies.py:58-134``) and of
``contact_calc/transformations.py`` (``res_contacts_xl`` :70-150, ``res_contacts`` :262-304,
``gen_counts`` :446-470, ``gen_frequencies`` :357-391).

Two sources, one table format:

* ``counts_to_table``      the device count tables (int32 [pairs], accumulated by the pair kernel) -> the INTENDED
                           counts: #frames in which any counted interaction joins two residues;
* ``shipped_counts``       a contact-list TSV -> the counts the reference script ACTUALLY produces.  The shipped
                           reader starts with ``current_frame = 0``, throws away the line on which the frame number
                           changes and never flushes the last frame (transformations.py:114-148); a drop-in has to
                           reproduce that to emit the same table from the same file;
* ``intended_counts_tsv``  the same file without those two quirks (equals the device table).
"""
from __future__ import annotations

from collections import defaultdict

import numpy as np

# get_contact_frequencies.py:107-109 -- note that hp/hplp/hpll, sblp/sbpl/sbll and hbll are NOT in the list
ALL_FREQ_ITYPES = ["sb", "pc", "ps", "ts", "vdw", "hb", "lhb", "hbbb", "hbsb", "hbss", "wb", "wb2", "hbls", "hblb",
                   "lwb", "lwb2", "br"]


class ParseError(ValueError):
    pass


def _residue_of(label: str) -> str:
    """First three colon fields of an atom label (transformations.py:293-294)."""
    return ":".join(label.split(":")[0:3])


def _tokens(line: str):
    tok = line.split("\t")
    try:
        frame = int(tok[0])
    except ValueError:
        raise ParseError("First column isn't a integer")
    if not 4 <= len(tok) <= 7:
        raise ParseError("Invalid number of tokens")
    return frame, tok


def _header_frames(line: str, current: int) -> int:
    """``# total_frames:20 beg:0 ...`` -> 20 (second blank-separated token, after the colon; :121-123)."""
    if "total_frames" in line:
        second = line.split(" ")[1]
        return int(second[second.find(":") + 1:])
    return current


def shipped_counts(lines, itypes=None, as_sets=False):
    """(total_frames, {(res1, res2): frame_count}) exactly as the reference script computes them
    (``as_sets``: the sets of frame numbers instead of their sizes, for label files)."""
    keep = None if itypes is None else set(itypes)
    per_pair = defaultdict(set)
    total_frames, open_frame, pending = 0, 0, []

    def flush():
        for fr, r1, r2 in pending:
            per_pair[(r1, r2)].add(fr)

    for raw in lines:
        line = raw.strip()
        total_frames = _header_frames(line, total_frames)
        if not line or line[0] == "#":
            continue
        frame, tok = _tokens(line)
        if frame == open_frame:
            if keep is None or tok[1] in keep:
                r1, r2 = _residue_of(tok[2]), _residue_of(tok[3])
                pending.append((frame, r2, r1) if r2 < r1 else (frame, r1, r2))
        else:                       # the frame number changed: this very line is lost
            flush()
            pending, open_frame = [], frame
    # the reference never flushes what is still pending at end of file
    return total_frames, (dict(per_pair) if as_sets else {k: len(v) for k, v in per_pair.items()})


def intended_counts_tsv(lines, itypes=None, as_sets=False):
    """Same reduction without the two shipped quirks (every counted line of every frame takes part)."""
    keep = None if itypes is None else set(itypes)
    per_pair = defaultdict(set)
    total_frames = 0
    for raw in lines:
        line = raw.strip()
        total_frames = _header_frames(line, total_frames)
        if not line or line[0] == "#":
            continue
        frame, tok = _tokens(line)
        if keep is None or tok[1] in keep:
            r1, r2 = _residue_of(tok[2]), _residue_of(tok[3])
            per_pair[(r2, r1) if r2 < r1 else (r1, r2)].add(frame)
    return total_frames, (dict(per_pair) if as_sets else {k: len(v) for k, v in per_pair.items()})


def native_counts(path: str, itypes=None, shipped: bool = True, threads: int | None = None):
    """(total_frames, {(res1, res2): frame_count}) of a contact-list TSV FILE through the library's threaded reader
    (``chc_count_tsv``, csrc/tsv_counts.h) -- the same table as ``shipped_counts`` / ``intended_counts_tsv``.  Returns
    None when the file's frame numbers are not strictly increasing (then only the line-by-line readers apply)."""
    import ctypes as C
    import os
    from . import _lib
    lib = _lib.load()
    handle = C.c_void_p()
    csv = ",".join(itypes).encode() if itypes is not None else None
    rc = lib.chc_count_tsv(os.fsencode(path), csv, int(bool(shipped)), int(threads or min(32, os.cpu_count() or 1)), C.byref(handle))
    try:
        if rc == -4:                                    # CHC_E_ORDER
            return None
        if rc == -1:
            raise ParseError(lib.chc_tsv_counts_error(handle).decode() or "malformed contact file")
        if rc != 0:
            raise OSError("cannot read %s (%d)" % (path, rc))
        nres, npair, nbytes, total = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        lib.chc_tsv_counts_sizes(handle, C.byref(nres), C.byref(npair), C.byref(nbytes), C.byref(total))
        text = np.zeros(max(nbytes.value, 1), np.uint8)
        off = np.zeros(nres.value + 1, np.int64)
        r1, r2 = np.zeros(max(npair.value, 1), np.int32), np.zeros(max(npair.value, 1), np.int32)
        cnt = np.zeros(max(npair.value, 1), np.int64)
        lib.chc_tsv_counts_get(handle, text.ctypes.data, off.ctypes.data, r1.ctypes.data, r2.ctypes.data, cnt.ctypes.data)
    finally:
        if handle:
            lib.chc_tsv_counts_free(handle)
    blob = text.tobytes()
    names = [blob[off[k]:off[k + 1]].decode() for k in range(nres.value)]
    n = npair.value
    return int(total.value), {(names[a], names[b]): int(c) for a, b, c in zip(r1[:n].tolist(), r2[:n].tolist(), cnt[:n].tolist())}


def counts_to_table(residue_label, counts: np.ndarray):
    """Device count table (upper triangle incl. diagonal, index = lo*R - lo*(lo-1)/2 + (hi-lo)) -> {(res1, res2): n}.
    ``residue_label`` is sorted as Python strings, so rank order is label order and res1 <= res2."""
    r = len(residue_label)
    idx = np.nonzero(counts)[0].astype(np.int64)
    # invert the triangular index: row lo starts at lo*r - lo*(lo-1)/2
    starts = np.arange(r, dtype=np.int64) * r - (np.arange(r, dtype=np.int64) * (np.arange(r, dtype=np.int64) - 1)) // 2
    lo = np.searchsorted(starts, idx, side="right") - 1
    hi = lo + (idx - starts[lo])
    return {(residue_label[a], residue_label[b]): int(c) for a, b, c in zip(lo.tolist(), hi.tolist(), counts[idx].tolist())}


def merge_counts(per_input):
    """``gen_frequencies``: [(frames, {pair: count})] -> (total_frames, {pair: (count, frequency)})."""
    total = sum(f for f, _ in per_input)
    acc = defaultdict(int)
    for _, table in per_input:
        for k, c in table.items():
            acc[k] += c
    return total, {k: (c, float(c) / total) for k, c in acc.items()}


def relabel_counts(residue_frames, labels):
    """Label files (``relabel`` :394-443) work on (frame, res1, res2) triples; only used with --label_file."""
    out = defaultdict(set)
    for (r1, r2), frames in residue_frames.items():
        if r1 in labels and r2 in labels:
            a, b = labels[r1], labels[r2]
            if a == b:
                continue
            if a > b:
                a, b = b, a
            out[(a, b)] |= frames
    return out


def parse_residuelabels(fh):
    """Two or three tab-separated columns: residue id, [tree.]label, colour (transformations.py:201-259)."""
    labels, colors, used = {}, {}, set()
    for line in fh:
        line = line.strip()
        if not line:
            continue
        col = line.split("\t")
        path = col[1] if len(col) > 1 else col[0]
        leaf = path.split(".")[-1]
        if leaf in used:
            raise AssertionError("Flare label '%s' used twice in %s" % (leaf, getattr(fh, "name", "<labels>")))
        used.add(leaf)
        labels[col[0]] = path
        colors[col[0]] = col[2] if len(col) > 2 else "white"
    return labels, colors


def write_frequency_table(fd, total_frames: int, itypes, freqs, order=None):
    """get_contact_frequencies.py:129-132.  ``freqs``: {(res1, res2): (count, frequency)}.  The reference writes in
    dict order (not deterministic across runs); rows are sorted here unless ``order`` is given."""
    fd.write("#\ttotal_frames:%d\tinteraction_types:%s\n" % (total_frames, ",".join(itypes)))
    fd.write("#\tColumns:\tresidue_1,\tresidue_2\tframe_count\tcontact_frequency\n")
    keys = order if order is not None else sorted(freqs)
    for r1, r2 in keys:
        fd.write("\t".join([r1, r2, "%.3f" % freqs[(r1, r2)][1]]) + "\n")


def write_count_table(fd, total_frames: int, itypes, counts, order=None):
    """Integer sidecar of the frequency table: ``res1 <TAB> res2 <TAB> frame_count``.  The reference
    only keeps ``count / total_frames`` printed with three decimals, which loses the counts when runs are merged or
    frame-weighted afterwards (chacra/scripts/process_hremd_output.py:205-222); with the counts the merge is exact."""
    fd.write("#\ttotal_frames:%d\tinteraction_types:%s\n" % (total_frames, ",".join(itypes)))
    fd.write("#\tColumns:\tresidue_1,\tresidue_2\tframe_count\n")
    keys = order if order is not None else sorted(counts)
    for r1, r2 in keys:
        fd.write("%s\t%s\t%d\n" % (r1, r2, counts[(r1, r2)]))



def read_count_table(path: str):
    """(total_frames, {(res1, res2): count}) from a file written by ``write_count_table``."""
    total, table = 0, {}
    with open(path) as fh:
        for line in fh:
            if line.startswith("#"):
                if "total_frames:" in line:
                    total = int(line.split("total_frames:")[1].split()[0])
                continue
            tok = line.rstrip("\n").split("\t")
            if len(tok) == 3:
                table[(tok[0], tok[1])] = int(tok[2])
    return total, table


def merge_runs_exact(count_files, output_file, itypes=None):
    """Frequency table of several runs of one state from their integer sidecars: sum of counts over sum of frames.
    The reference combines runs from the three-decimal frequency files weighted by frame counts
    (chacra/scripts/process_hremd_output.py:205-222), which keeps a rounding error of up to 5e-4 per run; this is exact."""
    per_input = [read_count_table(p) for p in count_files]
    total, freqs = merge_counts(per_input)
    with open(output_file, "w") as fd:
        write_frequency_table(fd, total, itypes or ALL_FREQ_ITYPES, freqs)
    return total, {k: c for k, (c, _) in freqs.items()}
