"""``compute_contacts``: the reference's Python entry point for the path, same signature and file output
(``external/getcontacts/contact_calc/compute_contacts.py:187-293``), with the frames processed by the CUDA
kernels behind ``libchacra_b200.so`` instead of VMD worker processes.

What stays on the host: argument handling, the one-off topology compilation (the reference's setup block
:226-238), reading trajectory fragments, and turning device records into text (:306-359).
"""
from __future__ import annotations

import math
import os
import time
import sys

import numpy as np

from .engine import ContactEngine, ChacraError, COMPUTE_ORDER, NEAR_KINDS, TRAJ_FRAG_SIZE
from .topology import read_pdb, compile_topology
from .trajectory import open_trajectory
from .tsv import ContactWriter, header_lines

# frames handed to the device per round trip: whole fragments, bounded by the record buffer
_RECORD_CAPACITY = 1 << 23


def compute_contacts(top, traj, output, itypes, geom_criterion_values, cores, beg, end, stride, distout,
                     ligand_sele, solv_sele, lipid_sele, sele1, sele2, device=0, frequency_output=None,
                     count_output=None, near_output=None, write_contacts=True):
    """Compute the contact list of ``traj`` and write it to ``output`` (contact-list TSV).

    ``cores`` is accepted for interface compatibility (the reference uses it as the number of VMD worker
    processes, :267); one GPU replaces them.  Extensions: ``frequency_output`` / ``count_output`` write the
    residue-pair frequency table / integer frame counts straight from the device count table (intended counts,
    SURVEY A.12 level L2); ``near_output`` lists every pair within 1e-4 A / 1e-3 degree of a cutoff;
    ``write_contacts=False`` skips the contact list altogether (no records leave the device, no text is written).
    """
    geom = dict(geom_criterion_values or {})
    itypes = list(itypes)
    sele2 = sele1 if sele2 is None else sele2
    atoms = read_pdb(top)
    trj = open_trajectory(traj)
    if hasattr(trj, "device"):
        trj.device = int(device)                          # readers that decode on the GPU (XTC)
    if trj.n_atoms != atoms.n:
        raise ChacraError("topology has %d atoms but the trajectory has %d" % (atoms.n, trj.n_atoms))
    sim_length = trj.n_frames
    xyz0 = trj.read(0, 0)[0]
    topo = compile_topology(atoms, xyz0, sele1, sele2, solv_sele or "", lipid_sele or "", ligand_sele or "",
                            vdw_res_diff=int(geom.get("VDW_RES_DIFF", 2)), hbond_res_diff=int(geom.get("HBOND_RES_DIFF", 1)))
    if not topo.sele1.any():                              # compute_contacts.py:240-245
        print("The --sele selection is empty!")
        sys.exit(1)
    if not topo.sele2.any():
        print("The --sele2 selection is empty!")
        sys.exit(1)
    end = sys.maxsize if end is None else end
    beg = max(min(beg, sim_length - 1), 0)                # :247-249
    end = min(max(end, beg), sim_length - 1)
    stride = max(1, stride)
    total_frames = math.ceil((end - beg + 1) / stride)    # what the header says (:309)
    num_fragments = math.ceil((end - beg + 1) / (TRAJ_FRAG_SIZE * stride))
    # The reference's fragments are never clamped to --end (:253-254): the last one runs to its own end or to the end of
    # the trajectory, so frames past --end are computed and written too.  Reproduced (the oracle does the same).
    last = min(beg + num_fragments * TRAJ_FRAG_SIZE * stride - 1, sim_length - 1)
    n_proc = len(range(beg, last + 1, stride))
    print("Processing %s (frame %d to %d with stride of %d) as %d fragments\n" % (traj, beg, end, stride, num_fragments))

    kernel_itypes = [t for t in COMPUTE_ORDER if t in itypes]
    track_near = near_output is not None
    eng = ContactEngine(topo, kernel_itypes, geom, n_states=1, emit_contacts=bool(write_contacts),
                        contact_capacity=_RECORD_CAPACITY, track_near=track_near, near_capacity=1 << 20, device=device)
    labels = atoms.atom_labels()
    near_fd = None
    if track_near:
        near_fd = open(near_output, "w")
        near_fd.write("# near-cutoff observations: |value - cutoff| < 1e-4 A (distances) or 1e-3 degree (angles); these pairs sit on the\n"
                      "# float32 boundary of a reference criterion and are listed apart from the contact list\n"
                      "# Columns: frame, criterion, atom_1, atom_2, value_minus_cutoff\n")
    near_total = 0

    def drain_near():
        nonlocal near_total
        rec, total = eng.drain_near()
        if total > len(rec):
            raise ChacraError("near-cutoff buffer overflow (%d observations)" % total)
        near_total += len(rec)
        rows = []
        for r in rec:                                    # atom_1 / atom_2 ordered like a contact line (compute_contacts.py:343-348)
            l1, l2 = labels[r["atom1"]], labels[r["atom2"]]
            rows.append((int(r["frame"]), int(r["kind"]), min(l1, l2), max(l1, l2), float(r["delta"])))
        for f, kind, l1, l2, delta in sorted(set(rows)):
            near_fd.write("%d\t%s\t%s\t%s\t%.3e\n" % (beg + f * stride, NEAR_KINDS[kind], l1, l2, delta))

    reader = getattr(trj, "read_device", trj.read)       # DCD and XTC frames are unpacked on the device and stay there
    stage = [0.0, 0.0, 0.0]                              # seconds: trajectory input, contacts + drain, TSV text
    done = 0                                             # selected frames processed so far
    if write_contacts:
        with open(output, "w") as fd:
            fd.write(header_lines(total_frames, beg, end, stride, itypes, distout))
            writer = ContactWriter(fd, labels, beg, stride, distout, TRAJ_FRAG_SIZE)
            per_frame = None                             # measured records per frame
            while done < n_proc:
                if per_frame is None:
                    n = min(TRAJ_FRAG_SIZE, n_proc - done)
                else:
                    room = int(0.5 * _RECORD_CAPACITY / max(per_frame, 1.0))
                    n = min(max(TRAJ_FRAG_SIZE, room // TRAJ_FRAG_SIZE * TRAJ_FRAG_SIZE), n_proc - done)
                first = beg + done * stride
                t0 = time.perf_counter()
                frames = reader(first, min(first + (n - 1) * stride, last), stride)
                t1 = time.perf_counter()
                eng.submit(frames, 0, done)
                eng.synchronize()
                rec = eng.drain_contacts(copy=False)      # reordered in place by the writer, gone before the next drain
                t2 = time.perf_counter()
                writer.write(rec)
                if track_near:
                    drain_near()
                t3 = time.perf_counter()
                stage[0] += t1 - t0; stage[1] += t2 - t1; stage[2] += t3 - t2
                per_frame = max(per_frame or 0.0, len(rec) / float(len(frames)))
                done += len(frames)
    else:
        # counts only: whole fragments in chunks of about 1 GiB of coordinates, nothing but the count table comes back
        per = max(TRAJ_FRAG_SIZE, (1 << 30) // (12 * atoms.n) // TRAJ_FRAG_SIZE * TRAJ_FRAG_SIZE)
        while done < n_proc:
            n = min(per, n_proc - done)
            first = beg + done * stride
            t0 = time.perf_counter()
            frames = reader(first, min(first + (n - 1) * stride, last), stride)
            t1 = time.perf_counter()
            eng.submit(frames, 0, done)
            eng.synchronize()
            if track_near:
                drain_near()
            stage[0] += t1 - t0; stage[1] += time.perf_counter() - t1
            done += len(frames)
    if near_fd is not None:
        near_fd.close()
    if os.environ.get("CHACRA_TIMING"):
        print("stages: trajectory input %.3f s, contacts + drain %.3f s, TSV text %.3f s" % tuple(stage), file=sys.stderr)
    if frequency_output or count_output:
        from .frequencies import ALL_FREQ_ITYPES, counts_to_table, write_count_table, write_frequency_table
        table = counts_to_table(topo.residue_label, eng.counts(0))
        if frequency_output:
            with open(frequency_output, "w") as fd:
                write_frequency_table(fd, total_frames, ALL_FREQ_ITYPES,
                                      {k: (c, float(c) / total_frames) for k, c in table.items()})
        if count_output:
            with open(count_output, "w") as fd:
                write_count_table(fd, total_frames, ALL_FREQ_ITYPES, table)
    eng.close()
    return None
