"""All thermodynamic states of an HREMD run on every GPU of the box: the multi-GPU product path.

The reference loops the states one after another (``chacra/scripts/process_hremd_output.py:134-175``) and, inside one
state, hands 100-frame fragments to ``--cores`` worker processes (``compute_contacts.py:250-293``).  Here the work
items are (state, fragment range): every rank (one process per GPU, ``torch.distributed``) takes a fragment-aligned
slice of EVERY state (``sharding.state_slices``), reads only those frames from the state trajectory, accumulates its
own ``int32 [states][pairs]`` table on its GPU, and the tables are merged with one all-reduce
(``sharding.merge_count_tables``; NCCL over NVLink).  Rank 0 writes ``freqs_state_{i}.tsv`` (+ integer counts, + the
near-cutoff list) in the formats ChACRA's consumers read.  With one process it is the single-GPU counts-only path.
"""
from __future__ import annotations

import math
import os
import sys

import numpy as np

from .engine import COMPUTE_ORDER, NEAR_KINDS, TRAJ_FRAG_SIZE, ChacraError, ContactEngine
from .sharding import merge_count_tables, state_slices


def _dist():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist, dist.get_rank(), dist.get_world_size()
    return None, 0, 1


def plan_states(frames_per_state, beg, end, stride, world, rank):
    """The reference's frame selection for every state (``compute_contacts.py:247-254``: clamped beg / end, fragments
    never clamped to --end) and this rank's share.  Returns (selected frames per state [(beg, n_proc, total_frames)],
    work [(state, first selected index, count)])."""
    sel = []
    for n in frames_per_state:
        b = max(min(beg, n - 1), 0)
        e = min(max(sys.maxsize if end is None else end, b), n - 1)
        nfrag = math.ceil((e - b + 1) / (TRAJ_FRAG_SIZE * stride))
        last = min(b + nfrag * TRAJ_FRAG_SIZE * stride - 1, n - 1)
        sel.append((b, len(range(b, last + 1, stride)), math.ceil((e - b + 1) / stride)))
    return sel, state_slices([s[1] for s in sel], world, rank)


def run_states(topology, trajectories, itypes=("all",), geometry=None, sele1="protein", sele2=None, solv="", lipid="", ligand="",
               beg=0, end=None, stride=1, device=None, track_near=False, chunk_bytes=1 << 30, compact=None, info=None):
    """Count table ``int32 [states][pairs]`` of all ``trajectories`` (one file per state), merged over the ranks of the
    initialised process group.  Returns (topo, counts tensor on this rank's GPU, [(beg, n_proc, total_frames)] per state,
    near records of THIS rank as a structured array or None)."""
    import torch
    from .topology import compile_topology, read_pdb
    from .trajectory import open_trajectory
    dist, rank, world = _dist()
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0")) if world > 1 else 0
    torch.cuda.set_device(device)
    geom = dict(geometry or {})
    atoms = read_pdb(topology)
    trjs = [open_trajectory(t) for t in trajectories]
    for t, path in zip(trjs, trajectories):
        if hasattr(t, "device"):
            t.device = int(device)
        if t.n_atoms != atoms.n:
            raise ChacraError("topology has %d atoms but %s has %d" % (atoms.n, path, t.n_atoms))
    xyz0 = trjs[0].read(0, 0)[0]                      # selections / disulfides / ligand ions from frame 0 of state 0
    topo = compile_topology(atoms, xyz0, sele1, sele1 if sele2 is None else sele2, solv or "", lipid or "", ligand or "",
                            vdw_res_diff=int(geom.get("VDW_RES_DIFF", 2)), hbond_res_diff=int(geom.get("HBOND_RES_DIFF", 1)))
    if not topo.sele1.any() or not topo.sele2.any():
        raise ChacraError("The --sele%s selection is empty!" % ("" if not topo.sele1.any() else "2"))
    kernel_itypes = list(COMPUTE_ORDER) if "all" in itypes else [t for t in COMPUTE_ORDER if t in itypes]
    sel, work = plan_states([t.n_frames for t in trjs], beg, end, max(1, stride), world, rank)
    stride = max(1, stride)
    eng = ContactEngine(topo, kernel_itypes, geom, n_states=len(trjs), emit_contacts=False, track_near=track_near,
                        near_capacity=1 << 20, device=device)
    per = max(TRAJ_FRAG_SIZE, int(chunk_bytes) // (12 * atoms.n) // TRAJ_FRAG_SIZE * TRAJ_FRAG_SIZE)
    near_parts = []
    for state, first, count in work:
        b = sel[state][0]
        reader = getattr(trjs[state], "read_device", trjs[state].read)
        done = 0
        while done < count:
            n = min(per, count - done)
            f0 = b + (first + done) * stride
            frames = reader(f0, f0 + (n - 1) * stride, stride)
            eng.submit(frames, state, first + done)
            eng.synchronize()
            if track_near:
                rec, total = eng.drain_near()
                if total > len(rec):
                    raise ChacraError("near-cutoff buffer overflow (%d observations)" % total)
                rec = rec.copy()
                rec["reserved"] = state
                near_parts.append(rec)
            done += len(frames)
    counts = eng.counts_tensor
    merge_count_tables(counts, compact=compact, info=info)
    near = np.concatenate(near_parts) if near_parts else None
    run_states.last_engine = eng                       # keeps the count tensor's context alive for the caller
    return topo, counts, sel, near


def write_state_tables(out_dir, topo, counts, sel, counts_too=False, prefix="freqs_state_"):
    """``freqs_state_{i}.tsv`` (frequency = frames / header total_frames, get_contact_frequencies.py:129-132) from the merged table."""
    from .frequencies import ALL_FREQ_ITYPES, counts_to_table, write_count_table, write_frequency_table
    os.makedirs(out_dir, exist_ok=True)
    host = counts.cpu().numpy()
    for s, (_, _, total) in enumerate(sel):
        table = counts_to_table(topo.residue_label, host[s])
        with open(os.path.join(out_dir, "%s%d.tsv" % (prefix, s)), "w") as fd:
            write_frequency_table(fd, total, ALL_FREQ_ITYPES, {k: (c, float(c) / total) for k, c in table.items()})
        if counts_too:
            with open(os.path.join(out_dir, "counts_state_%d.tsv" % s), "w") as fd:
                write_count_table(fd, total, ALL_FREQ_ITYPES, table)


def gather_near(near, labels, sel, stride, path):
    """Every rank's near-cutoff records -> one file written by rank 0 (state, frame, criterion, atom_1, atom_2, delta)."""
    dist, rank, world = _dist()
    parts = [near]
    if world > 1:
        parts = [None] * world
        dist.all_gather_object(parts, near)
    if rank != 0:
        return
    rows = set()
    for rec in parts:
        if rec is None:
            continue
        for r in rec:
            l1, l2 = labels[r["atom1"]], labels[r["atom2"]]
            s = int(r["reserved"])
            rows.add((s, sel[s][0] + int(r["frame"]) * stride, int(r["kind"]), min(l1, l2), max(l1, l2), float(r["delta"])))
    with open(path, "w") as fd:
        fd.write("# near-cutoff observations: |value - cutoff| < 1e-4 A (distances) or 1e-3 degree (angles)\n"
                 "# Columns: state, frame, criterion, atom_1, atom_2, value_minus_cutoff\n")
        for s, f, kind, l1, l2, delta in sorted(rows):
            fd.write("%d\t%d\t%s\t%s\t%s\t%.3e\n" % (s, f, NEAR_KINDS[kind], l1, l2, delta))
