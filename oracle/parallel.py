"""Test infrastructure: the numpy oracle (getcontacts_oracle.py) over many frames on several host processes.

Frames are independent except for the per-fragment state (hydrophobic set and vdW soft cutoff come from the first
frame of each 100-frame fragment), so every worker gets whole frame indices plus the fragment-opening
frame it needs.  The result has the layout of ``getcontacts_oracle.run`` (beg 0, stride 1).
"""
from __future__ import annotations

import math
import multiprocessing as mp
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import getcontacts_oracle as o  # noqa: E402

_G = {}


def _init(top_path, frames, itypes, geom, sel, collect_near):
    top = o.Topology(top_path)
    S = o.Setup(top, sel.get("sele1", "protein"), sel.get("sele2"), sel.get("solv", ""), sel.get("lipid", ""),
                sel.get("ligand", ""), xyz0=frames[0])
    _G.update(top=top, S=S, frames=frames, itypes=itypes, geom=geom, near=collect_near)


def _work(idx):
    top, S, frames = _G["top"], _G["S"], _G["frames"]
    out, near_all = [], []
    frags = {}
    for f in idx:
        k = f // o.TRAJ_FRAG_SIZE
        if k not in frags:
            frags[k] = o.FragmentState(xyz0=np.ascontiguousarray(frames[k * o.TRAJ_FRAG_SIZE], dtype=np.float32))
        near = [] if _G["near"] else None
        seen = set()
        for c in o.compute_frame_contacts(S, frames[f], _G["itypes"], _G["geom"], frags[k], near):
            text, rec = o.canonical(top, f, c)
            if text not in seen:
                seen.add(text)
                out.append(rec)
        if near:
            near_all += [(f,) + t for t in near]
    return out, near_all


def run_parallel(top_path, frames, itypes, procs=None, geom=None, collect_near=False, **sel):
    frames = np.ascontiguousarray(frames, dtype=np.float32)
    geom = dict(o.DEFAULT_GEOM if geom is None else geom)
    n = len(frames)
    procs = max(1, min(procs or (os.cpu_count() or 1), n))
    chunks = [list(range(k, n, procs)) for k in range(procs)]
    if procs == 1:
        _init(top_path, frames, itypes, geom, sel, collect_near)
        parts = [_work(chunks[0])]
    else:
        with mp.get_context("fork").Pool(procs, initializer=_init, initargs=(top_path, frames, itypes, geom, sel, collect_near)) as pool:
            parts = pool.map(_work, chunks)
        _init(top_path, frames[:1], itypes, geom, sel, collect_near)
    contacts = sorted((c for p in parts for c in p[0]), key=lambda c: c[0])
    near = sorted((t for p in parts for t in p[1]), key=lambda t: t[0])
    return dict(total_frames=n, beg=0, end=n - 1, stride=1, itypes=itypes, contacts=contacts, top=_G["top"], setup=_G["S"],
                frames=frames, near=near)
