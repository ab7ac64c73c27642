"""Replica -> state demultiplexing on the device.

femto's HREMD writes one trajectory per REPLICA (``r{j}.dcd``) plus a table ``replica_to_state_idx`` (one row per
cycle: the thermodynamic state every replica was in).  The reference first turns that into per-state frame lists
(``sort_replica_trajectories``, chacra/trajectories/process_hremd.py:56-94), then gathers the frames of a state
from every replica file, aligns them on CA atoms and writes ``state_{i}.xtc``
(``get_state_coordinates_from_replica`` :184-230, ``write_state_trajectory`` :233-307), which getcontacts reads
back.  Here the replica trajectories stay where they are in device memory and the binning kernel reads frame
``f`` of state ``s`` through a source index (``chc_submit_frames_indexed``): no state trajectory is written.

What is deliberately NOT reproduced: the CA alignment (a rigid motion; contacts depend on distances and angles
only) and the XTC quantisation to 0.001 nm of the intermediate file.  Contacts therefore are those of the
replica coordinates themselves; parity is defined against "gather the same frames, then run the contact path"
(bit-exact, tests/test_gpu_demux.py), and the host-side index logic against the reference's own function.
"""
from __future__ import annotations

import numpy as np


def _state_table(replica_to_state_idx, save_interval: int = 1, traj_len: int | None = None) -> np.ndarray:
    """[cycles][replicas] int table restricted to the cycles whose coordinates were saved (process_hremd.py:86-88)."""
    col = replica_to_state_idx
    if hasattr(col, "columns") or isinstance(col, dict):         # femto's state data frame: take the column
        col = col["replica_to_state_idx"]
    if hasattr(col, "values") and not isinstance(col, np.ndarray):
        col = col.values
    rows = [np.asarray(r) for r in col[::max(1, int(save_interval))]]
    if traj_len is not None:
        rows = rows[:int(traj_len)]
    if not rows:
        return np.zeros((0, 0), np.int64)
    table = np.vstack(rows).astype(np.int64)
    return table


def sort_replica_trajectories(replica_to_state_idx, save_interval: int = 1, traj_len: int | None = None) -> dict:
    """Same result as the reference's function of this name (process_hremd.py:56-94): ``out[replica][state]`` is the
    ascending array of saved-frame numbers at which ``replica`` was in ``state``."""
    table = _state_table(replica_to_state_idx, save_interval, traj_len)
    n = table.shape[1]
    out = {r: [] for r in range(n)}
    for r in range(n):
        col = table[:, r]
        order = np.argsort(col, kind="stable")
        bounds = np.searchsorted(col[order], np.arange(n + 1))
        for s in range(n):
            out[r].append(order[bounds[s]:bounds[s + 1]])
    return out


def state_sources(replica_to_state_idx, save_interval: int = 1, traj_len: int | None = None) -> np.ndarray:
    """int64 [states][frames]: the replica that holds frame f of state s (the inverse of every row of the table).
    Raises if a row is not a permutation — the reference would silently leave zero coordinates there
    (``coordinates = np.zeros(...)``, process_hremd.py:293)."""
    table = _state_table(replica_to_state_idx, save_interval, traj_len)
    f, n = table.shape
    if f == 0:
        return np.zeros((n, 0), np.int64)
    if table.min() < 0 or table.max() >= n or (np.sort(table, axis=1) != np.arange(n)).any():
        bad = int(np.nonzero((np.sort(table, axis=1) != np.arange(n)).any(axis=1) | (table.min(axis=1) < 0))[0][0])
        raise ValueError("replica_to_state_idx row %d is not a permutation of 0..%d" % (bad, n - 1))
    return np.ascontiguousarray(np.argsort(table, axis=1, kind="stable").T)


def demux_contacts(engine, replicas, replica_to_state_idx, save_interval: int = 1, states=None, first_index: int = 0):
    """Run the contact path for every state straight off the replica trajectories.

    ``replicas``: CUDA float32 tensor [R][F][N][3] (replica-major, the frames of ``r{j}.dcd`` in order).  For state s
    frame f is read from replica ``state_sources(...)[s][f]``; counts accumulate in ``engine`` under state s.
    Returns the int64 source table [S][F]."""
    import torch
    if not (isinstance(replicas, torch.Tensor) and replicas.is_cuda and replicas.dim() == 4 and replicas.is_contiguous()):
        raise ValueError("replicas must be a contiguous CUDA tensor [R][F][N][3]")
    n_rep, n_frames = int(replicas.shape[0]), int(replicas.shape[1])
    src = state_sources(replica_to_state_idx, save_interval, n_frames)
    if src.shape[0] != n_rep or src.shape[1] != n_frames:
        raise ValueError("state table gives %d states x %d saved frames, the trajectories hold %d replicas x %d frames"
                         % (src.shape[0], src.shape[1], n_rep, n_frames))
    flat = src * n_frames + np.arange(n_frames, dtype=np.int64)[None, :]        # frame number in the [R*F] pool
    pool = replicas.view(n_rep * n_frames, replicas.shape[2], 3)
    d_flat = torch.from_numpy(flat).to(replicas.device)
    for s in (range(n_rep) if states is None else states):
        engine.submit_indexed(pool, d_flat[s], state=int(s), first_index=first_index)
    return src


def load_state_table(source, save_interval: int = 1, traj_len: int | None = None) -> np.ndarray:
    """``replica_to_state_idx`` restricted to the saved cycles, from a femto ``samples.arrow`` stream file
    (``load_femto_data``, process_hremd.py:24-36), a ``.npy`` table [cycles][replicas], or anything ``state_sources``
    takes."""
    if isinstance(source, (str, bytes)) or hasattr(source, "__fspath__"):
        path = str(source)
        if path.endswith(".npy"):
            source = list(np.load(path))
        else:
            import pyarrow
            with pyarrow.OSFile(path, "rb") as fh:
                with pyarrow.RecordBatchStreamReader(fh) as reader:
                    source = reader.read_all().to_pandas()
    return _state_table(source, save_interval, traj_len)


def demux_contacts_from_files(engine, replica_paths, replica_to_state_idx, save_interval: int = 1, window: int = 2000,
                              fragment: int = 100):
    """Contacts of every state straight off the replica trajectory files ``r0.dcd, r1.dcd, ...`` (replica order): no
    state trajectory is written or held.  Frames are taken in windows of ``window`` frames (a multiple of the 100-frame
    fragment, so the per-fragment hydrophobic sets see the same first frames as a whole-trajectory run); a window holds
    R x window frames on the device.  Returns (n_frames, source table [S][F])."""
    import torch
    from .trajectory import open_trajectory
    readers = [open_trajectory(p) for p in replica_paths]
    n_rep = len(readers)
    n_frames = min(r.n_frames for r in readers)
    for r in readers:
        if hasattr(r, "device"):
            r.device = engine.device
        if r.n_atoms != engine.topo.n_atoms:
            raise ValueError("%s has %d atoms, the topology %d" % (getattr(r, "path", "trajectory"), r.n_atoms, engine.topo.n_atoms))
    src = state_sources(replica_to_state_idx, save_interval, n_frames)
    if src.shape[0] != n_rep:
        raise ValueError("state table has %d states, %d replica files were given" % (src.shape[0], n_rep))
    n_frames = src.shape[1]
    window = max(fragment, window // fragment * fragment)
    dev = torch.device("cuda", engine.device)
    for f0 in range(0, n_frames, window):
        f1 = min(f0 + window, n_frames)
        w = f1 - f0
        pool = torch.empty((n_rep, w, engine.topo.n_atoms, 3), dtype=torch.float32, device=dev)
        for r, reader in enumerate(readers):
            read = getattr(reader, "read_device", None)
            if read is not None:
                pool[r] = read(f0, f1 - 1)
            else:
                pool[r] = torch.from_numpy(np.ascontiguousarray(reader.read(f0, f1 - 1))).to(dev)
        flat = torch.from_numpy(src[:, f0:f1] * w + np.arange(w, dtype=np.int64)[None, :]).to(dev)
        flat_pool = pool.view(n_rep * w, engine.topo.n_atoms, 3)
        for s in range(n_rep):
            engine.submit_indexed(flat_pool, flat[s], state=s, first_index=f0)
        engine.synchronize()                              # the pool is released before the next window is read
    return n_frames, src
