#!/usr/bin/env python3
"""
Extension (no counterpart script in getcontacts): contact frequencies of ALL thermodynamic states on every GPU of the box.

Replaces the per-state loop of chacra/scripts/process_hremd_output.py:134-175 (get_dynamic_contacts + get_contact_frequencies
once per state, sequentially) for the frequency tables ChACRA's ContactFrequencies consumes.  One process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 get_state_frequencies.py \\
        --topology protein.pdb --trajectories state_0.xtc state_1.xtc ... --output_dir freqs/

Every rank takes a fragment-aligned slice of every state (compute_contacts.py:250-261 hands the same 100-frame fragments to
worker processes), the per-rank integer tables are merged with one all-reduce over NVLink, rank 0 writes freqs_state_{i}.tsv.
Run without torchrun it is the single-GPU counts-only path.  Counts are the intended ones (every frame counted).
"""
import argparse
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _path  # noqa: E402,F401


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--topology", required=True, help="structure file (PDB)")
    ap.add_argument("--trajectories", required=True, nargs="+", metavar="state_{i}.xtc|dcd",
                    help="one trajectory per state; sorted by the number before the suffix like process_hremd_output.py:182")
    ap.add_argument("--output_dir", required=True)
    ap.add_argument("--itypes", nargs="+", default=["all"])
    ap.add_argument("--sele", default="protein")
    ap.add_argument("--sele2", default=None)
    ap.add_argument("--solv", default="")
    ap.add_argument("--lipid", default="")
    ap.add_argument("--ligand", default="")
    ap.add_argument("--beg", type=int, default=0)
    ap.add_argument("--end", type=int, default=None)
    ap.add_argument("--stride", type=int, default=1)
    ap.add_argument("--counts", action="store_true", help="also write counts_state_{i}.tsv (integer frame counts)")
    ap.add_argument("--near_cutoff_output", default=None, metavar="PATH", help="list every pair within 1e-4 A / 1e-3 degree of a cutoff")
    args = ap.parse_args(argv)
    import torch
    import torch.distributed as dist
    from chacra_b200.engine import ChacraError
    from chacra_b200.states import gather_near, run_states, write_state_tables
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    started = False
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        started = True
    rank = dist.get_rank() if dist.is_initialized() else 0

    def key(path):
        m = re.split(r"_|\.", os.path.basename(path))
        try:
            return int(m[-2])
        except (ValueError, IndexError):
            return 0
    trajs = sorted(args.trajectories, key=key) if len({key(t) for t in args.trajectories}) == len(args.trajectories) else list(args.trajectories)
    try:
        info = {}
        topo, counts, sel, near = run_states(args.topology, trajs, args.itypes, None, args.sele, args.sele2, args.solv, args.lipid, args.ligand,
                                             args.beg, args.end, args.stride, device=local if world > 1 else 0,
                                             track_near=args.near_cutoff_output is not None, info=info)
    except ChacraError as exc:
        if "selection is empty" in str(exc):
            print(str(exc))
            sys.exit(1)
        raise
    if args.near_cutoff_output:
        gather_near(near, topo.atoms.atom_labels(), sel, max(1, args.stride), args.near_cutoff_output)
    if rank == 0:
        write_state_tables(args.output_dir, topo, counts, sel, counts_too=args.counts)
        print("wrote %d frequency tables to %s (%d rank(s)%s)" % (len(trajs), args.output_dir, world,
                                                                 ", compacted exchange" if info.get("compact") else ""))
    if started:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
