#!/usr/bin/env python3
"""
Extension (no counterpart script in getcontacts): residue-pair contact frequencies of EVERY thermodynamic state straight
from the femto HREMD replica trajectories, without writing state trajectories or contact lists.

Replaces, for the frequency tables, the chain  ReplicaHandler.write_state_trajectories -> state_{i}.xtc ->
get_dynamic_contacts -> cont_state_{i}.tsv -> get_contact_frequencies  of chacra/scripts/process_hremd_output.py:100-175
(SURVEY 8(f) rows X1-X3).  Output files freqs_state_{i}.tsv have the format ChACRA's ContactFrequencies loader reads.
The counts are the intended ones (every frame counted); the CA alignment and the XTC quantisation of the intermediate
state trajectories are not reproduced (contacts depend on distances and angles only).
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _path  # noqa: E402,F401


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--topology", required=True, help="structure file (PDB)")
    ap.add_argument("--replicas", required=True, nargs="+", metavar="r{j}.dcd", help="replica trajectories in replica order")
    ap.add_argument("--state_data", required=True, help="femto samples.arrow, or a .npy table [cycles][replicas] of state indices")
    ap.add_argument("--save_interval", type=int, default=1, help="cycles between saved frames (femto trajectory_interval)")
    ap.add_argument("--output_dir", required=True)
    ap.add_argument("--sele", default="protein")
    ap.add_argument("--sele2", default=None)
    ap.add_argument("--itypes", nargs="+", default=["all"])
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--counts", action="store_true", help="also write counts_state_{i}.tsv (integer frame counts)")
    ap.add_argument("--window", type=int, default=2000, help="frames per replica held on the device at a time")
    args = ap.parse_args(argv)
    from chacra_b200 import compile_topology, read_pdb
    from chacra_b200.demux import demux_contacts_from_files, load_state_table
    from chacra_b200.engine import COMPUTE_ORDER, ContactEngine
    from chacra_b200.frequencies import ALL_FREQ_ITYPES, counts_to_table, write_count_table, write_frequency_table
    from chacra_b200.trajectory import open_trajectory
    itypes = COMPUTE_ORDER if "all" in args.itypes else [t for t in COMPUTE_ORDER if t in args.itypes]
    atoms = read_pdb(args.topology)
    xyz0 = open_trajectory(args.replicas[0]).read(0, 0)[0]
    topo = compile_topology(atoms, xyz0, args.sele, args.sele2 or args.sele, "", "", "")
    if not topo.sele1.any() or not topo.sele2.any():
        print("The selection is empty!")
        sys.exit(1)
    table = load_state_table(args.state_data, args.save_interval)
    eng = ContactEngine(topo, itypes, n_states=len(args.replicas), device=args.device)
    n_frames, _ = demux_contacts_from_files(eng, args.replicas, list(table), 1, args.window)
    os.makedirs(args.output_dir, exist_ok=True)
    for s in range(len(args.replicas)):
        pairs = counts_to_table(topo.residue_label, eng.counts(s))
        with open(os.path.join(args.output_dir, "freqs_state_%d.tsv" % s), "w") as fd:
            write_frequency_table(fd, n_frames, ALL_FREQ_ITYPES, {k: (c, float(c) / n_frames) for k, c in pairs.items()})
        if args.counts:
            with open(os.path.join(args.output_dir, "counts_state_%d.tsv" % s), "w") as fd:
                write_count_table(fd, n_frames, ALL_FREQ_ITYPES, pairs)
    eng.close()
    print("wrote %d frequency tables (%d frames per state) to %s" % (len(args.replicas), n_frames, args.output_dir))


if __name__ == "__main__":
    main()
