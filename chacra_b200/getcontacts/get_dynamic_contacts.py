#!/usr/bin/env python3
"""
Computes the interaction network of a molecular-dynamics trajectory: salt bridges, pi-cation, pi-stacking,
T-stacking, van der Waals, hydrogen bonds (with water bridges) and hydrophobic contacts, frame by frame.

Drop-in for ``external/getcontacts/get_dynamic_contacts.py`` (same flags, same contact-list TSV): the frames are
processed by the sm_100a CUDA kernels of chacra_b200 instead of VMD worker processes.
"""
import datetime
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _path  # noqa: E402,F401
from contact_calc import argparsers as ap  # noqa: E402
from contact_calc.compute_contacts import compute_contacts  # noqa: E402

ALL_ITYPES = ["hp", "sb", "pc", "ps", "ts", "vdw", "hb"]      # header order of the reference (get_dynamic_contacts.py:68)


def process_geometric_criterion_args(args):
    """Flag values -> the cutoff dictionary the detectors read (get_dynamic_contacts.py:23-40)."""
    return {
        "SALT_BRIDGE_CUTOFF_DISTANCE": args.sb_cutoff_dist,
        "PI_CATION_CUTOFF_DISTANCE": args.pc_cutoff_dist, "PI_CATION_CUTOFF_ANGLE": args.pc_cutoff_ang,
        "PI_STACK_CUTOFF_DISTANCE": args.ps_cutoff_dist, "PI_STACK_CUTOFF_ANGLE": args.ps_cutoff_ang,
        "PI_STACK_PSI_ANGLE": args.ps_psi_ang,
        "T_STACK_CUTOFF_DISTANCE": args.ts_cutoff_dist, "T_STACK_CUTOFF_ANGLE": args.ts_cutoff_ang,
        "T_STACK_PSI_ANGLE": args.ts_psi_ang,
        "HBOND_CUTOFF_DISTANCE": args.hbond_cutoff_dist, "HBOND_CUTOFF_ANGLE": args.hbond_cutoff_ang,
        "HBOND_RES_DIFF": args.hbond_res_diff,
        "VDW_EPSILON": args.vdw_epsilon, "VDW_RES_DIFF": args.vdw_res_diff,
    }


def main(argv=None):
    parser = ap.build_getcontact_parser(__doc__, True)
    args = parser.parse_args(argv)
    if args.output is None and not args.no_contact_output:
        parser.error("the following arguments are required: --output")
    sele2 = args.sele if args.sele2 is None else args.sele2
    end = args.end if args.end else sys.maxsize
    if "all" in args.itypes:
        itypes = ALL_ITYPES
    else:
        for itype in args.itypes:
            if itype not in ALL_ITYPES:
                parser.error(itype + " is not a valid interaction type")
        itypes = args.itypes
    tic = datetime.datetime.now()
    compute_contacts(args.topology, args.trajectory, args.output, itypes, process_geometric_criterion_args(args),
                     args.cores, args.beg, end, args.stride, args.distout, args.ligand, args.solv, args.lipid,
                     args.sele, sele2, device=args.device, frequency_output=args.frequency_output, count_output=args.count_output,
                     near_output=args.near_cutoff_output, write_contacts=not args.no_contact_output)
    print("\nTotal computation time:", (datetime.datetime.now() - tic).total_seconds(), "seconds")
    for k, v in (("topology", args.topology), ("trajectory", args.trajectory), ("output", args.output),
                 ("cores", args.cores), ("solv", args.solv), ("sele", args.sele), ("sele2", sele2), ("stride", args.stride)):
        print("%s=%s" % (k, v))


if __name__ == "__main__":
    main()
