#!/usr/bin/env python3
"""
Computes the interaction network of a single structure (first model of a PDB file): salt bridges, pi-cation,
pi-stacking, T-stacking, van der Waals, hydrogen bonds with water bridges, hydrophobic contacts.

Drop-in for ``external/getcontacts/get_static_contacts.py`` (same flags; the hydrogen-bond angle defaults to 180
degrees here, ``argparsers.py:113``): the structure is its own one-frame trajectory.
"""
import datetime
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _path  # noqa: E402,F401
from contact_calc import argparsers as ap  # noqa: E402
from contact_calc.compute_contacts import compute_contacts  # noqa: E402
from get_dynamic_contacts import ALL_ITYPES, process_geometric_criterion_args  # noqa: E402


def main(argv=None):
    parser = ap.build_getcontact_parser(__doc__, False)
    args = parser.parse_args(argv)
    if args.output is None and not args.no_contact_output:
        parser.error("the following arguments are required: --output")
    sele2 = args.sele if args.sele2 is None else args.sele2
    if "all" in args.itypes:
        itypes = ALL_ITYPES
    else:
        for itype in args.itypes:
            if itype not in ALL_ITYPES:
                parser.error(itype + " is not a valid interaction type")
        itypes = args.itypes
    tic = datetime.datetime.now()
    compute_contacts(args.structure, args.structure, args.output, itypes, process_geometric_criterion_args(args), 1, 0, 0, 1,
                     args.distout, args.ligand, args.solv, args.lipid, args.sele, sele2, device=args.device,
                     frequency_output=args.frequency_output, count_output=args.count_output,
                     near_output=args.near_cutoff_output, write_contacts=not args.no_contact_output)
    print("\nTotal computation time: " + str((datetime.datetime.now() - tic).total_seconds()) + " seconds")
    for k, v in (("topology", args.structure), ("trajectory", args.structure), ("output", args.output), ("solv", args.solv),
                 ("sele", args.sele), ("sele2", sele2)):
        print("%s=%s" % (k, v))


if __name__ == "__main__":
    main()
