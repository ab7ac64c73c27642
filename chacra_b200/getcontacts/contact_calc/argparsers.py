"""Command-line flags of get_dynamic_contacts / get_static_contacts: the same names, defaults and error
behaviour as ``external/getcontacts/contact_calc/argparsers.py:25-127`` (usage + exit status 2 on a bad flag)."""
import argparse
import sys

_GEOMETRY = [
    # flag, type, default, help
    ("--sb_cutoff_dist", float, 4.0, "cutoff for distance between anion and cation atoms [default = 4.0 angstroms]"),
    ("--pc_cutoff_dist", float, 6.0, "cutoff for distance between cation and centroid of aromatic ring [default = 6.0 angstroms]"),
    ("--pc_cutoff_ang", float, 60, "cutoff for angle between normal vector projecting from aromatic plane and vector from "
                                   "aromatic center to cation atom [default = 60 degrees]"),
    ("--ps_cutoff_dist", float, 7.0, "cutoff for distance between centroids of two aromatic rings [default = 7.0 angstroms]"),
    ("--ps_cutoff_ang", float, 30, "cutoff for angle between the normal vectors projecting from each aromatic plane [default = 30 degrees]"),
    ("--ps_psi_ang", float, 45, "cutoff for angle between normal vector projecting from aromatic plane 1 and vector between "
                                "the two aromatic centroids [default = 45 degrees]"),
    ("--ts_cutoff_dist", float, 5.0, "cutoff for distance between centroids of two aromatic rings [default = 5.0 angstroms]"),
    ("--ts_cutoff_ang", float, 30, "cutoff for angle between the normal vectors projecting from each aromatic plane minus 90 "
                                   "degrees [default = 30 degrees]"),
    ("--ts_psi_ang", float, 45, "cutoff for angle between normal vector projecting from aromatic plane 1 and vector between "
                                "the two aromatic centroids [default = 45 degrees]"),
    ("--hbond_cutoff_dist", float, 3.5, "cutoff for distance between donor and acceptor atoms [default = 3.5 angstroms]"),
]


class _Parser(argparse.ArgumentParser):
    def error(self, message):
        self.print_help(sys.stderr)
        sys.stderr.write("\nError: %s\n" % message)
        sys.exit(2)


def build_getcontact_parser(program_description, trajectory):
    p = _Parser(description=program_description, formatter_class=argparse.RawTextHelpFormatter, add_help=False)
    req = p.add_argument_group("Required arguments")
    if trajectory:
        req.add_argument("--topology", type=str, metavar="PATH", required=True, help="path to topology file")
        req.add_argument("--trajectory", type=str, metavar="PATH", required=True, help="path to trajectory file")
    else:
        req.add_argument("--structure", type=str, metavar="PATH", required=True, help="path to structure file")
    req.add_argument("--output", type=str, metavar="PATH", required=False, default=None,
                     help="path to output file (required unless --no_contact_output is given)")
    req.add_argument("--itypes", type=str, nargs="+", required=True, metavar="STR",
                     help="Compute only these interaction types. Valid choices are:\n"
                          "* all, sb (salt-bridges), pc (pi-cation), ps (pi-stacking), ts (t-stacking),\n"
                          "* vdw (van der Waals), hb (hydrogen bonds), hp (hydrophobics)")
    opt = p.add_argument_group("Optional arguments")
    opt.add_argument("-h", "--help", action="help", default=argparse.SUPPRESS, help="show this help message and exit")
    opt.add_argument("--sele", type=str, metavar="STR", default="protein", help="VMD selection 1 [default = 'protein']")
    opt.add_argument("--sele2", type=str, metavar="STR", default=None, help="VMD selection 2 [default = <same as --sele>]")
    opt.add_argument("--ligand", type=str, metavar="STR", default="",
                     help="VMD selection of ligand atoms [default = 'not (protein or nucleic or solv or lipid)']")
    opt.add_argument("--solv", type=str, metavar="STR", default="", help="VMD selection of solvent atoms [default = 'resname HOH TIP3 T3P ..']")
    opt.add_argument("--lipid", type=str, metavar="STR", default="", help="VMD selection of lipid atoms [default = 'resname DLPE DMPC POPC ..']")
    if trajectory:
        opt.add_argument("--cores", type=int, metavar="INT", default=6, help="accepted for compatibility (one GPU does the work) [default = 6]")
        opt.add_argument("--beg", type=int, metavar="INT", default=0, help="first frame to read [default = 0]")
        opt.add_argument("--end", type=int, metavar="INT", default=None, help="last frame to read (unless end-of-trajectory is met first) [default = infty]")
        opt.add_argument("--stride", type=int, metavar="INT", default=1, help="steps between computed frames [default = 1]")
    opt.add_argument("--distout", default=False, action="store_true", help="write the atom_1-atom_2 distance as a last column")
    opt.add_argument("--device", type=int, metavar="INT", default=0, help="CUDA device ordinal (extension) [default = 0]")
    opt.add_argument("--frequency_output", type=str, metavar="PATH", default=None,
                     help="also write the residue-pair frequency table from the device count table (extension)")
    opt.add_argument("--count_output", type=str, metavar="PATH", default=None,
                     help="also write the integer residue-pair frame counts behind that table (extension)")
    opt.add_argument("--near_cutoff_output", type=str, metavar="PATH", default=None,
                     help="list every pair within 1e-4 A / 1e-3 degree of a cutoff in this file (extension)")
    opt.add_argument("--no_contact_output", default=False, action="store_true",
                     help="skip the contact list: with --frequency_output / --count_output nothing but the count table leaves the GPU (extension)")
    geo = p.add_argument_group("Geometric criteria")
    for flag, typ, default, text in _GEOMETRY:
        geo.add_argument(flag, type=typ, metavar="FLOAT", default=default, help=text)
    geo.add_argument("--hbond_cutoff_ang", type=float, metavar="FLOAT", default=70 if trajectory else 180,
                     help="cutoff for angle between donor hydrogen acceptor [default = %d degrees]" % (70 if trajectory else 180))
    geo.add_argument("--hbond_res_diff", type=int, metavar="INT", default=1,
                     help="minimum residue distance for which to consider computing hydrogen bond interactions [default = 1]")
    geo.add_argument("--vdw_epsilon", type=float, metavar="FLOAT", default=0.5,
                     help="amount of padding for calculating vanderwaals contacts [default = 0.5 angstroms]")
    geo.add_argument("--vdw_res_diff", type=int, metavar="INT", default=2,
                     help="minimum residue distance for which to consider computing vdw interactions [default = 2]")
    return p
