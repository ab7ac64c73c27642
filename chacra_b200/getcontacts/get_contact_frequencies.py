#!/usr/bin/env python3
"""
Take one or more contact-list files and compute, for every residue pair, the fraction of frames in which the two
residues are joined by at least one interaction of the requested types.

Drop-in for ``external/getcontacts/get_contact_frequencies.py`` (same flags, same table).  By default the
table is the one the reference script produces from the same file, including its two reader quirks (the line on
which the frame number changes is lost and the last frame is never counted, SURVEY A.12 level L3);
``--semantics intended`` counts every line of every frame (level L2, what the GPU count tables hold).
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _path  # noqa: E402,F401
from chacra_b200 import frequencies as fq  # noqa: E402


class _Parser(argparse.ArgumentParser):
    def error(self, message):
        self.print_help(sys.stderr)
        sys.stderr.write("\nError: %s\n" % message)
        sys.exit(2)


def main(argv=None):
    parser = _Parser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    parser.add_argument("--input_files", type=argparse.FileType("r"), required=True, nargs="+", metavar="FILE.tsv",
                        help="Path to one or more contact-file outputs")
    parser.add_argument("--label_file", type=argparse.FileType("r"), required=False, metavar="FILE.tsv",
                        help="A label file for standardizing residue names between different proteins")
    parser.add_argument("--output_file", type=argparse.FileType("w"), required=True, metavar="FILE.tsv",
                        help="Path to output file")
    parser.add_argument("--itypes", required=False, default="all", type=str, nargs="+", metavar="ITYPE",
                        help="Include only these interaction types in frequency computation (default: all)")
    parser.add_argument("--semantics", choices=["shipped", "intended"], default="shipped",
                        help="shipped = bit-compatible with the reference script (default); intended = every frame counts")
    args = parser.parse_args(argv)
    itypes = fq.ALL_FREQ_ITYPES if "all" in args.itypes else args.itypes
    labels = None
    if args.label_file is not None:
        labels, _ = fq.parse_residuelabels(args.label_file)
        args.label_file.close()
    reduce_ = fq.shipped_counts if args.semantics == "shipped" else fq.intended_counts_tsv
    per_input = []
    for fh in args.input_files:
        name = getattr(fh, "name", None)
        fast = None
        if labels is None and isinstance(name, str) and os.path.isfile(name):
            # a file on disk and no relabelling: the library's threaded reader (same table; None = not a file in frame order)
            fast = fq.native_counts(name, itypes, shipped=args.semantics == "shipped")
        if fast is not None:
            fh.close()
            per_input.append(fast)
            continue
        frames, sets = reduce_(fh, itypes, as_sets=True)
        fh.close()
        if labels is not None:
            sets = fq.relabel_counts(sets, labels)
        per_input.append((frames, {k: len(v) for k, v in sets.items()}))
    total, freqs = fq.merge_counts(per_input)
    fq.write_frequency_table(args.output_file, total, itypes, freqs)
    args.output_file.close()


if __name__ == "__main__":
    main()
