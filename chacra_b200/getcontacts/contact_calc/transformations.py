"""The frequency-side helpers of ``external/getcontacts/contact_calc/transformations.py`` under their names."""
try:
    from .. import _path  # noqa: F401
except ImportError:
    import _path  # noqa: F401

from chacra_b200.frequencies import (ParseError, parse_residuelabels, shipped_counts, intended_counts_tsv,  # noqa: E402,F401
                                     merge_counts as gen_frequencies)
