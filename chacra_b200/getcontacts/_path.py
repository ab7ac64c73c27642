"""Make ``chacra_b200`` importable when these scripts are run by path (``python .../get_dynamic_contacts.py``),
which is how ChACRA starts them (``chacra/scripts/process_hremd_output.py:139-175``)."""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
