"""``from contact_calc.compute_contacts import *`` -- the names the reference scripts pull in
(``external/getcontacts/get_dynamic_contacts.py:20``): ``compute_contacts`` plus the modules it relies on."""
import datetime  # noqa: F401  (re-exported like the reference module does)
import os  # noqa: F401
import sys  # noqa: F401

try:
    from .. import _path  # noqa: F401
except ImportError:  # imported as the top-level package ``contact_calc`` (script directory on sys.path)
    import _path  # noqa: F401

from chacra_b200.compute import compute_contacts  # noqa: E402,F401
from chacra_b200.engine import TRAJ_FRAG_SIZE  # noqa: E402,F401
from . import argparsers as ap  # noqa: E402,F401

__all__ = ["compute_contacts", "TRAJ_FRAG_SIZE", "ap", "datetime", "os", "sys"]
