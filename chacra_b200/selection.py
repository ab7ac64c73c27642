"""Atom-selection mini language (the subset of VMD's `atomselect` grammar that getcontacts and
ChACRA actually pass through ``--sele/--sele2/--solv/--lipid/--ligand``).

Replaces the Tcl ``atomselect top "<text>"`` round trips of the reference
(``external/getcontacts/contact_calc/contact_utils.py:276-289`` ``get_selection_indices`` and the
macro definitions at ``:411-579``).  The expression is tokenised, converted to postfix with a
shunting-yard pass and evaluated on boolean numpy masks, one mask per atom.

Supported: ``and or not ( )``, the flags ``all none protein nucleic water waters hydrogen noh ion
carbon sulfur backbone``, user macros (``solv lipid ligand``), keyword value lists
(``chain resname name resid index element segname residue type altloc insertion serial`` with
``a to b`` ranges for the numeric ones) and ``within R of <expr>``.
"""
from __future__ import annotations

import re

import numpy as np

_LEX = re.compile(r"""\s*(\(|\)|'[^']*'|"[^"]*"|[^\s()]+)""")
_STR_KEYS = {"chain": "chain", "resname": "resname", "name": "name", "element": "element", "segname": "segname",
             "type": "name", "altloc": "altloc", "insertion": "icode"}
_INT_KEYS = ("resid", "index", "residue", "serial")
_OPS = {"or": 1, "and": 2, "not": 3}

ION_RESNAMES = frozenset(
    "AL BA CA Ca CAL CD CES CLA CL Cl CO CS CU Cu CU1 CUA HG IN IOD K MG MN3 MO3 MO4 MO5 MO6 NA Na NAW OC7 PB POT PT "
    "RB SOD TB TL WO4 YB ZN ZN1 ZN2".split())
WATER_RESNAMES = frozenset("H2O HH0 OHH HOH OH2 SOL WAT TIP TIP2 TIP3 TIP4".split())


class SelectionError(ValueError):
    pass


def _isin(arr, values):
    return np.isin(arr, np.array(sorted(values), dtype=arr.dtype)) if len(values) else np.zeros(len(arr), bool)


class Selection:
    """Evaluate selection strings against an :class:`~chacra_b200.topology.AtomTable`."""

    def __init__(self, atoms, macros=None, xyz=None):
        self.atoms = atoms
        self.macros = dict(macros or {})
        self.xyz = atoms.xyz if xyz is None else xyz

    # -- leaves ------------------------------------------------------------------------------
    def _flag(self, word):
        a = self.atoms
        n = a.n
        if word in self.macros:
            m = self.macros[word]
            return m.copy() if isinstance(m, np.ndarray) else Selection(a, self.macros, self.xyz)(m)
        if word == "all":
            return np.ones(n, bool)
        if word == "none":
            return np.zeros(n, bool)
        if word == "protein":
            return a.is_protein.copy()
        if word == "nucleic":
            return a.is_nucleic.copy()
        if word in ("water", "waters"):
            return _isin(a.resname, WATER_RESNAMES)
        if word == "hydrogen":
            return a.is_hydrogen.copy()
        if word == "noh":
            return ~a.is_hydrogen
        if word == "ion":
            return _isin(a.resname, ION_RESNAMES)
        if word in ("carbon", "sulfur"):
            first = np.char.startswith(a.name, "C" if word == "carbon" else "S")
            return first & ~_isin(a.resname, ION_RESNAMES)
        if word == "backbone":
            return a.is_protein & _isin(a.name, ("N", "CA", "C", "O"))
        return None

    def _keyword(self, key, vals):
        a = self.atoms
        if not vals:
            raise SelectionError("keyword %r needs at least one value" % key)
        if key in _INT_KEYS:
            arr = {"resid": a.resid, "index": np.arange(a.n), "residue": a.residue, "serial": np.arange(a.n) + 1}[key]
            m = np.zeros(a.n, bool)
            k = 0
            while k < len(vals):
                if k + 2 < len(vals) and vals[k + 1] == "to":
                    m |= (arr >= int(vals[k])) & (arr <= int(vals[k + 2]))
                    k += 3
                else:
                    m |= arr == int(vals[k])
                    k += 1
            return m
        return _isin(getattr(a, _STR_KEYS[key]), set(vals))

    # -- parser --------------------------------------------------------------------------------
    def __call__(self, text: str) -> np.ndarray:
        toks = _LEX.findall(text)
        if not toks:
            raise SelectionError("empty selection")
        out, ops = [], []          # value stack is built lazily through a postfix program

        def emit(op):
            if op == "not":
                out.append(~out.pop())
            elif op[0] == "within":
                inner = out.pop()
                out.append(self._within(op[1], inner))
            else:
                b, a_ = out.pop(), out.pop()
                out.append(a_ | b if op == "or" else a_ & b)

        def prec(op):
            return 3 if isinstance(op, tuple) else _OPS[op]

        i, expect_value = 0, True
        while i < len(toks):
            t = toks[i]
            if t == "(":
                ops.append("(")
                i += 1
                expect_value = True
            elif t == ")":
                while ops and ops[-1] != "(":
                    emit(ops.pop())
                if not ops:
                    raise SelectionError("unbalanced ')' in %r" % text)
                ops.pop()
                i += 1
                expect_value = False
            elif t in ("and", "or"):
                while ops and ops[-1] != "(" and prec(ops[-1]) >= _OPS[t]:
                    emit(ops.pop())
                ops.append(t)
                i += 1
                expect_value = True
            elif t == "not":
                ops.append("not")
                i += 1
            elif t == "within":
                if i + 2 >= len(toks) or toks[i + 2] != "of":
                    raise SelectionError("expected 'within R of <selection>' in %r" % text)
                ops.append(("within", float(toks[i + 1])))
                i += 3
            elif t in _STR_KEYS or t in _INT_KEYS:
                j = i + 1
                vals = []
                while j < len(toks) and toks[j] not in ("and", "or", "not", "(", ")"):
                    v = toks[j]
                    vals.append(v[1:-1] if v[0] in "'\"" else v)
                    j += 1
                out.append(self._keyword(t, vals))
                i = j
                expect_value = False
            else:
                m = self._flag(t)
                if m is None:
                    raise SelectionError("unsupported selection keyword %r in %r" % (t, text))
                out.append(m)
                i += 1
                expect_value = False
            # unary operators bind to the value that was just completed
            while not expect_value and ops and (ops[-1] == "not" or isinstance(ops[-1], tuple)) \
                    and (i >= len(toks) or toks[i] != "("):
                emit(ops.pop())
        while ops:
            op = ops.pop()
            if op == "(":
                raise SelectionError("unbalanced '(' in %r" % text)
            emit(op)
        if len(out) != 1:
            raise SelectionError("cannot parse selection %r" % text)
        return out[0]

    def _within(self, radius, inner):
        from .topology import within_mask
        return within_mask(self.xyz, np.ones(self.atoms.n, bool), inner, radius)
