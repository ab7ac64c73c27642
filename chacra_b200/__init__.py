"""chacra_b200 -- B200-native contact detection behind getcontacts' interface (ChACRA's hot path).

Importing the package never touches CUDA; ``ContactEngine`` loads ``csrc/libchacra_b200.so`` and
fails loudly when it is missing (there is no CPU fallback).
"""
from .topology import read_pdb, compile_topology, AtomTable, CompiledTopology  # noqa: F401

__all__ = ["read_pdb", "compile_topology", "AtomTable", "CompiledTopology", "ContactEngine"]


def __getattr__(name):
    if name == "ContactEngine":
        from .engine import ContactEngine
        return ContactEngine
    raise AttributeError(name)
