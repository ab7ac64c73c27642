"""The C ABI: libchacra_b200.so builds for sm_100a, loads without a GPU, and exports every function that
include/chacra_b200.h declares; the ctypes table of chacra_b200/_lib.py binds exactly that set.  No device work here."""
import ctypes as C
import os
import re

import numpy as np

from conftest import ROOT
from chacra_b200 import _lib

HEADER = os.path.join(ROOT, "include", "chacra_b200.h")


def _declared():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)                  # comments mention names too
    body = text[text.index('extern "C"'):]
    names = set(re.findall(r"\b(chc_[a-z0-9_]+)\s*\(", body))
    assert len(names) > 25
    return names


def test_library_builds_and_exports_every_declared_function():
    path = _lib.build()
    assert os.path.isfile(path) and path.startswith(os.path.join(ROOT, "chacra_b200"))        # in-tree, not a JIT cache
    lib = C.CDLL(path)
    missing = [n for n in sorted(_declared()) if not hasattr(lib, n)]
    assert not missing, missing
    lib.chc_abi_version.restype = C.c_int
    assert lib.chc_abi_version() == 1


def test_ctypes_table_binds_exactly_the_header():
    assert set(_lib.SYMBOLS) == _declared()
    lib = _lib.load()
    for name, (res, args) in _lib.SYMBOLS.items():
        fn = getattr(lib, name)
        assert fn.restype == res and list(fn.argtypes) == list(args)


def test_record_layouts_match_the_header():
    from chacra_b200.engine import CONTACT_DTYPE, NEAR_DTYPE
    from chacra_b200.xtc import META_DTYPE
    text = open(HEADER).read()
    for struct, dtype, size in (("chc_contact", CONTACT_DTYPE, 32), ("chc_near", NEAR_DTYPE, 32), ("chc_xtc_frame", META_DTYPE, 64)):
        assert "} %s;" % struct in text and dtype.itemsize == size
    body = text[text.index("typedef struct chc_contact {"):text.index("} chc_contact;")]
    fields = re.findall(r"^\s*(?:int32_t|float)\s+(\w+);", body, flags=re.M)
    assert fields == ["frame", "itype", "atom1", "atom2", "water1", "water2", "distance", "reserved"] == list(CONTACT_DTYPE.names)


def test_host_only_entry_points_work_without_a_gpu(tmp_path):
    """The text stages are host code: they run here."""
    lib = _lib.load()
    buf = C.create_string_buffer(40)
    assert lib.chc_format_fixed3(C.c_float(2.0625), buf) == 5 and buf.value == b"2.062"     # tie to even, like printf
    p = tmp_path / "c.tsv"
    p.write_text("# total_frames:2 beg:0 end:1 stride:1 interaction_types:sb\n0\tsb\tA:ASP:1:OD1\tA:LYS:2:NZ\n1\tsb\tA:ASP:1:OD1\tA:LYS:2:NZ\n")
    from chacra_b200 import frequencies as fq
    assert fq.native_counts(str(p), ["sb"], shipped=False) == (2, {("A:ASP:1", "A:LYS:2"): 2})
    assert fq.native_counts(str(p), ["sb"], shipped=True) == (2, {("A:ASP:1", "A:LYS:2"): 1})


def test_device_entry_points_fail_loudly_without_a_gpu():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from chacra_b200.pca import run_permuted_pca
    from chacra_b200.trajectory import DCDFile
    from conftest import GOLDEN
    with pytest.raises(_lib.ChacraLibraryError):
        run_permuted_pca(np.zeros((3, 4)), 2, seed=1)
    with pytest.raises(_lib.ChacraLibraryError):
        DCDFile(os.path.join(GOLDEN, "5xnd_trajectory.dcd")).read_device()


def test_plain_c_client_links_and_runs(tmp_path):
    """The header compiles as C99 and a C program drives the host-only entry points through the shared library."""
    import subprocess
    lib_dir = os.path.dirname(_lib.build())
    exe = str(tmp_path / "abi_host")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "c", "abi_host.c"),
                    "-o", exe, "-L", lib_dir, "-lchacra_b200", "-Wl,-rpath," + lib_dir], check=True)
    out = subprocess.run([exe, str(tmp_path / "scratch.tsv")], capture_output=True, text=True)
    assert out.returncode == 0 and "abi_host ok" in out.stdout, out.stderr
