"""The ctypes mirrors in ``ops/_lib.py`` / ``ops/linlink.py`` must match the
C structs of ``include/multigrad_b200.h`` byte for byte.  Checked here without
a GPU by compiling a tiny C program against the header (the library repeats
the check against its own ``mgb_abi_sizes`` when it is loaded)."""
import ctypes as C
import shutil
import subprocess
from pathlib import Path

import pytest

from multigrad_b200.ops import _lib, linlink

ROOT = Path(__file__).resolve().parents[1]

MIRRORS = [("MgbPeerTable", _lib.MgbPeerTable),
           ("MgbErfBinModel", _lib.MgbErfBinModel),
           ("MgbStepArgs", _lib.MgbStepArgs),
           ("MgbGroupArgs", _lib.MgbGroupArgs),
           ("MgbLinModel", linlink.MgbLinModel),
           ("MgbLinArgs", linlink.MgbLinArgs)]


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs a C compiler")
def test_struct_sizes_and_offsets_match_the_header(tmp_path):
    lines = ['#include <stdio.h>', '#include <stddef.h>',
             '#include "multigrad_b200.h"', 'int main(void) {']
    for name, cls in MIRRORS:
        lines.append(f'  printf("{name} %zu\\n", sizeof({name}));')
        for field, _ in cls._fields_:
            lines.append(f'  printf("{name}.{field} %zu\\n", '
                         f'offsetof({name}, {field}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "abi.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "abi"
    subprocess.run(["gcc", "-std=c99", "-I", str(ROOT / "include"), "-o",
                    str(exe), str(src)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True,
                         text=True).stdout
    got = dict(line.split() for line in out.strip().splitlines())
    for name, cls in MIRRORS:
        assert int(got[name]) == C.sizeof(cls), name
        for field, _ in cls._fields_:
            assert int(got[f"{name}.{field}"]) == getattr(cls, field).offset, \
                f"{name}.{field}"


@pytest.mark.skipif(_lib.nvcc_path() is None, reason="needs nvcc to build")
def test_library_loads_without_a_gpu_driver():
    """``__graft_entry__.build()`` imports the library on a machine without
    ``libcuda.so.1``: the driver API must only be reached through
    ``cudaGetDriverEntryPoint`` (csrc/nvls.cu, csrc/linlink.cu), never through
    a link-time symbol."""
    path = _lib.build()
    if shutil.which("readelf"):
        dyn = subprocess.run(["readelf", "-d", str(path)], check=True,
                             capture_output=True, text=True).stdout
        assert "libcuda" not in dyn, dyn
    lib = _lib.load()                  # also runs the ABI size check
    assert lib.mgb_version() >= 100
    out = (C.c_uint32 * 8)()
    assert lib.mgb_abi_sizes(out, 8) == len(MIRRORS)
    assert [int(v) for v in out[:len(MIRRORS)]] == \
        [C.sizeof(cls) for _, cls in MIRRORS]
