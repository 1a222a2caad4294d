"""The C-ABI boundary without a GPU: libmmimo_b200.so loads, exports every function include/mmimo_b200.h declares,
the ctypes structures mirror the C structs (size and field offsets, checked against gcc), and entry points refuse
to run instead of computing anything when no device is present."""
import ctypes as C
import os
import re
import subprocess
import tempfile

import pytest

from conftest import ROOT
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _lib, _stage_abi

HEADER = os.path.join(ROOT, "include", "mmimo_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mmimo_[a-z0-9_]+)\s*\(", src)))


def test_header_functions_are_exported_and_bound():
    lib = _lib.load()
    names = declared_functions()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), "declared in the header but not exported: " + n
    assert sorted(_lib.EXPORTS) == names, set(names) ^ set(_lib.EXPORTS)
    assert set(_stage_abi.STAGE_EXPORTS) <= set(names)
    assert lib.mmimo_version().decode().startswith("mmimo_b200")


def test_every_entry_point_cites_the_reference():
    """Each declaration block names the reference file:line it replaces."""
    src = open(HEADER).read()
    assert len(re.findall(r"[a-z_]+\.py:\d+", src)) >= 40


STRUCTS = {"mmimo_link_config": _lib.LinkConfig, "mmimo_sim_request": _lib.SimRequest, "mmimo_sim_result": _lib.SimResult,
           "mmimo_modem_desc": _stage_abi.ModemDesc, "mmimo_pa_desc": _stage_abi.PaDesc}


def test_ctypes_structs_match_the_header():
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "mmimo_b200.h"', 'int main(void) {']
    for cname, ct in STRUCTS.items():
        lines.append('printf("%s %%zu\\n", sizeof(%s));' % (cname, cname))
        for fname, _ in ct._fields_:
            lines.append('printf("%s.%s %%zu\\n", offsetof(%s, %s));' % (cname, fname, cname, fname))
    lines.append("return 0; }")
    with tempfile.TemporaryDirectory() as d:
        src, exe = os.path.join(d, "t.c"), os.path.join(d, "t")
        open(src, "w").write("\n".join(lines))
        subprocess.run(["gcc", "-I", os.path.dirname(HEADER), src, "-o", exe], check=True)
        out = dict(l.split() for l in subprocess.run([exe], check=True, capture_output=True, text=True).stdout.splitlines())
    for cname, ct in STRUCTS.items():
        assert int(out[cname]) == C.sizeof(ct), cname
        for fname, _ in ct._fields_:
            assert int(out["%s.%s" % (cname, fname)]) == getattr(ct, fname).offset, (cname, fname)


def test_no_cpu_fallback():
    """Without a CUDA device nothing computes: creating a context fails loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.mmimo_create(C.byref(h), 0) != 0
    assert lib.mmimo_last_error()
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine, MmimoError
    with pytest.raises(MmimoError):
        LinkEngine(0)
    with pytest.raises(MmimoError):
        _stage_abi.get_ops()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "m_mimo_ofdm_with_nonlinear_pa_sim_b200")
    for base, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".inl")):
                txt = open(os.path.join(base, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), fn
                assert "link_oracle" not in txt, fn
