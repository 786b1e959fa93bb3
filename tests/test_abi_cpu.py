"""CPU-only: libvtn.so loads and exports every symbol include/vtn.h declares; ctypes signatures cover them all.
No compute call is made (no GPU here); vtn_ctx_create must fail loudly without a device."""
import ctypes as C
import os
import re

import helpers
from vt_b200 import _lib as L


def _declared():
    txt = open(os.path.join(helpers.ROOT, "include", "vtn.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(vtn_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    lib = L.load()
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), "libvtn.so does not export %s" % n
    assert sorted(L.SIGNATURES) == names


def test_ctx_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        return
    lib = L.load()
    h = C.c_void_p()
    rc = lib.vtn_ctx_create(C.byref(h), 0)
    assert rc < 0 and not h
    assert b"no CPU fallback" in lib.vtn_last_error(None)


def test_tool_library_exports_the_reference_entry_points():
    """libvtn_tools.so (include/vtn_tools.h): the reference's tool functions normalize / decompose / annotate_indels (+ uniq, decompose_blocksub)"""
    txt = open(os.path.join(helpers.ROOT, "include", "vtn_tools.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    names = sorted(set(re.findall(r"\b(vtn_tools?_[a-z0-9_]+)\s*\(", txt)))
    assert {"vtn_tool_normalize", "vtn_tool_decompose", "vtn_tool_annotate_indels", "vtn_tool_main"} <= set(names)
    lib = C.CDLL(os.path.join(helpers.ROOT, "vt_b200", "libvtn_tools.so"))
    for n in names:
        assert hasattr(lib, n), "libvtn_tools.so does not export %s" % n
