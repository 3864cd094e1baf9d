"""CPU: the C-ABI library loads and exports every symbol include/defl/*.h declares.
No decode call is made here (there is no GPU in this tier and no CPU path in the product)."""
import ctypes
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = []
    for h in ("infl.h", "infl_batch.h", "infl_wrap.h"):
        src = open(os.path.join(ROOT, "include", "defl", h)).read()
        names += re.findall(r"UNZ_EXPORT\s+[\w\s\*]+?\b(infl\w*)\s*\(", src)
    return sorted(set(names))


def test_library_exports_all_declared_symbols():
    import defl_b200
    L = defl_b200.lib()
    declared = _declared()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/defl but not exported"
    assert sorted(defl_b200.EXPORTED_SYMBOLS) == declared


def test_struct_layouts_match_headers():
    import defl_b200
    assert defl_b200.ITEM_DTYPE.itemsize == 32
    assert defl_b200.RESULT_DTYPE.itemsize == 32
    assert defl_b200.RESULT_DTYPE.fields["outlen"][1] == 8
    assert defl_b200.RESULT_DTYPE.fields["adler_stored"][1] == 24


def test_result_codes_are_the_reference_abi():
    import defl_b200 as d
    import errno
    # reference include/defl/common.h:80-90, :99
    assert (d.UNZ_UNFINISHED, d.UNZ_NOOP, d.UNZ_OK, d.UNZ_ERR) == (2, 1, 0, -1)
    assert d.UNZ_EFULL == -errno.ENOBUFS and d.INFL_ZLIB == 1
    hdr = open(os.path.join(ROOT, "include", "defl", "common.h")).read()
    for token in ("UNZ_UNFINISHED =  2", "UNZ_NOOP       =  1", "UNZ_OK         =  0", "UNZ_ERR        = -1",
                  "UNZ_EFOUND     = -1000", "-ENOMEM", "-EPERM", "-EBADF", "-ENOBUFS", "#define INFL_ZLIB 1"):
        assert token in hdr


def test_handle_api_without_device_is_safe():
    """Lifecycle calls that never reach the device behave as the reference's do."""
    import defl_b200 as d
    L = d.lib()
    L.infl_destroy(None)                                     # mem.c:337: NULL is a no-op
    assert L.infl_output_pos(None) == 0 and L.infl_input_pos(None) == 0
    out = np.zeros(16, dtype=np.uint8)
    h = L.infl_init(out.ctypes.data, 16, 0)
    assert h
    assert L.infl(h) == d.UNZ_NOOP                           # nothing included: infl.c:1544-1545
    L.infl_include(h, None, 0)
    assert L.infl(h) == d.UNZ_OK                             # only empty pieces: infl.c:1547-1550
    assert L.infl_resize_output(h, None, 4) == d.UNZ_ERR     # mem.c:282-283
    assert L.infl_resize_output(h, out.ctypes.data, 8) == d.UNZ_OK
    L.infl_reset(h, out.ctypes.data, 16, 1)
    assert L.infl(h) == d.UNZ_NOOP
    L.infl_destroy(h)
    assert d.infl_buf(b"", 4, 0)[0] == d.UNZ_OK              # empty input, no device needed


def test_no_cpu_fallback_in_product():
    """The product tree must not include, import, link or dlopen anything under oracle/ or tests/."""
    bad = re.compile(r'#\s*include\s*[<"][^>"]*(oracle|simt_emu|tests/)|liboracle|libdefl_ref|libemu|'
                     r'import\s+oracle|from\s+oracle|oracle_lib|dlopen')
    for dirpath, _, files in os.walk(os.path.join(ROOT, "defl_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not bad.search(txt), f
