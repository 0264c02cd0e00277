"""The C-ABI shared library loads without a GPU and exports exactly the entry points that
include/sdmm_b200.h declares (no compute calls here)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = set()
    for hdr in ("sdmm_b200.h", "sdmm_b200_render.h"):
        src = open(os.path.join(ROOT, "include", hdr)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names |= set(re.findall(r"\b(sdmmb_[a-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_header_declares_what_capi_binds():
    from sdmm_mitsuba_b200 import capi
    assert _declared() == sorted(capi.SYMBOLS)


def test_library_exports_every_declared_symbol():
    from sdmm_mitsuba_b200 import capi
    assert os.path.exists(capi.LIB_PATH), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    L = ctypes.CDLL(capi.LIB_PATH)
    for name in _declared():
        assert hasattr(L, name), name


def test_error_strings_and_argument_checks_without_gpu():
    """Entry points validate arguments before touching CUDA, so these run on a CPU-only box."""
    from sdmm_mitsuba_b200 import capi
    L = capi.lib()
    assert L.sdmmb_get_error_string(0) == b"ok"
    assert L.sdmmb_get_error_string(-7) == b"Data full."
    assert L.sdmmb_create(None, 0, 1, 16) == -1          # SDMMB_E_INVALID: null out pointer
    h = ctypes.c_void_p()
    assert L.sdmmb_create(ctypes.byref(h), 0, 0, 16) == -1
    assert L.sdmmb_create(ctypes.byref(h), 0, 4, 129) == -1   # K above SDMMB_MAX_K
    assert L.sdmmb_em_step(None, 0, 0, 1, 0, None, None) == -1
    assert L.sdmmb_destroy(None) == 0


def test_product_has_no_oracle_or_cpu_fallback():
    """Nothing under the product package imports the oracle (test infrastructure only)."""
    pkg = os.path.join(ROOT, "sdmm_mitsuba_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".inc", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "from oracle" not in text and "import oracle" not in text, f
                assert "sdmm_oracle" not in text and "liboracle" not in text, f
