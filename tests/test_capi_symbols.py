"""CPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol the header declares,
and fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    txt = open(os.path.join(ROOT, "include", "dso_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(dso_b200_[a-z_A-Z0-9]+)\s*\(", txt)))


def test_header_symbols_match_binding(pkg):
    syms = _header_symbols()
    assert len(syms) >= 25
    assert sorted(pkg.capi.SIGNATURES.keys()) == syms


def test_library_exports_every_symbol(pkg):
    L = pkg.capi.load()
    for s in _header_symbols():
        assert getattr(L, s) is not None


def test_header_is_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "dso_b200.h"\nint main(void){ dso_b200_track_result r; (void)r; return sizeof(dso_b200_precalc) == 112 ? 0 : 1; }\n')
    exe = tmp_path / "t"
    import subprocess
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    assert subprocess.call([str(exe)]) == 0


def test_struct_layouts_match(pkg, tmp_path):
    src = tmp_path / "s.c"
    src.write_text('#include <stdio.h>\n#include "dso_b200.h"\nint main(void){ printf("%zu %zu %zu\\n", sizeof(dso_b200_track_result), sizeof(dso_b200_track_opts), sizeof(dso_b200_precalc)); return 0; }\n')
    exe = tmp_path / "s"
    import subprocess
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    a, b, c = map(int, subprocess.check_output([str(exe)]).split())
    assert a == C.sizeof(pkg.capi.TrackResult) and b == C.sizeof(pkg.capi.TrackOpts) and c == C.sizeof(pkg.capi.Precalc)


def test_no_cpu_fallback(pkg):
    """without a GPU the context cannot be created and says so; with one this test is a no-op"""
    L = pkg.capi.load()
    h = C.c_void_p()
    rc = L.dso_b200_create(0, 640, 480, 2, C.byref(h))
    if rc == 0:
        L.dso_b200_destroy(h)
        pytest.skip("CUDA device present")
    assert rc == -2
    assert b"no CPU fallback" in L.dso_b200_last_error() or b"cuda" in L.dso_b200_last_error().lower()
    with pytest.raises(pkg.DsoError):
        pkg.Context(0, 640, 480, 2)


def test_product_does_not_touch_oracle():
    """the product tree must not include, link or import anything under oracle/"""
    bad = []
    pdir = os.path.join(ROOT, "dso-noted-in-chinese_b200")
    for dp, _, fs in os.walk(pdir):
        for f in fs:
            if f.endswith((".cu", ".cuh", ".h", ".cpp", ".py", "Makefile")):
                t = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"oracle_py|dso_oracle|liboracle|oracle/", t):
                    bad.append(os.path.join(dp, f))
    assert not bad, bad
