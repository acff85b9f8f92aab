"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/mtsb200.h
declares, and refuses to work without a GPU (no CPU fallback)."""
import ctypes
import os
import subprocess

import pytest


def test_library_exports_every_declared_symbol(pkg):
    L = pkg.binding.lib()
    names = pkg.binding.header_symbols()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, "declared in include/mtsb200.h but not exported: %s" % missing
    assert L.mtsb_abi_version() == 1


def test_header_compiles_as_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "mtsb200.h"\nint main(void){ mtsb_ray r; mtsb_hit h; (void)r; (void)h; return sizeof(mtsb_ray)==32 && sizeof(mtsb_hit)==16 ? 0 : 1; }\n')
    inc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include")
    exe = tmp_path / "t"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", inc, str(src), "-o", str(exe)])
    assert subprocess.call([str(exe)]) == 0


def test_struct_sizes_match_binding(pkg):
    B = pkg.binding
    assert ctypes.sizeof(B.BsdfDesc) == 64
    assert ctypes.sizeof(B.EmitterDesc) == 36
    assert ctypes.sizeof(B.Camera) == 152      # + aperture_radius, focus_distance (thinlens)
    assert ctypes.sizeof(B.FilmDesc) == 172
    assert ctypes.sizeof(B.IntegratorDesc) == 28


def test_no_cpu_fallback(pkg):
    """Without a CUDA device every entry point must fail loudly."""
    B = pkg.binding
    if B.device_count() > 0:
        pytest.skip("a GPU is present; the failure path is exercised on the CPU box")
    with pytest.raises(B.MtsbError) as e:
        B.Context(0)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_product_code_never_imports_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    bad = []
    for dp, _, files in os.walk(os.path.join(root, "mitsuba-im_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dp, f), errors="replace").read()
                if "import oracle" in text or "from oracle" in text or "liboracle" in text or "restate.c" in text or "oracle.py" in text:
                    bad.append(os.path.join(dp, f))
    assert not bad, "product files reference the oracle: %s" % bad
