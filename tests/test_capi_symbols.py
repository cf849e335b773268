"""The C-ABI library loads without a GPU, exports every symbol include/hgb200.h declares, and the ctypes table binds exactly those."""
import ctypes as C
import os
import re
import subprocess

import pytest

import harfang3d_b200
from harfang3d_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "hgb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return set(re.findall(r"\b(hgb_[a-z0-9_]+)\s*\(", src))


def test_header_table_and_library_agree():
    decl = header_symbols()
    assert decl == set(capi.SIGNATURES), (decl ^ set(capi.SIGNATURES))
    out = subprocess.check_output(["nm", "-D", "--defined-only", harfang3d_b200.library_path()], text=True)
    exported = set(re.findall(r"\bT (hgb_[a-z0-9_]+)", out))
    assert decl <= exported, decl - exported
    L = harfang3d_b200.lib()  # binds every symbol; no device needed
    assert L.hgb_last_error() is not None


def test_no_cpu_fallback_without_a_device():
    """hgb_create must fail loudly when there is no CUDA device (this container) instead of computing on the host"""
    L = harfang3d_b200.lib()
    if L.hgb_device_count() > 0:
        return
    try:
        harfang3d_b200.Context(0)
    except harfang3d_b200.HgbError as e:
        assert "no CPU path" in str(e) or "CUDA" in str(e)
    else:
        raise AssertionError("Context() succeeded without a CUDA device")


def test_product_does_not_touch_the_oracle():
    """nothing under harfang3d_b200/ may import, link or name oracle/ (the oracle is test infrastructure)"""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "harfang3d_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".cpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                for bad in ("hg_port", "libhg_ref", "oracle_lib", "from tests", "import tests"):
                    assert bad not in txt, (f, bad)
    out = subprocess.check_output(["ldd", harfang3d_b200.library_path()], text=True)
    assert "hg_port" not in out and "hg_ref" not in out


def test_no_contracted_float_math_in_the_kernels():
    """Bit-exactness with the reference's x86-64 build rests on FP32 multiplies and adds never being fused. -fmad=false covers the
    scalar code, but ptxas 12.9 contracts mul.rn.f32x2 -> add.rn.f32x2 into FFMA2 regardless, so the packed products of
    csrc/hgb_math.cuh must never feed a packed add directly. Guard: no FFMA2 anywhere in the library's SASS, and scalar FFMA only in
    the kernels that divide or take square roots (IEEE division and sqrt expand to FFMA sequences): the track samplers and the
    batched SetTransformWorld/LocalMatrix kernel (InverseFast, Decompose)."""
    import shutil
    import subprocess
    from harfang3d_b200 import capi
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(capi.library_path()) or not os.path.exists(cuobjdump):
        pytest.skip("library or cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", capi.library_path()], capture_output=True, text=True, check=True).stdout
    fn, ffma, packed = None, {}, 0
    for line in sass.splitlines():
        if "Function :" in line:
            fn = line.split("Function :")[1].strip()
        elif " FFMA2 " in line:
            raise AssertionError("FFMA2 (contracted packed multiply-add) in %s" % fn)
        elif " FFMA " in line or " FFMA." in line:
            ffma[fn] = ffma.get(fn, 0) + 1
        elif " FMUL2 " in line:
            packed += 1
    assert packed > 0, "the packed products are expected to compile to FMUL2"
    assert all("k_sample_tracks" in f or "k_set_matrices" in f for f in ffma), "scalar FFMA outside the dividing kernels: %s" % sorted(ffma)
