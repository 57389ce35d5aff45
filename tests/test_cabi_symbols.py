"""CPU: the C-ABI library loads and exports every symbol include/*.h declares (no compute)."""
import ctypes as C
import re
from pathlib import Path

from butterfly_b200 import capi

ROOT = Path(__file__).resolve().parent.parent


def _declared(header):
    txt = (ROOT / "include" / header).read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(bfCuda[A-Za-z0-9]+)\s*\(", txt)))


def test_libbfcuda_exports_everything_bf_cuda_h_declares(lib):
    names = _declared("bf_cuda.h") + _declared("bf_cuda_debug.h")
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert sorted(set(capi.EXPORTED_SYMBOLS + capi.DEBUG_EXPORTED_SYMBOLS)) == sorted(names)
    assert lib.bfCudaAbiVersion() == capi.ABI_VERSION


def test_attach_library_exports(lib):
    import pytest

    if not capi.ATTACH_LIB_PATH.exists():
        pytest.skip("attach layer is only built where the reference headers exist")
    nm = lambda p: __import__("subprocess").run(["nm", "-D", "--defined-only", str(p)], capture_output=True, text=True).stdout
    out = nm(capi.ATTACH_LIB_PATH)
    helm2 = nm(capi.LIB_DIR / "libbfcuda_helm2.so")
    for n in _declared("bf_cuda_attach.h"):
        # the recorder half of the header lives in libbfcuda_helm2.so (it must precede libbutterfly)
        where = helm2 if n in capi.HELM2_EXPORTED_SYMBOLS else out
        assert f" T {n}" in where, n
    for n in capi.ATTACH_EXPORTED_SYMBOLS:
        assert f" T {n}" in out, n
    for n in capi.HELM2_EXPORTED_SYMBOLS:
        assert f" T {n}" in helm2, n


def test_no_device_fails_loudly(lib):
    """On a machine without a GPU plan creation must fail with NO_DEVICE: there is no CPU fallback."""
    import torch

    if torch.cuda.is_available():
        return
    from butterfly_b200 import synth

    raw = synth.random_rect(50, 40, capi.REAL_F64, seed=1)
    d = raw.desc()
    h = C.c_void_p()
    assert lib.bfCudaPlanCreate(C.byref(d), C.byref(h)) == 5  # BF_CUDA_ERROR_NO_DEVICE
    assert b"no CPU fallback" in lib.bfCudaLastErrorMessage()
