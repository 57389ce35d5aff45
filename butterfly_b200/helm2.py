"""Host-side mirror of the Helmholtz value construction ABI (include/bf_cuda.h: BfCudaHelm2Recipe,
bfCudaPlanBuildHelm2, bfCudaHelm2Evaluate).  ctypes only; the values are computed by
csrc/bf_helm2.cu on the GPU.  `Recording` holds what libbfcuda_helm2.so (csrc/bf_helm2_record.c)
noted while the reference's own fac_helm2 laid out an operator: the recipes, the deduplicated point
pool and the leaf -> recipe map."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .rawops import RawOps  # noqa: F401 (re-exported)

HELM2_LIB_PATH = capi.LIB_DIR / "libbfcuda_helm2.so"

_recorder = None


def load_recorder() -> C.CDLL:
    """dlopen libbfcuda_helm2.so (csrc/bf_helm2_record.c) into the GLOBAL symbol scope.  It must be loaded
    BEFORE libbutterfly so that its bfHelm2GetKernelMatrix / bfHelm2GetReexpansionMatrix win, and it has no
    link-time dependency on any particular libbutterfly (the three bf* functions it calls are bound on first
    use) — exactly what `-lbfcuda_helm2 -lbutterfly` gives a C program.  ctypes.CDLL forces RTLD_NOW, which
    would need those symbols at load time, hence the direct dlopen with RTLD_LAZY."""
    global _recorder
    if _recorder is None:
        import os

        libc = C.CDLL(None)
        libc.dlopen.restype = C.c_void_p
        libc.dlopen.argtypes = [C.c_char_p, C.c_int]
        libc.dlerror.restype = C.c_char_p
        handle = libc.dlopen(str(HELM2_LIB_PATH).encode(), os.RTLD_LAZY | os.RTLD_GLOBAL)
        if not handle:
            raise OSError(f"dlopen({HELM2_LIB_PATH}): {(libc.dlerror() or b'').decode()}")
        _recorder = C.CDLL(str(HELM2_LIB_PATH), handle=handle)
    return _recorder


KERNEL, REEXPANSION = 0, 1
LAYER_POT_SINGLE, LAYER_POT_PV_DOUBLE, LAYER_POT_PV_NORMAL_DERIV_SINGLE, LAYER_POT_COMBINED_FIELD = 1, 2, 3, 5

RECIPE_DTYPE = np.dtype(
    [
        ("kind", "<i4"), ("layerPot", "<i4"), ("k", "<f8"), ("alpha", "<f8", (2,)), ("beta", "<f8", (2,)),
        ("rows", "<i8"), ("cols", "<i8"), ("src", "<i8"), ("numSrc", "<i8"), ("tgt", "<i8"), ("numTgt", "<i8"),
        ("equiv", "<i8"), ("numEquiv", "<i8"), ("srcNormals", "<i8"), ("tgtNormals", "<i8"), ("equivNormals", "<i8"),
    ],
    align=True,
)
assert RECIPE_DTYPE.itemsize == 136


class Build(C.Structure):
    """struct BfCudaHelm2Build"""

    _fields_ = [
        ("numPairs", C.c_int64), ("pairs", C.c_void_p),
        ("numRecipes", C.c_int64), ("recipes", C.c_void_p),
        ("numOps", C.c_int64), ("opRecipe", C.c_void_p),
        ("truncationScale", C.c_double),
        ("opRowOffset", C.c_void_p),
    ]


class Stats(C.Structure):
    """struct BfCudaHelm2Stats"""

    _fields_ = [(n, C.c_int64) for n in ("kernelRecipes", "kernelEntries", "reexpRecipes", "reexpEntries",
                                         "reexpSmall", "reexpLarge", "maxSweeps", "notConverged")] + [
        (n, C.c_double) for n in ("kernelMs", "reexpMs", "packMs")] + [("pinvClassMs", C.c_double * 6), ("pinvClassCount", C.c_int64 * 6)]

    def as_dict(self):
        return {n: (list(getattr(self, n)) if n.startswith("pinvClass") else getattr(self, n)) for n, _ in self._fields_}


class Recording:
    """Recipes + point pool + leaf map of one recorded operator (numpy arrays, so it can be saved)."""

    def __init__(self, pairs, recipes, op_recipe, truncation_scale: float = 0.0, op_row_offset=None):
        self.truncation_scale = float(truncation_scale)  # 0: the reference's rule (BfCudaHelm2Build)
        self.op_row_offset = None if op_row_offset is None else np.ascontiguousarray(op_row_offset, dtype=np.int64)
        self.pairs = np.ascontiguousarray(pairs, dtype=np.float64).reshape(-1, 2)
        self.recipes = np.ascontiguousarray(recipes, dtype=RECIPE_DTYPE)
        self.op_recipe = np.ascontiguousarray(op_recipe, dtype=np.int64)

    def build_struct(self) -> Build:
        b = Build()
        b.numPairs = len(self.pairs); b.pairs = self.pairs.ctypes.data
        b.numRecipes = len(self.recipes); b.recipes = self.recipes.ctypes.data
        b.numOps = len(self.op_recipe); b.opRecipe = self.op_recipe.ctypes.data
        b.truncationScale = self.truncation_scale
        b.opRowOffset = None if self.op_row_offset is None else self.op_row_offset.ctypes.data
        return b

    def remapped(self, op_index, row_offset) -> "Recording":
        """The same recipes for an op list derived from the recorded one: new op i is rows [row_offset[i], ...)
        of recorded op op_index[i] (rowshard.origin_of)."""
        op_index = np.asarray(op_index, dtype=np.int64)
        base = self.op_row_offset[op_index] if self.op_row_offset is not None else 0
        return Recording(self.pairs, self.recipes, self.op_recipe[op_index], self.truncation_scale, base + np.asarray(row_offset, np.int64))

    @classmethod
    def from_build(cls, b: Build) -> "Recording":
        """Deep copy of a BfCudaHelm2Build filled by bfCudaHelm2Resolve."""
        def arr(ptr, n, dt):
            if n == 0:
                return np.zeros(0, dt)
            raw = (C.c_char * (n * np.dtype(dt).itemsize)).from_address(ptr)
            return np.frombuffer(raw, dtype=dt).copy()
        return cls(arr(b.pairs, 2 * b.numPairs, np.float64), arr(b.recipes, b.numRecipes, RECIPE_DTYPE),
                   arr(b.opRecipe, b.numOps, np.int64))

    def save_into(self, d: dict):
        d.update(h2_pairs=self.pairs, h2_recipes=self.recipes, h2_op_recipe=self.op_recipe)

    @classmethod
    def load_from(cls, z) -> "Recording":
        return cls(z["h2_pairs"], z["h2_recipes"].astype(RECIPE_DTYPE), z["h2_op_recipe"])


def _lib():
    l = capi.load_library()
    if not getattr(l, "_helm2_bound", False):
        l.bfCudaPlanBuildHelm2.argtypes = [C.c_void_p, C.POINTER(Build), C.POINTER(Stats)]
        l.bfCudaPlanBuildHelm2.restype = C.c_int32
        l.bfCudaHelm2Evaluate.argtypes = [C.POINTER(Build), C.c_int64, C.c_void_p, C.c_int32]
        l.bfCudaHelm2Evaluate.restype = C.c_int32
        l._helm2_bound = True
    return l


def build_plan(plan, rec: Recording) -> dict:
    """bfCudaPlanBuildHelm2: compute the recorded leaves' values on the GPU into the plan's pool."""
    l = _lib()
    b = rec.build_struct()
    st = Stats()
    code = l.bfCudaPlanBuildHelm2(plan.handle, C.byref(b), C.byref(st))
    if code != capi.OK:
        raise capi.BfCudaError(code, "bfCudaPlanBuildHelm2", l.bfCudaLastErrorMessage().decode())
    return st.as_dict()


def evaluate(rec: Recording, index: int, device: int = -1) -> np.ndarray:
    """bfCudaHelm2Evaluate: the matrix of one recipe, computed on the GPU."""
    l = _lib()
    r = rec.recipes[index]
    out = np.zeros((int(r["rows"]), int(r["cols"])), np.complex128)
    b = rec.build_struct()
    code = l.bfCudaHelm2Evaluate(C.byref(b), int(index), out.ctypes.data, device)
    if code != capi.OK:
        raise capi.BfCudaError(code, "bfCudaHelm2Evaluate", l.bfCudaLastErrorMessage().decode())
    return out
