"""ctypes wrapper of the UNMODIFIED reference library (oracle/_ref/libbutterfly_ref.so, compiled
from /root/reference by oracle/Makefile) and of our thin C drivers around it
(oracle/ref_driver.c).  TEST INFRASTRUCTURE, NOT PRODUCT: only tests/, smoke() and bench.py's
cpu_baseline / --impl reference legs import this.  The prebuilt .so files travel to the GPU box;
/root/reference itself is never read at run time.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from butterfly_b200 import capi
from butterfly_b200.rawops import RawOps

HERE = Path(__file__).resolve().parent
REF_SO = HERE / "_ref" / "libbutterfly_ref.so"
DRV_SO = HERE / "_ref" / "libbf_oracle_drv.so"
REFERENCE = Path("/root/reference")  # only read where it exists (this container), never on the GPU box

_ref = _drv = _att = None


def available() -> bool:
    return REF_SO.exists() and DRV_SO.exists()


def attach_available() -> bool:
    return available() and capi.ATTACH_LIB_PATH.exists()


def load():
    global _ref, _drv
    if _drv is None:
        if not available():
            raise FileNotFoundError("oracle/_ref is not built (run `make -C oracle ref` where /root/reference exists)")
        _ref = C.CDLL(str(REF_SO), mode=C.RTLD_GLOBAL)
        d = C.CDLL(str(DRV_SO), mode=C.RTLD_GLOBAL)
        vp, i64, dbl = C.c_void_p, C.c_int64, C.c_double
        d.refHelm2Build.argtypes = [vp, i64, dbl]; d.refHelm2Build.restype = vp
        d.refHelm2GetMat.argtypes = [vp]; d.refHelm2GetMat.restype = vp
        d.refHelm2GetPerm.argtypes = [vp, vp]; d.refHelm2GetPerm.restype = None
        d.refHelm2DenseKernel.argtypes = [vp, vp]; d.refHelm2DenseKernel.restype = None
        d.refHelm2Free.argtypes = [vp]; d.refHelm2Free.restype = None
        d.refMatNumRows.argtypes = [vp]; d.refMatNumRows.restype = i64
        d.refMatNumCols.argtypes = [vp]; d.refMatNumCols.restype = i64
        d.refMatMulComplex.argtypes = [vp, vp, i64, i64, vp]; d.refMatMulComplex.restype = dbl
        d.refMatMulVecReal.argtypes = [vp, vp, i64, vp, C.c_int]; d.refMatMulVecReal.restype = dbl
        d.refPermuteRowsComplex.argtypes = [vp, i64, i64, vp]; d.refPermuteRowsComplex.restype = None
        d.refReversePerm.argtypes = [vp, i64, vp]; d.refReversePerm.restype = None
        d.refSeed.argtypes = [C.c_uint64]; d.refSeed.restype = None
        d.refComplexRandn.argtypes = [i64, vp]; d.refComplexRandn.restype = None
        d.refRealRandn.argtypes = [i64, vp]; d.refRealRandn.restype = None
        d.refOpenblasThreads.argtypes = [C.c_int]; d.refOpenblasThreads.restype = C.c_int
        d.refPrintBlocks.argtypes = [vp, C.c_char_p]; d.refPrintBlocks.restype = None
        d.refHelm2MakeSecondKind.argtypes = [vp, dbl]; d.refHelm2MakeSecondKind.restype = None
        d.refSolveGMRES.argtypes = [vp, vp, i64, dbl, i64, vp, vp]; d.refSolveGMRES.restype = dbl
        d.refCallGMRES.argtypes = [vp, vp, vp, i64, dbl, i64, vp, vp]; d.refCallGMRES.restype = dbl
        d.refMatRmulComplex.argtypes = [vp, vp, i64, i64, vp]; d.refMatRmulComplex.restype = dbl
        d.refMatMulVecRealStrided.argtypes = [vp, vp, i64, i64, vp, C.c_int]; d.refMatMulVecRealStrided.restype = dbl
        d.refMatMulVecComplexStrided.argtypes = [vp, vp, i64, i64, vp, C.c_int]; d.refMatMulVecComplexStrided.restype = dbl
        d.refMatGetView.argtypes = [vp]; d.refMatGetView.restype = vp
        d.refMatBlockDenseGetViewDirect.argtypes = [vp]; d.refMatBlockDenseGetViewDirect.restype = vp
        d.refMatCopy.argtypes = [vp]; d.refMatCopy.restype = vp
        d.refMatDelete.argtypes = [vp]; d.refMatDelete.restype = None
        d.refMatTranspose.argtypes = [vp]; d.refMatTranspose.restype = None
        _drv = d
    return _drv


def load_attach():
    """libbfcuda_attach.so: the C flattener + vtable attach, resolved against the loaded reference."""
    global _att
    if _att is None:
        load()
        capi.load_library()
        a = C.CDLL(str(capi.ATTACH_LIB_PATH), mode=C.RTLD_GLOBAL)
        a.bfCudaFlatten.argtypes = [C.c_void_p, C.POINTER(C.POINTER(capi.PlanDesc))]; a.bfCudaFlatten.restype = C.c_int32
        a.bfCudaFlattenEx.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.POINTER(capi.PlanDesc))]; a.bfCudaFlattenEx.restype = C.c_int32
        a.bfCudaFlattenFree.argtypes = [C.POINTER(capi.PlanDesc)]; a.bfCudaFlattenFree.restype = None

        class AttachOpts(C.Structure):
            _fields_ = [("device", C.c_int32), ("buildTranspose", C.c_int32), ("flattenFlags", C.c_uint32)]

        a.AttachOpts = AttachOpts
        a.bfCudaMatAttach.argtypes = [C.c_void_p, C.POINTER(AttachOpts)]; a.bfCudaMatAttach.restype = C.c_int32
        a.bfCudaMatDetach.argtypes = [C.c_void_p]; a.bfCudaMatDetach.restype = C.c_int32
        a.bfCudaMatGetPlan.argtypes = [C.c_void_p]; a.bfCudaMatGetPlan.restype = C.c_void_p
        a.bfCudaMatAttachSharded.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(AttachOpts)]; a.bfCudaMatAttachSharded.restype = C.c_int32
        a.bfCudaMatGetShardedPlan.argtypes = [C.c_void_p]; a.bfCudaMatGetShardedPlan.restype = C.c_void_p
        _att = a
    return _att


def circle_points(n: int) -> np.ndarray:
    """examples/simple/make_circle_test_data.py:12-15"""
    t = 2 * np.pi * np.arange(n) / n
    return np.ascontiguousarray(np.stack([np.cos(t), np.sin(t)], axis=1))


class Helm2:
    """bfFacHelm2MakeMultilevel on a point set, as examples/simple/bf_all_blocks.c:96-130."""

    def __init__(self, points: np.ndarray, k_wave: float):
        self.drv = load()
        pts = np.ascontiguousarray(points, dtype=np.float64)
        self.n = len(pts)
        self.h = self.drv.refHelm2Build(pts.ctypes.data, self.n, float(k_wave))
        self.mat = self.drv.refHelm2GetMat(self.h)

    def close(self):
        if self.h:
            self.drv.refHelm2Free(self.h)
            self.h = None

    def perm(self) -> np.ndarray:
        p = np.zeros(self.n, np.int64)
        self.drv.refHelm2GetPerm(self.h, p.ctypes.data)
        return p

    def dense_kernel(self) -> np.ndarray:
        K = np.zeros((self.n, self.n), np.complex128)
        self.drv.refHelm2DenseKernel(self.h, K.ctypes.data)
        return K

    def matmul(self, x: np.ndarray):
        """bfMatMul(A, X) through the object's CURRENT vtable (reference CPU path, or the GPU path
        once attached).  Returns (Y, seconds inside bfMatMul)."""
        x = np.ascontiguousarray(x, dtype=np.complex128)
        if x.ndim == 1:
            x = x[:, None]
        y = np.zeros((self.drv.refMatNumRows(self.mat), x.shape[1]), np.complex128)
        t = self.drv.refMatMulComplex(self.mat, x.ctypes.data, x.shape[0], x.shape[1], y.ctypes.data)
        return y, t

    def rmul(self, x: np.ndarray, mat=None):
        """bfMatRmul(A, X) = X A through the CURRENT vtable (src/mat_product.c:282-310); X is p x numRows(A)."""
        mat = mat or self.mat
        x = np.ascontiguousarray(x, dtype=np.complex128)
        y = np.zeros((x.shape[0], self.drv.refMatNumCols(mat)), np.complex128)
        t = self.drv.refMatRmulComplex(mat, x.ctypes.data, x.shape[0], x.shape[1], y.ctypes.data)
        return y, t

    def matmul_on(self, mat, x: np.ndarray):
        """bfMatMul on another BfMat* of the same shape (a view or a copy of this operator)."""
        x = np.ascontiguousarray(x, dtype=np.complex128)
        if x.ndim == 1:
            x = x[:, None]
        y = np.zeros((self.drv.refMatNumRows(mat), x.shape[1]), np.complex128)
        self.drv.refMatMulComplex(mat, x.ctypes.data, x.shape[0], x.shape[1], y.ctypes.data)
        return y

    def mulvec_complex(self, x: np.ndarray, stride: int = 1, transposed: bool = False):
        """bfMatMulVec / bfMatRmulVec with a BfVecComplex view of the given stride.  The reference's block
        types refuse the operand type: only valid on an attached object."""
        n = self.drv.refMatNumRows(self.mat) if transposed else self.drv.refMatNumCols(self.mat)
        buf = np.zeros(n * stride, np.complex128)
        buf[::stride] = x
        y = np.zeros(self.drv.refMatNumCols(self.mat) if transposed else self.drv.refMatNumRows(self.mat), np.complex128)
        t = self.drv.refMatMulVecComplexStrided(self.mat, buf.ctypes.data, n, stride, y.ctypes.data, int(transposed))
        if t < 0:
            raise RuntimeError("bfMatMulVec returned NULL")
        return y

    def transpose(self):
        """bfMatTranspose in place (for complex leaves: the conjugate transpose, src/mat_dense_complex.c:1475-1478)."""
        self.drv.refMatTranspose(self.mat)

    def get_view(self, direct: bool = False):
        """bfMatGetView(A) (what bfMatGet(A, BF_POLICY_VIEW) calls, multiple_scattering_context.c:616), or the
        typed constructor bfMatBlockDenseGetView called directly (bypasses the vtable)."""
        return self.drv.refMatBlockDenseGetViewDirect(self.mat) if direct else self.drv.refMatGetView(self.mat)

    def copy(self):
        return self.drv.refMatCopy(self.mat)

    def delete_mat(self, mat):
        self.drv.refMatDelete(mat)

    def make_second_kind(self, weight: float | None = None):
        """K <- K diag(w) + I/2 with the reference's own bfMatScaleCols / bfMatAddInplace
        (examples/multiple_scattering/multiple_scattering_context.c:456-487); w = trapezoid weight
        2 pi / n of the unit circle by default."""
        self.drv.refHelm2MakeSecondKind(self.h, float(2 * np.pi / self.n if weight is None else weight))

    def apply_kr_correction(self, order: int = 6):
        """bfHelm2ApplyKrCorrectionTree (src/helm2.c:398-414): Kapur-Rokhlin correction entries added to
        the operator in place; butterflied blocks they touch become MatSum [MatProduct, MatCooComplex]."""
        self.drv.refHelm2ApplyKrCorrection.argtypes = [C.c_void_p, C.c_int64]
        self.drv.refHelm2ApplyKrCorrection.restype = None
        self.drv.refHelm2ApplyKrCorrection(self.h, int(order))

    def sum_with_coo(self, rows: np.ndarray, cols: np.ndarray, values: np.ndarray):
        """A <- MatSum [A, C] for an n x n MatCooComplex C in tree order (the shape of a butterfly with
        sparse correction entries, src/mat_block_dense.c:493-499)."""
        r = np.ascontiguousarray(rows, dtype=np.int64)
        c = np.ascontiguousarray(cols, dtype=np.int64)
        v = np.ascontiguousarray(values, dtype=np.complex128)
        self.drv.refHelm2SumWithCoo.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        self.drv.refHelm2SumWithCoo.restype = None
        self.drv.refHelm2SumWithCoo(self.h, len(r), r.ctypes.data, c.ctypes.data, v.ctypes.data)
        self.mat = self.drv.refHelm2GetMat(self.h)

    def solve_gmres(self, b: np.ndarray, tol: float = 1e-12, max_iter: int = 50):
        """bfSolveGMRES(A, B, NULL, tol, maxNumIter, &numIter, NULL) (src/linalg.c:47) through the
        object's CURRENT vtable.  Returns (x, num_iter, seconds)."""
        b = np.ascontiguousarray(b, dtype=np.complex128).reshape(-1)
        x = np.zeros_like(b)
        it = C.c_int64()
        t = self.drv.refSolveGMRES(self.mat, b.ctypes.data, len(b), float(tol), int(max_iter), x.ctypes.data, C.byref(it))
        return x, it.value, t

    def make_block_jacobi(self, num_blocks: int = 8, weight: float | None = None):
        """The block Jacobi left preconditioner of multiple_scattering_context.c:544-600 for the
        second-kind operator: a MatBlockDiag of LU views (bfMatSolve = LU solves).  Returns a BfMat*."""
        self.drv.refHelm2MakeBlockJacobi.argtypes = [C.c_void_p, C.c_int64, C.c_double]
        self.drv.refHelm2MakeBlockJacobi.restype = C.c_void_p
        return self.drv.refHelm2MakeBlockJacobi(self.h, int(num_blocks), float(2 * np.pi / self.n if weight is None else weight))

    def solve_gmres_precond(self, b: np.ndarray, M, tol: float = 1e-12, max_iter: int = 50, cuda: bool = False):
        """bfSolveGMRES(A, B, NULL, tol, maxNumIter, &numIter, M) — or bfCudaSolveGMRESMat with the same
        arguments when `cuda` (operator attached).  Returns (x, num_iter, seconds)."""
        d = self.drv
        d.refCallGMRESPrecond.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_int64, C.c_void_p, C.POINTER(C.c_int64)]
        d.refCallGMRESPrecond.restype = C.c_double
        fn = C.cast(load_attach().bfCudaSolveGMRESMat if cuda else _ref.bfSolveGMRES, C.c_void_p)
        b = np.ascontiguousarray(b, dtype=np.complex128).reshape(-1)
        x = np.zeros_like(b)
        it = C.c_int64()
        t = d.refCallGMRESPrecond(fn, self.mat, M, b.ctypes.data, len(b), float(tol), int(max_iter), x.ctypes.data, C.byref(it))
        return x, it.value, t

    def solve_gmres_cuda(self, b: np.ndarray, tol: float = 1e-12, max_iter: int = 50):
        """bfCudaSolveGMRESMat (include/bf_cuda_attach.h): the bfSolveGMRES-shaped entry point of the
        attach layer, called with reference objects.  The operator must be attached."""
        a = load_attach()
        b = np.ascontiguousarray(b, dtype=np.complex128).reshape(-1)
        x = np.zeros_like(b)
        it = C.c_int64()
        fn = C.cast(a.bfCudaSolveGMRESMat, C.c_void_p)
        t = self.drv.refCallGMRES(fn, self.mat, b.ctypes.data, len(b), float(tol), int(max_iter), x.ctypes.data, C.byref(it))
        return x, it.value, t

    def flatten(self, flags: int = 0) -> RawOps:
        a = load_attach()
        d = C.POINTER(capi.PlanDesc)()
        st = a.bfCudaFlattenEx(self.mat, flags, C.byref(d))
        if st != 0:
            raise RuntimeError(f"bfCudaFlatten: status {st}")
        try:
            return RawOps.from_desc(d.contents)
        finally:
            a.bfCudaFlattenFree(d)

    def attach(self, build_transpose=0, flatten_flags=0):
        a = load_attach()
        o = a.AttachOpts(-1, build_transpose, flatten_flags)
        st = a.bfCudaMatAttach(self.mat, C.byref(o))
        if st != 0:
            raise capi.BfCudaError(st, "bfCudaMatAttach", capi.load_library().bfCudaLastErrorMessage().decode())

    def attach_sharded(self, comm, mode=capi.SHARD_BLOCK_ROWS, build_transpose=0):
        """bfCudaMatAttachSharded: the operator over the GPUs of `comm` (butterfly_b200.sharded.Comm), behind the
        same vtable."""
        a = load_attach()
        o = a.AttachOpts(-1, build_transpose, 0)
        st = a.bfCudaMatAttachSharded(self.mat, comm.handle, mode, C.byref(o))
        if st != 0:
            raise capi.BfCudaError(st, "bfCudaMatAttachSharded", capi.load_library().bfCudaLastErrorMessage().decode())

    def plan_handle(self):
        """bfCudaMatGetPlan: the plan behind the attached object (borrowed)."""
        h = load_attach().bfCudaMatGetPlan(self.mat)
        if not h:
            raise RuntimeError("object is not attached")
        return h

    def detach(self):
        st = load_attach().bfCudaMatDetach(self.mat)
        if st != 0:
            raise capi.BfCudaError(st, "bfCudaMatDetach")


def randn_complex(n: int, seed: int = 0) -> np.ndarray:
    """bfSeed(seed); bfComplexRandn(n) — the examples' RHS generator (src/rand.c:19-76)."""
    d = load()
    out = np.zeros(n, np.complex128)
    d.refSeed(seed)
    d.refComplexRandn(n, out.ctypes.data)
    return out


def permute_rows_complex(x: np.ndarray, perm: np.ndarray) -> np.ndarray:
    """bfMatPermuteRows on a copy (src/mat_dense_complex.c:741-774)."""
    d = load()
    y = np.ascontiguousarray(x if x.ndim == 2 else x[:, None], dtype=np.complex128).copy()
    p = np.ascontiguousarray(perm, dtype=np.int64)
    d.refPermuteRowsComplex(y.ctypes.data, y.shape[0], y.shape[1], p.ctypes.data)
    return y if x.ndim == 2 else y[:, 0]


def reverse_perm(perm: np.ndarray) -> np.ndarray:
    d = load()
    p = np.ascontiguousarray(perm, dtype=np.int64)
    r = np.zeros_like(p)
    d.refReversePerm(p.ctypes.data, len(p), r.ctypes.data)
    return r


def set_blas_threads(n: int) -> int:
    return load().refOpenblasThreads(n)


# ---- streamed real butterfly (LBO-style) -------------------------------
def fibonacci_sphere(n: int) -> np.ndarray:
    """n quasi-uniform points on the unit sphere (as tests/generate_data_for_test_linalg.py:29-32)."""
    i = np.arange(n) + 0.5
    phi = np.arccos(1 - 2 * i / n)
    theta = np.pi * (1 + 5 ** 0.5) * i
    return np.ascontiguousarray(np.stack([np.cos(theta) * np.sin(phi), np.sin(theta) * np.sin(phi), np.cos(phi)], axis=1))


def real_sph_harm_band(points: np.ndarray, degrees) -> tuple[np.ndarray, np.ndarray]:
    """Real spherical harmonics of the given degrees on unit-sphere points: the exact LBO
    eigenfunctions of the sphere (eigenvalue l(l+1), frequency sqrt(l(l+1))).  Returns (n x m band,
    m frequencies ascending)."""
    from scipy.special import sph_harm_y

    x, y, z = points.T
    polar = np.arccos(np.clip(z, -1, 1))
    azim = np.arctan2(y, x)
    cols, freqs = [], []
    for l in degrees:
        for m in range(-l, l + 1):
            Y = sph_harm_y(l, abs(m), polar, azim)
            if m == 0:
                c = Y.real
            elif m > 0:
                c = np.sqrt(2) * (-1) ** m * Y.real
            else:
                c = np.sqrt(2) * (-1) ** m * Y.imag
            cols.append(c)
            freqs.append(np.sqrt(l * (l + 1.0)))
    if not cols:
        return np.zeros((len(points), 0)), np.zeros(0)
    return np.ascontiguousarray(np.stack(cols, axis=1)), np.asarray(freqs)


class Stream:
    """bfFacStreamer fed with caller-supplied eigenbands (oracle/ref_driver.c refStream*: the body
    of bfLboFeedFacStreamerNextEigenband, src/lbo.c:96-137, minus the ARPACK call).  `band_fn(a, b,
    leftmost, rightmost)` returns (n x m band in original point order, m ascending frequencies in
    [a, b))."""

    def __init__(self, points: np.ndarray, freq_max: float, band_fn, freq_tree_depth: int = -1, tol: float = 1e-3,
                 row_tree_offset: int = 0, min_rows: int = 20, min_cols: int = 20):
        self.drv = d = load()
        vp, i64, dbl = C.c_void_p, C.c_int64, C.c_double
        d.refStreamBegin.argtypes = [vp, i64, dbl, i64, dbl, i64, i64, i64]; d.refStreamBegin.restype = vp
        d.refStreamRowTreeDepth.argtypes = [vp]; d.refStreamRowTreeDepth.restype = i64
        d.refStreamGetRowPerm.argtypes = [vp, vp]; d.refStreamGetRowPerm.restype = None
        d.refStreamIsDone.argtypes = [vp]; d.refStreamIsDone.restype = C.c_int
        d.refStreamCurrentBand.argtypes = [vp, vp, vp, vp, vp]; d.refStreamCurrentBand.restype = C.c_int
        d.refStreamFeed.argtypes = [vp, vp, i64, vp]; d.refStreamFeed.restype = C.c_int
        d.refStreamGetMat.argtypes = [vp]; d.refStreamGetMat.restype = vp
        d.refStreamNumFreqs.argtypes = [vp]; d.refStreamNumFreqs.restype = i64
        pts = np.ascontiguousarray(points, dtype=np.float64)
        self.n = len(pts)
        self.h = d.refStreamBegin(pts.ctypes.data, self.n, float(freq_max), int(freq_tree_depth), float(tol),
                                  int(row_tree_offset), int(min_rows), int(min_cols))
        self.bands = []
        while not d.refStreamIsDone(self.h):
            a, b = C.c_double(), C.c_double()
            lm, rm = C.c_int(), C.c_int()
            if d.refStreamCurrentBand(self.h, C.byref(a), C.byref(b), C.byref(lm), C.byref(rm)):
                raise RuntimeError("streamer has no current column leaf")
            band, freqs = band_fn(a.value, b.value, bool(lm.value), bool(rm.value))
            band = np.ascontiguousarray(band, dtype=np.float64)
            freqs = np.ascontiguousarray(freqs, dtype=np.float64)
            rc = d.refStreamFeed(self.h, band.ctypes.data, band.shape[1], freqs.ctypes.data)
            if rc:
                raise RuntimeError(f"refStreamFeed: {rc} (band [{a.value}, {b.value}) with {band.shape[1]} columns)")
            self.bands.append((a.value, b.value, band.shape[1]))
        self.mat = d.refStreamGetMat(self.h)
        self.num_rows = d.refMatNumRows(self.mat)
        self.num_cols = d.refMatNumCols(self.mat)

    def row_perm(self) -> np.ndarray:
        p = np.zeros(self.n, np.int64)
        self.drv.refStreamGetRowPerm(self.h, p.ctypes.data)
        return p

    def mulvec_strided(self, x: np.ndarray, stride: int, transposed: bool = False):
        """The same calls with a BfVecReal VIEW of stride `stride` (bfVecRealInitView)."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        buf = np.full(len(x) * stride, np.nan)
        buf[::stride] = x
        y = np.zeros(self.num_cols if transposed else self.num_rows)
        t = self.drv.refMatMulVecRealStrided(self.mat, buf.ctypes.data, len(x), stride, y.ctypes.data, int(transposed))
        return y, t

    def mulvec(self, x: np.ndarray, transposed: bool = False):
        """bfMatMulVec / bfMatRmulVec through the object's CURRENT vtable.  Returns (y, seconds)."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.zeros(self.num_cols if transposed else self.num_rows)
        t = self.drv.refMatMulVecReal(self.mat, x.ctypes.data, len(x), y.ctypes.data, int(transposed))
        return y, t

    flatten = Helm2.flatten
    attach = Helm2.attach
    detach = Helm2.detach
    plan_handle = Helm2.plan_handle
    attach_sharded = Helm2.attach_sharded


def sphere_stream(n: int = 4000, lmax: int = 31, tol: float = 1e-3, freq_tree_depth: int = -1) -> "Stream":
    """The survey's streamed-factorization probe: n Fibonacci-sphere points, real spherical harmonics
    of degree <= lmax grouped into the frequency tree's leaf bands."""
    pts = fibonacci_sphere(n)
    fmax = np.sqrt(lmax * (lmax + 1.0)) + 0.25

    def band_fn(a, b, leftmost, rightmost):
        degs = [l for l in range(lmax + 1) if (leftmost or np.sqrt(l * (l + 1.0)) >= a) and (rightmost or np.sqrt(l * (l + 1.0)) < b)]
        return real_sph_harm_band(pts, degs)

    s = Stream(pts, fmax, band_fn, freq_tree_depth=freq_tree_depth, tol=tol)
    s.points = pts
    return s


def sphere_fixture_stream(tol: float = 1e-3, freq_tree_depth: int = 2, min_rows: int = 20) -> "Stream":
    """bfFacStreamer fed with the reference's OWN LBO test fixture: the 500 vertices of tests/sphere.obj
    and the 32 eigenpairs tests/sphere_{Lam,Phi}.txt that tests/test_linalg.c:21-22 checks its ARPACK
    solve against.  Frequencies are sqrt(Lam) (src/lbo.c:117-125); the frequency-tree leaves below the
    lowest fixture frequency receive empty bands."""
    ref = Path(REFERENCE) / "tests"
    lam = np.loadtxt(ref / "sphere_Lam.txt")
    phi = np.loadtxt(ref / "sphere_Phi.txt")
    pts = np.array([[float(t) for t in ln.split()[1:4]] for ln in open(ref / "sphere.obj") if ln.startswith("v ")])
    freqs = np.sqrt(lam)

    def band_fn(a, b, leftmost, rightmost):
        sel = [i for i in range(len(freqs)) if (leftmost or freqs[i] >= a) and (rightmost or freqs[i] < b)]
        return phi[:, sel], freqs[sel]

    s = Stream(pts, float(freqs.max()) * 1.0001, band_fn, freq_tree_depth=freq_tree_depth, tol=tol, min_rows=min_rows, min_cols=1)
    s.points, s.phi, s.freqs = pts, phi, freqs
    return s
