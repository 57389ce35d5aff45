"""Lay out the reference's Helmholtz HODBF for N unit-circle points at wavenumber K INSIDE A RECORDING
(butterfly_b200/csrc/bf_helm2_record.c) and write the flattened structure together with the recorded
value recipes to an .npz file: what bfCudaPlanBuildHelm2 needs to compute the operator's values on the
GPU.

Run as a subprocess: libbfcuda_helm2.so must enter the global symbol scope BEFORE the reference
library so that its bfHelm2GetKernelMatrix / bfHelm2GetReexpansionMatrix are the ones fac_helm2
calls.  The reference's own quadtree + bfFacHelm2MakeMultilevel decide every block shape, offset, rank
and point set; no Hankel function or SVD is evaluated on the host.

usage: python tools/helm2_record.py N K OUT.npz
"""
import ctypes as C
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    n, k, out = int(sys.argv[1]), float(sys.argv[2]), sys.argv[3]
    import numpy as np

    from butterfly_b200 import capi, helm2

    rec_lib = helm2.load_recorder()  # first: its two builders win
    from oracle import reflib

    att = reflib.load_attach()
    rec_lib.bfCudaHelm2RecordEnd.argtypes = [C.POINTER(C.c_void_p)]
    rec_lib.bfCudaHelm2Resolve.argtypes = [C.c_void_p, C.POINTER(capi.PlanDesc), C.POINTER(helm2.Build)]
    rec_lib.bfCudaHelm2RecordingFree.argtypes = [C.c_void_p]

    t0 = time.time()
    assert rec_lib.bfCudaHelm2RecordBegin() == 0
    h = reflib.Helm2(reflib.circle_points(n), k)
    handle = C.c_void_p()
    assert rec_lib.bfCudaHelm2RecordEnd(C.byref(handle)) == 0
    t_layout = time.time() - t0

    d = C.POINTER(capi.PlanDesc)()
    assert att.bfCudaFlattenEx(h.mat, 0, C.byref(d)) == 0
    b = helm2.Build()
    st = rec_lib.bfCudaHelm2Resolve(handle, d, C.byref(b))
    assert st == 0, f"bfCudaHelm2Resolve: {st}"
    raw = helm2.RawOps.from_desc(d.contents)
    assert raw.values is None
    rec = helm2.Recording.from_build(b)
    assert (rec.op_recipe >= 0).all(), "a leaf of the operator was not produced by a recorded builder call"
    z = dict(dtype=raw.dtype, num_rows=raw.num_rows, num_cols=raw.num_cols, buf_rows=raw.buf_rows, ops=raw.ops,
             values=np.zeros(0), has_values=False, perm=h.perm(), seconds=time.time() - t0, layout_seconds=t_layout,
             points=reflib.circle_points(n), k_wave=k)
    rec.save_into(z)
    np.savez(out, **z)
    kinds = rec.recipes["kind"]
    print(f"helm2 recording N={n} k={k}: {raw.num_ops} leaves ({int((kinds == 0).sum())} kernel, "
          f"{int((kinds == 1).sum())} re-expansion recipes), {len(rec.pairs)} pooled points, "
          f"E={raw.num_entries} in {time.time() - t0:.1f}s")
    sys.stdout.flush()
    os._exit(0)  # the reference's teardown frees millions of tiny nodes; skip it


if __name__ == "__main__":
    main()
