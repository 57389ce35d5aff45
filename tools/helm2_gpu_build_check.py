"""The reference-facing flow of the on-device construction, end to end, in one process:

    bfCudaHelm2RecordBegin(); A = bfFacHelm2MakeMultilevel(...); bfCudaHelm2RecordEnd(&rec);
    bfCudaMatAttachHelm2(A, rec, ...);  Y = bfMatMul(A, X)

next to the unmodified reference building the same operator with its own value builders (the recorder
forwards to them outside a recording) and to the dense kernel matrix, which is what the reference's own
check compares against (examples/simple/bf_all_blocks.c:149-153).  Prints one JSON line.  A separate
process because libbfcuda_helm2.so has to enter the symbol scope before the reference library.

usage: python tools/helm2_gpu_build_check.py N K [--no-dense] [--no-reference] [--sample-rows R] [--truncation-scale S]
       [--geometry circle|ellipse|kite]   (default circle; the others are non-uniformly sampled closed curves)
"""
import ctypes as C
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    n, k = int(sys.argv[1]), float(sys.argv[2])
    import numpy as np

    from butterfly_b200 import capi, helm2

    rec_lib = helm2.load_recorder()
    from oracle import reflib

    att = reflib.load_attach()
    rec_lib.bfCudaHelm2RecordEnd.argtypes = [C.POINTER(C.c_void_p)]
    rec_lib.bfCudaHelm2RecordingFree.argtypes = [C.c_void_p]
    att.bfCudaMatAttachHelm2.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(helm2.Stats)]
    att.bfCudaMatAttachHelm2.restype = C.c_int32
    geom = sys.argv[sys.argv.index("--geometry") + 1] if "--geometry" in sys.argv else "circle"
    if geom == "circle":
        pts = reflib.circle_points(n)
    else:
        # non-uniform parameter spacing: clusters of points, uneven leaf boxes, unequal ranks
        t = 2 * np.pi * (np.arange(n) + 0.3 * np.sin(6 * np.pi * np.arange(n) / n) * n / (6 * np.pi)) / n
        if geom == "ellipse":
            pts = np.stack([1.5 * np.cos(t), 0.6 * np.sin(t)], axis=1)
        else:  # the "kite" of the boundary-integral literature
            pts = np.stack([np.cos(t) + 0.65 * np.cos(2 * t) - 0.65, 1.5 * np.sin(t)], axis=1)
        pts = np.ascontiguousarray(pts)
    out_geom = geom
    x = reflib.randn_complex(n, 0)
    X = np.stack([x, np.conj(x), 1j * np.roll(x, 3)], axis=1)
    out = {"n": n, "k": k, "geometry": out_geom}

    # create the CUDA context outside the timed build (a fixed per-process cost, reported on its own)
    tc = time.time()
    a, b = C.c_double(), C.c_double()
    capi.load_library().bfCudaMeasureFp64Peak(-1, C.byref(a), C.byref(b))
    out["cuda_context_s"] = time.time() - tc

    t0 = time.time()
    assert rec_lib.bfCudaHelm2RecordBegin() == 0
    h = reflib.Helm2(pts, k)                      # the reference's own layout code, no values
    handle = C.c_void_p()
    assert rec_lib.bfCudaHelm2RecordEnd(C.byref(handle)) == 0
    out["layout_s"] = time.time() - t0
    if "--truncation-scale" in sys.argv:
        rec_lib.bfCudaHelm2RecordingSetTruncationScale.argtypes = [C.c_void_p, C.c_double]
        out["truncation_scale"] = float(sys.argv[sys.argv.index("--truncation-scale") + 1])
        assert rec_lib.bfCudaHelm2RecordingSetTruncationScale(handle, out["truncation_scale"]) == 0
    st = helm2.Stats()
    t1 = time.time()
    code = att.bfCudaMatAttachHelm2(h.mat, handle, None, C.byref(st))
    if code != 0:
        raise SystemExit(f"bfCudaMatAttachHelm2: status {code}: {capi.load_library().bfCudaLastErrorMessage().decode()}")
    out["attach_s"] = time.time() - t1
    out["gpu_build_s"] = time.time() - t0
    out["stats"] = st.as_dict()
    rec_lib.bfCudaHelm2RecordingFree(handle)
    y_gpu, _ = h.matmul(x)                        # bfMatMul through the swapped vtable
    Y_gpu, _ = h.matmul(X)
    out["multi_rhs_consistent"] = float(np.linalg.norm(Y_gpu[:, 0] - y_gpu[:, 0]) / np.linalg.norm(y_gpu))

    rows = want = None
    if "--sample-rows" in sys.argv:
        # R rows of the dense kernel matrix in tree order (scipy's Hankel function): the reference's own
        # yardstick (bf_all_blocks.c:149-153) at sizes where the full matrix does not fit
        from scipy.special import hankel1

        R = int(sys.argv[sys.argv.index("--sample-rows") + 1])
        tp = pts[h.perm()]
        rows = np.random.default_rng(1).choice(n, R, replace=False)
        d = np.hypot(tp[rows, None, 0] - tp[None, :, 0], tp[rows, None, 1] - tp[None, :, 1])
        with np.errstate(invalid="ignore", divide="ignore"):
            Kr = 0.25j * hankel1(0, k * d)
        Kr[d == 0] = 0
        want = Kr @ x
        out["sample_rows"] = R
        out["gpu_built_vs_dense_rows"] = float(np.linalg.norm(y_gpu[rows, 0] - want) / np.linalg.norm(want))
    if "--no-reference" not in sys.argv:
        t2 = time.time()
        h_ref = reflib.Helm2(pts, k)              # not recording: the reference computes its own values
        out["reference_build_s"] = time.time() - t2
        y_ref, _ = h_ref.matmul(x)
        out["rel_l2_vs_reference"] = float(np.linalg.norm(y_gpu - y_ref) / np.linalg.norm(y_ref))
        if rows is not None:
            out["reference_vs_dense_rows"] = float(np.linalg.norm(y_ref[rows, 0] - want) / np.linalg.norm(want))
        if "--same-factors" in sys.argv:
            # the north star's own check at this N: the reference-built factors applied by the reference on the CPU
            # and, after bfCudaMatAttach of THAT object, on the GPU — 1 RHS and a 256-RHS block (the block's columns
            # checked against the CPU on a spread of 8 of them: 256 columns cost the reference about a minute)
            att.bfCudaMatDetach(h.mat)  # the GPU-built operator's plan is no longer needed
            rng = np.random.default_rng(9)
            X256 = rng.standard_normal((n, 256)) + 1j * rng.standard_normal((n, 256))
            probe = np.arange(5, 256, 32)
            cpu_t = [h_ref.matmul(x)[1] for _ in range(3)]
            Yp_ref, _ = h_ref.matmul(X256[:, probe].copy())
            ta = time.time()
            h_ref.attach()
            out["reference_operator_attach_s"] = time.time() - ta
            y_same, _ = h_ref.matmul(x)
            gpu_t = [h_ref.matmul(x)[1] for _ in range(10)]
            Y_same, t256 = h_ref.matmul(X256)
            h_ref.detach()
            rel = lambda a, b: float(np.linalg.norm(a - b) / np.linalg.norm(b))
            out["parity"] = {"rel_l2_same_factors_1rhs": rel(y_same, y_ref),
                             "rel_l2_same_factors_256rhs_max": max(rel(Y_same[:, c], Yp_ref[:, i]) for i, c in enumerate(probe)),
                             "columns_checked_of_256": int(len(probe)), "tolerance": 1e-12,
                             "what": "reference-built factors: bfMatMul on the CPU vs bfMatMul after bfCudaMatAttach of the same object"}
            out["same_config"] = {"n": n, "cpu_bfMatMul_ms": float(np.median(cpu_t)) * 1e3, "gpu_bfMatMul_ms": float(np.median(gpu_t[2:])) * 1e3,
                                  "gpu_bfMatMul_256rhs_ms": t256 * 1e3,
                                  "api": "bfMatMul through the swapped vtable: malloc'ed result, pageable operand, H2D + D2H inside"}
            out["same_config"]["speedup"] = out["same_config"]["cpu_bfMatMul_ms"] / out["same_config"]["gpu_bfMatMul_ms"]
        if "--no-dense" not in sys.argv:
            Kx = h_ref.dense_kernel() @ x
            out["reference_vs_dense"] = float(np.linalg.norm(y_ref[:, 0] - Kx) / np.linalg.norm(Kx))
            out["gpu_built_vs_dense"] = float(np.linalg.norm(y_gpu[:, 0] - Kx) / np.linalg.norm(Kx))
    print(json.dumps(out))
    sys.stdout.flush()
    import os
    os._exit(0)


if __name__ == "__main__":
    main()
