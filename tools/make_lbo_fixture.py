"""Build a REAL-path (configs[3]) operator with the reference's own streamer and persist it as a fixture.

The unmodified reference (oracle/_ref) runs bfFacStreamer (src/fac_streamer.c:386-518, mergeAndSplit src/fac.c:1080-1294)
on n Fibonacci-sphere points fed with the real spherical harmonics of degree <= lmax, grouped into the frequency
tree's leaf bands exactly as bfLboFeedFacStreamerNextEigenband (src/lbo.c:96-137) would feed ARPACK's eigenbands
(ARPACK is not in this image; on the sphere the Laplace-Beltrami eigenfunctions are known in closed form).  The
resulting BfMatProduct is flattened (bfCudaFlattenEx) and written as raw ops, together with seeded operands and the
REFERENCE's results for them (bfMatMulVec / bfMatRmulVec, src/mat_product.c) and its CPU time per apply, so that the
GPU box — which has neither /root/reference nor the minutes the streamer takes — can time the CUDA path on a
reference-structured real operator and check parity against outputs of the reference itself.

usage: python tools/make_lbo_fixture.py [--n 8000] [--lmax 45] [--tol 1e-3] [--out fixtures/lbo_sphere]
Writes <out>.raw.npz (RawOps.save) and <out>.ref.npz.  fixtures/ is git-ignored (it still travels to the GPU box)."""
import argparse
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=8000)
    ap.add_argument("--lmax", type=int, default=45)
    ap.add_argument("--tol", type=float, default=1e-3)
    ap.add_argument("--out", default=str(ROOT / "fixtures" / "lbo_sphere"))
    ap.add_argument("--vectors", type=int, default=4)
    args = ap.parse_args()
    from oracle import reflib  # test infrastructure: this script is a fixture generator, not the product path

    t0 = time.time()
    s = reflib.sphere_stream(n=args.n, lmax=args.lmax, tol=args.tol)
    build_s = time.time() - t0
    raw = s.flatten()
    rng = np.random.default_rng(2024)
    X = rng.standard_normal((s.num_cols, args.vectors))
    XT = rng.standard_normal((s.num_rows, args.vectors))
    Y = np.zeros((s.num_rows, args.vectors))
    YT = np.zeros((s.num_cols, args.vectors))
    tf, tt = [], []
    for j in range(args.vectors):
        for rep in range(3):  # (first call warms the caches; the minimum is the reference's figure)
            y, sec = s.mulvec(X[:, j])
            tf.append(sec)
            yt, sec = s.mulvec(XT[:, j], transposed=True)
            tt.append(sec)
        Y[:, j], YT[:, j] = y, yt
    # the streamer's own accuracy, as its checkRelError reports it (src/fac_streamer.c:286-301): butterfly vs the dense
    # eigenvector block, here on the whole matrix
    pts = s.points
    Phi, _ = reflib.real_sph_harm_band(pts, list(range(args.lmax + 1)))
    perm = s.row_perm()
    dense = np.empty_like(Phi)
    dense[perm] = Phi  # dst[perm[i]] = src[i] (src/mat_dense_real.c permute semantics)
    err_fwd = np.linalg.norm(Y - dense @ X) / np.linalg.norm(dense @ X)
    alt = np.linalg.norm(Y - Phi[perm] @ X) / np.linalg.norm(Phi[perm] @ X)
    out = Path(args.out)
    out.parent.mkdir(parents=True, exist_ok=True)
    raw.save(str(out) + ".raw.npz")
    np.savez_compressed(str(out) + ".ref.npz", X=X, Y=Y, XT=XT, YT=YT, n=args.n, lmax=args.lmax, tol=args.tol,
                        num_rows=s.num_rows, num_cols=s.num_cols, reference_build_s=build_s,
                        cpu_forward_ms=1e3 * min(tf), cpu_transpose_ms=1e3 * min(tt), cpu_cores=1,
                        butterfly_vs_dense_rel_l2=min(err_fwd, alt), entries=raw.num_entries, ops=raw.num_ops)
    print(f"n={args.n} lmax={args.lmax}: {s.num_rows} x {s.num_cols}, ops {raw.num_ops}, entries {raw.num_entries} "
          f"({raw.num_entries / (s.num_rows * s.num_cols):.3f} of dense), streamer {build_s:.1f} s, "
          f"reference MulVec {1e3 * min(tf):.2f} ms, RmulVec {1e3 * min(tt):.2f} ms, butterfly vs dense {min(err_fwd, alt):.2e}")


if __name__ == "__main__":
    main()
