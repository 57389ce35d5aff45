"""Generate the golden fixtures under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE
(oracle/_ref, built from /root/reference).  Needs /root/reference; run here, commit the output.

  helm2_n2048_meta.npz   bfFacHelm2MakeMultilevel on the 2048-point unit circle, k = 50 (the
                         example's points-per-wavelength ratio, run_bf_all_blocks_test.sh:8-9):
                         tree permutation, structural fingerprint of the flattened tree, E,
                         x = bfComplexRandn after bfSeed(0), y = bfMatMul(A, x) and the permuted
                         variants produced by bfMatPermuteRows.  No factor values (too large).
  helm2_block.npz        ONE butterfly of that operator (a BfMatProduct [BlockDiag, BlockCoo...,
                         BlockDiag] from bfFacHelm2Make) with its factor values, an operand
                         supported on its source box and the reference's bfMatMul result on its
                         target box.
  lbo_stream_n600.npz    the REAL path: bfFacStreamer (src/fac_streamer.c:386) fed with the real
                         spherical harmonics of degree <= 11 on 600 Fibonacci-sphere points through
                         the restated body of bfLboFeedFacStreamerNextEigenband (src/lbo.c:96-137,
                         oracle/ref_driver.c refStream*), tol 1e-3; the flattened bfFacSpanGetMat
                         object WITH its factor values, a seeded operand pair and the reference's
                         bfMatMulVec / bfMatRmulVec results.
  lbo_sphere_fixture.npz the same streamer fed with the reference's own LBO test fixture
                         (tests/sphere.obj + tests/sphere_{Lam,Phi}.txt, the eigenpairs
                         tests/test_linalg.c:21-22 checks against): flattened operator, seeded
                         operands, the reference's bfMatMulVec / bfMatRmulVec results and the
                         products with the dense eigenvector matrix the factorization approximates.
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from butterfly_b200 import capi  # noqa: E402
from butterfly_b200.rawops import RawOps  # noqa: E402
from oracle import reflib  # noqa: E402

OUT = ROOT / "tests" / "golden"


def extract_product(raw: RawOps, which: int = 0):
    """Sub-operator made of the `which`-th group of ops connected through temp buffers."""
    ops = raw.ops
    first_temp_ops = np.nonzero(ops["outBuf"] >= 2)[0]
    seen_bufs, groups = set(), []
    for i in first_temp_ops:
        b = int(ops["outBuf"][i])
        if b in seen_bufs:
            continue
        # a product's temps are consecutive buffers: walk forward while ops chain through them
        comp = {b}
        grew = True
        while grew:
            grew = False
            for j in np.nonzero(np.isin(ops["inBuf"], list(comp)) | np.isin(ops["outBuf"], list(comp)))[0]:
                for bb in (int(ops["inBuf"][j]), int(ops["outBuf"][j])):
                    if bb >= 2 and bb not in comp:
                        comp.add(bb)
                        grew = True
        seen_bufs |= comp
        groups.append(sorted(comp))
        if len(groups) > which:
            break
    comp = groups[which]
    sel = np.nonzero(np.isin(ops["inBuf"], comp) | np.isin(ops["outBuf"], comp))[0]
    sub = ops[sel].copy()
    rd = sub[sub["inBuf"] == 0]
    wr = sub[sub["outBuf"] == 1]
    c0, c1 = int(rd["inOff"].min()), int((rd["inOff"] + rd["k"]).max())
    r0, r1 = int(wr["outOff"].min()), int((wr["outOff"] + wr["m"]).max())
    remap = {0: 0, 1: 1}
    buf_rows = [c1 - c0, r1 - r0]
    for b in comp:
        remap[b] = len(buf_rows)
        buf_rows.append(int(raw.buf_rows[b]))
    sub["inOff"][sub["inBuf"] == 0] -= c0
    sub["outOff"][sub["outBuf"] == 1] -= r0
    sub["inBuf"] = [remap[int(b)] for b in sub["inBuf"]]
    sub["outBuf"] = [remap[int(b)] for b in sub["outBuf"]]
    # re-pack the values
    e = 2 if raw.dtype == capi.COMPLEX_F64 else 1
    sizes = (sub["m"] * sub["k"] * e).astype(np.int64)
    offs = np.concatenate([[0], np.cumsum(sizes)])
    vals = np.zeros(int(offs[-1]))
    for i in range(len(sub)):
        o = int(sub["val"][i])
        vals[offs[i]:offs[i + 1]] = raw.values[o:o + sizes[i]]
        sub["val"][i] = offs[i]
    return RawOps(raw.dtype, r1 - r0, c1 - c0, np.array(buf_rows, np.int64), sub, vals), (r0, r1, c0, c1)


def main():
    OUT.mkdir(parents=True, exist_ok=True)
    n, k = 2048, 50.0
    h = reflib.Helm2(reflib.circle_points(n), k)
    perm = h.perm()
    rev = reflib.reverse_perm(perm)
    x = reflib.randn_complex(n, 0)
    y, _ = h.matmul(x)
    x3 = np.stack([x, np.roll(x, 7) * 1j, np.conj(x)], axis=1)
    y3, _ = h.matmul(x3)
    raw = h.flatten()
    np.savez_compressed(
        OUT / "helm2_n2048_meta.npz", n=n, k=k, perm=perm, rev_perm=rev, x=x, y=y[:, 0], x3=x3, y3=y3,
        fingerprint=raw.fingerprint(), buf_rows=raw.buf_rows, num_entries=raw.num_entries,
        x_permuted=reflib.permute_rows_complex(x, rev), y_permuted=reflib.permute_rows_complex(y[:, 0], perm),
    )
    sub, (r0, r1, c0, c1) = extract_product(raw, 0)
    xs = np.zeros(n, np.complex128)
    xs[c0:c1] = x[c0:c1]
    ys, _ = h.matmul(xs)
    sub.save(OUT / "helm2_block.npz")
    np.savez_compressed(OUT / "helm2_block_io.npz", x=xs[c0:c1], y=ys[r0:r1, 0], box=np.array([r0, r1, c0, c1]))
    print("full: ops", raw.num_ops, "E", raw.num_entries, "| block:", sub.num_ops, "ops, E", sub.num_entries,
          "rows", (r0, r1), "cols", (c0, c1), "factors", len(sub.buf_rows) - 1)
    h.close()

    # ---- real (streamed) path ----
    st = reflib.sphere_stream(600, 11)
    raw_s = st.flatten()
    rng = np.random.default_rng(0)
    xs_ = rng.standard_normal(st.num_cols)
    vs_ = rng.standard_normal(st.num_rows)
    ys_, _ = st.mulvec(xs_)
    zs_, _ = st.mulvec(vs_, transposed=True)
    raw_s.save(OUT / "lbo_stream_n600.npz")
    np.savez_compressed(OUT / "lbo_stream_n600_io.npz", x=xs_, y=ys_, v=vs_, z=zs_, row_perm=st.row_perm(),
                        bands=np.array(st.bands), fingerprint=raw_s.fingerprint())
    print("stream: ops", raw_s.num_ops, "E", raw_s.num_entries, "shape", (st.num_rows, st.num_cols), "bands", st.bands)

    # ---- real path on the reference's own fixture ----
    sf = reflib.sphere_fixture_stream()
    raw_f = sf.flatten()
    xf = rng.standard_normal(sf.num_cols)
    vf = rng.standard_normal(sf.num_rows)
    yf, _ = sf.mulvec(xf)
    zf, _ = sf.mulvec(vf, transposed=True)
    pf = sf.row_perm()
    raw_f.save(OUT / "lbo_sphere_fixture.npz")
    np.savez_compressed(OUT / "lbo_sphere_fixture_io.npz", x=xf, y=yf, v=vf, z=zf, row_perm=pf, bands=np.array(sf.bands),
                        y_dense=sf.phi[pf] @ xf, z_dense=sf.phi[pf].T @ vf, fingerprint=raw_f.fingerprint())
    print("fixture stream: ops", raw_f.num_ops, "E", raw_f.num_entries, "shape", (sf.num_rows, sf.num_cols), "bands", sf.bands,
          "err vs dense Phi", np.linalg.norm(yf - sf.phi[pf] @ xf) / np.linalg.norm(yf))


if __name__ == "__main__":
    main()
