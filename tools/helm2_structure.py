"""Write the flattened, VALUE-LESS structure of the reference's Helmholtz HODBF for N points on the
unit circle (examples/simple/make_circle_test_data.py:12-15) and wavenumber k to an .npz file.

Run as a subprocess (it must load tools/shape_stub.c's library BEFORE the reference so that the
kernel-matrix builders are interposed; a process that already loaded the real reference cannot do
that).  The reference's own quadtree + bfFacHelm2MakeMultilevel decide every block shape, offset
and rank; only the O(N^1.4) value computation is skipped.  Workload generation for bench.py — the
oracle and the parity tests never use it.

usage: python tools/helm2_structure.py N K OUT.npz
"""
import ctypes as C
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    n, k, out = int(sys.argv[1]), float(sys.argv[2]), sys.argv[3]
    stub = ROOT / "oracle" / "_ref" / "libbf_shape_stub.so"
    C.CDLL(str(stub), mode=C.RTLD_GLOBAL)  # first in the global scope: its two symbols win
    import numpy as np

    from oracle import reflib

    t0 = time.time()
    h = reflib.Helm2(reflib.circle_points(n), k)
    raw = h.flatten()
    assert raw.values is None, "stub was not interposed: the reference computed values"
    np.savez(out, dtype=raw.dtype, num_rows=raw.num_rows, num_cols=raw.num_cols, buf_rows=raw.buf_rows,
             ops=raw.ops, values=np.zeros(0), has_values=False, perm=h.perm(), seconds=time.time() - t0)
    print(f"helm2 structure N={n} k={k}: {raw.num_ops} leaves, E={raw.num_entries} in {time.time() - t0:.1f}s")
    # the reference's teardown frees millions of tiny nodes; skip it
    sys.stdout.flush()
    import os
    os._exit(0)


if __name__ == "__main__":
    main()
