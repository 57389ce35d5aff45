"""CPU: the Hankel functions the on-device Helmholtz construction evaluates (csrc/bf_hankel.inl), compiled
for the host from the same source text and compared with scipy over the argument range a butterfly sees
(k r from 1e-6 to 3e4).  CUDA's own j0 / y0 are only good to 5e-12 absolute beyond |x| = 8, which is why
the kernels carry these."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent


def test_hankel_series_match_scipy(tmp_path):
    from scipy.special import hankel1

    so = tmp_path / "libhankel_host.so"
    subprocess.run(["g++", "-O2", "-fPIC", "-shared", "-I" + str(ROOT / "butterfly_b200" / "csrc"), "-o", str(so),
                    str(ROOT / "tests" / "hankel_host.cpp")], check=True)
    l = C.CDLL(str(so))
    rng = np.random.default_rng(0)
    x = np.concatenate([10 ** rng.uniform(-6, 0, 5000), rng.uniform(0, 8, 20000), rng.uniform(3.9, 5.1, 5000),
                        rng.uniform(4, 60, 20000), rng.uniform(50, 30000, 20000), [4.0, 5.0, 2.9e-8]])
    h0, h1 = np.zeros(2 * len(x)), np.zeros(2 * len(x))
    l.bfHankelHost(x.ctypes.data_as(C.c_void_p), C.c_long(len(x)), h0.ctypes.data_as(C.c_void_p), h1.ctypes.data_as(C.c_void_p))
    for got, order in ((h0.view(np.complex128), 0), (h1.view(np.complex128), 1)):
        want = hankel1(order, x)
        assert (np.abs(got - want) <= 4e-15 * np.abs(want)).all(), order
