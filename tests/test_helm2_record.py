"""CPU: the recorder of the reference's Helmholtz value builders (csrc/bf_helm2_record.c).

An operator laid out INSIDE a recording must have exactly the structure the unmodified reference
builds, and every recorded recipe must describe the leaf it stands for: evaluated here with scipy's
Hankel functions (an independent restatement of src/helm2.c:93-125 and :321-365), a kernel recipe
reproduces the reference's leaf values, and a re-expansion recipe's Z_equiv, Z_orig are solved by the
reference's own leaf to the accuracy of its truncated SVD.  No GPU involved: this pins the host half of
the on-device construction."""
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

from butterfly_b200 import capi, helm2

ROOT = Path(__file__).resolve().parent.parent
N, K_WAVE = 2048, 50.0


def _need():
    from oracle import reflib

    if not (reflib.attach_available() and helm2.HELM2_LIB_PATH.exists()):
        pytest.skip("needs oracle/_ref and libbfcuda_helm2.so (built where the reference sources exist)")
    return reflib


@pytest.fixture(scope="module")
def recorded(tmp_path_factory):
    _need()
    out = tmp_path_factory.mktemp("helm2") / "rec.npz"
    # a subprocess: the recorder has to enter the symbol scope before the reference does
    subprocess.run([sys.executable, str(ROOT / "tools" / "helm2_record.py"), str(N), repr(K_WAVE), str(out)], check=True,
                   stdout=subprocess.DEVNULL)
    z = np.load(out)
    raw = helm2.RawOps(int(z["dtype"]), int(z["num_rows"]), int(z["num_cols"]), z["buf_rows"], z["ops"].astype(capi.RAWOP_DTYPE), None)
    return raw, helm2.Recording.load_from(z), z["perm"]


@pytest.fixture(scope="module")
def reference_built():
    reflib = _need()
    h = reflib.Helm2(reflib.circle_points(N), K_WAVE)
    return h, h.flatten()


def hankel_kernel(pairs, rc, src_key="src"):
    """i/4 H0(k |x_t - x_s|), 0 where the points coincide (src/helm2.c:93-125)."""
    from scipy.special import hankel1

    s = pairs[rc[src_key]: rc[src_key] + rc["num" + src_key.capitalize()]]
    t = pairs[rc["tgt"]: rc["tgt"] + rc["numTgt"]]
    r = np.hypot(s[None, :, 0] - t[:, None, 0], s[None, :, 1] - t[:, None, 1])
    with np.errstate(invalid="ignore", divide="ignore"):
        z = 0.25j * hankel1(0, rc["k"] * r)
    z[r == 0] = 0
    return z


def test_recorded_structure_is_the_references(recorded, reference_built):
    raw, rec, perm = recorded
    h, ref_raw = reference_built
    assert raw.values is None and ref_raw.values is not None
    assert np.array_equal(raw.fingerprint(), ref_raw.fingerprint())
    assert np.array_equal(raw.buf_rows, ref_raw.buf_rows)
    assert np.array_equal(perm, h.perm())
    assert (rec.op_recipe >= 0).all() and len(np.unique(rec.op_recipe)) == raw.num_ops
    r = rec.recipes[rec.op_recipe]
    assert np.array_equal(r["rows"] * r["cols"], raw.ops["m"] * raw.ops["k"])
    assert set(np.unique(r["kind"])) == {helm2.KERNEL, helm2.REEXPANSION}
    assert (r["layerPot"] == helm2.LAYER_POT_SINGLE).all() and (r["k"] == K_WAVE).all()


def _leaf(ref_raw, i, rows, cols):
    o = ref_raw.ops[i]
    return ref_raw.values[int(o["val"]): int(o["val"]) + 2 * rows * cols].view(np.complex128).reshape(rows, cols)


def test_kernel_recipes_reproduce_the_reference_leaves(recorded, reference_built):
    raw, rec, _ = recorded
    _, ref_raw = reference_built
    idx = [i for i in range(raw.num_ops) if rec.recipes[rec.op_recipe[i]]["kind"] == helm2.KERNEL]
    assert idx
    for i in idx[:: max(1, len(idx) // 25)]:
        rc = rec.recipes[rec.op_recipe[i]]
        ref = _leaf(ref_raw, i, int(rc["rows"]), int(rc["cols"]))
        mine = hankel_kernel(rec.pairs, rc)
        assert np.abs(mine - ref).max() <= 1e-13 * np.abs(ref).max()


def test_reexpansion_recipes_are_solved_by_the_reference_leaves(recorded, reference_built):
    raw, rec, _ = recorded
    _, ref_raw = reference_built
    idx = [i for i in range(raw.num_ops) if rec.recipes[rec.op_recipe[i]]["kind"] == helm2.REEXPANSION]
    assert idx
    for i in idx[:: max(1, len(idx) // 25)]:
        rc = rec.recipes[rec.op_recipe[i]]
        x_ref = _leaf(ref_raw, i, int(rc["rows"]), int(rc["cols"]))
        z_equiv = hankel_kernel(rec.pairs, rc, "equiv")
        z_orig = hankel_kernel(rec.pairs, rc)
        assert z_equiv.shape == (rc["numTgt"], rc["numEquiv"]) and z_orig.shape == (rc["numTgt"], rc["numSrc"])
        assert np.linalg.norm(z_equiv @ x_ref - z_orig) <= 1e-11 * np.linalg.norm(z_orig)


def test_point_pool_is_deduplicated(recorded):
    _, rec, _ = recorded
    # the proxy circles are shared by many leaves: far fewer pooled points than recipe operands
    operands = int((rec.recipes["numSrc"] + rec.recipes["numTgt"] + rec.recipes["numEquiv"]).sum())
    assert len(rec.pairs) < operands / 2


def test_recipes_follow_the_leaves_through_sharding_cuts(recorded):
    """rowshard cuts tall leaves into row slabs; origin_of / Recording.remapped carry every slab back to its
    recipe and its first row inside the recipe's matrix (BfCudaHelm2Build.opRowOffset)."""
    from butterfly_b200 import rowshard

    raw, rec, _ = recorded
    cut = rowshard.presplit_tall_leaves(raw, max_rows=40)
    idx, off = rowshard.origin_of(cut)
    assert cut.num_ops > raw.num_ops and len(idx) == cut.num_ops
    o, c = raw.ops, cut.ops
    assert np.array_equal(c["outOff"] - o["outOff"][idx], off)
    assert np.array_equal(c["inOff"], o["inOff"][idx]) and np.array_equal(c["k"], o["k"][idx])
    assert (off >= 0).all() and (off + c["m"] <= o["m"][idx]).all()
    covered = np.zeros(raw.num_ops, np.int64)
    np.add.at(covered, idx, c["m"])
    assert np.array_equal(covered, o["m"])  # the slabs tile their leaf exactly
    # a second round of cuts composes
    cut2 = rowshard.presplit_tall_leaves(cut, max_rows=16)
    idx2, off2 = rowshard.origin_of(cut2)
    assert np.array_equal(cut2.ops["outOff"] - o["outOff"][idx2], off2)
    masked = rowshard.mask_for_rank(cut2, np.zeros(cut2.num_ops, np.int64), 0)
    assert np.array_equal(rowshard.origin_of(masked)[1], off2)
    r2 = rec.remapped(idx2, off2)
    assert np.array_equal(r2.op_recipe, rec.op_recipe[idx2]) and np.array_equal(r2.op_row_offset, off2)
    rows = rec.recipes["rows"][r2.op_recipe]
    assert (r2.op_row_offset + cut2.ops["m"] <= rows).all()


def test_recorder_forwards_outside_a_recording():
    """Outside a recording the two builders defined by libbfcuda_helm2.so hand the call to the reference's own
    definitions (csrc/bf_helm2_record.c: forwardTarget).  The library is linked against no libbutterfly, so
    when it comes in through dlopen the target is found among the loaded objects: an operator built THROUGH
    the recorder must be bit-identical to the golden output of the plain reference."""
    _need()
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r)\n"
        "from butterfly_b200 import helm2; helm2.load_recorder()\n"
        "from oracle import reflib\n"
        "h = reflib.Helm2(reflib.circle_points(2048), 50.0)\n"
        "y, _ = h.matmul(reflib.randn_complex(2048, 0))\n"
        "m = np.load(%r)\n"
        "assert np.array_equal(y[:, 0], m['y']), 'forwarded build differs from the reference build'\n"
    ) % (str(ROOT), str(ROOT / "tests" / "golden" / "helm2_n2048_meta.npz"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-1000:] + r.stderr[-2000:]
