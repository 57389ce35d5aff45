"""GMRES over the block-row sharded operator (configs[4]: repeated applies inside an iterative solve
across GPUs).  The algorithm is bfSolveGMRES (src/linalg.c:47-318) for one right-hand side, no
preconditioner — modified Gram-Schmidt, the reference's Givens rotations
(src/vec_complex.c:155-168,230-290), convergence on |s[j+1]| / beta < tol and the iterate rebuilt from
the first j basis vectors — distributed by rows: every rank keeps ITS rows of the Krylov basis in HBM,
dots are local partial sums + one all_reduce of a scalar, and the one full-length vector an iteration
needs (the operand of the next apply) is all_gathered from the row owners.

orthogonalization="mgs" follows the reference step for step (j + 2 scalar all_reduces in iteration j);
"cgs" projects against all basis vectors at once (2 all_reduces per iteration, the north star's count)
at the price of classical Gram-Schmidt's weaker orthogonality.

torch is plumbing here (device vectors, torch.distributed); the operator applies are the hand-written
stage kernels behind RowShardedOperator.apply.  The single-GPU, fully fused variant is
bfCudaSolveGMRES (csrc/bf_gmres.cu).
"""
from __future__ import annotations

import cmath
import math


def _givens(a: complex, b: complex):
    """bfVecComplexGetGivensRotation, src/vec_complex.c:276-287."""
    if abs(b) == 0:
        return 1.0 + 0j, 0j
    if abs(b) > abs(a):
        t = -a / b
        s = 1.0 / math.sqrt(1 + abs(t) ** 2)
        return t * s, s + 0j
    t = -b / a
    c = 1.0 / math.sqrt(1 + abs(t) ** 2)
    return c + 0j, t * c


def _rot(c, s, z0, z1):
    """mulInplace_givensComplex, src/vec_complex.c:155-168."""
    return c.conjugate() * z0 - s * z1, s * z0 + c * z1


def solve_gmres_sharded(op, b, x0=None, tol: float = 1e-12, max_iter: int = 50, orthogonalization: str = "mgs"):
    """Solve A x = b with A = `op` (RowShardedOperator), b replicated on every rank (torch tensor of n
    complex entries on the operator's device).  Returns (x_full, num_iter, rel_residual, num_applies);
    x_full is replicated."""
    import torch
    import torch.distributed as dist

    world = op.world
    r0, r1 = op.y_range
    n = b.numel()
    dev, dt = b.device, b.dtype
    b = b.reshape(n, 1)
    ybuf = torch.zeros(n, 1, dtype=dt, device=dev)

    def allsum(t):
        if world > 1:
            tr = torch.view_as_real(t) if t.is_complex() else t
            dist.all_reduce(tr, group=op.group)
        return t

    def full_from_local(v_loc):
        if world == 1:
            return v_loc.reshape(n, 1)
        ybuf[r0:r1, 0] = v_loc
        return op.gather_rows(ybuf)

    applies = 0
    if x0 is not None:
        x0 = x0.reshape(n, 1)
        op.apply(x0.contiguous(), ybuf)
        applies += 1
        r_loc = (b[r0:r1, 0] - ybuf[r0:r1, 0]).clone()
    else:
        r_loc = b[r0:r1, 0].clone()
    beta = math.sqrt(float(allsum((r_loc.conj() * r_loc).real.sum().reshape(1))[0]))
    V = torch.empty(max_iter + 1, r1 - r0, dtype=dt, device=dev)
    H = [[0j] * (max_iter + 1) for _ in range(max_iter)]
    S = [0j] * (max_iter + 1)
    S[0] = beta + 0j
    J = []
    residual, converged, j = 0.0, not beta > 0.0, 0
    if beta > 0.0:
        V[0] = r_loc / beta
    while j < max_iter and not converged:
        op.apply(full_from_local(V[j]).contiguous(), ybuf)  # W = A V[j], src/linalg.c:155
        applies += 1
        w = ybuf[r0:r1, 0].clone()
        hcol = torch.zeros(j + 2, dtype=dt, device=dev)
        if orthogonalization == "mgs":  # src/linalg.c:172-186
            for i in range(j + 1):
                h = allsum(torch.vdot(V[i], w).reshape(1))
                hcol[i] = h[0]
                w -= h[0] * V[i]
        else:
            h = allsum(V[: j + 1].conj() @ w)
            hcol[: j + 1] = h
            w -= V[: j + 1].T @ h
        nrm = torch.sqrt(allsum((w.conj() * w).real.sum().reshape(1)))
        hcol[j + 1] = nrm[0]
        V[j + 1] = w / nrm[0]
        hh = [complex(v) for v in hcol.cpu().tolist()]
        for i in range(j):
            hh[i], hh[i + 1] = _rot(J[i][0], J[i][1], hh[i], hh[i + 1])
        c, s = _givens(hh[j], hh[j + 1])
        J.append((c, s))
        hh[j], hh[j + 1] = _rot(c, s, hh[j], hh[j + 1])
        S[j], S[j + 1] = _rot(c, s, S[j], S[j + 1])
        for i in range(j + 1):
            H[j][i] = hh[i]
        residual = abs(S[j + 1]) / beta
        if residual < tol:  # the reference leaves the loop before counting this column
            converged = True
            break
        j += 1
    y = [0j] * max(j, 1)
    for r in range(j - 1, -1, -1):  # back substitution, src/linalg.c:240-266
        acc = S[r]
        for c_ in range(r + 1, j):
            acc -= H[c_][r] * y[c_]
        y[r] = acc / H[r][r]
    x_loc = torch.zeros(r1 - r0, dtype=dt, device=dev)
    if j > 0:
        x_loc = V[:j].T @ torch.tensor(y[:j], dtype=dt, device=dev)
    if x0 is not None:
        x_loc = x_loc + x0[r0:r1, 0]
    x_full = full_from_local(x_loc).reshape(n).clone()
    return x_full, j, residual, applies
