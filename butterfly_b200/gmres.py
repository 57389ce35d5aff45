"""GMRES over the block-row sharded operator (configs[4]: repeated applies inside an iterative solve
across GPUs).  The algorithm is bfSolveGMRES (src/linalg.c:47-318) for one right-hand side, no
preconditioner — modified Gram-Schmidt, the reference's Givens rotations
(src/vec_complex.c:155-168,230-290), convergence on |s[j+1]| / beta < tol and the iterate rebuilt from
the first j basis vectors — distributed by rows: every rank keeps ITS rows of the Krylov basis in HBM,
dots are local partial sums + one all_reduce of a scalar, and the one full-length vector an iteration
needs (the operand of the next apply) is all_gathered from the row owners.

orthogonalization="mgs" follows the reference step for step (j + 2 scalar all_reduces in iteration j);
"cgs" projects against all basis vectors at once (2 all_reduces per iteration, the north star's count)
at the price of classical Gram-Schmidt's weaker orthogonality; "cgs2" repeats that pass once (4
all_reduces), which restores orthogonality to working precision.

On CUDA tensors the local halves of every pass are the kernels of csrc/bf_gmres.cu
(bfCudaGmresProject / Subtract / Scale: all projections of a pass in one sweep, fixed-order sums) and
torch is plumbing only (buffers, all_reduce / all_gather over NCCL).  On CPU tensors (the gloo tests of
the host logic) the same steps are spelled with torch ops.  The single-GPU, fully fused variant is
bfCudaSolveGMRES.
"""
from __future__ import annotations

import ctypes as C
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


class _NativeGramSchmidt:
    """The CUDA halves of a Gram-Schmidt pass on this rank's rows (csrc/bf_gmres.cu)."""

    def __init__(self, V, max_iter):
        import torch

        from . import capi

        self.lib = capi.load_library()
        self.check = capi.check
        dev = V.device
        self.V, self.nloc = V, V.shape[1]
        self.scratch = torch.empty(int(self.lib.bfCudaGmresScratchBytes(max_iter + 1)), dtype=torch.uint8, device=dev)
        self.h = torch.zeros(max_iter + 3, dtype=V.dtype, device=dev)    # projections (+ |w|^2 slot)
        self.h2 = torch.zeros(max_iter + 3, dtype=V.dtype, device=dev)   # second pass of cgs2
        self.nrm2 = torch.zeros(1, dtype=V.dtype, device=dev)
        self.stream = lambda: C.c_void_p(torch.cuda.current_stream(dev).cuda_stream or 1)

    def _ptr(self, t, offset=0):
        return C.c_void_p(t.data_ptr() + 16 * offset)

    def project(self, first, m, w, out, out_offset=0):
        """out[out_offset + i] = <V[first + i], w> (local part), i < m; out[out_offset + m] = |w|^2 (local)."""
        self.check(self.lib.bfCudaGmresProject(self._ptr(self.V, first * self.nloc), self.nloc, m, self._ptr(w), self.nloc,
                                               self._ptr(out, out_offset), self._ptr(self.scratch), self.stream()), "bfCudaGmresProject")

    def subtract(self, first, m, h, h_offset, w):
        """w -= sum_i h[h_offset + i] V[first + i]; nrm2 = |w|^2 (local part)."""
        self.check(self.lib.bfCudaGmresSubtract(self._ptr(self.V, first * self.nloc), self.nloc, m, self._ptr(h, h_offset), self._ptr(w),
                                                self.nloc, self._ptr(self.nrm2), self._ptr(self.scratch), self.stream()), "bfCudaGmresSubtract")

    def scale(self, w, dst, dst2=None):
        self.check(self.lib.bfCudaGmresScale(self._ptr(w), self.nloc, self._ptr(self.nrm2), self._ptr(dst),
                                             self._ptr(dst2) if dst2 is not None else None, self.stream()), "bfCudaGmresScale")


def solve_gmres_sharded(op, b, x0=None, tol: float = 1e-12, max_iter: int = 50, orthogonalization: str = "mgs"):
    """Solve A x = b with A = `op` (RowShardedOperator), b replicated on every rank (torch tensor of n
    complex entries on the operator's device).  Returns (x_full, num_iter, rel_residual, num_applies);
    x_full is replicated."""
    import torch
    import torch.distributed as dist

    if orthogonalization not in ("mgs", "cgs", "cgs2"):
        raise ValueError(f"orthogonalization = {orthogonalization!r}")
    world = op.world
    r0, r1 = op.y_range
    nloc = r1 - r0
    n = b.numel()
    dev, dt = b.device, b.dtype
    b = b.reshape(n, 1)
    ybuf = torch.zeros(n, 1, dtype=dt, device=dev)
    xfull = torch.zeros(n, 1, dtype=dt, device=dev)
    equal_rows = all(op.y_cuts[i + 1] - op.y_cuts[i] == nloc for i in range(world))
    native = b.is_cuda
    if native and dt != torch.complex128:
        raise TypeError("the CUDA Gram-Schmidt kernels are complex128")

    def allsum(t):
        if world > 1:
            tr = torch.view_as_real(t) if t.is_complex() else t
            dist.all_reduce(tr, group=op.group)
        return t

    def full_from_local(v_loc):
        """Every rank's rows of a distributed vector -> the replicated operand of the next apply."""
        if world == 1:
            return v_loc.reshape(n, 1)
        if equal_rows and native:
            dist.all_gather_into_tensor(torch.view_as_real(xfull).reshape(-1), torch.view_as_real(v_loc.contiguous()).reshape(-1), group=op.group)
            return xfull
        ybuf[r0:r1, 0] = v_loc
        return op.gather_rows(ybuf, out=xfull)

    applies = 0
    if x0 is not None:
        x0 = x0.reshape(n, 1)
        op.apply(x0.contiguous(), ybuf)
        applies += 1
        r_loc = (b[r0:r1, 0] - ybuf[r0:r1, 0]).clone()
    else:
        r_loc = b[r0:r1, 0].clone()
    beta = math.sqrt(float(allsum((r_loc.conj() * r_loc).real.sum().reshape(1))[0]))
    V = torch.empty(max_iter + 1, nloc, dtype=dt, device=dev)
    gs = _NativeGramSchmidt(V, max_iter) if native else None
    H = [[0j] * (max_iter + 1) for _ in range(max_iter)]
    S = [0j] * (max_iter + 1)
    S[0] = beta + 0j
    J = []
    residual, converged, j = 0.0, not beta > 0.0, 0
    if beta > 0.0:
        V[0] = r_loc / beta
    def close_column(jj, hh):
        """Rotations, new Givens pair, rotated right-hand side for Hessenberg column jj (src/linalg.c:206-231);
        True when the residual test fires (the reference then leaves the loop without counting jj)."""
        nonlocal residual
        for i in range(jj):
            hh[i], hh[i + 1] = _rot(J[i][0], J[i][1], hh[i], hh[i + 1])
        c, s = _givens(hh[jj], hh[jj + 1])
        J.append((c, s))
        hh[jj], hh[jj + 1] = _rot(c, s, hh[jj], hh[jj + 1])
        S[jj], S[jj + 1] = _rot(c, s, S[jj], S[jj + 1])
        for i in range(jj + 1):
            H[jj][i] = hh[i]
        residual = abs(S[jj + 1]) / beta
        return residual < tol

    if native:
        # The Hessenberg column of iteration j is read on the host while the GPU already runs iteration
        # j + 1 (the host never drains the queue); a convergence found one iteration late costs one
        # discarded apply and changes nothing in the result.
        pin = torch.empty(2, max_iter + 2, dtype=dt).pin_memory()
        events = [torch.cuda.Event(), torch.cuda.Event()]

        def read_column(jj):
            events[jj & 1].synchronize()
            host = pin[jj & 1, : jj + 2].tolist()
            return [complex(v) for v in host[: jj + 1]] + [complex(math.sqrt(host[jj + 1].real))]

        pending = None
        while j < max_iter and not converged:
            op.apply(full_from_local(V[j]).contiguous(), ybuf)  # W = A V[j], src/linalg.c:155
            applies += 1
            m = j + 1
            w = ybuf[r0:r1, 0]  # orthogonalised in place; the next apply overwrites it
            if orthogonalization == "mgs":  # src/linalg.c:172-186, one all_reduce per basis vector
                for i in range(m):
                    gs.project(i, 1, w, gs.h, i)
                    allsum(gs.h[i:i + 1])
                    gs.subtract(i, 1, gs.h, i, w)
                hdev = gs.h[:m]
            else:
                gs.project(0, m, w, gs.h)
                allsum(gs.h[:m])
                gs.subtract(0, m, gs.h, 0, w)
                hdev = gs.h[:m]
                if orthogonalization == "cgs2":
                    gs.project(0, m, w, gs.h2)
                    allsum(gs.h2[:m])
                    gs.subtract(0, m, gs.h2, 0, w)
                    hdev = gs.h[:m] + gs.h2[:m]
            allsum(gs.nrm2)
            gs.scale(w, V[j + 1])
            pin[j & 1, : m + 1].copy_(torch.cat([hdev, gs.nrm2]), non_blocking=True)
            events[j & 1].record()
            if pending is not None and close_column(pending, read_column(pending)):
                converged, j = True, pending
                break
            pending = j
            j += 1
        if not converged and pending is not None and close_column(pending, read_column(pending)):
            converged, j = True, pending
    while not native and j < max_iter and not converged:
        op.apply(full_from_local(V[j]).contiguous(), ybuf)
        applies += 1
        m = j + 1
        w = ybuf[r0:r1, 0].clone()
        hcol = torch.zeros(j + 2, dtype=dt, device=dev)
        if orthogonalization == "mgs":
            for i in range(m):
                h = allsum(torch.vdot(V[i], w).reshape(1))
                hcol[i] = h[0]
                w -= h[0] * V[i]
        else:
            for _ in range(2 if orthogonalization == "cgs2" else 1):
                h = allsum(V[:m].conj() @ w)
                hcol[:m] += h
                w -= V[:m].T @ h
        nrm = torch.sqrt(allsum((w.conj() * w).real.sum().reshape(1)))
        hcol[j + 1] = nrm[0]
        V[j + 1] = w / nrm[0]
        if close_column(j, [complex(v) for v in hcol.cpu().tolist()]):
            converged = True
            break
        j += 1
    y = [0j] * max(j, 1)
    for r in range(j - 1, -1, -1):  # back substitution, src/linalg.c:240-266
        acc = S[r]
        for c_ in range(r + 1, j):
            acc -= H[c_][r] * y[c_]
        y[r] = acc / H[r][r]
    x_loc = torch.zeros(nloc, dtype=dt, device=dev)
    if j > 0:
        x_loc = V[:j].T @ torch.tensor(y[:j], dtype=dt, device=dev)
    if x0 is not None:
        x_loc = x_loc + x0[r0:r1, 0]
    x_full = full_from_local(x_loc).reshape(n).clone()
    return x_full, j, residual, applies
