"""The sparse direct solve as a numpy model on the CPU: multifrontal LU on the plan of opticalflow_b200/ndplan.py - the same
front layout, padding, extend-add through ``parent_slot`` and triangular sweeps as csrc/direct_solver.cuh - against SciPy's
SuperLU solve of the reference matrix (oracle/flow_oracle.py).  It holds the host-side plan to what it is for: a plan
with a wrong slot, a missed border pixel or a separator the stencil reaches across gives a wrong solution here, without a GPU."""
import numpy as np
import pytest
import scipy.linalg
import scipy.sparse.linalg as spla

from opticalflow_b200 import ndplan, synthetic
from oracle import flow_oracle as fo


def interior_system(P, C, alpha, beta):
    """The reference system (:836-1072) with its boundary rows eliminated: Schur complement on the interior unknowns, ordered as
    the device orders them (three planes of interior pixels, pixel p = (i - 1) * (Ny - 2) + (j - 1))."""
    Nx, Ny = P.shape
    A = fo.assemble_csr(P, alpha, beta).tocsc()
    b = fo.planes_to_interleaved(fo.build_rhs(P, C))
    ii, jj = np.meshgrid(np.arange(1, Nx - 1), np.arange(1, Ny - 1), indexing='ij')
    inner = np.concatenate([(3 * Ny * ii + 3 * jj + q).ravel() for q in range(3)])
    ghost = np.setdiff1d(np.arange(3 * Nx * Ny), inner)
    assert not b[ghost].any()
    A_II, A_IG = A[inner][:, inner], A[inner][:, ghost]
    A_GI, A_GG = A[ghost][:, inner], A[ghost][:, ghost]
    S = A_II - A_IG @ spla.spsolve(A_GG.tocsc(), A_GI.tocsc())
    return S.tocsr(), b[inner], A, b, inner


def multifrontal_solve(S, b, levels, M):
    var = lambda pix: (pix[None, :] + M * np.arange(3)[:, None]).T.reshape(-1)      # (pixel, field) pixel-major, like the fronts
    fac, schur = [None] * len(levels), None
    for d in range(len(levels) - 1, -1, -1):                                         # numeric factorisation, leaves to root
        lv = levels[d]
        nf, no = 3 * lv.n_front, 3 * lv.n_own
        F = np.zeros((lv.n_nodes, nf, nf))
        for n in range(lv.n_nodes):
            pix = np.concatenate([lv.own[n], lv.border[n]])
            ok = np.repeat(pix >= 0, 3)
            v = var(np.where(pix >= 0, pix, 0))
            A = S[v][:, v].toarray()
            A[~ok] = 0; A[:, ~ok] = 0
            A[no:, no:] = 0                                   # border x border entries belong to fronts higher up
            pad = np.flatnonzero(~ok[:no])
            A[pad, pad] = 1.0                                 # padded owned variables: identity
            F[n] = A
        if schur is not None:                                 # extend-add of the children's Schur complements
            ch = levels[d + 1]
            for c in range(ch.n_nodes):
                slot = ch.parent_slot[c]
                okp = slot >= 0
                vs = (3 * slot[okp][:, None] + np.arange(3)[None, :]).reshape(-1)
                sel = np.flatnonzero(np.repeat(okp, 3))
                F[c // 2][np.ix_(vs, vs)] += schur[c][np.ix_(sel, sel)]
        lus, Xs, F21s, schur = [], [], [], []
        for n in range(lv.n_nodes):
            lu = scipy.linalg.lu_factor(F[n, :no, :no])
            X = scipy.linalg.lu_solve(lu, F[n, :no, no:])
            schur.append(F[n, no:, no:] - F[n, no:, :no] @ X)
            lus.append(lu); Xs.append(X); F21s.append(F[n, no:, :no].copy())
        fac[d] = (lus, Xs, F21s)
    w = b.copy()
    for d in range(len(levels) - 1, -1, -1):                  # forward sweep
        lv = levels[d]
        lus, Xs, F21s = fac[d]
        for n in range(lv.n_nodes):
            ok_o, ok_b = np.repeat(lv.own[n] >= 0, 3), np.repeat(lv.border[n] >= 0, 3)
            vo, vb = var(np.where(lv.own[n] >= 0, lv.own[n], 0)), var(np.where(lv.border[n] >= 0, lv.border[n], 0))
            z = scipy.linalg.lu_solve(lus[n], np.where(ok_o, w[vo], 0.0))
            w[vo[ok_o]] = z[ok_o]
            np.subtract.at(w, vb[ok_b], (F21s[n] @ z)[ok_b])
    for d in range(len(levels)):                              # backward sweep
        lv = levels[d]
        lus, Xs, F21s = fac[d]
        for n in range(lv.n_nodes):
            ok_o, ok_b = np.repeat(lv.own[n] >= 0, 3), np.repeat(lv.border[n] >= 0, 3)
            vo, vb = var(np.where(lv.own[n] >= 0, lv.own[n], 0)), var(np.where(lv.border[n] >= 0, lv.border[n], 0))
            xo = np.where(ok_o, w[vo], 0.0) - Xs[n] @ np.where(ok_b, w[vb], 0.0)
            w[vo[ok_o]] = xo[ok_o]
    return w


@pytest.mark.parametrize('shape,alpha', [((9, 11), 1e5), ((16, 13), 1e3), ((23, 31), 1e5), ((23, 31), 1e3), ((12, 40), 2.15e4)])
def test_multifrontal_model_on_the_plan_matches_superlu(shape, alpha):
    Nx, Ny = shape
    movie = fo.blur_movie(synthetic.rotating_shearing_movie(Nx, Ny, 2, seed=2), 1.5)
    P, C = movie
    S, b, A, b_full, inner = interior_system(P, C, alpha, 1e4)
    x = multifrontal_solve(S, b, ndplan.build(Nx, Ny), (Nx - 2) * (Ny - 2))
    want = spla.spsolve(A, b_full)[inner]                     # the reference's direct branch (:1146-1147), interior unknowns
    assert np.linalg.norm(x - want) / np.linalg.norm(want) < 1e-7            # two direct solves of an ill-conditioned system; a wrong plan gives O(1)
    assert np.linalg.norm(S @ x - b) / np.linalg.norm(b) < 1e-10
