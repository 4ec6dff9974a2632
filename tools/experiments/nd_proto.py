"""numpy prototype of the multifrontal nested-dissection solve on the plan of opticalflow_b200/ndplan.py, against SciPy's
direct solve of the same interior system (test-only: imports oracle/).   python tools/experiments/nd_proto.py Nx Ny alpha"""
import os, sys, time
import numpy as np, scipy.linalg, scipy.sparse.linalg as spla
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from spectrum import interior_matrix
from oracle import flow_oracle as fo
from opticalflow_b200 import ndplan, synthetic

Nx, Ny, alpha = int(sys.argv[1]), int(sys.argv[2]), float(sys.argv[3])
movie = fo.blur_movie(synthetic.rotating_shearing_movie(Nx, Ny, 2, seed=2), 2.5)
P, C = movie
S = interior_matrix(P, alpha, 1e4).tocsr()
mx, my = Nx - 2, Ny - 2; M = mx * my
b = fo.build_rhs(P, C)[:, 1:-1, 1:-1].reshape(-1)
levels = ndplan.build(Nx, Ny)
print('levels', len(levels), [(l.n_nodes, l.n_own, l.n_border) for l in levels], 'front memory %.1f MB' % (ndplan.memory_bytes(levels) / 1e6))
var = lambda pix: (pix[None, :] + M * np.arange(3)[:, None]).T.reshape(-1)        # variable ids of pixels, (pixel, field) pixel-major
t0 = time.time()
fac = [None] * len(levels); schur = None
for d in range(len(levels) - 1, -1, -1):
    lv = levels[d]; nf = 3 * lv.n_front; no = 3 * lv.n_own
    F = np.zeros((lv.n_nodes, nf, nf))
    for n in range(lv.n_nodes):
        pix = np.concatenate([lv.own[n], lv.border[n]]); ok = np.repeat(pix >= 0, 3)
        v = var(np.where(pix >= 0, pix, 0))
        A = S[v][:, v].toarray(); A[~ok] = 0; A[:, ~ok] = 0
        A[no:, no:] = 0                                   # border x border entries belong to nodes higher up
        pad = np.flatnonzero(~ok[:no]); A[pad, pad] = 1.0
        F[n] = A
    if schur is not None:                                  # extend-add the children's Schur complements
        ch = levels[d + 1]
        for c in range(ch.n_nodes):
            slot = ch.parent_slot[c]; okp = slot >= 0
            vs = (3 * slot[okp][:, None] + np.arange(3)[None, :]).reshape(-1)
            sel = np.flatnonzero(np.repeat(okp, 3))
            F[c // 2][np.ix_(vs, vs)] += schur[c][np.ix_(sel, sel)]
    lus, Xs, F21s, schur_new = [], [], [], []
    for n in range(lv.n_nodes):
        lu = scipy.linalg.lu_factor(F[n, :no, :no])
        X = scipy.linalg.lu_solve(lu, F[n, :no, no:])
        schur_new.append(F[n, no:, no:] - F[n, no:, :no] @ X)
        lus.append(lu); Xs.append(X); F21s.append(F[n, no:, :no].copy())
    fac[d] = (lus, Xs, F21s); schur = schur_new
print('factorised in %.1f s' % (time.time() - t0))
w = b.copy()
for d in range(len(levels) - 1, -1, -1):                   # forward
    lv = levels[d]; lus, Xs, F21s = fac[d]
    for n in range(lv.n_nodes):
        ok_o = np.repeat(lv.own[n] >= 0, 3); ok_b = np.repeat(lv.border[n] >= 0, 3)
        vo = var(np.where(lv.own[n] >= 0, lv.own[n], 0)); vb = var(np.where(lv.border[n] >= 0, lv.border[n], 0))
        t = np.where(ok_o, w[vo], 0.0)
        z = scipy.linalg.lu_solve(lus[n], t)
        w[vo[ok_o]] = z[ok_o]
        u = F21s[n] @ z
        np.subtract.at(w, vb[ok_b], u[ok_b])
for d in range(len(levels)):                               # backward
    lv = levels[d]; lus, Xs, F21s = fac[d]
    for n in range(lv.n_nodes):
        ok_o = np.repeat(lv.own[n] >= 0, 3); ok_b = np.repeat(lv.border[n] >= 0, 3)
        vo = var(np.where(lv.own[n] >= 0, lv.own[n], 0)); vb = var(np.where(lv.border[n] >= 0, lv.border[n], 0))
        xb = np.where(ok_b, w[vb], 0.0)
        xo = np.where(ok_o, w[vo], 0.0) - Xs[n] @ xb
        w[vo[ok_o]] = xo[ok_o]
xd = spla.spsolve(S.tocsc(), b)
print('N %dx%d alpha %.3g: |x_nd - x_spsolve| / |x| = %.3e   residual %.3e' % (Nx, Ny, alpha, np.linalg.norm(w - xd) / np.linalg.norm(xd),
      np.linalg.norm(S @ w - b) / np.linalg.norm(b)))
