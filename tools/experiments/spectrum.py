"""Is the reference's linear system definite?  Counts the eigenvalues with positive real part of the interior operator S
(the reference matrix with its boundary rows eliminated - what a Krylov solver iterates on) for the d-3/d-4 images.

  python tools/experiments/spectrum.py            -> table on stdout (profiles/r02_indefinite_regime.txt)

Test-only (imports oracle/); not part of the product.  S is built by probing oracle/mg_model.apply_eliminated with a
3 x 3 colouring (27 operator applications).  Small frames: dense eigenvalues.  Larger frames: shift-invert ARPACK around a
positive shift, which returns the eigenvalues nearest to it - all the positive ones once k exceeds their number."""
import os
import sys
import time

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import flow_oracle as fo, mg_model as mg  # noqa: E402
from opticalflow_b200 import synthetic  # noqa: E402


def interior_matrix(P, alpha, beta):
    Nx, Ny = P.shape
    mx, my = Nx - 2, Ny - 2
    M = mx * my
    ii, jj = np.meshgrid(np.arange(mx), np.arange(my), indexing='ij')
    rows, cols, vals = [], [], []
    for q in range(3):
        for ci in range(3):
            for cj in range(3):
                x = np.zeros((3, Nx, Ny))
                x[q, 1:-1, 1:-1][((ii % 3) == ci) & ((jj % 3) == cj)] = 1.0
                y = mg.apply_eliminated(P, x, alpha, beta)
                for qr in range(3):
                    yy = y[qr, 1:-1, 1:-1]
                    i, j = np.nonzero(yy)
                    si = i + ((ci - i % 3 + 1) % 3 - 1)            # the probed column within distance 1 of row (i, j)
                    sj = j + ((cj - j % 3 + 1) % 3 - 1)
                    ok = (si >= 0) & (si < mx) & (sj >= 0) & (sj < my)
                    rows.append(qr * M + i[ok] * my + j[ok]); cols.append(q * M + si[ok] * my + sj[ok]); vals.append(yy[i, j][ok])
    return sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(3 * M, 3 * M)).tocsc()


def positive_eigenvalues(S, dense_limit=7000, k=200, shift=60.0):
    n = S.shape[0]
    if n <= dense_limit:
        ev = np.linalg.eigvals(S.toarray())
        return ev[ev.real > 0], 'dense'
    ev = spla.eigs(S, k=min(k, n - 2), sigma=shift, which='LM', return_eigenvectors=False, tol=1e-8)
    pos = ev[ev.real > 0]
    note = 'shift-invert k=%d' % k + (' (ALL returned are positive: a lower bound)' if len(pos) == len(ev) else '')
    return pos, note


def main():
    sizes = [int(v) for v in sys.argv[1:]] or [32, 48, 96, 128]
    alphas = [1e3, 2.15e3, 4.64e3, 1e4, 2.15e4, 4.64e4, 1e5]
    print('# eigenvalues of the interior operator S with POSITIVE real part (the rest of the spectrum is negative, down to ~ -8 (alpha + I^2));')
    print('# image: first frame of the d-3 movie (rotating_shearing_movie, seed 2) blurred with sigma 2.5, remodelling_alpha 1e4')
    print('%-6s %-9s %-10s %-12s %-14s %s' % ('N', 'unknowns', 'alpha', 'positive', 'largest real', 'method'))
    for N in sizes:
        P = fo.blur_movie(synthetic.rotating_shearing_movie(N, N, 1, seed=2), 2.5)[0]
        for alpha in alphas:
            t0 = time.time()
            S = interior_matrix(P, alpha, 1e4)
            pos, how = positive_eigenvalues(S)
            print('%-6d %-9d %-10.3g %-12d %-14.4g %s  (%.0f s)' % (N, S.shape[0], alpha, len(pos), pos.real.max() if len(pos) else 0.0, how,
                                                                  time.time() - t0), flush=True)
    print('# max I^2 of these images is ~3.4e4: the operator is indefinite (and non-symmetric) for alpha below ~max I^2 / 3, with a number of')
    print('# positive modes that grows with the frame area; they are smooth, so every coarse level of a multigrid hierarchy inherits them.')


if __name__ == '__main__':
    main()
