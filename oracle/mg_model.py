"""TEST INFRASTRUCTURE ONLY - numpy model of the solver the CUDA path implements.

This is an executable specification of the *eliminated* operator (ghost ring folded in
by mirror-on-load), the level hierarchy, the block-Jacobi smoother and the V-cycle
preconditioned BiCGStab.  It exists so that design decisions (coarsening alignment,
smoother damping, cycle shape) can be evaluated on the CPU and so that the CUDA kernels
have a line-by-line numpy counterpart to be tested against.  It is not part of the
reference; the reference solves with PETSc (optical_flow.py:1081-1132).
"""
import numpy as np


def fill_ghosts(f, corner_sum=True):
    """Ghost ring from the boundary rows of the reference system (optical_flow.py:965-1070):
    q(0,j)=q(2,j), q(N-1,j)=q(N-3,j), same for columns; corner rows read
    q(0,0)=q(2,0)+q(0,2) (two off-diagonals) -> 2*q(2,2) once the edges are mirrored.
    corner_sum=False gives the plain even reflection q(0,0)=q(2,2) (used for transfers)."""
    f[..., 0, :] = f[..., 2, :]
    f[..., -1, :] = f[..., -3, :]
    f[..., :, 0] = f[..., :, 2]
    f[..., :, -1] = f[..., :, -3]
    if corner_sum:
        f[..., 0, 0] *= 2; f[..., 0, -1] *= 2; f[..., -1, 0] *= 2; f[..., -1, -1] *= 2
    return f


def zero_ghosts(f):
    f[..., 0, :] = 0; f[..., -1, :] = 0; f[..., :, 0] = 0; f[..., :, -1] = 0
    return f


def level_coefficients(P, hinv):
    """Interior coefficient fields at grid spacing h=1/hinv (fine level hinv=1 == reference :812-827)."""
    I = P[1:-1, 1:-1]
    Ix = (P[2:, 1:-1] - P[:-2, 1:-1]) * (0.5 * hinv)
    Iy = Ix                                            # reference 'dy' == 'dx' (:698-699)
    Ixx = (P[2:, 1:-1] + P[:-2, 1:-1] - 2 * I) * hinv * hinv
    Iyy = (P[1:-1, 2:] + P[1:-1, :-2] - 2 * I) * hinv * hinv
    Ixy = (P[2:, 2:] - P[2:, :-2] - P[:-2, 2:] + P[:-2, :-2]) * (0.25 * hinv * hinv)
    return I, Ix, Iy, Ixx, Iyy, Ixy


def _s(f, di, dj):
    Nx, Ny = f.shape
    return f[1 + di:Nx - 1 + di, 1 + dj:Ny - 1 + dj]


def apply_eliminated(P, x, alpha, beta, hinv=1.0):
    """y = S x: interior rows with the ghost ring substituted; ghost rows of y are 0."""
    x = fill_ghosts(x.copy())
    ux, uy, g = x
    I, Ix, Iy, Ixx, Iyy, Ixy = level_coefficients(P, hinv)
    h2 = hinv * hinv
    a, be = alpha * h2, beta * h2
    II = I * I * h2
    cross = lambda f: _s(f, -1, -1) + _s(f, 1, 1) - _s(f, -1, 1) - _s(f, 1, -1)
    y = np.zeros_like(x)
    y[0, 1:-1, 1:-1] = ((I * Ixx - 2 * II - 4 * a) * _s(ux, 0, 0) + I * Ixy * _s(uy, 0, 0)
                        + (II - I * Ix * hinv + a) * _s(ux, -1, 0) + (II + I * Ix * hinv + a) * _s(ux, 1, 0)
                        + a * (_s(ux, 0, -1) + _s(ux, 0, 1))
                        + (I * Ix * hinv / 2) * (_s(uy, 0, 1) - _s(uy, 0, -1))
                        + (I * Iy * hinv / 2) * (_s(uy, 1, 0) - _s(uy, -1, 0))
                        + (II / 4) * cross(uy) + (I * hinv / 2) * (_s(g, -1, 0) - _s(g, 1, 0)))
    y[1, 1:-1, 1:-1] = ((I * Iyy - 2 * II - 4 * a) * _s(uy, 0, 0) + I * Ixy * _s(ux, 0, 0)
                        + (II - I * Iy * hinv + a) * _s(uy, 0, -1) + (II + I * Iy * hinv + a) * _s(uy, 0, 1)
                        + a * (_s(uy, -1, 0) + _s(uy, 1, 0))
                        + (I * Iy * hinv / 2) * (_s(ux, 1, 0) - _s(ux, -1, 0))
                        + (I * Ix * hinv / 2) * (_s(ux, 0, 1) - _s(ux, 0, -1))
                        + (II / 4) * cross(ux) + (I * hinv / 2) * (_s(g, 0, -1) - _s(g, 0, 1)))
    y[2, 1:-1, 1:-1] = ((-1 - 4 * be) * _s(g, 0, 0) + Ix * _s(ux, 0, 0) + Iy * _s(uy, 0, 0)
                        + be * (_s(g, -1, 0) + _s(g, 1, 0) + _s(g, 0, -1) + _s(g, 0, 1))
                        + (I * hinv / 2) * (_s(ux, 1, 0) - _s(ux, -1, 0))
                        + (I * hinv / 2) * (_s(uy, 0, 1) - _s(uy, 0, -1)))
    return y


def block_diagonal_inverse(P, alpha, beta, hinv=1.0):
    """Inverse of the pointwise 3x3 diagonal block [[d0,c,0],[c,d1,0],[Ix,Iy,d2]] (interior)."""
    I, Ix, Iy, Ixx, Iyy, Ixy = level_coefficients(P, hinv)
    h2 = hinv * hinv
    d0 = I * Ixx - 2 * I * I * h2 - 4 * alpha * h2
    d1 = I * Iyy - 2 * I * I * h2 - 4 * alpha * h2
    c = I * Ixy
    d2 = (-1 - 4 * beta * h2) * np.ones_like(I)
    return d0, d1, c, d2, Ix, Iy


def apply_block_jacobi(P, r, alpha, beta, hinv=1.0):
    """z = D^-1 r on the interior (ghost ring 0)."""
    d0, d1, c, d2, Ix, Iy = block_diagonal_inverse(P, alpha, beta, hinv)
    r0, r1, r2 = r[0, 1:-1, 1:-1], r[1, 1:-1, 1:-1], r[2, 1:-1, 1:-1]
    det = d0 * d1 - c * c
    z = np.zeros_like(r)
    z0 = (d1 * r0 - c * r1) / det
    z1 = (d0 * r1 - c * r0) / det
    z[0, 1:-1, 1:-1] = z0
    z[1, 1:-1, 1:-1] = z1
    z[2, 1:-1, 1:-1] = (r2 - Ix * z0 - Iy * z1) / d2
    return z


# ---------------------------------------------------------------- grid transfers
def coarse_size(N):
    """Coarse node K <-> fine node 2K-1 (both Neumann mirror nodes K=1 / i=1 coincide)."""
    return (N + 3) // 2


def _fw_axis(f, axis, Nc, clamp):
    """Full weighting 1/4,1/2,1/4 around fine 2K-1 along one axis, for K=0..Nc-1."""
    N = f.shape[axis]
    K = np.arange(Nc)
    out = 0
    for off, w in ((-1, 0.25), (0, 0.5), (1, 0.25)):
        idx = 2 * K - 1 + off
        idx = np.clip(idx, 0, N - 1) if clamp else np.clip(idx, 0, N - 1)
        out = out + w * np.take(f, idx, axis=axis)
    return out


def restrict_image(P):
    Nx, Ny = P.shape
    return _fw_axis(_fw_axis(P, 0, coarse_size(Nx), True), 1, coarse_size(Ny), True)


def restrict_residual(r):
    """Full weighting of a residual (ghost ring evenly reflected first); coarse ghost ring 0."""
    r = fill_ghosts(r.copy(), corner_sum=False)
    Nx, Ny = r.shape[-2:]
    out = _fw_axis(_fw_axis(r, r.ndim - 2, coarse_size(Nx), True), r.ndim - 1, coarse_size(Ny), True)
    return zero_ghosts(out)


def prolong(e, Nx, Ny):
    """Bilinear interpolation coarse -> fine (coarse ghost ring evenly reflected first)."""
    e = fill_ghosts(e.copy(), corner_sum=False)

    def axis_interp(f, axis, N):
        Nc = f.shape[axis]
        i = np.arange(N)
        lo = np.clip((i + 1) // 2, 0, Nc - 1)
        hi = np.clip((i + 2) // 2, 0, Nc - 1)     # odd i: lo==hi==K ; even i: K, K+1
        return 0.5 * (np.take(f, lo, axis=axis) + np.take(f, hi, axis=axis))

    out = axis_interp(axis_interp(e, e.ndim - 2, Nx), e.ndim - 1, Ny)
    return zero_ghosts(out)


# ---------------------------------------------------------------- multigrid
class Hierarchy:
    def __init__(self, P, alpha, beta, min_size=8, max_levels=12):
        self.alpha, self.beta = float(alpha), float(beta)
        self.images = [np.asarray(P, dtype=np.float64)]
        self.hinv = [1.0]
        while len(self.images) < max_levels:
            Nx, Ny = self.images[-1].shape
            if min(coarse_size(Nx), coarse_size(Ny)) < min_size or coarse_size(Nx) >= Nx:
                break
            self.images.append(restrict_image(self.images[-1]))
            self.hinv.append(self.hinv[-1] * 0.5)

    def apply(self, l, x):
        return apply_eliminated(self.images[l], x, self.alpha, self.beta, self.hinv[l])

    def jacobi(self, l, r):
        return apply_block_jacobi(self.images[l], r, self.alpha, self.beta, self.hinv[l])

    def vcycle(self, l, b, nu1=2, nu2=2, omega=0.8, coarse_sweeps=30, x=None):
        L = len(self.images) - 1
        if x is None:
            x = omega * self.jacobi(l, b)
            sweeps = nu1 - 1
        else:
            sweeps = nu1
        if l == L:
            sweeps = coarse_sweeps
        for _ in range(sweeps):
            x = x + omega * self.jacobi(l, b - self.apply(l, x))
        if l == L:
            return x
        r = b - self.apply(l, x)
        ec = self.vcycle(l + 1, restrict_residual(r), nu1, nu2, omega, coarse_sweeps)
        x = x + prolong(ec, *x.shape[-2:])
        for _ in range(nu2):
            x = x + omega * self.jacobi(l, b - self.apply(l, x))
        return x


def bicgstab(apply_A, b, x0, precond, rtol=1e-6, max_it=1000, callback=None):
    """Right-preconditioned BiCGStab, convergence on the true-recurrence residual ||r||/||b||."""
    x = x0.copy()
    r = b - apply_A(x)
    rhat = r.copy()
    bnorm = np.linalg.norm(b)
    rho = alpha = omega = 1.0
    v = np.zeros_like(b); p = np.zeros_like(b)
    history = [np.linalg.norm(r) / bnorm]
    if history[-1] <= rtol:
        return x, 0, history
    for it in range(1, max_it + 1):
        rho_new = np.vdot(rhat, r)
        beta = (rho_new / rho) * (alpha / omega)
        rho = rho_new
        p = r + beta * (p - omega * v)
        ph = precond(p)
        v = apply_A(ph)
        alpha = rho / np.vdot(rhat, v)
        s = r - alpha * v
        sh = precond(s)
        t = apply_A(sh)
        omega = np.vdot(t, s) / np.vdot(t, t)
        x = x + alpha * ph + omega * sh
        r = s - omega * t
        history.append(np.linalg.norm(r) / bnorm)
        if callback:
            callback(it, history[-1])
        if history[-1] <= rtol:
            return x, it, history
    return x, max_it, history
