"""numpy model of the GPU hierarchy (averaged coefficient planes)."""
import sys
sys.path.insert(0, '/root/repo')
import numpy as np, scipy.linalg
from oracle import mg_model as mg, flow_oracle as fo

def mirror(i, N):
    i = np.clip(i, 0, N - 1)
    return np.where(i == 0, 2, np.where(i == N - 1, N - 3, i))

def coefs_from_image(P):
    """7 planes at every pixel whose 3x3 neighbourhood exists (interior); ghost ring filled by reflection."""
    Nx, Ny = P.shape
    c = np.zeros((7, Nx, Ny))
    I = P[1:-1, 1:-1]
    Ix = 0.5 * (P[2:, 1:-1] - P[:-2, 1:-1])
    Ixx = P[2:, 1:-1] + P[:-2, 1:-1] - 2 * I
    Iyy = P[1:-1, 2:] + P[1:-1, :-2] - 2 * I
    Ixy = 0.25 * (P[2:, 2:] - P[2:, :-2] - P[:-2, 2:] + P[:-2, :-2])
    for q, v in enumerate((I * I, I * Ix, I * Ixx, I * Iyy, I * Ixy, I, Ix)):
        c[q, 1:-1, 1:-1] = v
    return c

def fw_coef(c, scale=True):
    """full weighting around fine 2K-1 with reflect-clamped indices, all coarse nodes incl. ghosts."""
    _, Nx, Ny = c.shape
    Ncx, Ncy = mg.coarse_size(Nx), mg.coarse_size(Ny)
    out = np.zeros((c.shape[0], Ncx, Ncy))
    K = np.arange(Ncx); L = np.arange(Ncy)
    for di, wi in ((-1, .25), (0, .5), (1, .25)):
        ii = mirror(2 * K - 1 + di, Nx)
        for dj, wj in ((-1, .25), (0, .5), (1, .25)):
            jj = mirror(2 * L - 1 + dj, Ny)
            out += wi * wj * c[:, ii][:, :, jj]
    if scale:
        out *= np.array([.25, .5, 1, 1, 1, .5, 1])[:, None, None]
    return out

def apply_c(c, x, a, be):
    """eliminated operator with coefficient planes (pre-scaled); a, be already scaled by h2."""
    x = mg.fill_ghosts(x.copy())
    ux, uy, g = x
    s = mg._s
    II, IIx, IIxx, IIyy, IIxy, Ih, Ix = [p[1:-1, 1:-1] for p in c]
    cross = lambda f: s(f, -1, -1) + s(f, 1, 1) - s(f, -1, 1) - s(f, 1, -1)
    y = np.zeros_like(x)
    y[0, 1:-1, 1:-1] = ((IIxx - 2 * II - 4 * a) * s(ux, 0, 0) + IIxy * s(uy, 0, 0) + (II - IIx + a) * s(ux, -1, 0) + (II + IIx + a) * s(ux, 1, 0)
                        + a * (s(ux, 0, -1) + s(ux, 0, 1)) + 0.5 * IIx * (s(uy, 0, 1) - s(uy, 0, -1) + s(uy, 1, 0) - s(uy, -1, 0))
                        + 0.25 * II * cross(uy) + 0.5 * Ih * (s(g, -1, 0) - s(g, 1, 0)))
    y[1, 1:-1, 1:-1] = ((IIyy - 2 * II - 4 * a) * s(uy, 0, 0) + IIxy * s(ux, 0, 0) + (II - IIx + a) * s(uy, 0, -1) + (II + IIx + a) * s(uy, 0, 1)
                        + a * (s(uy, -1, 0) + s(uy, 1, 0)) + 0.5 * IIx * (s(ux, 1, 0) - s(ux, -1, 0) + s(ux, 0, 1) - s(ux, 0, -1))
                        + 0.25 * II * cross(ux) + 0.5 * Ih * (s(g, 0, -1) - s(g, 0, 1)))
    y[2, 1:-1, 1:-1] = ((-1 - 4 * be) * s(g, 0, 0) + Ix * (s(ux, 0, 0) + s(uy, 0, 0)) + be * (s(g, -1, 0) + s(g, 1, 0) + s(g, 0, -1) + s(g, 0, 1))
                        + 0.5 * Ih * (s(ux, 1, 0) - s(ux, -1, 0) + s(uy, 0, 1) - s(uy, 0, -1)))
    return y

def jacobi_c(c, r, a, be):
    II, IIx, IIxx, IIyy, IIxy, Ih, Ix = [p[1:-1, 1:-1] for p in c]
    d0 = IIxx - 2 * II - 4 * a; d1 = IIyy - 2 * II - 4 * a; d2 = -1 - 4 * be
    det = d0 * d1 - IIxy ** 2
    r0, r1, r2 = r[0, 1:-1, 1:-1], r[1, 1:-1, 1:-1], r[2, 1:-1, 1:-1]
    z = np.zeros_like(r)
    z0 = (d1 * r0 - IIxy * r1) / det; z1 = (d0 * r1 - IIxy * r0) / det
    z[0, 1:-1, 1:-1] = z0; z[1, 1:-1, 1:-1] = z1; z[2, 1:-1, 1:-1] = (r2 - Ix * (z0 + z1)) / d2
    return z

class GH:
    def __init__(self, P, alpha, beta, min_size=6, coef_mode='avg'):
        self.alpha, self.beta = alpha, beta
        c0 = coefs_from_image(P)
        self.c = [c0]; self.h2 = [1.0]; self.shapes = [P.shape]
        img = P
        while True:
            Nx, Ny = self.shapes[-1]
            cx, cy = mg.coarse_size(Nx), mg.coarse_size(Ny)
            if cx < min_size or cy < min_size or cx >= Nx: break
            if coef_mode == 'avg':
                if len(self.c) == 1:
                    # level 1: FW of level-0 coefficient fields evaluated at reflect-clamped fine pixels
                    c0g = c0.copy()
                    self.c.append(fw_coef(c0g))
                else:
                    self.c.append(fw_coef(self.c[-1]))
            else:
                img = mg.restrict_image(img)
                hinv = np.sqrt(self.h2[-1]) * 0.5
                cc = coefs_from_image(img)
                cc *= np.array([hinv**2, hinv * hinv, hinv**2, hinv**2, hinv**2, hinv, hinv])[:, None, None]
                self.c.append(cc)
            self.h2.append(self.h2[-1] * 0.25); self.shapes.append((cx, cy))
    def apply(self, l, x): return apply_c(self.c[l], x, self.alpha * self.h2[l], self.beta * self.h2[l])
    def jacobi(self, l, r): return jacobi_c(self.c[l], r, self.alpha * self.h2[l], self.beta * self.h2[l])
