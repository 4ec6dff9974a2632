import sys, time
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.abspath(__file__)))
import gpuh
from gpuh import *
from opticalflow_b200 import synthetic
N = int(sys.argv[1]); fix = int(sys.argv[2]); gam = [int(c) for c in sys.argv[3]] if len(sys.argv) > 3 else [1]*12
alpha, beta = float(__import__('os').environ.get('ALPHA', 1e5)), 1e4
movie = fo.blur_movie(synthetic.rotating_shearing_movie(N, N, 2, seed=2), 2.5)
P, C = movie
b = mg.zero_ghosts(fo.build_rhs(P, C))
import os
H = GH(P, alpha, beta, coef_mode=os.environ.get('COEF_MODE', 'avg'))
shapes = H.shapes
if os.environ.get('COEF_ROUND'):
    for _l in range(1, len(H.c)):
        if os.environ['COEF_ROUND'] == 'f16':
            H.c[_l] = H.c[_l].astype(np.float16).astype(np.float64)
        else:      # bfloat16: keep 8 significant bits
            _v = np.ascontiguousarray(H.c[_l].astype(np.float32)).view(np.uint32); _v = ((_v + 0x8000) & 0xFFFF0000).astype(np.uint32); H.c[_l] = _v.view(np.float32).astype(np.float64)
# far-side rule per level and axis: 0 = node mirror (ghost = q(N-3)), 1 = midpoint mirror (ghost = q(N-2))
far = [(0, 0)]
pos = [(0.0, 0.0)]   # delta (axis offset beyond last interior node, in units of this level's cells)
for l in range(1, len(shapes)):
    rule = []; dl = []
    for ax in range(2):
        Mf = shapes[l - 1][ax] - 2; df = pos[l - 1][ax]
        d = df / 2 if Mf % 2 == 1 else (df + 1) / 2
        dl.append(d); rule.append((1 if d >= 0.25 else 0) if fix == 1 else (min(2 * d, 1.0) if fix == 2 else 0))
    far.append(tuple(rule)); pos.append(tuple(dl))
print('shapes', shapes); print('delta', pos); print('far rule', far)
cur = {'l': 0}
def fill(f, l, corner_sum=True):
    f[..., 0, :] = f[..., 2, :]
    t = far[l][0]; f[..., -1, :] = (1 - t) * f[..., -3, :] + t * f[..., -2, :]
    f[..., :, 0] = f[..., :, 2]
    t = far[l][1]; f[..., :, -1] = (1 - t) * f[..., :, -3] + t * f[..., :, -2]
    if corner_sum:
        f[..., 0, 0] *= 2; f[..., 0, -1] *= 2; f[..., -1, 0] *= 2; f[..., -1, -1] *= 2
    return f
def apply(l, x):
    saved = mg.fill_ghosts
    mg.fill_ghosts = lambda f, corner_sum=True: fill(f, l, corner_sum)
    try: return H.apply(l, x)
    finally: mg.fill_ghosts = saved
def restrict(r, l):   # fine level l -> l+1
    r = fill(r.copy(), l, corner_sum=False)
    Nx, Ny = r.shape[-2:]
    return mg.zero_ghosts(mg._fw_axis(mg._fw_axis(r, r.ndim - 2, shapes[l + 1][0], True), r.ndim - 1, shapes[l + 1][1], True))
def prolong(e, l):    # coarse level l+1 -> l
    e = fill(e.copy(), l + 1, corner_sum=False)
    Nx, Ny = shapes[l]
    def ai(f, axis, Nn):
        Nc = f.shape[axis]; i = np.arange(Nn)
        lo = np.clip((i + 1) // 2, 0, Nc - 1); hi = np.clip((i + 2) // 2, 0, Nc - 1)
        return 0.5 * (np.take(f, lo, axis=axis) + np.take(f, hi, axis=axis))
    return mg.zero_ghosts(ai(ai(e, e.ndim - 2, Nx), e.ndim - 1, Ny))
def dense(l):
    Nx, Ny = shapes[l]; n = 3 * Nx * Ny; M = np.zeros((n, n)); e = np.zeros((3, Nx, Ny))
    for k in range(n):
        e.flat[k] = 1.0; M[:, k] = apply(l, mg.zero_ghosts(e.copy())).ravel(); e.flat[k] = 0.0
    g = np.ones((3, Nx, Ny), bool); g[:, 1:-1, 1:-1] = False; idx = np.flatnonzero(g.ravel()); M[idx, idx] = 1.0
    return M
LU = {l: scipy.linalg.lu_factor(dense(l)) for l in range(len(shapes) - 1, len(shapes))}
def smooth(l, x, bb, ws):
    for w in ws: x = x + w * H.jacobi(l, bb - apply(l, x))
    return x
def cycle(l, bb, pre, post, exact_at, gamma=1):
    if l == exact_at: return mg.zero_ghosts(scipy.linalg.lu_solve(LU[l], bb.ravel()).reshape(bb.shape))
    x = smooth(l, pre[0] * H.jacobi(l, bb), bb, pre[1:])
    rc = restrict(bb - apply(l, x), l)
    ec = cycle(l + 1, rc, pre, post, exact_at, gamma)
    for _ in range(gamma[l] - 1):
        ec = ec + cycle(l + 1, rc - apply(l + 1, ec), pre, post, exact_at, gamma)
    return smooth(l, x + prolong(ec, l), bb, post)
def radius(pre, post, ex, gamma=1, its=30):
    e = mg.zero_ghosts(np.random.default_rng(1).standard_normal(b.shape))
    for it in range(its):
        en = e - cycle(0, apply(0, e), pre, post, ex, gamma); lam = np.linalg.norm(en) / np.linalg.norm(e); e = en / np.linalg.norm(en)
    return lam
pre, post = [1.0], [1.3, 0.5]
x, it, hist = mg.bicgstab(lambda v: apply(0, v), b, np.zeros_like(b), lambda r: cycle(0, r, pre, post, len(shapes) - 1, gam), rtol=1e-6, max_it=60)
import math
print('N', N, 'fix', fix, 'gamma', gam, 'BiCGStab iterations to 1e-6:', it, ' last relres %.2e' % hist[-1], ' per-cycle factor %.3f' % (hist[-1] ** (1.0 / (2 * it))), flush=True)
