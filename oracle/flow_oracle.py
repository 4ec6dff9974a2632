"""TEST INFRASTRUCTURE ONLY - CPU restatement (numpy/scipy, float64) of the hot path
``variational_optical_flow`` of kursawe/OpticalFlow (source/optical_flow.py:715-1210).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import this module; the product path
(``opticalflow_b200``) never does and fails loudly without its CUDA library.

Parity pinning: the reference ships NO golden vectors or asserting tests for this path
(SURVEY.md section 4).  This restatement is pinned instead against outputs of the
reference's own code, imported unmodified in the build container (oracle/ref_import.py)
and run through its ``use_direct_solver=True`` branch; those outputs are committed as
``tests/golden/*.npz`` by ``oracle/make_golden.py``.  The reference's PETSc KSP branch
(:1081-1145) and real ``skimage`` are not installed anywhere we can run: parity of the
iterative branch and of the blur against real skimage is "unpinned" (see DESIGN.md).

Each function cites the reference lines it restates.
"""
import numpy as np
import scipy.sparse
import scipy.sparse.linalg

FIELD_NAMES = ('ux', 'uy', 'remodelling')


# --------------------------------------------------------------------------------------
# blur  (optical_flow.py:282-306 -> skimage.filters.gaussian -> scipy.ndimage.gaussian_filter,
#        mode='nearest', truncate=4.0; separable correlate1d along axis 0 then axis 1)
# --------------------------------------------------------------------------------------
def gaussian_weights(sigma, truncate=4.0):
    radius = int(truncate * float(sigma) + 0.5)
    k = np.arange(-radius, radius + 1, dtype=np.float64)
    w = np.exp(-0.5 / (float(sigma) * float(sigma)) * k * k)
    return w / w.sum()


def gaussian_blur(frame, sigma):
    frame = np.asarray(frame, dtype=np.float64)
    w = gaussian_weights(sigma)
    r = (len(w) - 1) // 2
    out = frame
    for axis in (0, 1):
        pad = [(0, 0), (0, 0)]
        pad[axis] = (r, r)
        padded = np.pad(out, pad, mode='edge')
        acc = np.zeros_like(out)
        n = out.shape[axis]
        for k in range(2 * r + 1):
            sl = [slice(None), slice(None)]
            sl[axis] = slice(k, k + n)
            acc += w[k] * padded[tuple(sl)]
        out = acc
    return out


def blur_movie(movie, smoothing_sigma):
    """optical_flow.py:282-306."""
    blurred = np.zeros_like(movie, dtype='double')
    for index in range(movie.shape[0]):
        blurred[index] = gaussian_blur(movie[index], smoothing_sigma)
    return blurred


# --------------------------------------------------------------------------------------
# helpers (optical_flow.py:676-713, :1241-1316)
# --------------------------------------------------------------------------------------
def numerical_derivative(m, rule):
    """optical_flow.py:676-713 - including the 'dy' == 'dx' copy-paste (:698-699)."""
    if rule in ('dx', 'dy'):
        return (m[2:, 1:-1] - m[:-2, 1:-1]) / 2
    if rule in ('dxy', 'dyx'):
        return (m[2:, 2:] - m[2:, :-2] - m[:-2, 2:] + m[:-2, :-2]) / 4
    if rule == 'dxx':
        return m[2:, 1:-1] + m[:-2, 1:-1] - 2 * m[1:-1, 1:-1]
    if rule == 'dyy':
        return m[1:-1, 2:] + m[1:-1, :-2] - 2 * m[1:-1, 1:-1]
    if rule == 'bar_x':
        return m[2:, 1:-1] + m[:-2, 1:-1]
    if rule == 'bar_y':
        return m[1:-1, 2:] + m[1:-1, :-2]
    if rule == 'bar':
        return m[:-2, 1:-1] + m[2:, 1:-1] + m[1:-1, 2:] + m[1:-1, :-2]
    raise ValueError(rule)


def apply_constant_boundary_condition(image):
    """optical_flow.py:1304-1316 (in place, this order)."""
    image[0, :] = image[2, :]
    image[-1, :] = image[-3, :]
    image[:, 0] = image[:, 2]
    image[:, -1] = image[:, -3]


def get_index_set(N_i, N_j, i_offset, j_offset, quantity, include_boundaries=False):
    """optical_flow.py:1241-1302: idx = 3*N_j*(i+di) + 3*(j+dj) + q."""
    if include_boundaries:
        i0, i1, j0, j1 = 0, N_i, 0, N_j
    else:
        i0, i1, j0, j1 = 1, N_i - 1, 1, N_j - 1
    ii, jj = np.meshgrid(np.arange(i0, i1, dtype=np.int32), np.arange(j0, j1, dtype=np.int32),
                         indexing='ij')
    q = FIELD_NAMES.index(quantity)
    return (3 * N_j * (ii + i_offset) + 3 * (jj + j_offset) + q).flatten()


def planes_to_interleaved(planes):
    """[3,Nx,Ny] SoA planes -> reference DOF vector (idx = 3*Ny*i + 3*j + q)."""
    return np.ascontiguousarray(np.moveaxis(planes, 0, -1)).reshape(-1)


def interleaved_to_planes(vector, Nx, Ny):
    return np.ascontiguousarray(np.moveaxis(vector.reshape(Nx, Ny, 3), -1, 0))


# --------------------------------------------------------------------------------------
# derivative block (optical_flow.py:812-827)
# --------------------------------------------------------------------------------------
def image_coefficients(P):
    """Interior coefficient fields of the previous frame. dIdy := dIdx (reference :813 + :698)."""
    I = P[1:-1, 1:-1]
    Ix = numerical_derivative(P, 'dx')
    Iy = numerical_derivative(P, 'dy')
    Ixx = numerical_derivative(P, 'dxx')
    Iyy = numerical_derivative(P, 'dyy')
    Ixy = numerical_derivative(P, 'dyx')
    return I, Ix, Iy, Ixx, Iyy, Ixy


def build_rhs(P, C):
    """RHS as 3 SoA planes [3,Nx,Ny]; boundary rows are 0 (optical_flow.py:889,:938,:962)."""
    Nx, Ny = P.shape
    b = np.zeros((3, Nx, Ny))
    I = P[1:-1, 1:-1]
    Ixt = (C[2:, 1:-1] - C[:-2, 1:-1] - P[2:, 1:-1] + P[:-2, 1:-1]) / 2      # :815-816
    Iyt = (C[1:-1, 2:] - C[1:-1, :-2] - P[1:-1, 2:] + P[1:-1, :-2]) / 2      # :818-819
    It = (C - P)[1:-1, 1:-1]                                                # :821-823
    b[0, 1:-1, 1:-1] = -I * Ixt
    b[1, 1:-1, 1:-1] = -I * Iyt
    b[2, 1:-1, 1:-1] = -It
    return b


# --------------------------------------------------------------------------------------
# the linear operator, matrix-free (optical_flow.py:836-1070)
# --------------------------------------------------------------------------------------
def _shift(f, di, dj):
    """f(i+di, j+dj) on the interior."""
    Nx, Ny = f.shape
    return f[1 + di:Nx - 1 + di, 1 + dj:Ny - 1 + dj]


def apply_operator(P, x, speed_alpha, remodelling_alpha):
    """y = A x for SoA planes x[3,Nx,Ny]; rows exactly as assembled at :843-1070."""
    a, be = float(speed_alpha), float(remodelling_alpha)
    ux, uy, g = x[0], x[1], x[2]
    Nx, Ny = P.shape
    I, Ix, Iy, Ixx, Iyy, Ixy = image_coefficients(P)
    y = np.zeros_like(x)
    cross = lambda f: _shift(f, -1, -1) + _shift(f, 1, 1) - _shift(f, -1, 1) - _shift(f, 1, -1)
    # ux rows :843-887
    y[0, 1:-1, 1:-1] = ((I * (Ixx - 2 * I) - 4 * a) * _shift(ux, 0, 0) + I * Ixy * _shift(uy, 0, 0)
                        + (I * (I - Ix) + a) * _shift(ux, -1, 0) + (I * (I + Ix) + a) * _shift(ux, 1, 0)
                        + a * _shift(ux, 0, -1) + a * _shift(ux, 0, 1)
                        + (I * Ix / 2) * (_shift(uy, 0, 1) - _shift(uy, 0, -1))
                        + (I * Iy / 2) * (_shift(uy, 1, 0) - _shift(uy, -1, 0))
                        + (I * I / 4) * cross(uy)
                        + (I / 2) * (_shift(g, -1, 0) - _shift(g, 1, 0)))
    # uy rows :892-936
    y[1, 1:-1, 1:-1] = ((I * (Iyy - 2 * I) - 4 * a) * _shift(uy, 0, 0) + I * Ixy * _shift(ux, 0, 0)
                        + (I * (I - Iy) + a) * _shift(uy, 0, -1) + (I * (I + Iy) + a) * _shift(uy, 0, 1)
                        + a * _shift(uy, -1, 0) + a * _shift(uy, 1, 0)
                        + (I * Iy / 2) * (_shift(ux, 1, 0) - _shift(ux, -1, 0))
                        + (I * Ix / 2) * (_shift(ux, 0, 1) - _shift(ux, 0, -1))
                        + (I * I / 4) * cross(ux)
                        + (I / 2) * (_shift(g, 0, -1) - _shift(g, 0, 1)))
    # remodelling rows :942-960
    y[2, 1:-1, 1:-1] = ((-1 - 4 * be) * _shift(g, 0, 0) + Ix * _shift(ux, 0, 0) + Iy * _shift(uy, 0, 0)
                        + be * (_shift(g, -1, 0) + _shift(g, 1, 0) + _shift(g, 0, -1) + _shift(g, 0, 1))
                        + (I / 2) * (_shift(ux, 1, 0) - _shift(ux, -1, 0))
                        + (I / 2) * (_shift(uy, 0, 1) - _shift(uy, 0, -1)))
    # boundary rows :965-972, :1009-1016, :1053-1070.  Corners are in two sets: diagonal stays 1
    # (assignment, not accumulation) and both off-diagonals are present.
    for q in range(3):
        f = x[q]
        yb = y[q]
        yb[0, :] = f[0, :] - f[2, :]
        yb[-1, :] = f[-1, :] - f[-3, :]
        yb[1:-1, 0] = f[1:-1, 0] - f[1:-1, 2]
        yb[1:-1, -1] = f[1:-1, -1] - f[1:-1, -3]
        yb[0, 0] -= f[0, 2]
        yb[0, -1] -= f[0, -3]
        yb[-1, 0] -= f[-1, 2]
        yb[-1, -1] -= f[-1, -3]
    return y


def assemble_csr(P, speed_alpha, remodelling_alpha):
    """The same matrix as the reference's LIL assembly (:836-1072), built by vectorised COO.

    Ordering is the reference's interleaved DOF numbering (get_index_set)."""
    a, be = float(speed_alpha), float(remodelling_alpha)
    Nx, Ny = P.shape
    n = 3 * Nx * Ny
    I, Ix, Iy, Ixx, Iyy, Ixy = [np.ascontiguousarray(f).ravel() for f in image_coefficients(P)]
    one = np.ones_like(I)
    rows, cols, vals = [], [], []

    def put(rq, cq, di, dj, v):
        rows.append(get_index_set(Nx, Ny, 0, 0, rq))
        cols.append(get_index_set(Nx, Ny, di, dj, cq))
        vals.append(np.broadcast_to(v, I.shape).astype(np.float64))

    # ux rows
    put('ux', 'ux', 0, 0, I * (Ixx - 2 * I) - 4 * a)
    put('ux', 'uy', 0, 0, I * Ixy)
    put('ux', 'ux', -1, 0, I * (I - Ix) + a)
    put('ux', 'ux', 1, 0, I * (I + Ix) + a)
    put('ux', 'ux', 0, -1, a * one)
    put('ux', 'ux', 0, 1, a * one)
    put('ux', 'uy', 0, -1, -I * Ix / 2)
    put('ux', 'uy', 0, 1, I * Ix / 2)
    put('ux', 'uy', -1, 0, -I * Iy / 2)
    put('ux', 'uy', 1, 0, I * Iy / 2)
    put('ux', 'uy', -1, -1, I * I / 4)
    put('ux', 'uy', 1, 1, I * I / 4)
    put('ux', 'uy', -1, 1, -I * I / 4)
    put('ux', 'uy', 1, -1, -I * I / 4)
    put('ux', 'remodelling', -1, 0, I / 2)
    put('ux', 'remodelling', 1, 0, -I / 2)
    # uy rows
    put('uy', 'uy', 0, 0, I * (Iyy - 2 * I) - 4 * a)
    put('uy', 'ux', 0, 0, I * Ixy)
    put('uy', 'uy', 0, -1, I * (I - Iy) + a)
    put('uy', 'uy', 0, 1, I * (I + Iy) + a)
    put('uy', 'uy', -1, 0, a * one)
    put('uy', 'uy', 1, 0, a * one)
    put('uy', 'ux', -1, 0, -I * Iy / 2)
    put('uy', 'ux', 1, 0, I * Iy / 2)
    put('uy', 'ux', 0, -1, -I * Ix / 2)
    put('uy', 'ux', 0, 1, I * Ix / 2)
    put('uy', 'ux', -1, -1, I * I / 4)
    put('uy', 'ux', 1, 1, I * I / 4)
    put('uy', 'ux', -1, 1, -I * I / 4)
    put('uy', 'ux', 1, -1, -I * I / 4)
    put('uy', 'remodelling', 0, -1, I / 2)
    put('uy', 'remodelling', 0, 1, -I / 2)
    # remodelling rows
    put('remodelling', 'remodelling', 0, 0, (-1 - 4 * be) * one)
    put('remodelling', 'ux', 0, 0, Ix)
    put('remodelling', 'uy', 0, 0, Iy)
    for di, dj in ((-1, 0), (1, 0), (0, -1), (0, 1)):
        put('remodelling', 'remodelling', di, dj, be * one)
    put('remodelling', 'ux', -1, 0, -I / 2)
    put('remodelling', 'ux', 1, 0, I / 2)
    put('remodelling', 'uy', 0, -1, -I / 2)
    put('remodelling', 'uy', 0, 1, I / 2)
    # boundary rows (unique diagonal + one mirror entry per set the pixel belongs to)
    ii, jj = np.meshgrid(np.arange(Nx), np.arange(Ny), indexing='ij')
    on_edge = (ii == 0) | (ii == Nx - 1) | (jj == 0) | (jj == Ny - 1)
    for q in range(3):
        idx = lambda i, j: 3 * Ny * i + 3 * j + q
        bi, bj = ii[on_edge], jj[on_edge]
        rows.append(idx(bi, bj)); cols.append(idx(bi, bj)); vals.append(np.ones(bi.size))
        for mask, mi, mj in ((ii == 0, 2, 0), (ii == Nx - 1, -2, 0), (jj == 0, 0, 2), (jj == Ny - 1, 0, -2)):
            bi, bj = ii[mask], jj[mask]
            rows.append(idx(bi, bj)); cols.append(idx(bi + mi, bj + mj)); vals.append(-np.ones(bi.size))
    A = scipy.sparse.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))),
                                shape=(n, n))
    return A.tocsr()


# --------------------------------------------------------------------------------------
# epilogue (optical_flow.py:1159-1183)
# --------------------------------------------------------------------------------------
def functionals(P, C, v_x, v_y, remodelling, speed_alpha, remodelling_alpha):
    """L1 / speed / remodelling functionals of one pair, pixel units, 'dy'=='dx' (:1167-1183)."""
    I, Ix, Iy, _, _, _ = image_coefficients(P)
    It = (C - P)[1:-1, 1:-1]
    dvx_dx = numerical_derivative(v_x, 'dx'); dvx_dy = numerical_derivative(v_x, 'dy')
    dvy_dx = numerical_derivative(v_y, 'dx'); dvy_dy = numerical_derivative(v_y, 'dy')
    dr_dx = numerical_derivative(remodelling, 'dx'); dr_dy = numerical_derivative(remodelling, 'dy')
    L1 = np.sum(np.power(It + v_x[1:-1, 1:-1] * Ix + v_y[1:-1, 1:-1] * Iy
                         + I * dvx_dx + I * dvy_dy - remodelling[1:-1, 1:-1], 2))
    speed_f = speed_alpha * np.sum(dvx_dx ** 2 + dvx_dy ** 2 + dvy_dx ** 2 + dvy_dy ** 2)
    rem_f = remodelling_alpha * np.sum(dr_dx ** 2 + dr_dy ** 2)
    return L1, speed_f, rem_f


def solve_pair(P, C, speed_alpha, remodelling_alpha):
    """Direct solve of one pair (the reference's spsolve branch :1146-1147).

    Returns raw solution planes [3,Nx,Ny] and the true relative residual (:1151)."""
    Nx, Ny = P.shape
    A = assemble_csr(P, speed_alpha, remodelling_alpha)
    b = planes_to_interleaved(build_rhs(P, C))
    x = scipy.sparse.linalg.spsolve(A, b)
    relres = np.linalg.norm(A @ x - b) / np.linalg.norm(b)
    return interleaved_to_planes(x, Nx, Ny), relres


def variational_optical_flow(movie, delta_x=1.0, delta_t=1.0, speed_alpha=1.0,
                             remodelling_alpha=1000.0, smoothing_sigma=None,
                             initial_v_x=0.0, initial_v_y=0.0, initial_remodelling=0.0,
                             use_direct_solver=True):
    """Restatement of optical_flow.py:715-1210 with the direct-solver branch.

    initial_* are accepted for signature parity; a direct solve does not depend on them."""
    movie = np.asarray(movie).astype(np.float64)
    analysed = blur_movie(movie, smoothing_sigma) if smoothing_sigma is not None else movie
    T, Nx, Ny = movie.shape
    all_v_x = np.zeros((T - 1, Nx, Ny)); all_v_y = np.zeros((T - 1, Nx, Ny))
    all_rem = np.zeros((T - 1, Nx, Ny))
    L1 = np.zeros(T); rem_f = np.zeros(T); speed_f = np.zeros(T)
    residuals = []
    for k in range(1, T):
        P, C = analysed[k - 1], analysed[k]
        x, relres = solve_pair(P, C, speed_alpha, remodelling_alpha)
        residuals.append(relres)
        v_x, v_y, rem = all_v_x[k - 1], all_v_y[k - 1], all_rem[k - 1]
        v_x[:], v_y[:], rem[:] = x[0], x[1], x[2]
        for f in (v_x, v_y, rem):
            apply_constant_boundary_condition(f)
        L1[k], speed_f[k], rem_f[k] = functionals(P, C, v_x, v_y, rem, speed_alpha, remodelling_alpha)
    all_v_x *= delta_x / delta_t
    all_v_y *= delta_x / delta_t
    result = dict(v_x=all_v_x, v_y=all_v_y, speed=np.sqrt(all_v_x ** 2 + all_v_y ** 2),
                  remodelling=all_rem, original_data=movie, delta_x=delta_x, delta_t=delta_t,
                  blurred_data=analysed, converged=bool(max(residuals) < 1e-6),
                  L1_functional=np.sum(L1), remodelling_functional=np.sum(rem_f),
                  speed_functional=np.sum(rem_f))      # :1205 assigns the remodelling functional
    result['relative_residuals'] = residuals
    return result
