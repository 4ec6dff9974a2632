"""TEST INFRASTRUCTURE ONLY - CPU restatement (numpy) of the callers either side of the hot path
(SURVEY.md section 8f NEXT-3 / NEXT-4) in kursawe/OpticalFlow ``source/optical_flow.py``:

* ``liu_shen_optical_flow_jit``            :426-673   (2-field Jacobi iteration, deprecated driver :1318-1529)
* ``apply_adaptive_threshold``             :308-338   (-> cv2.adaptiveThreshold, MEAN_C / THRESH_BINARY)
* ``apply_clahe``                          :340-374   (-> cv2.createCLAHE(...).apply on uint16 frames)
* background subtraction                   :195-198   (inline in ``conduct_optical_flow``)

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module.

Parity pinning: numba 0.65 and opencv-python 4.13 ARE installed in the build container, so the reference's own
functions run there unmodified (``oracle/ref_import.py``); ``oracle/make_golden.py`` committed their outputs as
``tests/golden/legacy_*.npz`` and ``tests/test_oracle_cpu.py`` holds this restatement to them (Liu-Shen: 1e-12
relative; threshold and CLAHE: bit-exact).  The OpenCV algorithms are restated from OpenCV 4.x
``modules/imgproc/src/thresh.cpp`` (adaptiveThreshold), ``box_filter`` (normalised uint8 box mean, replicate
border) and ``clahe.cpp`` (16-bit path: 65536 bins, clip + redistribute, float LUT scaling, bilinear LUT blend).
"""
import numpy as np


# --------------------------------------------------------------------------------------
# Liu-Shen Jacobi iteration, optical_flow.py:426-673
# --------------------------------------------------------------------------------------
def _mirror(a):
    """apply_constant_boundary_condition (:1304-1316) on a frame padded by one ring: ghost = mirror about the edge pixel."""
    return np.pad(a, 1, mode='reflect')


def liu_shen_coefficients(previous, current):
    """Per-pixel image terms of :529-560 (previous/current already carry the mirrored ring)."""
    P, C = previous, current
    c = (slice(1, -1), slice(1, -1))
    up, dn = (slice(2, None), slice(1, -1)), (slice(0, -2), slice(1, -1))          # i+1, i-1
    rt, lt = (slice(1, -1), slice(2, None)), (slice(1, -1), slice(0, -2))          # j+1, j-1
    out = dict()
    out['I'] = P[c]
    out['Ix'] = (P[up] - P[dn]) / 2
    out['Iy'] = (P[rt] - P[lt]) / 2
    out['Ixt'] = (C[up] - C[dn] - P[up] + P[dn]) / 2
    out['Iyt'] = (C[rt] - C[lt] - P[rt] + P[lt]) / 2
    out['Ixx'] = P[up] + P[dn] - 2 * P[c]
    out['Iyy'] = P[rt] + P[lt] - 2 * P[c]
    out['Ixy'] = (P[2:, 2:] - P[2:, :-2] - P[:-2, 2:] + P[:-2, :-2]) / 4
    return out


def liu_shen_jacobi_sweep(vx, vy, coef, alpha):
    """One pass of the pixel loop :512-648 over all pixels (Jacobi: reads the old iterate only)."""
    Nx, Ny = vx.shape
    mx, my = _mirror(vx), _mirror(vy)                       # :509-511, derivative terms see mirrored ghosts
    zx, zy = np.pad(vx, 1), np.pad(vy, 1)                   # :522-538, the 8-neighbour sum sees zeros outside
    c = (slice(1, -1), slice(1, -1))
    up, dn = (slice(2, None), slice(1, -1)), (slice(0, -2), slice(1, -1))
    rt, lt = (slice(1, -1), slice(2, None)), (slice(1, -1), slice(0, -2))

    def d_x(m): return (m[up] - m[dn]) / 2
    def d_y(m): return (m[rt] - m[lt]) / 2
    def d_xy(m): return (m[2:, 2:] - m[2:, :-2] - m[:-2, 2:] + m[:-2, :-2]) / 4
    def bar8(z): return (z[dn] + z[up] + z[rt] + z[lt] + z[:-2, :-2] + z[:-2, 2:] + z[2:, :-2] + z[2:, 2:])

    I, Ix, Iy = coef['I'], coef['Ix'], coef['Iy']
    F0 = (-I * coef['Ixt'] - I * (2 * Ix * d_x(mx) + Iy * d_x(my) + Ix * d_y(my))
          - I ** 2 * (mx[up] + mx[dn] + d_xy(my)) - alpha * bar8(zx))
    F1 = (-I * coef['Iyt'] - I * (2 * Iy * d_y(my) + Ix * d_y(mx) + Iy * d_x(mx))
          - I ** 2 * (my[rt] + my[lt] + d_xy(mx)) - alpha * bar8(zy))
    # boundary_prefactor :619-629: 8 inside, 5 on an edge, 3 in a corner
    edge_i = np.zeros((Nx, 1), bool)
    edge_i[0] = edge_i[-1] = True
    edge_j = np.zeros((1, Ny), bool)
    edge_j[0, 0] = edge_j[0, -1] = True
    prefactor = np.where(edge_i & edge_j, 3.0, np.where(edge_i | edge_j, 5.0, 8.0))
    a00 = I * coef['Ixx'] - 2 * I ** 2 - prefactor * alpha
    a01 = I * coef['Ixy']
    a11 = I * coef['Iyy'] - 2 * I ** 2 - prefactor * alpha
    det = a00 * a11 - a01 * a01
    return (a11 * F0 - a01 * F1) / det, (a00 * F1 - a01 * F0) / det


def liu_shen_optical_flow(movie, delta_x=1.0, delta_t=1.0, alpha=100, remodelling_alpha=1.0, initial_v_x=None,
                          initial_v_y=None, initial_remodelling=None, max_iterations=10, tolerance=1e-9,
                          include_remodelling=True):
    """optical_flow.py:426-673.  Returns (v_x, v_y, speed, remodelling, last_iteration_index) like the reference."""
    movie = np.asarray(movie, dtype=np.float64)
    T, Nx, Ny = movie.shape
    zeros = np.zeros((Nx, Ny))
    initial_v_x = zeros if initial_v_x is None else np.asarray(initial_v_x, float)
    initial_v_y = zeros if initial_v_y is None else np.asarray(initial_v_y, float)
    initial_remodelling = zeros if initial_remodelling is None else np.asarray(initial_remodelling, float)
    all_vx = np.zeros((T - 1, Nx, Ny))
    all_vy = np.zeros((T - 1, Nx, Ny))
    all_rem = np.zeros((T - 1, Nx, Ny))
    for k in range(1, T):
        coef = liu_shen_coefficients(_mirror(movie[k - 1]), _mirror(movie[k]))
        vx = initial_v_x * delta_t / delta_x                 # :497-498 (every pair restarts from the guess)
        vy = initial_v_y * delta_t / delta_x
        for _ in range(max_iterations):
            vx, vy = liu_shen_jacobi_sweep(vx, vy, coef, alpha)
        all_vx[k - 1], all_vy[k - 1], all_rem[k - 1] = vx, vy, initial_remodelling
    all_vx *= delta_x / delta_t                              # :669-670
    all_vy *= delta_x / delta_t
    return all_vx, all_vy, np.sqrt(all_vx ** 2 + all_vy ** 2), all_rem, max_iterations - 1


# --------------------------------------------------------------------------------------
# adaptive threshold, optical_flow.py:308-338 -> cv2.adaptiveThreshold(src, 1.0, MEAN_C, THRESH_BINARY, k, C)
# --------------------------------------------------------------------------------------
def to_uint8_range(movie):
    """:328  np.array(movie/np.max(movie)*255.0, dtype='uint8') (truncation; defined for non-negative movies)."""
    movie = np.asarray(movie)
    return np.array(movie / np.max(movie) * 255.0, dtype='uint8')


def box_mean_u8(frame_u8, window):
    """OpenCV normalised box filter on uint8, BORDER_REPLICATE: integer window sum, scaled by 1/k^2, rounded to
    nearest (k odd => k^2 odd => a tie cannot occur, so the rounding mode does not matter)."""
    r = window // 2
    padded = np.pad(frame_u8.astype(np.int64), r, mode='edge')
    integral = np.zeros((padded.shape[0] + 1, padded.shape[1] + 1), np.int64)
    integral[1:, 1:] = padded.cumsum(0).cumsum(1)
    Nx, Ny = frame_u8.shape
    s = (integral[window:window + Nx, window:window + Ny] - integral[:Nx, window:window + Ny]
         - integral[window:window + Nx, :Ny] + integral[:Nx, :Ny])
    k2 = window * window
    return (2 * s + k2) // (2 * k2)


def apply_adaptive_threshold(movie, window_size=51, threshold=0.0):
    """:308-338.  dst = 1 where src > mean - ceil(threshold) (thresh.cpp: idelta = cvCeil(delta) for THRESH_BINARY,
    tab[src - mean + 255] with tab[i] = (i - 255 > -idelta))."""
    transformed = to_uint8_range(movie)
    idelta = int(np.ceil(threshold))
    out = np.zeros(transformed.shape, dtype=bool)
    for index in range(transformed.shape[0]):
        frame = transformed[index]
        mean = box_mean_u8(frame, window_size)
        out[index] = (frame.astype(np.int64) - mean) > -idelta
    return out


# --------------------------------------------------------------------------------------
# CLAHE, optical_flow.py:340-374 -> cv2.createCLAHE(clipLimit, (tile_number, round(tile_number*Ny/Nx))).apply(uint16)
# --------------------------------------------------------------------------------------
HIST_SIZE = 65536


def clahe_geometry(Nx, Ny, tile_number):
    """tileGridSize = (tilesX, tilesY): tilesX splits the columns (axis 2), tilesY the rows (:364-366)."""
    tiles_x = int(tile_number)
    tiles_y = int(round(tile_number * (Ny / Nx)))
    if Ny % tiles_x == 0 and Nx % tiles_y == 0:
        ext_x, ext_y = Nx, Ny
    else:                                                     # clahe.cpp: copyMakeBorder(..., BORDER_REFLECT_101)
        ext_x, ext_y = Nx + (tiles_y - Nx % tiles_y), Ny + (tiles_x - Ny % tiles_x)
    return tiles_x, tiles_y, ext_x, ext_y, ext_y // tiles_x, ext_x // tiles_y     # ..., tile width, tile height


def clahe_luts(frame_u16, clip_limit, tile_number):
    Nx, Ny = frame_u16.shape
    tiles_x, tiles_y, ext_x, ext_y, tw, th = clahe_geometry(Nx, Ny, tile_number)
    ext = np.pad(frame_u16, ((0, ext_x - Nx), (0, ext_y - Ny)), mode='reflect') if (ext_x, ext_y) != (Nx, Ny) else frame_u16
    area = tw * th
    lut_scale = np.float32(HIST_SIZE - 1) / np.float32(area)
    clip = 0
    if clip_limit > 0.0:
        clip = max(int(clip_limit * area / HIST_SIZE), 1)
    luts = np.zeros((tiles_y, tiles_x, HIST_SIZE), np.uint16)
    for ty in range(tiles_y):
        for tx in range(tiles_x):
            tile = ext[ty * th:(ty + 1) * th, tx * tw:(tx + 1) * tw]
            hist = np.bincount(tile.ravel(), minlength=HIST_SIZE).astype(np.int64)
            if clip > 0:
                clipped = int(np.sum(np.maximum(hist - clip, 0)))
                hist = np.minimum(hist, clip)
                batch = clipped // HIST_SIZE
                residual = clipped - batch * HIST_SIZE
                hist += batch
                if residual != 0:
                    step = max(HIST_SIZE // residual, 1)
                    idx = np.arange(0, HIST_SIZE, step)[:residual]
                    hist[idx] += 1
            cum = np.cumsum(hist)
            scaled = cum.astype(np.float32) * lut_scale              # float multiply, then cvRound (half to even)
            luts[ty, tx] = np.clip(np.rint(scaled), 0, 65535).astype(np.uint16)
    return luts, (tiles_x, tiles_y, tw, th)


def clahe_frame(frame_u16, clip_limit, tile_number):
    Nx, Ny = frame_u16.shape
    luts, (tiles_x, tiles_y, tw, th) = clahe_luts(frame_u16, clip_limit, tile_number)
    f32 = np.float32
    inv_tw, inv_th = f32(1.0) / f32(tw), f32(1.0) / f32(th)
    txf = np.arange(Ny, dtype=np.int32).astype(f32) * inv_tw - f32(0.5)
    tx1 = np.floor(txf).astype(np.int32)
    xa = (txf - tx1.astype(f32)).astype(f32)
    xa1 = (f32(1.0) - xa).astype(f32)
    tx2 = np.minimum(tx1 + 1, tiles_x - 1)
    tx1 = np.maximum(tx1, 0)
    tyf = np.arange(Nx, dtype=np.int32).astype(f32) * inv_th - f32(0.5)
    ty1 = np.floor(tyf).astype(np.int32)
    ya = (tyf - ty1.astype(f32)).astype(f32)[:, None]
    ya1 = (f32(1.0) - ya).astype(f32)
    ty2 = np.minimum(ty1 + 1, tiles_y - 1)
    ty1 = np.maximum(ty1, 0)
    v = frame_u16.astype(np.int64)
    l11 = luts[ty1[:, None], tx1[None, :], v].astype(f32)
    l12 = luts[ty1[:, None], tx2[None, :], v].astype(f32)
    l21 = luts[ty2[:, None], tx1[None, :], v].astype(f32)
    l22 = luts[ty2[:, None], tx2[None, :], v].astype(f32)
    res = (l11 * xa1 + l12 * xa) * ya1 + (l21 * xa1 + l22 * xa) * ya
    return np.clip(np.rint(res), 0, 65535).astype(np.uint16)


def apply_clahe(movie, clipLimit=50000, tile_number=10):
    """:340-374 (returns float64 like the reference)."""
    converted = np.asarray(movie).astype(np.uint16)
    out = np.zeros(converted.shape, dtype='double')
    for index in range(converted.shape[0]):
        out[index] = clahe_frame(converted[index], clipLimit, tile_number)
    return out


# --------------------------------------------------------------------------------------
# background subtraction, optical_flow.py:195-198
# --------------------------------------------------------------------------------------
def subtract_background(movie, blurred_sigma10, background):
    out = np.zeros_like(blurred_sigma10)
    mask = blurred_sigma10 > background
    out[mask] = np.asarray(movie)[mask] - background
    return out
