"""Seeded synthetic movies for the five BASELINE.json configs.

Host-side numpy only.  The same bytes are fed to the oracle and to the GPU path.
``gaussian_hat_pair`` restates the reference's only synthetic generator
(optical_flow.py:376-423, used by analysis/analyse_variational_optical_flow.py:26-32).
"""
import numpy as np


def _smooth_periodic(noise, sigma):
    """Gaussian low-pass with periodic wrap (== scipy gaussian_filter(mode='wrap') up to truncation)."""
    Nx, Ny = noise.shape
    fx = np.fft.fftfreq(Nx)[:, None]
    fy = np.fft.rfftfreq(Ny)[None, :]
    transfer = np.exp(-2.0 * (np.pi * sigma) ** 2 * (fx * fx + fy * fy))
    return np.fft.irfft2(np.fft.rfft2(noise) * transfer, s=noise.shape)


def textured_frame(Nx, Ny, seed=0, sigma_texture=2.0, lo=22.0, hi=220.0):
    """Band-limited random texture mapped affinely to [lo, hi]."""
    rng = np.random.default_rng(seed)
    base = _smooth_periodic(rng.standard_normal((Nx, Ny)), sigma_texture)
    base = (base - base.min()) / (base.max() - base.min())
    return lo + (hi - lo) * base


def fourier_shift(frame, shift_x, shift_y):
    """Translate a (periodic) frame by (shift_x, shift_y) pixels along axis 0 / axis 1."""
    Nx, Ny = frame.shape
    fx = np.fft.fftfreq(Nx)[:, None]
    fy = np.fft.rfftfreq(Ny)[None, :]
    phase = np.exp(-2j * np.pi * (fx * shift_x + fy * shift_y))
    return np.fft.irfft2(np.fft.rfft2(frame) * phase, s=frame.shape)


def translating_movie(Nx, Ny, n_frames=2, shift=(1.5, -0.7), seed=0):
    """Config d-1 family: textured frame translated by t*shift pixels."""
    base = textured_frame(Nx, Ny, seed=seed)
    return np.stack([fourier_shift(base, t * shift[0], t * shift[1]) for t in range(n_frames)])


def actin_like_movie(Nx=512, Ny=512, n_frames=100, seed=1, n_filaments=400, shift=(0.8, 0.5)):
    """Config d-2: random line filaments, blurred, noisy, uniformly translating."""
    rng = np.random.default_rng(seed)
    frame = np.zeros((Nx, Ny))
    for _ in range(n_filaments):
        length = rng.uniform(40, 200)
        angle = rng.uniform(0, np.pi)
        x0, y0 = rng.uniform(0, Nx), rng.uniform(0, Ny)
        intensity = rng.uniform(50, 255)
        s = np.arange(0, length, 0.5)
        xi = np.round(x0 + s * np.cos(angle)).astype(int) % Nx
        yi = np.round(y0 + s * np.sin(angle)).astype(int) % Ny
        frame[xi, yi] = np.maximum(frame[xi, yi], intensity)
    frame = _smooth_periodic(frame, 1.5) + 10.0
    movie = np.stack([fourier_shift(frame, t * shift[0], t * shift[1]) for t in range(n_frames)])
    movie += rng.normal(0.0, 2.0, movie.shape)
    return movie


def rotating_shearing_frames(Nx, Ny, indices, seed=2, omega=1e-3, shear=5e-4):
    """Frames ``indices`` of the config d-3 movie (any subset, any order): a frame depends on frame 0 only, so a pool of
    processes can generate a long movie in pieces (bench.py)."""
    base = textured_frame(Nx, Ny, seed=seed)
    ii, jj = np.meshgrid(np.arange(Nx, dtype=np.float64), np.arange(Ny, dtype=np.float64), indexing='ij')
    cx, cy = (Nx - 1) / 2.0, (Ny - 1) / 2.0
    vx = -omega * (jj - cy) + shear * (jj - cy)
    vy = omega * (ii - cx)
    return np.stack([base if t == 0 else _bicubic_sample(base, ii - t * vx, jj - t * vy) for t in indices])


def rotating_shearing_movie(Nx=1024, Ny=1024, n_frames=500, seed=2, omega=1e-3, shear=5e-4):
    """Config d-3: frame t = frame 0 warped by t * v(r), v = omega z x (r-c) + shear (y-cy) x_hat.

    Warp is a backward bicubic (Catmull-Rom) lookup with clamped coordinates."""
    return rotating_shearing_frames(Nx, Ny, range(n_frames), seed, omega, shear)


def _bicubic_sample(img, x, y):
    Nx, Ny = img.shape
    x0 = np.floor(x).astype(np.int64); y0 = np.floor(y).astype(np.int64)
    tx = x - x0; ty = y - y0

    def w(t):
        return (((-t + 2) * t - 1) * t / 2, ((3 * t - 5) * t * t + 2) / 2,
                ((-3 * t + 4) * t + 1) * t / 2, (t - 1) * t * t / 2)

    wx, wy = w(tx), w(ty)
    out = np.zeros_like(x)
    for a in range(4):
        xi = np.clip(x0 + a - 1, 0, Nx - 1)
        for b in range(4):
            yi = np.clip(y0 + b - 1, 0, Ny - 1)
            out += wx[a] * wy[b] * img[xi, yi]
    return out


def gaussian_hat_frame(x_position, y_position, sigma=1.0, width=20.0, dimension=1000):
    """exp(-((x-x0)^2+(y-y0)^2)/sigma^2) on linspace(0,width,dimension) (optical_flow.py:412-418)."""
    x = np.linspace(0, width, dimension)
    frame = np.exp((-(x[:, None] - x_position) ** 2 - (x[None, :] - y_position) ** 2) / sigma ** 2)
    return frame, x[1] - x[0]


def gaussian_hat_pair(dimension=50, v_x=0.1, v_y=0.2, brightening=0.05):
    """Inputs of simple_test_with_data_on_boundary (analysis/analyse_variational_optical_flow.py:27-32)."""
    first, delta_x = gaussian_hat_frame(2.5, 2.5, sigma=3, width=5, dimension=dimension)
    second, _ = gaussian_hat_frame(2.5 + v_x, 2.5 + v_y, sigma=3, width=5, dimension=dimension)
    return np.stack((first, second + brightening)), delta_x
