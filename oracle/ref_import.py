"""TEST INFRASTRUCTURE ONLY - never imported by the product path.

Imports the *unmodified* reference module ``/root/reference/source/optical_flow.py``
in THIS container so that its own derivative / assembly / spsolve / boundary /
functional code can be executed (``use_direct_solver=True`` branch,
optical_flow.py:1146-1147).  Three third-party imports of the reference are not
installed here (matplotlib :4-12, skimage :13, petsc4py :17); inert stub modules
are injected for them first.  ``skimage.filters.gaussian`` is stubbed by its
documented delegation ``scipy.ndimage.gaussian_filter(mode='nearest',
truncate=4.0)`` - parity of the blur against real skimage is therefore unpinned.

``/root/reference`` does not exist on the GPU box.  ``oracle/build_ref.py`` (run by
``__graft_entry__.build()`` in the build container) places an untouched copy of the one
reference module this path consists of in ``oracle/_ref/`` - a build output, git-ignored,
never committed - which travels to the box with the snapshot; this importer falls back to
it.  Used by ``oracle/make_golden.py`` (fixture generator), by the ``not gpu`` tests that
validate ``oracle/flow_oracle.py`` against the reference, and by ``bench.py``'s reference
arm / ``cpu_baseline`` leg (the reference's own CPU code timed on the box's host cores).
"""
import contextlib
import io
import os
import sys
import types

import numpy as np

_CANDIDATES = ('/root/reference/source', os.path.join(os.path.dirname(os.path.abspath(__file__)), '_ref'))
REFERENCE_SOURCE_DIR = next((d for d in _CANDIDATES if os.path.isfile(os.path.join(d, 'optical_flow.py'))), _CANDIDATES[0])


def reference_available():
    return os.path.isfile(os.path.join(REFERENCE_SOURCE_DIR, 'optical_flow.py'))


class _AnyModule(types.ModuleType):
    """Attribute-tolerant dummy module (plt.rc(...) is called at import time, :7-9)."""

    def __getattr__(self, key):
        if key.startswith('__'):
            raise AttributeError(key)
        sub = _AnyModule(self.__name__ + '.' + key)
        setattr(self, key, sub)
        return sub

    def __call__(self, *args, **kwargs):
        return None


class _InertPetscObject:
    is_converged = True

    def __getattr__(self, key):
        return lambda *a, **kw: self


class _KSP:
    class NormType:
        NORM_UNPRECONDITIONED = 0

    def create(self):
        return _InertPetscObject()


class _PETSc:
    KSP = _KSP

    def Options(self):
        return _InertPetscObject()

    def Vec(self):
        return _InertPetscObject()

    def Mat(self):
        return _InertPetscObject()


_reference_module = None


def import_reference():
    """Return the reference ``optical_flow`` module, imported unmodified behind stubs."""
    global _reference_module
    if _reference_module is not None:
        return _reference_module
    if not reference_available():
        raise RuntimeError('reference sources are not present at ' + REFERENCE_SOURCE_DIR)
    import scipy.ndimage
    import scipy.sparse.linalg  # noqa: F401  (the reference uses it at :1147 without importing it)

    for name in ['matplotlib', 'matplotlib.pyplot', 'matplotlib.ticker', 'matplotlib.animation',
                 'skimage', 'skimage.filters', 'petsc4py']:
        if name not in sys.modules:
            sys.modules[name] = _AnyModule(name)
    sys.modules['matplotlib.animation'].FuncAnimation = object
    sys.modules['skimage.filters'].gaussian = (
        lambda image, sigma=1, preserve_range=False, **kw:
        scipy.ndimage.gaussian_filter(np.asarray(image, float), sigma, mode='nearest', truncate=4.0))
    sys.modules['skimage'].filters = sys.modules['skimage.filters']
    sys.modules['petsc4py'].PETSc = _PETSc()
    if REFERENCE_SOURCE_DIR not in sys.path:
        sys.path.insert(0, REFERENCE_SOURCE_DIR)
    with contextlib.redirect_stdout(io.StringIO()):
        import optical_flow
    _reference_module = optical_flow
    return optical_flow


def reference_solve(movie, **kwargs):
    """Reference ``variational_optical_flow(..., use_direct_solver=True)`` with stdout captured.

    Returns (result_dict, list_of_true_relative_residuals_per_pair, captured_stdout).
    ``result['converged']`` comes from the inert PETSc stub and is overwritten here
    from the independently calculated residual the reference prints (:1150-1151).
    """
    ref = import_reference()
    kwargs = dict(kwargs)
    kwargs['use_direct_solver'] = True
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        result = ref.variational_optical_flow(movie, **kwargs)
    lines = buf.getvalue().splitlines()
    residuals = [float(lines[k + 1]) for k, line in enumerate(lines)
                 if line.startswith('Final residual norm independently calculated')]
    result['converged'] = bool(residuals and residuals[-1] < 1e-6)
    return result, residuals, buf.getvalue()


def reference_system(previous_frame, current_frame, speed_alpha, remodelling_alpha):
    """Capture the reference-assembled CSR matrix and RHS for one frame pair.

    ``scipy.sparse.linalg.spsolve`` is looked up by attribute at call time (:1147), so
    rebinding it lets us copy (A, b) out of the reference's own assembly code.
    """
    import scipy.sparse.linalg
    ref = import_reference()
    captured = {}
    original = scipy.sparse.linalg.spsolve

    def capture(A, b, *args, **kwargs):
        captured['A'] = A.copy().tocsr()
        captured['b'] = np.array(b, copy=True)
        return np.zeros_like(b)

    scipy.sparse.linalg.spsolve = capture
    try:
        with contextlib.redirect_stdout(io.StringIO()), np.errstate(all='ignore'):
            ref.variational_optical_flow(np.stack([previous_frame, current_frame]),
                                         speed_alpha=speed_alpha,
                                         remodelling_alpha=remodelling_alpha,
                                         use_direct_solver=True)
    finally:
        scipy.sparse.linalg.spsolve = original
    return captured['A'], captured['b']
