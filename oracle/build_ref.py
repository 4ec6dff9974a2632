"""TEST INFRASTRUCTURE - recipe that makes the reference's own code available on the GPU box.

The hot path of kursawe/OpticalFlow is one pure-Python module (source/optical_flow.py; numba-jitted helpers, SciPy
assembly and SuperLU solve).  There is nothing to compile: this recipe copies that module, byte for byte, from where it
lies under /root/reference into ``oracle/_ref/`` (a build output: git-ignored, not gpurun-ignored, so it travels with the
snapshot like the built .so files) and records its SHA-256.  ``oracle/ref_import.py`` imports it behind stubs for the
three packages the reference imports and this image lacks (matplotlib, skimage, petsc4py).

  python oracle/build_ref.py        (also run by __graft_entry__.build() when /root/reference is present)
"""
import hashlib
import os
import shutil
import sys

SOURCE = '/root/reference/source/optical_flow.py'
TARGET_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), '_ref')


def build(verbose=False):
    if not os.path.isfile(SOURCE):
        return False
    os.makedirs(TARGET_DIR, exist_ok=True)
    target = os.path.join(TARGET_DIR, 'optical_flow.py')
    shutil.copyfile(SOURCE, target)
    digest = hashlib.sha256(open(target, 'rb').read()).hexdigest()
    with open(os.path.join(TARGET_DIR, 'SHA256'), 'w') as f:
        f.write('%s  source/optical_flow.py (unmodified copy of %s)\n' % (digest, SOURCE))
    if verbose:
        print('oracle/_ref/optical_flow.py', digest)
    return True


if __name__ == '__main__':
    sys.exit(0 if build(verbose=True) else 1)
