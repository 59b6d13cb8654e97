"""Build recipe of the CUDA library (sm_100a only; nvcc cross-compiles without a GPU).

``python -m epde_b200.build`` or ``__graft_entry__.build()``.  The shared object is written
in-tree (``epde_b200/libepde_b200.so``) so it travels with the repo snapshot; it is git-ignored.
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(PKG, 'csrc', 'epde_b200.cu')
HDR = os.path.join(os.path.dirname(PKG), 'include', 'epde_b200.h')
LIB = os.path.join(PKG, 'libepde_b200.so')

NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC', '-shared',
              '--split-compile=0']      # ptxas works on the ~200 template instantiations in parallel (build 2.7 -> 1.5 min)


def find_nvcc():
    for cand in (os.environ.get('NVCC'), shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError('nvcc not found: the epde_b200 CUDA library cannot be built')


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.exists(p) and os.path.getmtime(p) > t for p in (SRC, HDR))


def build(force=False, verbose=False):
    """Compile csrc/epde_b200.cu -> libepde_b200.so.  Returns the library path."""
    if not force and not is_stale():
        return LIB
    cmd = [find_nvcc()] + NVCC_FLAGS + ([] if not verbose else ['-Xptxas', '-v']) + [SRC, '-o', LIB + '.tmp']
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError('nvcc failed:\n' + ' '.join(cmd) + '\n' + res.stdout + res.stderr)
    os.replace(LIB + '.tmp', LIB)
    if verbose:
        sys.stderr.write(res.stderr)
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
