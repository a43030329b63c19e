"""In-tree build of the native core: ``nvcc`` -> ``_native/libpdmnist_b200.so``.

One target, one architecture: ``-gencode arch=compute_100a,code=sm_100a``
(the ``a`` suffix is required for tcgen05/TMEM).  nvcc cross-compiles without
a GPU, so this runs on the CPU build box; the resulting ``.so`` is git-ignored
but travels with the repo snapshot to the GPU box.

Usage: ``python -m pytorch_distributed_mnist_b200.ops._build [--force] [--verbose]``
"""
from __future__ import annotations

import concurrent.futures
import glob
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REPO_DIR = os.path.dirname(PKG_DIR)
CSRC_DIR = os.path.join(PKG_DIR, 'csrc')
INCLUDE_DIR = os.path.join(REPO_DIR, 'include')
NATIVE_DIR = os.path.join(PKG_DIR, '_native')
OBJ_DIR = os.path.join(REPO_DIR, 'build', 'obj')
LIB_NAME = 'libpdmnist_b200.so'
LIB_PATH = os.path.join(NATIVE_DIR, LIB_NAME)

ARCH_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a']
NVCC_FLAGS = ['-O3', '-std=c++17', '-lineinfo', '--expt-relaxed-constexpr',
              '-Xcompiler', '-fPIC,-O3,-Wall,-Wno-unused-function', '-Xptxas', '-v']


def find_nvcc():
    for cand in (os.environ.get('NVCC'), shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.isfile(cand):
            return cand
    raise RuntimeError('nvcc not found (set NVCC or put it on PATH)')


def sources():
    return sorted(glob.glob(os.path.join(CSRC_DIR, '*.cu')) + glob.glob(os.path.join(CSRC_DIR, '*.cpp')))


def _deps_mtime():
    hdrs = glob.glob(os.path.join(CSRC_DIR, '*.cuh')) + glob.glob(os.path.join(CSRC_DIR, '*.h')) + \
        glob.glob(os.path.join(INCLUDE_DIR, '*.h'))
    return max([os.path.getmtime(h) for h in hdrs] + [0.0])


def _compile_one(nvcc, src, obj, verbose):
    cmd = [nvcc] + ARCH_FLAGS + NVCC_FLAGS + ['-I', INCLUDE_DIR, '-I', CSRC_DIR, '-x', 'cu', '-c', src, '-o', obj]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log = proc.stdout + proc.stderr
    with open(obj + '.log', 'w') as f:
        f.write(' '.join(cmd) + '\n' + log)
    if proc.returncode != 0:
        raise RuntimeError('nvcc failed for {}:\n{}'.format(src, log))
    if verbose:
        print(log)
    return obj


def build(force=False, verbose=False):
    """Compile every source under csrc/ for sm_100a and link the shared library.
    Returns the library path.  No-op when the library is newer than all inputs."""
    srcs = sources()
    if not srcs:
        raise RuntimeError('no sources under ' + CSRC_DIR)
    newest = max([os.path.getmtime(s) for s in srcs] + [_deps_mtime(), os.path.getmtime(__file__)])

    def up_to_date():
        return os.path.isfile(LIB_PATH) and os.path.getmtime(LIB_PATH) >= newest

    if not force and up_to_date():
        return LIB_PATH
    os.makedirs(OBJ_DIR, exist_ok=True)
    os.makedirs(NATIVE_DIR, exist_ok=True)
    # Every rank of a spawn job may get here at once (first run on a fresh box): one of them builds, the
    # others wait on the lock and then find the library up to date.  Objects are shared paths under
    # build/obj, so two concurrent builds would link each other's half-written files.
    import fcntl
    with open(os.path.join(OBJ_DIR, '.build.lock'), 'w') as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and up_to_date():
                return LIB_PATH
            return _build_locked(srcs, force, verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(srcs, force, verbose):
    nvcc = find_nvcc()
    hdr_mtime = max(_deps_mtime(), os.path.getmtime(__file__))
    jobs = []
    objs = []
    for s in srcs:
        obj = os.path.join(OBJ_DIR, os.path.basename(s) + '.o')
        objs.append(obj)
        if force or not os.path.isfile(obj) or os.path.getmtime(obj) < max(os.path.getmtime(s), hdr_mtime):
            jobs.append((s, obj))
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as pool:
        futs = [pool.submit(_compile_one, nvcc, s, o, verbose) for s, o in jobs]
        for f in futs:
            f.result()
    tmp = LIB_PATH + '.tmp.{}'.format(os.getpid())
    link = [nvcc] + ARCH_FLAGS + ['-shared', '-o', tmp] + objs + ['-lpthread', '-ldl', '-lrt']
    proc = subprocess.run(link, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError('link failed:\n' + proc.stdout + proc.stderr)
    os.replace(tmp, LIB_PATH)
    return LIB_PATH


if __name__ == '__main__':
    path = build(force='--force' in sys.argv, verbose='--verbose' in sys.argv)
    print(path)
