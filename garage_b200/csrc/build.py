"""Build libb2g.so (sm_100a only) with nvcc, in-tree, no torch headers.

    python -m garage_b200.csrc.build [--force] [--verbose]

Object files land in `garage_b200/csrc/_obj/` (git-ignored), the library in
`garage_b200/libb2g.so`.  Only stale objects are recompiled.
"""
import glob
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
LIB = os.path.join(PKG, 'libb2g.so')
OBJ = os.path.join(HERE, '_obj')
NVCC = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
FLAGS = [
    '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
    '-Xcompiler', '-fPIC',
    '--expt-relaxed-constexpr', '-Xptxas', '-v', *os.environ.get('B2G_NVCC_EXTRA', '').split(),
]
# precise math is nvcc's default; fast-math is never enabled (fp32 parity)


def _deps():
    return glob.glob(os.path.join(HERE, '*.cuh')) + [
        os.path.join(os.path.dirname(PKG), 'include', 'b2g.h'), __file__]


def _stale(src, obj):
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    return any(os.path.getmtime(p) > t for p in [src] + _deps())


def _compile(src, obj, verbose):
    cmd = [NVCC] + FLAGS + ['-c', src, '-o', obj]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = res.stdout + res.stderr
    with open(obj + '.log', 'w') as f:
        f.write(log)
    if res.returncode != 0:
        raise RuntimeError('nvcc failed for {}:\n{}'.format(src, log))
    if verbose:
        print(log)
    return obj


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(HERE, '*.cu')))
    objs = [os.path.join(OBJ, os.path.basename(s)[:-3] + '.o') for s in srcs]
    todo = [(s, o) for s, o in zip(srcs, objs) if force or _stale(s, o)]
    if todo:
        with ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            list(ex.map(lambda so: _compile(so[0], so[1], verbose), todo))
    if todo or not os.path.exists(LIB):
        cmd = [NVCC, '-shared', '-o', LIB] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a',
                                                     '-lcudart_static', '-lpthread', '-ldl', '-lrt']
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError('link failed:\n' + res.stdout + res.stderr)
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='--verbose' in sys.argv))
