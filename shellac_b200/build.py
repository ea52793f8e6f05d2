"""Build libshellac_b200.so in-tree with nvcc for sm_100a (no torch, no cmake).

    python -m shellac_b200.build            # build if stale
    python -m shellac_b200.build --force

The .so lands next to this file so it travels with the source tree (it is git-ignored).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OUT = os.path.join(HERE, 'libshellac_b200.so')
SOURCES = ['api.cu']
DEPS = ['api.cu', 'parse_kernel.cuh', 'bitmap_walk.cuh', 'frame_kernel.cuh', 'shl_device.cuh', 'corpus.cuh',
        os.path.join('..', '..', 'include', 'shellac_b200.h')]
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
              '-Xcompiler', '-fPIC', '-shared', '-cudart', 'static', '-ldl']


def stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build(force=False, verbose=False):
    if not force and not stale():
        return OUT
    nvcc = os.environ.get('NVCC', 'nvcc')
    cmd = [nvcc] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + \
          ['-o', OUT] + [os.path.join(CSRC, s) for s in SOURCES]
    subprocess.check_call(cmd)
    return OUT


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
