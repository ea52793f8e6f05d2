#!/usr/bin/env python
"""Build named variants of the library with extra -D flags into build/variants/ (git-ignored, but
shipped to the GPU box) for A/B timing:  build_variants.py name:-DX=1,-DY=0 name2:...
Then on the box: SHL_LIB=build/variants/lib_<name>.so python tools/quick_kernel_time.py"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from shellac_b200 import build as B  # noqa: E402

out_dir = os.path.join(ROOT, 'build', 'variants')
os.makedirs(out_dir, exist_ok=True)


def one(spec):
    name, _, flags = spec.partition(':')
    out = os.path.join(out_dir, 'lib_%s.so' % name)
    cmd = ['nvcc'] + B.NVCC_FLAGS + [f for f in flags.split(',') if f] + ['-o', out, os.path.join(B.CSRC, 'api.cu')]
    subprocess.check_call(cmd)
    return out


with ThreadPoolExecutor(4) as ex:
    for o in ex.map(one, sys.argv[1:]):
        print(o)
