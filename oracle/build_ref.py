#!/usr/bin/env python
"""oracle/build_ref.py — TEST / BENCH INFRASTRUCTURE: the recipe that builds oracle/_ref/.

The reference's hot path is two pure-Python files.  Where /root/reference is mounted (the build
container; never the GPU box) this compiles them, UNMODIFIED and from where they lie, to CPython
bytecode: oracle/_ref/HttpParser.bc and StreamBuf.bc (.pyc content; the extension keeps file-syncing tools that skip *.pyc from dropping them).  Bytecode is the reference's "binary" for
this interpreter: oracle/_ref/ is git-ignored (no reference source enters the repository) but is
not gpurun-ignored, so it travels to the GPU box like the built .so files do, and
`bench.py --impl reference` can then time the reference itself (kind "reference") instead of the
CPython port (kind "port").  Also builds the host corpus generator (oracle/_build/libcorpus.so).

    python oracle/build_ref.py        # prints what it built / why it could not
"""
import os
import py_compile
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.environ.get('SHL_REFERENCE', '/root/reference/src/python/shellac/server')
OUT = os.path.join(HERE, '_ref')
BUILD = os.path.join(HERE, '_build')
FILES = ('HttpParser', 'StreamBuf')


def build_corpus_lib(force=False):
    src = os.path.join(HERE, 'corpus_host.cpp')
    dep = os.path.join(HERE, '..', 'shellac_b200', 'csrc', 'corpus.cuh')
    lib = os.path.join(BUILD, 'libcorpus.so')
    if force or not os.path.exists(lib) or os.path.getmtime(lib) < max(os.path.getmtime(src), os.path.getmtime(dep)):
        os.makedirs(BUILD, exist_ok=True)
        subprocess.check_call(['g++', '-O2', '-std=c++17', '-shared', '-fPIC', '-o', lib, src])
    return lib


def build_ref():
    """Returns the list of .pyc written, [] when the reference is not mounted here."""
    if not os.path.isdir(REF_DIR):
        return []
    os.makedirs(OUT, exist_ok=True)
    done = []
    for name in FILES:
        src = os.path.join(REF_DIR, name + '.py')
        dst = os.path.join(OUT, name + '.bc')
        if not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            py_compile.compile(src, cfile=dst, doraise=True)
        done.append(dst)
    return done


if __name__ == '__main__':
    print(build_corpus_lib())
    got = build_ref()
    print('\n'.join(got) if got else 'reference not mounted at %s: oracle/_ref not (re)built' % REF_DIR)
    sys.exit(0)
