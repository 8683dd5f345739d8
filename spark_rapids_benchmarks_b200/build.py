"""Builds libndsb200.so in-tree with nvcc for sm_100a (no JIT cache, no torch dependency)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libndsb200.so')
SOURCES = ['capi.cu', 'exec.cu', 'tables.cu', 'ops.cu', 'decode.cu', 'decode_fast.cu', 'fused_scan.cu', 'noalloc.cu', 'parquet_meta.cc', 'nccl_dyn.cc']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC', '-Xcompiler', '-fvisibility=hidden']


def _newest_source_mtime():
    m = 0.0
    for root in (CSRC, os.path.join(HERE, '..', 'include')):
        for f in os.listdir(root):
            m = max(m, os.path.getmtime(os.path.join(root, f)))
    return m


def build(force=False, verbose=False):
    """Compile every CUDA source to objects (in parallel) and link the shared library."""
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= _newest_source_mtime():
        return LIB
    nvcc = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
    objdir = os.path.join(HERE, 'build')
    os.makedirs(objdir, exist_ok=True)
    procs = []
    objs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src + '.o')
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + os.environ.get('NDS_NVCC_EXTRA', '').split() + ['-c', os.path.join(CSRC, src), '-o', obj]   # NDS_NVCC_EXTRA: tuning experiments
        if verbose:
            print(' '.join(cmd))
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out.decode())
            raise RuntimeError('nvcc failed on %s' % src)
        if verbose and out:
            print(out.decode())
    cmd = [nvcc, '-shared', '-o', LIB] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a', '-ldl']
    subprocess.check_call(cmd)
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose=True))
