"""Builds libtnt_b200.so (the C-ABI library) in-tree with nvcc for sm_100a.  No torch involved."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OUT = os.path.join(HERE, 'libtnt_b200.so')
SOURCES = ['core.cu', 'gemm.cu', 'gemm_splitk.cu', 'gemm_ln.cu', 'dynconv.cu', 'elementwise.cu', 'context.cu', 'pointer.cu', 'backward.cu', 'attention.cu', 'attention_tc.cu', 'attention_step.cu', 'attention_short.cu', 'attention_bwd.cu', 'attention_bwd_tc.cu', 'adasoftmax.cu',
           'engine_load.cu', 'engine_fwd.cu', 'engine_gen.cu', 'engine_beam.cu', 'engine_train.cu', 'optimizer.cu']
FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
         '-Xcompiler', '-fPIC', '-DNDEBUG']


def _stale(src_paths, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    deps = list(src_paths) + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.cuh', '.h'))]
    deps.append(os.path.join(os.path.dirname(HERE), 'include', 'tnt_b200.h'))
    return any(os.path.getmtime(p) > t for p in deps)


def build(force=False, verbose=False):
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    objdir = os.path.join(HERE, 'build')
    os.makedirs(objdir, exist_ok=True)
    objs = [os.path.join(objdir, os.path.basename(s)[:-3] + '.o') for s in srcs]
    todo = [(s, o) for s, o in zip(srcs, objs) if force or _stale([s], o)]

    def cc(so):
        s, o = so
        cmd = ['nvcc'] + FLAGS + ['-c', s, '-o', o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError('nvcc failed for %s:\n%s' % (s, r.stderr))
        if verbose and r.stderr.strip():
            print(r.stderr)
    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(cc, todo))
    if todo or not os.path.exists(OUT):
        cmd = ['nvcc', '-shared', '-o', OUT] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a', '-lcudart']
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError('link failed:\n%s' % r.stderr)
    return OUT


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose=True))
