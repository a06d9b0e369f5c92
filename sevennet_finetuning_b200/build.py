"""Builds libsevennet_b200.so in-tree with nvcc for sm_100a (no other architecture)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libsevennet_b200.so')
DRIVER = os.path.join(HERE, 'b200_pair_driver')   # stand-alone C++ caller of the C-ABI (csrc/pair_driver.cpp)
SOURCES = ['capi.cu', 'conv.cu', 'conv_pair.cu', 'edge.cu', 'radial.cu', 'radial_tc.cu', 'radial_table.cu', 'linear.cu', 'linear_tc.cu', 'node.cu', 'graph.cu', 'engine.cu']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
              '--use_fast_math=false', '-Xcompiler', '-fPIC', '-Xcompiler', '-O2']


def _stale():
    if not os.path.exists(LIB) or not os.path.exists(DRIVER):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [
        os.path.join(os.path.dirname(HERE), 'include', 'sevennet_b200.h')]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, 'build'), exist_ok=True)
    for s in SOURCES:
        o = os.path.join(HERE, 'build', s.replace('.cu', '.o'))
        objs.append(o)
        cmd = [nvcc] + [f for f in NVCC_FLAGS if f != '--use_fast_math=false'] + (
            ['-Xptxas', '-v'] if verbose else []) + ['-c', os.path.join(CSRC, s), '-o', o]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    fail = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f'--- {s}\n{out}\n')
        fail |= p.returncode != 0
    if fail:
        raise RuntimeError('nvcc failed')
    subprocess.check_call([nvcc, '-shared', '-o', LIB] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a',
                                                                   '-lcudart'])
    # the native caller: plain C++ against include/sevennet_b200.h, no torch, no Python
    subprocess.check_call([nvcc, '-O2', '-std=c++17', os.path.join(CSRC, 'pair_driver.cpp'), '-o', DRIVER, '-L' + HERE,
                           '-lsevennet_b200', '-lcudart', '-Xlinker', '-rpath', '-Xlinker', '$ORIGIN'])
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
