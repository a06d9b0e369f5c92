"""Builds libsevennet_b200.so in-tree with nvcc for sm_100a (no other architecture)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libsevennet_b200.so')
DRIVER = os.path.join(HERE, 'b200_pair_driver')   # stand-alone C++ caller of the C-ABI (csrc/pair_driver.cpp)
SOURCES = ['capi.cu', 'conv.cu', 'conv_pair.cu', 'edge.cu', 'radial.cu', 'radial_tc.cu', 'radial_table.cu', 'table_train.cu', 'linear.cu', 'linear_tc.cu', 'node.cu', 'graph.cu', 'engine.cu']
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
        build_torch_ops()
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
    build_torch_ops()
    return LIB


TORCH_OPS = os.path.join(HERE, 'libb200_torch_ops.so')   # torch.ops.b200.* for TorchScript deployment (csrc/torch_ops.cpp)


def build_torch_ops(force=False):
    """g++ against the torch headers; links libsevennet_b200.so (rpath $ORIGIN) and the torch libraries."""
    src = os.path.join(CSRC, 'torch_ops.cpp')
    if not force and os.path.exists(TORCH_OPS) and os.path.getmtime(TORCH_OPS) > max(os.path.getmtime(src), os.path.getmtime(LIB)):
        return TORCH_OPS
    import torch
    from torch.utils import cpp_extension as ce
    inc = [f'-I{p}' for p in ce.include_paths()] + ['-I/usr/local/cuda/include']
    libdir = os.path.join(os.path.dirname(torch.__file__), 'lib')
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cmd = ['g++', '-O2', '-std=c++17', '-fPIC', '-shared', f'-D_GLIBCXX_USE_CXX11_ABI={abi}', src, '-o', TORCH_OPS] + inc + [
        f'-L{HERE}', '-lsevennet_b200', f'-L{libdir}', '-lc10', '-lc10_cuda', '-ltorch_cpu', '-ltorch_cuda', '-ltorch',
        '-L/usr/local/cuda/lib64', '-lcudart', '-Wl,-rpath,$ORIGIN', f'-Wl,-rpath,{libdir}']
    subprocess.check_call(cmd)
    return TORCH_OPS


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
