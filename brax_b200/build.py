"""Builds brax_b200/libbrax_b200.so (the C ABI + sm_100a kernels) in-tree.

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the
GPU box with the snapshot.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libbrax_b200.so')
SOURCES = ['bx_api.cu']  # single translation unit (includes bx_step_kernel.cu)
DEPS = ['bx_api.cu', 'bx_step_kernel.cu', 'bx_math.cuh', 'bx_tables.h',
        os.path.join('..', '..', 'include', 'brax_b200.h')]

NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3',
              '-std=c++17', '-shared', '-Xcompiler', '-fPIC']


def stale() -> bool:
  if not os.path.exists(LIB):
    return True
  t = os.path.getmtime(LIB)
  return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build(force: bool = False, verbose: bool = False, extra=()) -> str:
  if force or stale():
    nvcc = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
    cmd = [nvcc] + NVCC_FLAGS + list(extra) + (
        ['-Xptxas', '-v'] if verbose else []) + [
            os.path.join(CSRC, s) for s in SOURCES] + ['-o', LIB]
    subprocess.check_call(cmd)
  return LIB


if __name__ == '__main__':
  print(build(force=True, verbose=True))
