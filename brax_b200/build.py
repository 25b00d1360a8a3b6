"""Builds brax_b200/libbrax_b200.so (the C ABI + sm_100a kernels) in-tree.

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the
GPU box with the snapshot.
"""
import os
import subprocess
import time

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libbrax_b200.so')
DEPS = ['bx_api.cu', 'bx_reset.cu', 'bx_kernels.cu', 'bx_step_kernel.cuh', 'bx_params.h', 'bx_math.cuh', 'bx_tables.h',
        os.path.join('..', '..', 'include', 'brax_b200.h')]
# (launch bound threads, min blocks per SM): one kernel translation unit each
BOUNDS = [(256, 2), (320, 2), (384, 2), (512, 1)]
# units that also instantiate the capsule-box contact class (the others would only add build time)
MESH_BOUNDS = [(320, 2), (512, 1)]
# extra units: (name, threads, min blocks per SM)
EXTRA_UNITS = [('320x3', 320, 3)]

NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3',
              '-std=c++17', '-Xcompiler', '-fPIC']


def stale(lib=None) -> bool:
  lib = lib or LIB
  if not os.path.exists(lib):
    return True
  t = os.path.getmtime(lib)
  return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build(force: bool = False, verbose: bool = False, extra=(), out: str = None) -> str:
  """Compiles bx_api.cu + one bx_kernels.cu object per launch bound (in
  parallel) and links them into the shared library."""
  out = out or LIB
  if not (force or stale(out)):
    return out
  nvcc = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
  import tempfile
  objdir = os.path.join(tempfile.gettempdir(), 'brax_b200_obj_' + os.path.basename(out))
  os.makedirs(objdir, exist_ok=True)
  common = [nvcc] + NVCC_FLAGS + list(extra) + (['-Xptxas', '-v'] if verbose else [])
  jobs = [(common + ['-c', os.path.join(CSRC, 'bx_api.cu'), '-o', os.path.join(objdir, 'bx_api.o')])]
  objs = [os.path.join(objdir, 'bx_api.o'), os.path.join(objdir, 'bx_reset.o')]
  jobs.append(common + ['-c', os.path.join(CSRC, 'bx_reset.cu'), '-o', os.path.join(objdir, 'bx_reset.o')])
  # one object per (launch bound, dynamics mode, env epilogue): independent nvcc jobs, run in parallel below
  for mode in (0, 1):
    for env in (0, 1):
      tu = ['-DBX_TU_MODE=%d' % mode, '-DBX_TU_ENV=%d' % env]
      for maxt, minb in BOUNDS:
        o = os.path.join(objdir, 'bx_kernels_%d_m%de%d.o' % (maxt, mode, env))
        objs.append(o)
        mesh = ['-DBX_TU_MESH=1'] if (maxt, minb) in MESH_BOUNDS else []
        jobs.append(common + mesh + tu + ['-DBX_TU_MAXT=%d' % maxt, '-DBX_TU_MINB=%d' % minb, '-c',
                                          os.path.join(CSRC, 'bx_kernels.cu'), '-o', o])
      for name, maxt, minb in EXTRA_UNITS:
        o = os.path.join(objdir, 'bx_kernels_%s_m%de%d.o' % (name, mode, env))
        objs.append(o)
        jobs.append(common + tu + ['-DBX_TU_MAXT=%d' % maxt, '-DBX_TU_MINB=%d' % minb, '-DBX_TU_NAME=%s' % name, '-c',
                                   os.path.join(CSRC, 'bx_kernels.cu'), '-o', o])
  # incremental: an object is recompiled only if one of ITS sources is newer (same flags: the object
  # directory is keyed by the output name; `extra` flags force everything)
  kernel_deps = ['bx_kernels.cu', 'bx_step_kernel.cuh', 'bx_params.h', 'bx_math.cuh', 'bx_tables.h',
                 os.path.join('..', '..', 'include', 'brax_b200.h')]
  deps_of = {'bx_api.o': ['bx_api.cu', 'bx_params.h', 'bx_tables.h', os.path.join('..', '..', 'include', 'brax_b200.h')],
             'bx_reset.o': ['bx_reset.cu', 'bx_math.cuh', os.path.join('..', '..', 'include', 'brax_b200.h')]}
  stamp = os.path.join(objdir, 'flags.txt')
  flags_now = ' '.join(common)
  same_flags = os.path.exists(stamp) and open(stamp).read() == flags_now

  def fresh(job):
    o = job[job.index('-o') + 1]
    if not same_flags or not os.path.exists(o):
      return False
    deps = deps_of.get(os.path.basename(o), kernel_deps)
    return all(os.path.getmtime(os.path.join(CSRC, d)) <= os.path.getmtime(o) for d in deps)

  jobs = [j for j in jobs if not fresh(j)]
  # at most one nvcc per core at a time (the mesh-class objects are the long ones: they go first)
  jobs.sort(key=lambda j: '-DBX_TU_MESH=1' not in j)
  limit = max(1, os.cpu_count() or 1)
  running, rcs = [], []
  for j in jobs:
    while len(running) >= limit:
      done = [p for p in running if p.poll() is not None]
      if not done:
        time.sleep(0.2)
        continue
      for p in done:
        running.remove(p)
        rcs.append(p.returncode)
    running.append(subprocess.Popen(j))
  rcs += [p.wait() for p in running]
  if any(rcs):
    raise subprocess.CalledProcessError(max(rcs), 'nvcc')
  with open(stamp, 'w') as f:
    f.write(flags_now)
  subprocess.check_call([nvcc, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', out] + objs)
  return out


def build_ffi(verbose: bool = True):
  """Compiles csrc/bx_ffi.cc (the XLA FFI handler) next to the library -- only where jaxlib's headers
  resolve.  Returns the .so path, or None after printing why it was skipped."""
  try:
    import jax.ffi  # noqa: F401
    inc = jax.ffi.include_dir()
  except Exception as e:  # jax is not installed on this box (and cannot be: no network)
    if verbose:
      print('brax_b200: bx_ffi.cc not built -- jaxlib headers unavailable on this box (%s: %s)' % (type(e).__name__, e))
    return None
  out = os.path.join(HERE, 'libbrax_b200_ffi.so')
  subprocess.check_call(['g++', '-std=c++17', '-O2', '-fPIC', '-shared', '-I' + inc, '-I/usr/local/cuda/include',
                         os.path.join(CSRC, 'bx_ffi.cc'), '-o', out, '-L' + HERE, '-lbrax_b200', '-Wl,-rpath,$ORIGIN'])
  return out


if __name__ == '__main__':
  print(build(force=True, verbose=True))
  build_ffi()
