"""A compiled caller of the C ABI: examples/c_driver.c (plain C, no Python / torch in the
process) creates the system from the raw bytes of the descriptor structs and steps it through
`bx_step_host` -- the drop-in boundary exactly as a cgo / JNI / XLA-FFI binding would use it
(SURVEY.md 8(b)).  CPU: it compiles against include/brax_b200.h and links against the built
library.  GPU: its outputs pass the same oracle check as the Python mirror's and equal them bit
for bit."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from tests import util

SRC = os.path.join(util.ROOT, 'examples', 'c_driver.c')


def _build(tmp_path):
  from brax_b200 import build as bx_build
  lib = bx_build.build()
  exe = str(tmp_path / 'c_driver')
  libdir = os.path.dirname(lib)
  subprocess.check_call(['gcc', '-std=c99', '-O2', '-Wall', '-Werror', '-I', os.path.join(util.ROOT, 'include'), SRC,
                         '-o', exe, '-L', libdir, '-lbrax_b200', '-Wl,-rpath,' + libdir,
                         '-Wl,--allow-shlib-undefined'])
  return exe


def _system_blob(desc, env, path):
  """bx_system_desc + its arrays as raw bytes (the layout IS the ABI)."""
  d, keep = desc.to_c(env)
  bodies, joints, acts, forces, pairs, cenv, creset = keep[:7]
  with open(path, 'wb') as f:
    f.write(bytes(d))
    for arr, cnt in ((bodies, d.num_bodies), (joints, d.num_joints), (acts, d.num_actuators),
                     (forces, d.num_forces), (pairs, d.num_pairs)):
      f.write(bytes(arr)[:cnt * C.sizeof(arr._type_)])
    if cenv is not None:
      f.write(bytes(cenv))
    if creset is not None:
      f.write(bytes(creset))


def test_c_driver_compiles_and_links(tmp_path):
  exe = _build(tmp_path)
  r = subprocess.run([exe], capture_output=True, text=True)
  assert r.returncode == 2 and 'usage' in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize('name,envname', [('ant', 'ant'), ('humanoid_spring', None), ('fetch', None)])
def test_c_driver_steps_like_the_python_mirror(tmp_path, name, envname):
  from brax_b200 import system as bsys
  from tests.test_gpu_parity import _flat, check_step
  exe = _build(tmp_path)
  desc = util.load_desc(name)
  env = None
  if envname:
    from tests.test_env_oracle import VARIANTS
    env = VARIANTS[envname][1]
  g = util.load_golden(name)
  ins = [np.ascontiguousarray(g[k].astype(np.float32)) for k in util.QP_KEYS]
  b, n, t_n = ins[0].shape[0], desc.num_bodies, 3
  rng = np.random.default_rng(3)
  acts = rng.uniform(-1, 1, (t_n, b, desc.act_size)).astype(np.float32)
  acts[0] = g['act'].astype(np.float32)
  _system_blob(desc, env, tmp_path / 'system.blob')

  def run(steps):
    with open(tmp_path / 'io.blob', 'wb') as f:
      f.write(np.asarray([b, steps], np.int64).tobytes())
      for x in ins:
        f.write(x.tobytes())
      f.write(acts[:steps].tobytes())
    r = subprocess.run([exe, str(tmp_path / 'system.blob'), str(tmp_path / 'io.blob'), str(tmp_path / 'out.blob')],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert '%d launches' % steps in r.stdout, r.stdout
    raw = np.fromfile(tmp_path / 'out.blob', np.float32)
    s = bsys.System(desc, 0, env)
    k, o, m = s.num_contacts, (s.obs_size if env else 0), (s.num_metrics if env else 0)
    shapes = [('pos', (b, n, 3)), ('rot', (b, n, 4)), ('vel', (b, n, 3)), ('ang', (b, n, 3))]
    shapes += [(key, (b, n, 3)) for key in util.INFO_KEYS[:6]]
    shapes += [('contact_pos', (b, k, 3)), ('contact_normal', (b, k, 3)), ('contact_penetration', (b, k))]
    if env:
      shapes += [('obs', (b, o)), ('reward', (b,)), ('done', (b,)), ('metrics', (b, m))]
    out, off = {}, 0
    for key, shp in shapes:
      cnt = int(np.prod(shp))
      out[key] = raw[off:off + cnt].reshape(shp)
      off += cnt
    assert off == raw.size
    return s, out

  # one step: the oracle check of the parity tests
  s, one = run(1)
  check_step(name, ins, acts[0], one)
  # three chained steps: bit-identical to the Python mirror on the same library
  s, got = run(t_n)
  qp = bsys.QP(*ins)
  for t in range(t_n):
    res = s.step(qp, acts[t], env=env is not None)
    qp = res[0]
  want = _flat(res[0], res[1])
  for key in util.QP_KEYS + util.INFO_KEYS:
    np.testing.assert_array_equal(got[key], want[key], err_msg=key)
  if env:
    for key in ('obs', 'reward', 'done', 'metrics'):
      np.testing.assert_array_equal(got[key], res[2][key], err_msg=key)
