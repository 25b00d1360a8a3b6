"""bx_step_host_static: the host-buffer step that allocates nothing.

The entry point has no counterpart in the reference
(include/brax_b200.h): the device staging area is a file-scope `static __device__`
array of the library, looked up with cudaGetSymbolAddress once per device.  Checked here,
through the C ABI, against the existing entry point `bx_step_host` on the same host buffers:
the two launch the same kernel build on the same 32-env tiles, so the results are compared
bit for bit (the tolerance of tests/test_gpu_parity.py::test_host_path_pinned_buffers_and_chunking),
on two calls in a row (the second call finds the first call's contents in the array).

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest

from brax_b200 import lib as bx_lib
from brax_b200.desc import CStepIO
from brax_b200.system import _OUT, System
from tests import util


def _torch():
  return pytest.importorskip('torch')


def _host_io(s, b, rng, with_info=True):
  """Pageable numpy inputs near the default pose + fresh output arrays, and the CStepIO over them."""
  qp0 = s.default_qp(batch=b)
  ins = {
      'pos': qp0.pos + rng.uniform(-.02, .02, qp0.pos.shape).astype(np.float32),
      'rot': np.ascontiguousarray(qp0.rot),
      'vel': rng.uniform(-.5, .5, qp0.vel.shape).astype(np.float32),
      'ang': rng.uniform(-.5, .5, qp0.ang.shape).astype(np.float32),
      'act': rng.uniform(-1, 1, (b, s.act_size)).astype(np.float32),
  }
  dims = {'n': s.num_bodies, 'k': s.num_contacts}

  def outputs():
    out = {}
    for name, w, dim in (_OUT if with_info else _OUT[:4]):
      shape = (b, dims[dim], w) if w else (b, dims[dim])
      out[name] = np.full(shape, np.nan, dtype=np.float32)  # every element must be written
    return out

  def io_of(out):
    io = CStepIO()
    for k, v in ins.items():
      setattr(io, k, v.ctypes.data)
    for k, v in out.items():
      setattr(io, k, v.ctypes.data if v.size else None)
    return io

  return ins, outputs, io_of


@pytest.mark.gpu
@pytest.mark.parametrize('name', ['ant', 'humanoid_spring'])
def test_static_scratch_step_equals_step_host_two_calls(name):
  torch = _torch()
  lib = bx_lib.load()
  s = System(util.load_desc(name))
  stream = torch.cuda.Stream()
  rng = np.random.default_rng(11)
  for call, b in enumerate((257, 96)):   # ragged last tile; the second call re-uses the array
    ins, outputs, io_of = _host_io(s, b, rng)
    want, got = outputs(), outputs()
    io_w, io_g = io_of(want), io_of(got)
    assert lib.bx_step_host(s._h, b, C.byref(io_w)) == 0, bx_lib.last_error()
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = lib.bx_step_host_static(s._h, b, C.byref(io_g), C.c_void_p(stream.cuda_stream))
    assert rc == 0, bx_lib.last_error()
    if call > 0:  # (the first call may load the library's CUDA module, which holds the array)
      assert torch.cuda.mem_get_info()[0] >= free_before, 'device memory was allocated'
    for k in want:
      assert np.isfinite(got[k]).all(), (call, k)
      np.testing.assert_array_equal(got[k], want[k], err_msg='call %d: %s' % (call, k))
    assert np.abs(got['pos_o'] - ins['pos']).max() > 0  # a step was taken


@pytest.mark.gpu
def test_static_scratch_default_stream_and_qp_only():
  """NULL stream = default stream; Info pointers may be NULL like in bx_step_host."""
  lib = bx_lib.load()
  s = System(util.load_desc('ant'))
  rng = np.random.default_rng(5)
  b = 64
  _, outputs, io_of = _host_io(s, b, rng, with_info=False)
  want, got = outputs(), outputs()
  io_w, io_g = io_of(want), io_of(got)
  assert lib.bx_step_host(s._h, b, C.byref(io_w)) == 0, bx_lib.last_error()
  assert lib.bx_step_host_static(s._h, b, C.byref(io_g), None) == 0, bx_lib.last_error()
  for k in want:
    np.testing.assert_array_equal(got[k], want[k], err_msg=k)


@pytest.mark.gpu
def test_static_scratch_rejects_batches_beyond_the_array():
  lib = bx_lib.load()
  s = System(util.load_desc('ant'))
  cap = int(lib.bx_static_scratch_bytes())
  assert cap == 32 << 20
  per_env = 4 * (13 * s.num_bodies + s.act_size + 13 * s.num_bodies)   # inputs + the QP outputs
  b = cap // per_env + 64
  rng = np.random.default_rng(1)
  _, outputs, io_of = _host_io(s, b, rng, with_info=False)
  out = outputs()
  io = io_of(out)
  assert lib.bx_step_host_static(s._h, b, C.byref(io), None) != 0
  msg = bx_lib.last_error()
  assert 'static scratch' in msg and str(cap) in msg, msg
  assert np.isnan(out['pos_o']).all()   # nothing ran
  # the same batch goes through the allocating entry point
  assert lib.bx_step_host(s._h, b, C.byref(io)) == 0, bx_lib.last_error()
  assert np.isfinite(out['pos_o']).all()
