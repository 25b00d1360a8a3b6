"""The C-ABI library loads and exports every symbol include/brax_b200.h declares.

CPU-only checks (no compute calls): symbol table vs the header, struct sizes of
the ctypes mirror vs the compiled library, and the construction-time errors of
`bx_system_create` that are raised before any device is touched
(SURVEY.md 8(b) "Errors": fail loudly, never fall back).
"""
import ctypes as C
import os
import re
import subprocess

import pytest

from brax_b200 import build as bx_build
from brax_b200 import desc as bx_desc
from brax_b200 import envs as bx_envs
from brax_b200 import lib as bx_lib
from tests import util

HEADER = os.path.join(util.ROOT, 'include', 'brax_b200.h')


def _declared():
  src = open(HEADER).read()
  src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
  return sorted(set(re.findall(r'\b(bx_[a-z_0-9]+)\s*\(', src)))


@pytest.fixture(scope='module')
def lib():
  bx_build.build()
  return bx_lib.load()


def test_header_symbols_exported(lib):
  names = _declared()
  assert 'bx_step' in names and 'bx_system_create' in names
  for n in names:
    assert hasattr(lib, n), 'libbrax_b200.so does not export %s' % n
  assert set(names) == set(bx_lib.SYMBOLS)


def test_sm100a_code_is_embedded(lib):
  out = subprocess.run(['cuobjdump', '-lelf', bx_lib.LIB_PATH],
                       capture_output=True, text=True).stdout
  assert 'sm_100a' in out


def test_struct_sizes_match_header():
  """ctypes mirror vs the C compiler's view of include/brax_b200.h."""
  prog = r'''
#include <stdio.h>
#include "brax_b200.h"
int main(void) {
  printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %d\n", sizeof(bx_body), sizeof(bx_joint),
         sizeof(bx_actuator), sizeof(bx_force), sizeof(bx_pair),
         sizeof(bx_system_desc), sizeof(bx_step_io), sizeof(bx_env_desc), sizeof(bx_reset_joint),
         sizeof(bx_reset_probe), sizeof(bx_reset_desc), BX_ABI_VERSION);
  return 0;
}'''
  import tempfile
  with tempfile.TemporaryDirectory() as td:
    src = os.path.join(td, 's.c')
    open(src, 'w').write(prog)
    exe = os.path.join(td, 's')
    subprocess.check_call(['gcc', '-I', os.path.dirname(HEADER), src, '-o', exe])
    got = [int(x) for x in subprocess.check_output([exe]).split()]
  want = [C.sizeof(c) for c in (bx_desc.CBody, bx_desc.CJoint, bx_desc.CActuator,
                                bx_desc.CForce, bx_desc.CPair,
                                bx_desc.CSystemDesc, bx_desc.CStepIO, bx_envs.CEnvDesc,
                                bx_desc.CResetJoint, bx_desc.CResetProbe, bx_desc.CResetDesc)]
  assert got[:-1] == want
  assert got[-1] == bx_desc.ABI_VERSION


def _create(lib, cdesc):
  h = C.c_void_p()
  rc = lib.bx_system_create(C.byref(cdesc), 0, C.byref(h))
  msg = bx_lib.last_error()
  if rc == 0:
    lib.bx_system_destroy(h)
  return rc, msg


def test_create_rejects_abi_skew(lib):
  cdesc, keep = util.load_desc('ant').to_c()
  cdesc.abi_version = 1
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'abi_version' in msg
  cdesc, keep = util.load_desc('ant').to_c()
  cdesc.struct_size = 8
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'size mismatch' in msg


def test_create_rejects_uncovered_features(lib):
  cdesc, keep = util.load_desc('ant').to_c()
  cdesc.pairs[0].kind = 7  # e.g. mesh / heightmap / hull-hull
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'out of scope' in msg
  cdesc, keep = util.load_desc('ant').to_c()
  cdesc.joints[0].kind = bx_desc.JOINT_KIND['spring_revolute']
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'dynamics_mode' in msg
  cdesc, keep = util.load_desc('ant').to_c()
  cdesc.joints[0].body_c = 99
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'out of range' in msg


def test_create_rejects_bad_env_desc(lib):
  """The optional env epilogue (bx_env_desc) is validated before any device is touched."""
  env = bx_envs.EnvDesc('ant', use_contact_forces=True)
  cdesc, keep = util.load_desc('ant').to_c(env)
  keep[5].struct_size = 12
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'bx_env_desc size mismatch' in msg
  cdesc, keep = util.load_desc('ant').to_c(env)
  keep[5].kind = 9
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'unsupported env kind' in msg
  cdesc, keep = util.load_desc('ant').to_c(env)
  keep[5].joint_free_dofs[0] = 5
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'joint_free_dofs' in msg


def test_create_rejects_bad_reset_desc(lib):
  """The optional reset tables (bx_reset_desc) are validated before any device is touched."""
  cdesc, keep = util.load_desc('ant').to_c()
  keep[6].struct_size = 12
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'bx_reset_desc size mismatch' in msg
  cdesc, keep = util.load_desc('ant').to_c()
  keep[6].joints[3].body_c = 99
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'joint body index' in msg
  cdesc, keep = util.load_desc('ant').to_c()
  keep[6].joints[0].take[1] = 40
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'take index' in msg
  cdesc, keep = util.load_desc('ant').to_c()
  keep[6].probes[2].tree = 7
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'probe index' in msg


def test_tile_bytes_with_env_epilogue(lib):
  """The env rows grow the tile; ant and humanoid still fit two tiles per SM."""
  from brax_b200 import system as bx_system
  for name, env in (('ant', bx_envs.EnvDesc('ant', use_contact_forces=True)),
                    ('humanoid', bx_envs.EnvDesc('humanoid'))):
    d = util.load_desc(name)
    plain, _ = bx_system.tile_bytes(d)
    with_env, _ = bx_system.tile_bytes(d, env)
    assert plain <= with_env
    assert 2 * (with_env + 1024) <= 233472, (name, with_env)


def test_no_cpu_fallback(lib):
  """Without a CUDA device the library refuses to create a system."""
  import torch
  if torch.cuda.is_available():
    pytest.skip('a GPU is present')
  cdesc, keep = util.load_desc('ant').to_c()
  rc, msg = _create(lib, cdesc)
  assert rc != 0 and 'no CPU fallback' in msg


def test_product_does_not_import_oracle():
  """brax_b200/ never imports or links anything under oracle/."""
  pkg = os.path.join(util.ROOT, 'brax_b200')
  for dp, _, fs in os.walk(pkg):
    for f in fs:
      if f.endswith(('.py', '.cu', '.cuh', '.h')):
        text = open(os.path.join(dp, f)).read()
        assert not re.search(r'^\s*(from|import)\s+oracle', text, re.M), f
        assert 'liboracle' not in text, f


def test_xla_ffi_handler_is_a_real_source_and_its_build_is_gated():
  """The JAX side of the boundary exists as source (csrc/bx_ffi.cc + jax_ffi.py) and its build step is
  skipped with a printed reason where jaxlib's headers are missing (this box: no jax, no network)."""
  src = open(os.path.join(os.path.dirname(bx_build.CSRC), 'csrc', 'bx_ffi.cc')).read()
  assert 'XLA_FFI_DEFINE_HANDLER_SYMBOL(bx_step_ffi' in src and src.count('.Ret<F32>()') == 13 and src.count('.Arg<F32>()') == 5
  import brax_b200.jax_ffi as jf   # imports without jax
  try:
    import jax  # noqa: F401
    have = True
  except Exception:
    have = False
  out = bx_build.build_ffi(verbose=False)
  assert (out is not None) == have
  if not have:
    with pytest.raises(Exception):
      jf.register()
