"""TEST INFRASTRUCTURE ONLY: build reference `brax.System`s on numpy.

Puts oracle/jaxstub and /root/reference on sys.path so the UNMODIFIED
reference physics (brax/physics/*.py) runs on its own numpy float64 backend
(SURVEY.md F5).  Only usable in the build container (the GPU box has no
/root/reference); used by tools/gen_systems.py and oracle/gen_golden.py to
produce the committed descriptors and golden vectors.

Env configs are pulled out of brax/envs/<env>.py by regex instead of importing
`brax.envs` (which needs gym/dm_env/host_callback).
"""
import os
import re
import sys
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = os.environ.get('BRAX_REFERENCE', '/root/reference')


def setup():
  if not os.path.isdir(REFERENCE):
    raise RuntimeError('reference tree %s is not present' % REFERENCE)
  for p in (REFERENCE, os.path.join(HERE, 'jaxstub')):
    if p not in sys.path:
      sys.path.insert(0, p)
  warnings.simplefilter('ignore')
  import brax  # noqa: F401
  return brax


def env_config_text(env: str, var: str = '_SYSTEM_CONFIG') -> str:
  src = open(os.path.join(REFERENCE, 'brax', 'envs', env + '.py')).read()
  m = re.search(r'^%s\s*=\s*"""(.*?)"""' % var, src, re.S | re.M)
  if not m:
    raise KeyError('%s has no %s' % (env, var))
  return m.group(1)


def parse(text: str):
  brax = setup()
  from google.protobuf import text_format
  return text_format.Parse(text, brax.Config())


def env_system(env: str, var: str = '_SYSTEM_CONFIG'):
  brax = setup()
  return brax.System(parse(env_config_text(env, var)))
