"""The committed descriptors (brax_b200/systems/*.json) ARE what `brax_b200.extract.describe`
reads off the reference's constructed `brax.System` -- the drop-in route `System.from_brax`
(INTEGRATION.md 2).  Needs the reference tree (build container only; skipped elsewhere)."""
import json
import os
import sys

import pytest

from tests import util

sys.path.insert(0, os.path.join(util.ROOT, 'oracle'))
import refsys  # noqa: E402

pytestmark = pytest.mark.skipif(not os.path.isdir(refsys.REFERENCE), reason='reference tree not present')


@pytest.mark.parametrize('name,env,var,extra', [
    ('ant', 'ant', '_SYSTEM_CONFIG', 0), ('humanoid_spring', 'humanoid', '_SYSTEM_CONFIG_SPRING', 0),
    ('grasp', 'grasp', '_SYSTEM_CONFIG', 3), ('fetch', 'fetch', '_SYSTEM_CONFIG', 0)])
def test_committed_descriptor_is_a_fresh_extraction(name, env, var, extra):
  from brax_b200.extract import describe
  s = refsys.env_system(env, var)
  d = describe(s, name, act_size=int(s.num_joint_dof + s.num_forces_dof) + extra)
  d.validate()
  want = json.load(open(os.path.join(util.SYSTEMS, name + '.json')))
  assert json.loads(d.to_json()) == want
  # the reset tables describe the same default state the reference returns
  assert d.reset['num_dof'] == len(s.default_angle())
