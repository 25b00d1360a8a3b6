"""Shared helpers for the tests (oracle = checker only)."""
import os

import numpy as np

from brax_b200.desc import SystemDesc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')
SYSTEMS = os.path.join(ROOT, 'brax_b200', 'systems')

QP_KEYS = ('pos', 'rot', 'vel', 'ang')
INFO_KEYS = ('contact_vel', 'contact_ang', 'joint_vel', 'joint_ang',
             'actuator_vel', 'actuator_ang', 'contact_pos', 'contact_normal',
             'contact_penetration')


def load_desc(name: str) -> SystemDesc:
  return SystemDesc.load(os.path.join(SYSTEMS, name + '.json'))


def load_golden(name: str):
  return dict(np.load(os.path.join(GOLDEN, name + '_steps.npz')))


def golden_names():
  return sorted(f[:-len('_steps.npz')] for f in os.listdir(GOLDEN)
                if f.endswith('_steps.npz'))
