"""TEST INFRASTRUCTURE ONLY (checker of the CUDA reset kernel; the product never imports it).

numpy restatement of `System.default_qp(joint_angle=..., joint_velocity=...)`
(brax/physics/system.py:112-242), batched over envs, on the static tables of
`SystemDesc.reset` (brax_b200/extract.py:_reset_desc).  Pinned to the reference itself by
tests/golden/reset.npz (oracle/gen_golden_reset.py).
"""
import numpy as np


def _rotate(v, q):
  # math.rotate (brax/math.py:25-40), same operation order
  s, u = q[..., 0:1], q[..., 1:]
  r = 2 * (np.sum(u * v, -1, keepdims=True) * u) + (s * s - np.sum(u * u, -1, keepdims=True)) * v
  return r + 2 * s * np.cross(u, v)


def _quat_mul(u, v):
  # math.quat_mul (brax/math.py:130-145)
  return np.stack([
      u[..., 0] * v[..., 0] - u[..., 1] * v[..., 1] - u[..., 2] * v[..., 2] - u[..., 3] * v[..., 3],
      u[..., 0] * v[..., 1] + u[..., 1] * v[..., 0] + u[..., 2] * v[..., 3] - u[..., 3] * v[..., 2],
      u[..., 0] * v[..., 2] - u[..., 1] * v[..., 3] + u[..., 2] * v[..., 0] + u[..., 3] * v[..., 1],
      u[..., 0] * v[..., 3] + u[..., 1] * v[..., 2] - u[..., 2] * v[..., 1] + u[..., 3] * v[..., 0]], -1)


def _quat_rot_axis(axis, angle):
  # math.quat_rot_axis (brax/math.py:173-187)
  s = np.sin(angle / 2)[..., None]
  return np.concatenate([np.cos(angle / 2)[..., None], axis * s], -1)


def default_qp(desc, joint_angle, joint_velocity, dtype=np.float32):
  """joint_angle / joint_velocity [B,D] -> dict pos [B,N,3] rot [B,N,4] vel ang [B,N,3]."""
  r = desc.reset
  dt = np.dtype(dtype).type
  ja = np.asarray(joint_angle, dtype)
  jv = np.asarray(joint_velocity, dtype)
  b = ja.shape[0]
  qp = {k: np.tile(np.asarray(r['base'][k], dtype), (b, 1, 1)) for k in ('pos', 'rot', 'vel', 'ang')}
  ja = np.concatenate([ja, np.zeros((b, 1), dtype)], 1)      # system.py:163
  jv = np.concatenate([jv, np.zeros((b, 1), dtype)], 1)
  eye = np.eye(3, dtype=dtype)
  for j in r['joints']:
    angles, vels = ja[:, j['take']], jv[:, j['take']]
    rot, ref = np.asarray(j['rot'], dtype), np.asarray(j['ref'], dtype)
    axes = np.stack([_rotate(eye[i], rot) for i in range(3)])               # system.py:182
    ang = (axes.T[None] * vels[:, None, :]).sum(-1)                          # dot(axes.T, vels) :183
    lrot = np.tile(ref, (b, 1))
    for i in range(3):                                                       # system.py:185-189
      axis = _rotate(np.tile(axes[i], (b, 1)), lrot)
      lrot = _quat_mul(_quat_rot_axis(axis, angles[:, i]), lrot)
    p, c = j['body_p'], j['body_c']
    rot_p, pos_p = qp['rot'][:, p].copy(), qp['pos'][:, p].copy()
    world_rot = _quat_mul(rot_p, lrot)                                       # system.py:204
    local_pos = np.asarray(j['off_p'], dtype) - _rotate(np.tile(np.asarray(j['off_c'], dtype), (b, 1)), lrot)
    qp['pos'][:, c] = pos_p + _rotate(local_pos, rot_p)                      # system.py:206
    qp['rot'][:, c] = world_rot
    qp['ang'][:, c] = _rotate(ang.astype(dtype), rot_p)                      # system.py:207
  for g in r['groups']:                                                      # system.py:211-240
    zs = [np.zeros(b, dtype)] if g['zero'] else []
    for pr in g['probes']:
      pt = _rotate(np.tile(np.asarray(pr['point'], dtype), (b, 1)), qp['rot'][:, pr['body']])
      zs.append(qp['pos'][:, pr['body'], 2] + pt[:, 2] - dt(pr['radius']))   # bodies.py:71-93
    min_z = np.min(np.stack(zs), 0)
    for c in g['bodies']:
      qp['pos'][:, c, 2] = qp['pos'][:, c, 2] - min_z
  return {k: v.astype(dtype) for k, v in qp.items()}
