"""Host-side helpers of the pose-error path (the parts of bop_toolkit_lib/misc.py the driver
needs).  Symmetry sets are computed once per object on the host and uploaded as [S,12] float64
(SURVEY.md §8 a12); everything per-pair lives in the CUDA kernels."""
import datetime
import math
import os
import sys

import numpy as np


def log(s):
  """UTC-timestamped log line (same format as bop_toolkit_lib/misc.py:18-31, without pytz)."""
  now = datetime.datetime.now(datetime.timezone.utc)
  sys.stdout.write('{}/{}|{:02d}:{:02d}:{:02d}: {}\n'.format(
    now.month, now.day, now.hour, now.minute, now.second, s))
  sys.stdout.flush()


def ensure_dir(path):
  if path and not os.path.exists(path):
    os.makedirs(path, exist_ok=True)


def _axis_angle_matrix(angle, axis):
  """Rodrigues rotation, term order as in transform.rotation_matrix (transform.py:326-335) so
  that the result is bit-identical: cos*I + (1-cos) a a^T + sin [a]x."""
  a = np.array(axis[:3], dtype=np.float64)
  a /= math.sqrt(float(np.dot(a, a)))
  c, s = math.cos(angle), math.sin(angle)
  R = np.diag([c, c, c])
  R += np.outer(a, a) * (1.0 - c)
  a = a * s
  R += np.array([[0.0, -a[2], a[1]], [a[2], 0.0, -a[0]], [-a[1], a[0], 0.0]])
  return R


def get_symmetry_transformations(model_info, max_sym_disc_step):
  """Symmetry set of an object model -- same contract as misc.get_symmetry_transformations
  (bop_toolkit_lib/misc.py:43-90): list of {'R': 3x3, 't': 3x1}.

  Discrete part: identity + every 4x4 in 'symmetries_discrete'.  Continuous part: for each
  axis, ceil(pi / step) rotations per turn, i = 1 .. count-1 (the identity rotation itself is
  NOT part of the set when a continuous symmetry exists).  Result = continuous o discrete.
  """
  discrete = [(np.eye(3), np.zeros((3, 1)))]
  for flat in model_info.get('symmetries_discrete', []):
    m = np.reshape(flat, (4, 4))
    discrete.append((m[:3, :3], m[:3, 3].reshape((3, 1))))
  continuous = []
  for sym in model_info.get('symmetries_continuous', []):
    axis = np.array(sym['axis'])
    offset = np.array(sym['offset']).reshape((3, 1))
    count = int(np.ceil(np.pi / max_sym_disc_step))
    step = 2.0 * np.pi / count
    for i in range(1, count):
      R = _axis_angle_matrix(i * step, axis)
      continuous.append((R, -R.dot(offset) + offset))
  if not continuous:
    return [{'R': R, 't': t} for R, t in discrete]
  out = []
  for Rd, td in discrete:
    for Rc, tc in continuous:
      out.append({'R': Rc.dot(Rd), 't': Rc.dot(td) + tc})
  return out


def overlapping_sphere_projections(radius, p1, p2):
  """Host twin of the device gate (vsd.cu sphere_projections_overlap); contract of
  bop_toolkit_lib/misc.py:309-331."""
  p1 = np.asarray(p1, dtype=np.float64).reshape(3)
  p2 = np.asarray(p2, dtype=np.float64).reshape(3)
  if p1[2] == 0 or p2[2] == 0:
    return False
  d = p1[:2] / p1[2] - p2[:2] / p2[2]
  return bool(math.sqrt(d[0] * d[0] + d[1] * d[1]) < radius * (1.0 / p1[2] + 1.0 / p2[2]))


def get_error_signature(error_type, n_top, **kwargs):
  """Directory-name signature of an error setting (contract of misc.py:334-350)."""
  sign = 'error=' + error_type + '_ntop=' + str(n_top)
  if error_type == 'vsd':
    tau = kwargs['vsd_tau']
    tau_str = 'inf' if tau == float('inf') else '{:.3f}'.format(tau)
    sign += '_delta={:.3f}_tau={}'.format(kwargs['vsd_delta'], tau_str)
  return sign


def get_score_signature(correct_th, visib_gt_min):
  """File-name signature of a score setting (contract of misc.py:353-361)."""
  return 'th=' + '-'.join('{:.3f}'.format(t) for t in correct_th) + \
    '_min-visib={:.3f}'.format(visib_gt_min)
