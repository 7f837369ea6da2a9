"""CPU: the C-ABI library builds, loads and exports every symbol include/bop_b200.h declares;
host-side logic that needs no GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

from bop_toolkit_b200 import _lib, inout, misc, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
  hdr = open(os.path.join(ROOT, 'include', 'bop_b200.h')).read()
  hdr = re.sub(r'/\*.*?\*/', '', hdr, flags=re.S)
  return sorted(set(re.findall(r'\b(bop_[a-z0-9_]+)\s*\(', hdr)))


def test_library_exports_every_declared_symbol():
  _lib.build()
  lib = ctypes.CDLL(_lib.SO_PATH)
  declared = _declared_symbols()
  assert len(declared) >= 20
  for name in declared:
    assert hasattr(lib, name), 'missing export: ' + name
  assert sorted(_lib.EXPORTED_SYMBOLS) == declared     # the ctypes table covers the header
  assert _lib.lib().bop_abi_version() == _lib.ABI_VERSION == 6


def test_struct_layouts_match_header():
  assert ctypes.sizeof(_lib.ModelDev) == 64
  assert ctypes.sizeof(_lib.VsdParams) == 8 * _lib.BOP_MAX_TAUS + 8 + 8 * 4


def test_no_cpu_fallback_without_device():
  lib = _lib.lib()
  if lib.bop_device_count() > 0:
    pytest.skip('a GPU is present')
  with pytest.raises(_lib.BopCudaError):
    _lib.require_cuda()
  out = np.zeros(1)
  pose = np.zeros(12)
  rc = lib.bop_re_te_host(pose.ctypes.data, pose.ctypes.data, 1, out.ctypes.data, None)
  assert rc != 0 and b'no CUDA device' in lib.bop_last_error()
  from bop_toolkit_b200 import pose_error
  with pytest.raises(_lib.BopCudaError):
    pose_error.te(np.zeros(3), np.ones(3))


def test_error_signature_and_json_layout(tmp_path):
  assert misc.get_error_signature('vsd', -1, vsd_delta=15.0, vsd_tau=0.05) == \
    'error=vsd_ntop=-1_delta=15.000_tau=0.050'
  assert misc.get_error_signature('mssd', 1) == 'error=mssd_ntop=1'
  recs = [{'im_id': 3, 'obj_id': 1, 'est_id': 0, 'score': 0.5, 'errors': {0: [float('inf')], 2: [0.25]}}]
  p = str(tmp_path / 'errors_000001.json')
  inout.save_json(p, recs)
  txt = open(p).read()
  assert txt == '[\n  {"errors": {"0": [Infinity], "2": [0.25]}, "est_id": 0, "im_id": 3, "obj_id": 1, "score": 0.5}\n]'
  back = inout.load_json(p, keys_to_int=True)
  assert back[0]['errors'][0] == [float('inf')] and back[0]['errors'][2] == [0.25]


def test_bop_results_roundtrip(tmp_path):
  R_e, t_e, _, _ = synth.make_pose_pairs(5, seed=1)
  res = [{'scene_id': 1, 'im_id': i, 'obj_id': 2, 'score': 0.1 * i, 'R': R_e[i], 't': t_e[i].reshape(3, 1),
          'time': 0.5} for i in range(5)]
  p = str(tmp_path / 'x_lmo-test.csv')
  inout.save_bop_results(p, res)
  back = inout.load_bop_results(p)
  assert len(back) == 5
  for a, b in zip(res, back):
    assert np.array_equal(a['R'], b['R']) and np.array_equal(a['t'], b['t']) and a['score'] == b['score']


def test_overlap_gate_host_twin(golden):
  g = golden
  got = [misc.overlapping_sphere_projections(75.0, g['t_e'][i], g['t_g'][i]) for i in range(len(g['t_e']))]
  assert np.array_equal(np.array(got), g['sphere_proj_overlap'])


def test_closed_and_oriented_mesh_check():
  """Host-side gate of the rasteriser's back-face culling."""
  import numpy as np
  from bop_toolkit_b200 import batch, synth
  pts, f = synth.make_mesh(12, 24, 100.0, seed=1)
  n = len(pts)
  assert batch.closed_and_oriented(f, n)
  assert batch.closed_and_oriented(f[:, ::-1], n)                   # inside-out is still consistent
  assert not batch.closed_and_oriented(f[:-1], n)                   # hole
  assert not batch.closed_and_oriented(np.vstack([f[:5], f[5:, ::-1]]), n)   # mixed winding
  assert not batch.closed_and_oriented(np.vstack([f, f[:1]]), n)    # duplicate face
  assert not batch.closed_and_oriented(np.zeros((0, 3), int), n)
  tet = np.array([[0, 1, 2], [0, 3, 1], [1, 3, 2], [0, 2, 3]])
  assert batch.closed_and_oriented(tet, 4)
