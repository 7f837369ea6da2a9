"""Golden error files of the DRIVER (scripts/eval_calc_errors.py of the reference).

Run here (the reference does not exist on the GPU box):  python tests/golden/make_golden_driver.py

A small synthetic dataset in BOP layout is written with bop_toolkit_b200.synth.write_bop_dataset
(deterministic: seeded numpy + the float64 C oracle rasteriser for the depth images), then the
REAL reference script is executed unmodified (runpy) for every error type, with the usual shim
(stub pytz / imageio / png, numpy.float) and `renderer.create_renderer` patched to return the
oracle rasteriser -- the reference's GL renderers cannot run here.  The oracle renders with
window positions snapped to 1/256 px, as OpenGL implementations do (GL_SUBPIXEL_BITS = 8 on the
GPUs the reference ran on): with ~400-pixel silhouettes of the small test models a single edge
pixel moves VSD by 2.5e-3, so the comparison needs the same snapping rule on both sides.  The text of every errors_*.json
it writes is stored in tests/golden/driver_golden.json together with a checksum of the dataset,
so the tests can regenerate the same dataset on the GPU box and compare file by file.
"""
import hashlib
import json
import os
import runpy
import shutil
import sys
import tempfile
import types
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from bop_toolkit_b200 import synth            # noqa: E402
from oracle import renderer_oracle            # noqa: E402

DATASET_KW = dict(dataset='lmo', split='test', scene_ids=(2, 3), n_images=3, obj_ids=(1, 5, 6, 8, 9, 10, 11, 12),
                  objs_per_image=4, ests_per_target=3, seed=7)
RUNS = [('vsd', 1), ('vsd', -1), ('mssd', 1), ('mssd', 0), ('mspd', 1), ('mspd', -1), ('ad', 1),
        ('add', 1), ('adi', 1), ('proj', 1), ('cus', 1), ('rete', 1), ('re', 1), ('te', 0)]


def oracle_render_fn(pts, faces, R, t, K):
  return renderer_oracle.render_depth(pts, faces, R, t, K[0, 0], K[1, 1], K[0, 2], K[1, 2],
                                      synth.IM_H, synth.IM_W)


def dataset_checksum(info):
  """sha256 over the decoded depth images, the results CSV and the targets (byte-exact inputs)."""
  import cv2
  h = hashlib.sha256()
  base = os.path.join(info['datasets_path'], 'lmo')
  for scene_id in info['scene_ids']:
    ddir = os.path.join(base, 'test', '%06d' % scene_id, 'depth')
    for name in sorted(os.listdir(ddir)):
      h.update(cv2.imread(os.path.join(ddir, name), cv2.IMREAD_UNCHANGED).tobytes())
  for path in (os.path.join(info['results_path'], info['result_filename']),
               os.path.join(base, 'test_targets_bop19.json')):
    h.update(open(path, 'rb').read())
  return h.hexdigest()


def install_reference_shim():
  import cv2
  pytz = types.ModuleType('pytz')
  pytz.utc = types.SimpleNamespace(localize=lambda dt: dt)
  sys.modules['pytz'] = pytz
  imageio = types.ModuleType('imageio')
  imageio.imread = lambda path: cv2.imread(path, cv2.IMREAD_UNCHANGED)
  sys.modules['imageio'] = imageio
  sys.modules['png'] = types.ModuleType('png')
  np.float = float
  sys.path.insert(0, '/root/reference')
  warnings.filterwarnings('ignore')
  from bop_toolkit_lib import inout as ref_inout
  from bop_toolkit_lib import renderer as ref_renderer

  class RefOracleRenderer(renderer_oracle.RendererOracle):
    def add_object(self, obj_id, model_path=None, **kwargs):
      m = ref_inout.load_ply(model_path)     # the reference's own PLY loader
      self.models[obj_id] = (m['pts'], m['faces'].astype(np.int32))

  ref_renderer.create_renderer = lambda w, h, renderer_type='cpp', mode='rgb+depth', **kw: \
    RefOracleRenderer(w, h, mode, subpixel_bits=8)


def main():
  renderer_oracle.build()
  tmp = tempfile.mkdtemp(prefix='bop_golden_')
  try:
    info = synth.write_bop_dataset(tmp, oracle_render_fn, **DATASET_KW)
    install_reference_shim()
    script = '/root/reference/scripts/eval_calc_errors.py'
    files = {}
    for error_type, n_top in RUNS:
      eval_path = os.path.join(tmp, 'eval_%s_%d' % (error_type, n_top))
      argv = [script, '--error_type=' + error_type, '--n_top=%d' % n_top,
              '--result_filenames=' + info['result_filename'],
              '--results_path=' + info['results_path'], '--eval_path=' + eval_path,
              '--datasets_path=' + info['datasets_path'], '--renderer_type=vispy']
      old = sys.argv
      sys.argv = argv
      try:
        runpy.run_path(script, run_name='__main__')
      finally:
        sys.argv = old
      for dirpath, _, names in os.walk(eval_path):
        for name in names:
          full = os.path.join(dirpath, name)
          rel = os.path.relpath(full, eval_path)
          files['%s|%d|%s' % (error_type, n_top, rel)] = open(full).read()
    out = {'dataset_kw': {k: (list(v) if isinstance(v, tuple) else v) for k, v in DATASET_KW.items()},
           'dataset_sha256': dataset_checksum(info), 'runs': [list(r) for r in RUNS], 'files': files}
    with open(os.path.join(HERE, 'driver_golden.json'), 'w') as f:
      json.dump(out, f, indent=0, sort_keys=True)
    print('wrote %d files, dataset sha256 %s' % (len(files), out['dataset_sha256']))
  finally:
    shutil.rmtree(tmp, ignore_errors=True)


if __name__ == '__main__':
  main()
