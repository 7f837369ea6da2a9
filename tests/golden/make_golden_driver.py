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

DATASET_KW = dict(dataset='lm', split='test', scene_ids=(2, 3), n_images=3, obj_ids=tuple(range(1, 16)),
                  objs_per_image=4, ests_per_target=3, seed=7)
# (error type, n_top) whose error files are also scored; score settings = extra flags
SCORE_RUNS = [('vsd', 1), ('vsd', -1), ('mssd', 1), ('mssd', 0), ('mspd', -1), ('ad', 1), ('rete', 1), ('cus', 1)]
SCORE_SETTINGS = [[], ['--visib_gt_min=0.3'], ['--correct_th_vsd=0.15', '--correct_th_mssd=0.1',
                                               '--correct_th_mspd=5', '--correct_th_ad=0.05',
                                               '--correct_th_rete=10,3', '--correct_th_cus=0.3']]
RUNS = [('vsd', 1), ('vsd', -1), ('mssd', 1), ('mssd', 0), ('mspd', 1), ('mspd', -1), ('ad', 1),
        ('add', 1), ('adi', 1), ('proj', 1), ('cus', 1), ('rete', 1), ('re', 1), ('te', 0)]


def oracle_render_fn(pts, faces, R, t, K):
  return renderer_oracle.render_depth(pts, faces, R, t, K[0, 0], K[1, 1], K[0, 2], K[1, 2],
                                      synth.IM_H, synth.IM_W)


def dataset_checksum(info):
  """sha256 over the decoded depth images, the results CSV and the targets (byte-exact inputs)."""
  import cv2
  h = hashlib.sha256()
  base = os.path.join(info['datasets_path'], DATASET_KW['dataset'])
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


def run_script(script, argv):
  old = sys.argv
  sys.argv = [script] + argv
  try:
    runpy.run_path(script, run_name='__main__')
  finally:
    sys.argv = old


def collect(root):
  out = {}
  for dirpath, _, names in os.walk(root):
    for name in names:
      full = os.path.join(dirpath, name)
      out[os.path.relpath(full, root)] = open(full).read()
  return out


def patch_dataset_params(info):
  """calc_gt_info.py / calc_gt_masks.py have no flags (dataset and paths are edited in their `p`
  dict): route their dataset_params calls to the synthetic dataset instead."""
  from bop_toolkit_lib import dataset_params as ref_dp
  orig_split, orig_model = ref_dp.get_split_params, ref_dp.get_model_params
  ds, split = DATASET_KW['dataset'], DATASET_KW['split']
  ref_dp.get_split_params = lambda path, name, sp, st=None: orig_split(info['datasets_path'], ds, split, None)
  ref_dp.get_model_params = lambda path, name, mt=None: orig_model(info['datasets_path'], ds, mt)
  return lambda: (setattr(ref_dp, 'get_split_params', orig_split), setattr(ref_dp, 'get_model_params', orig_model))


def main():
  import cv2
  renderer_oracle.build()
  tmp = tempfile.mkdtemp(prefix='bop_golden_')
  try:
    info = synth.write_bop_dataset(tmp, oracle_render_fn, **DATASET_KW)
    install_reference_shim()
    sys.modules['imageio'].imwrite = lambda path, im, **kw: cv2.imwrite(path, im)
    ds_dir = os.path.join(info['datasets_path'], DATASET_KW['dataset'])
    split_dir = os.path.join(ds_dir, DATASET_KW['split'])

    # ---- scene_gt_info.json by the reference's calc_gt_info.py ---------------------------
    undo = patch_dataset_params(info)
    run_script('/root/reference/scripts/calc_gt_info.py', [])
    gt_info = {str(s): open(os.path.join(split_dir, '%06d' % s, 'scene_gt_info.json')).read()
               for s in info['scene_ids']}
    # ---- masks by the reference's calc_gt_masks.py (it only visits scene 0: link it) --------
    first = info['scene_ids'][0]
    os.symlink(os.path.join(split_dir, '%06d' % first), os.path.join(split_dir, '000000'))
    run_script('/root/reference/scripts/calc_gt_masks.py', [])
    os.unlink(os.path.join(split_dir, '000000'))
    undo()
    masks = {}
    for kind in ('mask', 'mask_visib'):
      mdir = os.path.join(split_dir, '%06d' % first, kind)
      for name in sorted(os.listdir(mdir)):
        m = cv2.imread(os.path.join(mdir, name), cv2.IMREAD_UNCHANGED)
        assert m.dtype == np.uint8 and set(np.unique(m)) <= {0, 255}
        masks['%s/%s' % (kind, name)] = np.packbits(m > 0)
    np.savez_compressed(os.path.join(HERE, 'gt_masks_golden.npz'), scene_id=first, **masks)

    # ---- error files by the reference's eval_calc_errors.py ------------------------------
    files = {}
    eval_paths = {}
    for error_type, n_top in RUNS:
      eval_path = os.path.join(tmp, 'eval_%s_%d' % (error_type, n_top))
      eval_paths[(error_type, n_top)] = eval_path
      run_script('/root/reference/scripts/eval_calc_errors.py', [
        '--error_type=' + error_type, '--n_top=%d' % n_top,
        '--result_filenames=' + info['result_filename'], '--results_path=' + info['results_path'],
        '--eval_path=' + eval_path, '--datasets_path=' + info['datasets_path'],
        '--renderer_type=vispy'])
      for rel, text in collect(eval_path).items():
        files['%s|%d|%s' % (error_type, n_top, rel)] = text

    # ---- scores / matches by the reference's eval_calc_scores.py -------------------------
    scores = {}
    for error_type, n_top in SCORE_RUNS:
      eval_path = eval_paths[(error_type, n_top)]
      result_dir = os.path.join(eval_path, os.listdir(eval_path)[0])
      dirs = sorted(os.listdir(result_dir))
      if error_type == 'vsd':
        dirs = [d for d in dirs if d.endswith('tau=0.200') or d.endswith('tau=0.500')]
      for si, extra in enumerate(SCORE_SETTINGS):
        before = set(collect(eval_path))
        run_script('/root/reference/scripts/eval_calc_scores.py', [
          '--error_dir_paths=' + ','.join(os.path.join(os.path.basename(result_dir), d) for d in dirs),
          '--eval_path=' + eval_path, '--datasets_path=' + info['datasets_path']] + extra)
        after = collect(eval_path)
        for rel in sorted(set(after) - before):
          scores['%s|%d|%d|%s' % (error_type, n_top, si, rel)] = after[rel]
    # ---- the whole BOP19 evaluation by the reference's eval_bop19_pose.py -----------------
    # It shells out to eval_calc_errors.py / eval_calc_scores.py 123 times; the children need
    # the same shim, so subprocess.call is routed to runpy in this process.  The children take
    # the dataset folder from config.datasets_path (BOP_PATH), patched here.
    import subprocess
    from bop_toolkit_lib import config as ref_config
    ref_config.datasets_path = info['datasets_path']
    real_call = subprocess.call

    def call_in_process(cmd):
      assert cmd[0] == 'python'
      run_script(cmd[1], list(cmd[2:]))
      return 0
    subprocess.call = call_in_process
    bop19_eval = os.path.join(tmp, 'eval_bop19')
    try:
      run_script('/root/reference/scripts/eval_bop19_pose.py', [
        '--result_filenames=' + info['result_filename'], '--results_path=' + info['results_path'],
        '--eval_path=' + bop19_eval, '--renderer_type=vispy'])
    finally:
      subprocess.call = real_call
    bop19 = {'scores': {}, 'matches_sha256': {}, 'n_error_files': 0}
    for rel, text in collect(bop19_eval).items():
      name = os.path.basename(rel)
      if name.startswith('scores_'):
        bop19['scores'][rel] = text
      elif name.startswith('matches_'):
        bop19['matches_sha256'][rel] = hashlib.sha256(text.encode()).hexdigest()
      else:
        bop19['n_error_files'] += 1

    # ---- the reference's dataset tables (fields the drivers read) ------------------------
    from bop_toolkit_lib import dataset_params as ref_dp
    tables = {}
    for ds in ['lm', 'lmo', 'tless', 'tudl', 'tyol', 'ruapc', 'icmi', 'icbin', 'itodd', 'hbs', 'hb', 'ycbv', 'hope']:
      tables['model|%s' % ds] = ref_dp.get_model_params('/d', ds, 'eval')
      for split in ['train', 'val', 'test']:
        for st in [None, 'pbr', 'primesense', 'kinect', 'real', 'synt', 'render']:
          try:
            r = ref_dp.get_split_params('/d', ds, split, st)
          except Exception:
            continue
          tables['split|%s|%s|%s' % (ds, split, st)] = {
            k: (list(r[k]) if isinstance(r[k], tuple) else r[k]) for k in r if k in
            ('name', 'split', 'split_type', 'base_path', 'split_path', 'im_size', 'scene_ids',
             'scene_camera_tpath', 'scene_gt_tpath', 'scene_gt_info_tpath', 'depth_tpath',
             'mask_tpath', 'mask_visib_tpath')}
    out = {'tables': tables, 'dataset_kw': {k: (list(v) if isinstance(v, tuple) else v) for k, v in DATASET_KW.items()},
           'dataset_sha256': dataset_checksum(info), 'runs': [list(r) for r in RUNS], 'files': files,
           'gt_info': gt_info, 'score_runs': [list(r) for r in SCORE_RUNS],
           'score_settings': SCORE_SETTINGS, 'scores': scores, 'bop19': bop19}
    with open(os.path.join(HERE, 'driver_golden.json'), 'w') as f:
      json.dump(out, f, indent=0, sort_keys=True)
    print('wrote %d error files, %d score files, %d masks, %d bop19 score files, dataset sha256 %s'
          % (len(files), len(scores), len(masks), len(bop19['scores']), out['dataset_sha256']))
  finally:
    shutil.rmtree(tmp, ignore_errors=True)


if __name__ == '__main__':
  main()
