"""GT info and GT masks -- batched drop-ins for scripts/calc_gt_info.py and
scripts/calc_gt_masks.py of the reference.

The reference renders every GT instance one at a time (on a 3x canvas for the info), converts
three full frames to distance images and reduces them with numpy; here all instances of a
scene go through one launch of the fused raster + visibility kernel (batch.gt_info_batch) and
only counts, boxes and (for the masks) the two 8-bit images leave the GPU.  Outputs:
`scene_gt_info.json` with the reference's fields (calc_gt_info.py:165-175) and
`mask/`, `mask_visib/` PNGs (calc_gt_masks.py:122-133).

The reference scripts are configured by editing a dict; the same keys are flags here.
"""
import argparse
import os

import numpy as np

from . import dataset_params, inout, misc


def default_params():
  return {'dataset': 'lm', 'dataset_split': 'test', 'dataset_split_type': None, 'delta': 15,
          'renderer_type': 'cuda',
          'datasets_path': os.environ.get('BOP_PATH', '/path/to/bop/datasets')}


def parse_args(argv=None):
  p = default_params()
  ap = argparse.ArgumentParser()
  for k, v in p.items():
    ap.add_argument('--' + k, default=v)
  a = ap.parse_args(argv)
  for k in p:
    p[k] = getattr(a, k)
  p['delta'] = float(p['delta'])
  if p['dataset_split_type'] in ('', 'None'):
    p['dataset_split_type'] = None
  return p


def present_scene_ids(dp_split):
  """Scene folders that exist on disk (dataset_params.get_present_scene_ids, :649-658)."""
  if not os.path.isdir(dp_split['split_path']):
    return []
  return sorted(int(d) for d in os.listdir(dp_split['split_path'])
                if d.isdigit() and os.path.isdir(os.path.join(dp_split['split_path'], d)))


def _setup(p, model_type):
  from . import batch
  dp_split = dataset_params.get_split_params(p['datasets_path'], p['dataset'], p['dataset_split'],
                                             p['dataset_split_type'])
  if p['dataset'] == 'tless' and model_type is None:
    model_type = 'cad'
  dp_model = dataset_params.get_model_params(p['datasets_path'], p['dataset'], model_type)
  ms = batch.ModelSet(p.get('device'))
  for obj_id in dp_model['obj_ids']:
    path = dp_model['model_tpath'].format(obj_id=obj_id)
    if os.path.exists(path):
      m = inout.load_ply(path)
      ms.add(obj_id, m['pts'], m['faces'])
  return batch, dp_split, dp_model, ms


def _scene_batch(p, dp_split, scene_id, load_depth_stack, device=None):
  """Flattens the GT instances of a scene: (im_ids, depth tensor, scales, K rows, per-instance
  image position / obj ids / poses, scene_gt)."""
  scene_camera = inout.load_scene_camera(dp_split['scene_camera_tpath'].format(scene_id=scene_id))
  scene_gt = inout.load_scene_gt(dp_split['scene_gt_tpath'].format(scene_id=scene_id))
  im_ids = sorted(scene_gt.keys())
  w, h = dp_split['im_size']
  paths = []
  for im_id in im_ids:
    path = dp_split['depth_tpath'].format(scene_id=scene_id, im_id=im_id)
    if not os.path.exists(path):
      path = path.replace('.tif', '.png')      # calc_gt_info.py:104-105
    paths.append(path)
  depth, sc = load_depth_stack(paths, [scene_camera[i]['depth_scale'] for i in im_ids], h, w,
                              device=device)
  K = np.stack([np.asarray(scene_camera[i]['cam_K'], np.float64).reshape(9) for i in im_ids])
  im_pos, obj_ids, poses = [], [], []
  for pos, im_id in enumerate(im_ids):
    for gt in scene_gt[im_id]:
      im_pos.append(pos)
      obj_ids.append(gt['obj_id'])
      poses.append(np.concatenate([gt['cam_R_m2c'].reshape(9), gt['cam_t_m2c'].reshape(3)]))
  return im_ids, depth, sc, K, np.asarray(im_pos, np.int32), obj_ids, \
    np.asarray(poses, np.float64).reshape(-1, 12), scene_gt


def calc_gt_info(p, log=misc.log, write=True, images_per_chunk=512):
  """-> {scene_id: {im_id: [info dicts]}}; writes scene_gt_info.json per scene."""
  from .eval_calc_errors import load_depth_stack
  batch, dp_split, dp_model, ms = _setup(p, None)
  out = {}
  for scene_id in present_scene_ids(dp_split):
    log('Calculating GT info - dataset: {} ({}, {}), scene: {}'.format(
      p['dataset'], p['dataset_split'], p['dataset_split_type'], scene_id))
    im_ids, depth, sc, K, im_pos, obj_ids, poses, scene_gt = _scene_batch(p, dp_split, scene_id, load_depth_stack, ms.device)
    res = batch.gt_info_batch(ms, ms.indices(obj_ids), poses, depth, im_pos, K, p['delta'], sc,
                              margin=None, zero_copy=(not depth.is_cuda) and depth.is_pinned())
    counts = res['counts'].cpu().numpy()
    fract = res['visib_fract'].cpu().numpy()
    bbox = res['bbox'].cpu().numpy()
    info, k = {}, 0
    for im_id in im_ids:
      info[im_id] = []
      for _ in scene_gt[im_id]:
        info[im_id].append({
          'px_count_all': int(counts[k, 0]), 'px_count_valid': int(counts[k, 1]),
          'px_count_visib': int(counts[k, 2]), 'visib_fract': float(fract[k]),
          'bbox_obj': [int(e) for e in bbox[k, :4]], 'bbox_visib': [int(e) for e in bbox[k, 4:]]})
        k += 1
    out[scene_id] = info
    if write:
      path = dp_split['scene_gt_info_tpath'].format(scene_id=scene_id)
      misc.ensure_dir(os.path.dirname(path))
      inout.save_json(path, info)
  return out


def calc_gt_masks(p, log=misc.log, instances_per_chunk=256):
  """Writes mask/ and mask_visib/ PNGs (0 / 255, 8-bit) for every GT instance."""
  import cv2
  from .eval_calc_errors import load_depth_stack
  batch, dp_split, dp_model, ms = _setup(p, None)
  n_written = 0
  for scene_id in present_scene_ids(dp_split):
    if not os.path.exists(dp_split['scene_gt_tpath'].format(scene_id=scene_id)):
      log('Skip scene {} (GT file not found).'.format(scene_id))
      continue
    log('Calculating masks - dataset: {} ({}, {}), scene: {}'.format(
      p['dataset'], p['dataset_split'], p['dataset_split_type'], scene_id))
    im_ids, depth, sc, K, im_pos, obj_ids, poses, scene_gt = _scene_batch(p, dp_split, scene_id, load_depth_stack, ms.device)
    for tp in ('mask_tpath', 'mask_visib_tpath'):
      misc.ensure_dir(os.path.dirname(dp_split[tp].format(scene_id=scene_id, im_id=0, gt_id=0)))
    names = [(im_id, g) for im_id in im_ids for g in range(len(scene_gt[im_id]))]
    oi = ms.indices(obj_ids)
    for c0 in range(0, len(names), instances_per_chunk):
      sl = slice(c0, c0 + instances_per_chunk)
      res = batch.gt_info_batch(ms, oi[sl], poses[sl], depth, im_pos[sl], K, p['delta'], sc,
                                margin=(0, 0), want_masks=True, zero_copy=(not depth.is_cuda) and depth.is_pinned())
      mask, visib = res['mask'].cpu().numpy(), res['mask_visib'].cpu().numpy()
      for k, (im_id, g) in enumerate(names[sl]):
        cv2.imwrite(dp_split['mask_tpath'].format(scene_id=scene_id, im_id=im_id, gt_id=g), mask[k])
        cv2.imwrite(dp_split['mask_visib_tpath'].format(scene_id=scene_id, im_id=im_id, gt_id=g), visib[k])
        n_written += 2
  return n_written


def main(argv=None):
  calc_gt_info(parse_args(argv))


if __name__ == '__main__':
  main()
