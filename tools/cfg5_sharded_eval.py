#!/usr/bin/env python
"""BASELINE.json configs[4] in miniature: a synthetic YCB-V-layout test set evaluated end to end
by the drop-in drivers, sharded over the GPUs of one node.

  python tools/cfg5_sharded_eval.py --images 40                       (1 GPU)
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 \
      tools/cfg5_sharded_eval.py --images 40                          (2 GPUs)

Stages (all through the package's public drivers): write the dataset (depth rendered with the
CUDA rasteriser), calc_gt_info -> scene_gt_info.json, eval_calc_errors for vsd / mssd / mspd with
n_top = -1 (images sharded over the ranks, error records all-gathered over NCCL), then the BOP19
average recalls with the in-process score stage.  With more than one rank, rank 0 also evaluates
everything unsharded and checks that the gathered records are bit-identical.  Prints one JSON line.
"""
import argparse
import json
import os
import shutil
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--images', type=int, default=20, help='images per scene')
  ap.add_argument('--config', default='cfg5', choices=['cfg5', 'cfg3'],
                  help='cfg5: YCB-V layout, 12 scenes, 21 models, 5 targets/image, n_top=-1; '
                       'cfg3: LM-O layout, 1 scene, 8 models in every image, n_top=1 (BASELINE.json configs)')
  ap.add_argument('--root', default=None)
  ap.add_argument('--keep', action='store_true')
  ap.add_argument('--host_decode_too', action='store_true',
                  help='also time the VSD pass with OpenCV decoding the depth PNGs on the host')
  ap.add_argument('--profile', action='store_true', help='cProfile of the error + score stages (rank 0)')
  args = ap.parse_args()

  import torch
  import torch.distributed as dist
  from bop_toolkit_b200 import batch, calc_gt_info, eval_calc_errors as ece, eval_calc_scores as ecs, synth

  world = int(os.environ.get('WORLD_SIZE', '1'))
  rank = int(os.environ.get('RANK', '0'))
  local = int(os.environ.get('LOCAL_RANK', '0'))
  torch.cuda.set_device(local)
  if world > 1:
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
  if args.config == 'cfg5':
    ds = dict(dataset='ycbv', scene_ids=tuple(range(48, 60)), obj_ids=tuple(range(1, 22)), objs_per_image=5,
              ests_per_target=2)
    n_top = -1
  else:
    ds = dict(dataset='lmo', scene_ids=(2,), obj_ids=(1, 5, 6, 8, 9, 10, 11, 12), objs_per_image=8,
              ests_per_target=2)
    n_top = 1
  dataset = ds['dataset']
  root = args.root or os.path.join(tempfile.gettempdir(), 'bop_%s_%d' % (args.config, args.images))
  quiet = lambda s: None
  t = {}

  # ---- dataset (rank 0) ----------------------------------------------------------------
  if rank == 0:
    t0 = time.time()
    if os.path.exists(root):
      shutil.rmtree(root)
    cache = {}

    def render_fn(pts, faces, R, t_, K):
      key = id(pts)
      if key not in cache:
        ms = batch.ModelSet('cuda:%d' % local)
        ms.add(0, pts, faces)
        cache[key] = ms
      cam = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]])
      return batch.render_batch(cache[key], 0, batch.pack_poses(R[None], t_[None]), cam,
                                synth.IM_H, synth.IM_W)[0].cpu().numpy()
    info = synth.write_bop_dataset(root, render_fn, split='test', n_images=args.images, seed=3,
                                   mesh_res=(50, 100), **ds)
    t['write_dataset_s'] = time.time() - t0
    t0 = time.time()
    p = calc_gt_info.parse_args(['--dataset=' + dataset, '--dataset_split=test',
                                 '--datasets_path=' + info['datasets_path']])
    p['device'] = 'cuda:%d' % local
    calc_gt_info.calc_gt_info(p, log=quiet)
    t['calc_gt_info_s'] = time.time() - t0
    with open(os.path.join(root, 'info.json'), 'w') as f:
      json.dump(info, f)
  if world > 1:
    dist.barrier()
  info = json.load(open(os.path.join(root, 'info.json')))

  # ---- errors: sharded over the ranks ------------------------------------------------------
  def params(error_type):
    return ece.parse_args(['--error_type=' + error_type, '--n_top=%d' % n_top,
                           '--result_filenames=' + info['result_filename'],
                           '--results_path=' + info['results_path'], '--eval_path=' + info['eval_path'],
                           '--datasets_path=' + info['datasets_path']])
  records, n_pairs = {}, {}
  prof = None
  if args.profile and rank == 0:
    import cProfile
    prof = cProfile.Profile()
    prof.enable()
  for et in ('vsd', 'mssd', 'mspd'):
    torch.cuda.synchronize()
    t0 = time.time()
    records[et] = ece.calc_errors(params(et), log=quiet, write=(rank == 0))['synth_%s-test' % dataset]
    torch.cuda.synchronize()
    t['errors_%s_s' % et] = time.time() - t0
    n_pairs[et] = sum(len(r['errors']) for recs in records[et].values() for r in recs)

  if args.host_decode_too:
    # the same VSD pass (no files written, caches warm) with the depth PNGs decoded by OpenCV on
    # the host threads (the r01 path) and by the GPU decoder
    for mode in ('host', 'gpu'):
      os.environ['BOP_B200_DEPTH_DECODE'] = mode
      torch.cuda.synchronize()
      t0 = time.time()
      rec = ece.calc_errors(params('vsd'), log=quiet, write=False)['synth_%s-test' % dataset]
      torch.cuda.synchronize()
      t['errors_vsd_%s_png_decode_nowrite_s' % mode] = time.time() - t0
      if mode == 'host':
        rec_host = rec
    os.environ.pop('BOP_B200_DEPTH_DECODE', None)
  out = None
  if args.host_decode_too:
    t['vsd_records_equal_host_decode'] = bool(rec_host == records['vsd'])
  if rank == 0:
    # ---- sharded == unsharded (bit-identical records) ----------------------------------
    identical = None
    if world > 1:
      identical = True
      real_info = ece._dist_info
      ece._dist_info = lambda: (0, 1)
      try:
        for et in ('vsd', 'mssd', 'mspd'):
          single = ece.calc_errors(params(et), log=quiet, write=False)['synth_%s-test' % dataset]
          identical = identical and (single == records[et])
      finally:
        ece._dist_info = real_info
    # ---- BOP19 average recalls (eval_bop19_pose.py:47-66): in-process score stage --------
    t0 = time.time()
    ths = {'vsd': [round(x, 2) for x in np.arange(0.05, 0.51, 0.05)],
           'mssd': [round(x, 2) for x in np.arange(0.05, 0.51, 0.05)],
           'mspd': [float(x) for x in np.arange(5, 51, 5)]}
    ar, n_scores, ctx = {}, 0, {}
    p_vsd = params('vsd')
    from bop_toolkit_b200 import score_sweep
    for et in ('vsd', 'mssd', 'mspd'):
      recalls = []
      ps = ecs.parse_args(['--datasets_path=' + info['datasets_path']])
      sweep = score_sweep.ThresholdSweep(ps, et, n_top, dataset, 'test', None, records[et], ctx)
      for col in range(len(p_vsd['vsd_taus']) if et == 'vsd' else 1):
        for sc, _ in sweep.run(ths[et], col=col, want_matches=False):
          recalls.append(sc['recall'])
          n_scores += 1
      ar[et] = float(np.mean(recalls))
    t['scores_s'] = time.time() - t0
    if prof is not None:
      import pstats
      prof.disable()
      pstats.Stats(prof, stream=sys.stderr).sort_stats('cumulative').print_stats(45)
    out = {'config': 'synthetic %s-test layout (%s): %d scene(s) x %d images, %d models of 4902 verts, n_top=%d'
                     % (dataset, args.config, len(ds['scene_ids']), args.images, len(ds['obj_ids']), n_top),
           'n_gpus': world, 'pairs': n_pairs, 'seconds': {k: (round(v, 3) if isinstance(v, float) else v) for k, v in t.items()},
           'pairs_per_s_driver': {et: round(n_pairs[et] / t['errors_%s_s' % et], 1) for et in n_pairs},
           'AR': ar, 'AR_mean': float(np.mean(list(ar.values()))), 'score_settings_evaluated': n_scores,
           'sharded_equals_unsharded': identical}
    print(json.dumps(out))
    if not args.keep:
      shutil.rmtree(root, ignore_errors=True)
  if world > 1:
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
  main()
