"""BOP Challenge 2019 evaluation (6D localization) -- drop-in for scripts/eval_bop19_pose.py.

The reference launches `eval_calc_errors.py` three times and `eval_calc_scores.py` 120 times as
subprocesses, each re-reading targets, GT and the error JSON (eval_bop19_pose.py:147-230).  Here
the same work runs in ONE process: the three error types go through the batched CUDA driver
(eval_calc_errors.calc_errors), and the 120 (error setting x threshold) scores are computed from
the in-memory records (eval_calc_scores.score_errors).  Every file the reference leaves behind is
still written -- errors_*.json, scores_*.json, matches_*.json and scores_bop19.json -- with the
same names and layout.  Same flags (:88-103) plus --datasets_path (the reference takes it from
the BOP_PATH environment variable only).
"""
import argparse
import os
import time

import numpy as np

from . import eval_calc_errors as ece
from . import eval_calc_scores as ecs
from . import inout, misc, score_sweep


def default_params():
  return {
    'errors': [
      {'n_top': -1, 'type': 'vsd', 'vsd_deltas': dict(ece.VSD_DELTAS),
       'vsd_taus': list(np.arange(0.05, 0.51, 0.05)), 'vsd_normalized_by_diameter': True,
       'correct_th': [[th] for th in np.arange(0.05, 0.51, 0.05)]},
      {'n_top': -1, 'type': 'mssd', 'correct_th': [[th] for th in np.arange(0.05, 0.51, 0.05)]},
      {'n_top': -1, 'type': 'mspd', 'correct_th': [[th] for th in np.arange(5, 51, 5)]},
    ],
    'visib_gt_min': -1,
    'max_sym_disc_step': 0.01,
    'renderer_type': 'cuda',
    'result_filenames': ['/relative/path/to/csv/with/results'],
    'results_path': os.environ.get('BOP_RESULTS_PATH', '/path/to/folder/with/results'),
    'eval_path': os.environ.get('BOP_EVAL_PATH', '/path/to/eval/folder'),
    'datasets_path': os.environ.get('BOP_PATH', '/path/to/bop/datasets'),
    'targets_filename': 'test_targets_bop19.json',
  }


def parse_args(argv=None):
  p = default_params()
  ap = argparse.ArgumentParser()
  ap.add_argument('--renderer_type', default=p['renderer_type'])
  ap.add_argument('--result_filenames', default=','.join(p['result_filenames']),
                  help='Comma-separated names of files with results.')
  for k in ('results_path', 'eval_path', 'targets_filename', 'datasets_path'):
    ap.add_argument('--' + k, default=p[k])
  a = ap.parse_args(argv)
  p['renderer_type'] = str(a.renderer_type)
  p['result_filenames'] = a.result_filenames.split(',')
  for k in ('results_path', 'eval_path', 'targets_filename', 'datasets_path'):
    p[k] = str(getattr(a, k))
  return p


def average_time_per_image(ests):
  """Mean run time over images; -1 if any estimate carries a negative time; all estimates of an
  image must report the same time within 1 ms (eval_bop19_pose.py:122-143)."""
  times = {}
  for est in ests:
    key = '{:06d}_{:06d}'.format(est['scene_id'], est['im_id'])
    if est['time'] < 0:
      return -1.0
    if key in times:
      if abs(times[key] - est['time']) > 0.001:
        raise ValueError('The running time for scene {} and image {} is not the same for all '
                         'estimates.'.format(est['scene_id'], est['im_id']))
    else:
      times[key] = est['time']
  return np.mean(list(times.values()))


def evaluate(p, log=misc.log, device=None):
  """-> {result_name: final scores dict}; writes the reference's files (rank 0)."""
  rank, _ = ece._dist_info()
  out = {}
  for result_filename in p['result_filenames']:
    log('===========')
    log('EVALUATING: {}'.format(result_filename))
    log('===========')
    t0 = time.time()
    result_name, _, dataset, split, split_type = ece.parse_result_filename(result_filename)
    ests = inout.load_bop_results(os.path.join(p['results_path'], result_filename), version='bop19')
    avg_time = average_time_per_image(ests)
    average_recalls, ctx = {}, {}
    for error in p['errors']:
      pe = ece.default_params()
      pe.update(n_top=error['n_top'], error_type=error['type'], result_filenames=[result_filename],
                renderer_type=p['renderer_type'], results_path=p['results_path'],
                eval_path=p['eval_path'], datasets_path=p['datasets_path'],
                targets_filename=p['targets_filename'], max_sym_disc_step=p['max_sym_disc_step'],
                skip_missing=True)
      if error['type'] == 'vsd':
        # the reference round-trips these through str() on the command line (:165-169)
        pe.update(vsd_deltas={k: float(v) for k, v in error['vsd_deltas'].items()},
                  vsd_taus=[float(str(t)) for t in error['vsd_taus']],
                  vsd_normalized_by_diameter=bool(str(error['vsd_normalized_by_diameter'])))
      records = ece.calc_errors(pe, device=device, log=log)[result_name]
      if rank != 0:
        continue
      if error['type'] == 'vsd':
        settings = [(misc.get_error_signature('vsd', error['n_top'], vsd_delta=pe['vsd_deltas'][dataset],
                                              vsd_tau=tau), tau_id)
                    for tau_id, tau in enumerate(pe['vsd_taus'])]
      else:
        settings = [(misc.get_error_signature(error['type'], error['n_top']), None)]
      recalls = []
      ps = ecs.default_params()
      ps.update(eval_path=p['eval_path'], datasets_path=p['datasets_path'],
                targets_filename=p['targets_filename'], visib_gt_min=float(p['visib_gt_min']))
      # the reference round-trips the thresholds through str() on the command line (:205-207)
      correct_ths = [[float(str(t)) for t in correct_th] for correct_th in error['correct_th']]
      # One vectorised pass per error type over every threshold (and tau) when the thresholds are
      # scalars (they are in BOP19); otherwise one score_errors call per setting.
      sweep = None
      if all(len(c) == 1 for c in correct_ths):
        sweep = score_sweep.ThresholdSweep(ps, error['type'], error['n_top'], dataset, split, split_type,
                                           records, ctx)
      for error_sign, tau_id in settings:
        if sweep is not None:
          results = sweep.run([c[0] for c in correct_ths], col=tau_id or 0)
        else:
          recs = records if tau_id is None else {s: ece.records_for_tau(r, tau_id) for s, r in records.items()}
          results = []
          for correct_th in correct_ths:
            ps['correct_th'][error['type']] = correct_th
            results.append(ecs.score_errors(ps, error['type'], error['n_top'], dataset, split,
                                            split_type, recs, do_print=False, ctx=ctx))
        base = os.path.join(p['eval_path'], result_name, error_sign)
        misc.ensure_dir(base)
        for correct_th, (scores, matches) in zip(correct_ths, results):
          sign = misc.get_score_signature(correct_th, ps['visib_gt_min'])
          inout.save_json(os.path.join(base, 'scores_{}.json'.format(sign)), scores)
          inout.save_json(os.path.join(base, 'matches_{}.json'.format(sign)), matches)
          recalls.append(scores['recall'])
      average_recalls[error['type']] = np.mean(recalls)
      log('Recall scores: {}'.format(' '.join(map(str, recalls))))
      log('Average recall: {}'.format(average_recalls[error['type']]))
    if rank != 0:
      continue
    log('Evaluation of {} took {}s.'.format(result_filename, time.time() - t0))
    final = {'bop19_average_recall_{}'.format(e['type']): average_recalls[e['type']] for e in p['errors']}
    final['bop19_average_recall'] = np.mean([average_recalls['vsd'], average_recalls['mssd'],
                                             average_recalls['mspd']])
    final['bop19_average_time_per_image'] = avg_time
    inout.save_json(os.path.join(p['eval_path'], result_name, 'scores_bop19.json'), final)
    log('FINAL SCORES:')
    for k, v in final.items():
      log('- {}: {}'.format(k, v))
    out[result_name] = final
  log('Done.')
  return out


def main(argv=None):
  p = parse_args(argv)
  world = int(os.environ.get('WORLD_SIZE', '1'))
  if world > 1:
    import torch
    import torch.distributed as dist
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
  evaluate(p)
  if world > 1:
    dist.destroy_process_group()


if __name__ == '__main__':
  main()
