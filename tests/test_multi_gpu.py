"""N > 1 on real GPUs (BASELINE.json configs[4]): the sharded drivers launched the
way a user launches them -- one process per GPU under torch.distributed.run -- must produce the
bytes of a single-process run.  world_size = the number of GPUs on the box, at least 2: on a box
with one GPU two ranks share it and the record gather goes through gloo
(BOP_B200_DIST_BACKEND=gloo), so the shard / LPT / unequal-gather logic still runs on CUDA kernels.
"""
import json
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
from test_driver import dataset, driver_golden, _params  # noqa: F401  (fixtures)


def _free_port():
  s = socket.socket()
  s.bind(('127.0.0.1', 0))
  port = s.getsockname()[1]
  s.close()
  return port


def _launch(world, argv, timeout=900):
  import torch
  env = dict(os.environ)
  if torch.cuda.device_count() < world:
    env['BOP_B200_DIST_BACKEND'] = 'gloo'
  cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(world),
         '--master-addr', '127.0.0.1', '--master-port', str(_free_port())] + argv
  r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
  assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
  return r.stdout


def _world():
  """GPUs on the box, at least 2 (BOP_TEST_WORLD overrides: more ranks than GPUs share them over gloo)."""
  import torch
  if os.environ.get('BOP_TEST_WORLD'):
    return max(2, int(os.environ['BOP_TEST_WORLD']))
  return max(2, torch.cuda.device_count())


def _tree(path):
  out = {}
  for dirpath, _, names in os.walk(path):
    for name in names:
      full = os.path.join(dirpath, name)
      out[os.path.relpath(full, path)] = open(full, 'rb').read()
  return out


@pytest.mark.gpu
@pytest.mark.parametrize('error_type,n_top', [('vsd', -1), ('mssd', 1), ('mspd', 0)])
def test_sharded_driver_writes_the_single_process_files(dataset, tmp_path, error_type, n_top):
  """scripts/eval_calc_errors.py under torchrun (images sharded over the ranks, records gathered)
  vs the same script in one process: identical error files."""
  from bop_toolkit_b200 import eval_calc_errors as ece
  single = str(tmp_path / 'single')
  ece.calc_errors(_params(dataset, error_type, n_top, single), log=lambda s: None)
  sharded = str(tmp_path / 'sharded')
  _launch(_world(), [os.path.join('scripts', 'eval_calc_errors.py'), '--error_type=' + error_type,
                     '--n_top=%d' % n_top, '--result_filenames=' + dataset['result_filename'],
                     '--results_path=' + dataset['results_path'], '--eval_path=' + sharded,
                     '--datasets_path=' + dataset['datasets_path']])
  a, b = _tree(single), _tree(sharded)
  assert sorted(a) == sorted(b) and len(a) > 0
  for k in a:
    assert a[k] == b[k], k


@pytest.mark.gpu
def test_cfg5_leg_is_world_size_invariant():
  """The strong-scaling leg of bench.py (tools/cfg5_leg.py) at a miniature size: gathered records
  and the 120 recalls hash the same at world_size 1 and N; shards are unequal (LPT on pairs per
  image with 3..7 instances per image)."""
  one = json.loads(_launch(1, [os.path.join('tools', 'cfg5_leg.py'), '--images', '96']).strip().splitlines()[-1])
  w = _world()
  many = json.loads(_launch(w, [os.path.join('tools', 'cfg5_leg.py'), '--images', '96']).strip().splitlines()[-1])
  assert many['n_gpus'] == w and one['n_gpus'] == 1
  assert one['pairs_total'] == many['pairs_total'] == sum(many['per_rank_pairs']) > 0
  assert one['records_sha256'] == many['records_sha256']
  assert one['recalls_sha256'] == many['recalls_sha256']
  assert one['AR'] == many['AR']
  assert 0.0 < one['AR_mean'] < 1.0
