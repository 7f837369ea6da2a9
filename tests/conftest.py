import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)


def pytest_configure(config):
  config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def golden():
  path = os.path.join(ROOT, 'tests', 'golden', 'pose_error_golden.npz')
  return dict(np.load(path))


@pytest.fixture(scope='session')
def oracle_built():
  from oracle import renderer_oracle
  renderer_oracle.build()
  return True


def syms_from_array(arr):
  return [{'R': a[:9].reshape(3, 3).copy(), 't': a[9:].reshape(3, 1).copy()} for a in arr]
