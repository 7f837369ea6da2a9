#!/usr/bin/env python
"""Drop-in for the reference's scripts/eval_calc_errors.py (same flags, same output files):
  python scripts/eval_calc_errors.py --error_type=vsd --result_filenames=method_lmo-test.csv ...
  torchrun --nproc-per-node 8 scripts/eval_calc_errors.py ...      (images sharded over the GPUs)
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bop_toolkit_b200.eval_calc_errors import main  # noqa: E402

if __name__ == '__main__':
  main()
