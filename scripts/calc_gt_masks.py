#!/usr/bin/env python
"""Drop-in for the reference's scripts/calc_gt_masks.py (its `p` dict keys are flags here)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bop_toolkit_b200 import calc_gt_info as m  # noqa: E402

if __name__ == '__main__':
  m.calc_gt_masks(m.parse_args())
