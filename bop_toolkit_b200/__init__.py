"""bop_toolkit_b200 -- B200-native implementation of the BOP toolkit's pose-error hot path
(bop_toolkit_lib/pose_error.py driven by scripts/eval_calc_errors.py of gist-ailab/bop_toolkit).

  pose_error   drop-in vsd / mssd / mspd / add / adi / re / te / proj
  renderer     Renderer interface + create_renderer(..., 'cuda')
  batch        batched fast path (ModelSet, *_batch)
  misc, inout  host-side helpers and file formats of the path
"""
__version__ = '0.1.0'
