"""ctypes binding of libbop_b200.so (the C ABI declared in include/bop_b200.h).

There is no CPU fallback: if the shared library is missing this module raises, and every
compute entry point fails when no CUDA device is usable.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, 'csrc')
SO_PATH = os.path.join(CSRC, 'libbop_b200.so')
SOURCES = ['api.cu', 'pose_points.cu', 'vsd.cu', 'png_depth.cu', 'error_json.cpp']
HEADERS = ['common.cuh', 'raster.cuh', 'png_inflate.cuh', os.path.join('..', '..', 'include', 'bop_b200.h')]

BOP_MAX_TAUS = 16
ABI_VERSION = 6            # BOP_B200_ABI_VERSION of include/bop_b200.h
STATIC_MAX_PAIRS = 4096   # BOP_STATIC_MAX_PAIRS (bop_vsd_from_depth_static_dev)
COST_TYPES = {'step': 0, 'tlinear': 1}
VISIB_MODES = {'bop19': 0, 'bop18': 1}
DEPTH_F32, DEPTH_U16 = 0, 1


class ModelDev(ctypes.Structure):
  """struct bop_model_dev (include/bop_b200.h)."""
  _fields_ = [
    ('pts', ctypes.c_void_p), ('verts4', ctypes.c_void_p), ('faces4', ctypes.c_void_p),
    ('syms', ctypes.c_void_p), ('diameter', ctypes.c_double), ('n_pts', ctypes.c_int32),
    ('n_faces', ctypes.c_int32), ('n_syms', ctypes.c_int32), ('pts_f64', ctypes.c_int32),
    ('cull', ctypes.c_int32), ('cull_min_dist', ctypes.c_float)]


class PngInfo(ctypes.Structure):
  """struct bop_png_info (include/bop_b200.h)."""
  _fields_ = [
    ('width', ctypes.c_int32), ('height', ctypes.c_int32), ('bit_depth', ctypes.c_int32),
    ('color_type', ctypes.c_int32), ('interlace', ctypes.c_int32), ('status', ctypes.c_int32),
    ('idat_bytes', ctypes.c_int64)]


PNG_UNSUPPORTED, PNG_STAGING_TOO_SMALL = -104, -105


class VsdParams(ctypes.Structure):
  """struct bop_vsd_params (include/bop_b200.h)."""
  _fields_ = [
    ('taus', ctypes.c_double * BOP_MAX_TAUS), ('delta', ctypes.c_double),
    ('n_taus', ctypes.c_int32), ('normalized_by_diameter', ctypes.c_int32),
    ('cost_type', ctypes.c_int32), ('visib_mode', ctypes.c_int32), ('height', ctypes.c_int32),
    ('width', ctypes.c_int32), ('gate', ctypes.c_int32), ('depth_dtype', ctypes.c_int32)]


def _source_hash(paths, cmd):
  import hashlib
  h = hashlib.sha256(' '.join(cmd).encode())
  for p in paths:
    with open(p, 'rb') as f:
      h.update(os.path.basename(p).encode() + b'\0' + f.read())
  return h.hexdigest()


def build(force=False, verbose=False):
  """Compiles the CUDA sources for sm_100a into csrc/libbop_b200.so (in-tree).  Whether the
  library is current is decided by a hash of the sources and the compile command kept beside it
  (libbop_b200.so.src): file times do not survive copying the tree to another machine."""
  srcs = [os.path.join(CSRC, s) for s in SOURCES]
  deps = srcs + [os.path.join(CSRC, h) for h in HEADERS]
  extra = os.environ.get('BOP_NVCC_FLAGS', '').split()     # e.g. -DBOP_DEBUG_CHECKS (device-side range checks)
  cmd = ['nvcc', '-O3', '-std=c++17', '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo'] + extra + \
        ['-Xcompiler', '-fPIC', '-shared', '-o', SO_PATH] + srcs
  want = _source_hash(deps, [c for c in cmd if c != SO_PATH and c not in srcs] + SOURCES)
  stamp = SO_PATH + '.src'
  if not force and os.path.exists(SO_PATH) and os.path.exists(stamp):
    with open(stamp) as f:
      if f.read().strip() == want:
        return SO_PATH
  import shutil
  if shutil.which('nvcc') is None and os.path.exists(SO_PATH) and not force:
    return SO_PATH          # a machine without the toolkit runs what was built elsewhere
  if verbose:
    cmd.insert(1, '-Xptxas=-v')
  if os.path.exists(stamp):
    os.remove(stamp)
  subprocess.check_call(cmd, cwd=CSRC)
  with open(stamp, 'w') as f:
    f.write(want)
  return SO_PATH


_P = ctypes.c_void_p
_I = ctypes.c_int
_SIGNATURES = {
  'bop_abi_version': (_I, []),
  'bop_last_error': (ctypes.c_char_p, []),
  'bop_device_count': (_I, []),
  'bop_kernel_launch_count': (ctypes.c_uint64, []),
  'bop_re_te_dev': (_I, [_P, _P, _I, _P, _P, _P]),
  'bop_re_te_host': (_I, [_P, _P, _I, _P, _P]),
  'bop_mssd_dev': (_I, [_P, _P, _P, _P, _I, _I, _I, _P, _P]),
  'bop_mspd_dev': (_I, [_P, _P, _P, _P, _P, _I, _I, _P, _P]),
  'bop_add_dev': (_I, [_P, _P, _P, _P, _I, _I, _I, _P, _P]),
  'bop_proj_dev': (_I, [_P, _P, _P, _P, _P, _I, _I, _P, _P]),
  'bop_adi_dev': (_I, [_P, _P, _P, _P, _I, _I, _I, _P, _P]),
  'bop_render_workspace_bytes': (ctypes.c_size_t, [_I, _I, _I, _I]),
  'bop_render_depth_dev': (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P, _P, ctypes.c_size_t, _P]),
  'bop_vsd_workspace_bytes': (ctypes.c_size_t, [_I, _I, _I, _I]),
  'bop_vsd_from_depth_dev': (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _P, _P, _P, _P,
                                   ctypes.c_size_t, _P]),
  'bop_vsd_from_depth_static_dev': (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _P, _P, _P, _P]),
  'bop_vsd_dev': (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P, ctypes.c_size_t,
                       _P]),
  'bop_vsd_split_dev': (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P, ctypes.c_size_t,
                             _P, _P]),
  'bop_vsd_grouped_dev': (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _P, _I, _I, _P, _P, _P, _P,
                               ctypes.c_size_t, _P]),
  'bop_mask_overlap_dev': (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P,
                                ctypes.c_size_t, _P]),
  'bop_mask_stats_workspace_bytes': (ctypes.c_size_t, [_I]),
  'bop_mask_overlap_from_depth_dev': (_I, [_P, _P, _I, _I, _I, _P, _P, _P, _P, _P, ctypes.c_size_t, _P]),
  'bop_gt_info_workspace_bytes': (ctypes.c_size_t, [_I, _I, _I, _I, _I, _I]),
  'bop_gt_info_dev': (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, ctypes.c_double, _I, _I,
                           _P, _P, _P, _P, _P, _P, ctypes.c_size_t, _P]),
  'bop_model_upload': (_I, [_P, _I, _P, _I, _P, _I, ctypes.c_double, _P, _P]),
  'bop_model_free': (_I, [_P, _P]),
  'bop_mssd_host': (_I, [_P, _I, _P, _P, _I, _I, _P]),
  'bop_mspd_host': (_I, [_P, _I, _P, _P, _P, _I, _P]),
  'bop_add_host': (_I, [_P, _I, _P, _P, _I, _I, _P]),
  'bop_adi_host': (_I, [_P, _I, _P, _P, _I, _I, _P]),
  'bop_render_depth_host': (_I, [_P, _P, _P, _I, _I, _I, _P]),
  'bop_vsd_host': (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _P, _P]),
  'bop_vsd_set_tuning': (_I, [_I, _I, _I, _I]),
  'bop_vsd_set_exact': (_I, [_I]),
  'bop_vsd_depth_bytes_fetched': (_I, [_P, _I, _P]),
  'bop_points_set_exact': (_I, [_I]),
  'bop_vsd_debug_triangle_list': (_I, [_P, _I, _I, _I, _I, _I, _I, _P, _I, _P]),
  'bop_profile_enable': (_I, [_I]),
  'bop_profile_reset': (_I, []),
  'bop_profile_sum_ms': (_I, [_I, _I, _P, _P]),
  'bop_png_stage_files_host': (_I, [_P, _I, _P, ctypes.c_size_t, _P, _P, _P, _I, _P]),
  'bop_png_scan_host': (_I, [_P, ctypes.c_size_t, _P, _P]),
  'bop_png_decode_workspace_bytes': (ctypes.c_size_t, [_I, _I, _I]),
  'bop_png_decode_dev': (_I, [_P, _P, _P, _I, _I, _I, _P, _P, _P, ctypes.c_size_t, _P]),
  'bop_png_decode_selftest_host': (_I, [_P, ctypes.c_int32, _I, _I, _P]),
  'bop_write_error_file_host': (_I, [ctypes.c_char_p, ctypes.c_int64, _P, _P, _P, _P, ctypes.c_int64,
                                     _P, _P, _P, _I, _I]),
  'bop_format_double_host': (_I, [ctypes.c_double, ctypes.c_char_p]),
  'bop_parse_results_csv_host': (_I, [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int64, _P, _P, _P, _P, _P, _P]),
}
EXPORTED_SYMBOLS = sorted(_SIGNATURES)

_lib = None


def lib():
  """Returns the loaded library; raises ImportError if it has not been built."""
  global _lib
  if _lib is None:
    if not os.path.exists(SO_PATH):
      raise ImportError(
        'bop_toolkit_b200: %s is missing. Build it with `python -c "import __graft_entry__ as g; '
        'g.build()"` (needs nvcc). There is no CPU fallback.' % SO_PATH)
    l = ctypes.CDLL(SO_PATH)
    stale = 'bop_toolkit_b200: %s is older than these sources (%%s); rebuild it with `python -c "import ' \
      '__graft_entry__ as g; g.build()"`' % SO_PATH
    for name, (res, args) in _SIGNATURES.items():
      try:
        fn = getattr(l, name)
      except AttributeError:
        raise ImportError(stale % ('no symbol %s' % name))
      fn.restype = res
      fn.argtypes = args
    if l.bop_abi_version() != ABI_VERSION:
      raise ImportError(stale % ('ABI %d, expected %d' % (l.bop_abi_version(), ABI_VERSION)))
    _lib = l
  return _lib


class BopCudaError(RuntimeError):
  pass


def check(rc):
  if rc != 0:
    msg = lib().bop_last_error().decode('utf-8', 'replace')
    if msg in ('Unknown pixel matching cost.', 'Unknown visibility mode.'):
      raise ValueError(msg)  # same exception types as the reference (pose_error.py:88)
    raise BopCudaError('libbop_b200 call failed (rc=%d): %s' % (rc, msg))


def require_cuda():
  if lib().bop_device_count() <= 0:
    raise BopCudaError('bop_toolkit_b200 needs a CUDA device (sm_100a); there is no CPU fallback.')
