"""Oracle renderer -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Implements the reference's Renderer interface (bop_toolkit_lib/renderer.py:11-74) on top of
oracle/raster_oracle.c so that the reference's own pose_error.vsd -- or oracle.pose_error_oracle.vsd
-- runs unmodified on top of it.  Rendered pixels: parity UNPINNED (no runnable reference renderer).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
  so = os.path.join(_HERE, 'liboracle.so')
  srcs = [os.path.join(_HERE, f) for f in ('raster_oracle.c', 'raycast_oracle.c')]
  if force or not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(f) for f in srcs):
    subprocess.check_call(['make', '-C', _HERE, '-s', 'liboracle.so'])
  return so


def _lib():
  global _LIB
  if _LIB is None:
    so = os.path.join(_HERE, 'liboracle.so')
    if not os.path.exists(so):
      build()
    _LIB = ctypes.CDLL(so)
    _LIB.oracle_render_depth.restype = ctypes.c_int
    _LIB.oracle_render_depth.argtypes = [
      ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
      ctypes.c_void_p, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_double,
      ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    _LIB.oracle_render_depth_sub.restype = ctypes.c_int
    _LIB.oracle_render_depth_sub.argtypes = [
      ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
      ctypes.c_void_p, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_double,
      ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    _LIB.raycast_render_depth.restype = ctypes.c_int
    _LIB.raycast_render_depth.argtypes = _LIB.oracle_render_depth.argtypes
  return _LIB


def raycast_depth(pts, faces, R, t, fx, fy, cx, cy, height, width):
  """The second restatement (oracle/raycast_oracle.c): per-pixel ray / triangle intersection in
  camera space.  Exact window positions (no sub-pixel snapping)."""
  pts = np.ascontiguousarray(pts, dtype=np.float64)
  faces = np.ascontiguousarray(faces, dtype=np.int32)
  R = np.ascontiguousarray(R, dtype=np.float64).reshape(9)
  t = np.ascontiguousarray(t, dtype=np.float64).reshape(3)
  depth = np.empty((height, width), dtype=np.float32)
  rc = _lib().raycast_render_depth(
    pts.ctypes.data, pts.shape[0], faces.ctypes.data, faces.shape[0], R.ctypes.data,
    t.ctypes.data, float(fx), float(fy), float(cx), float(cy), int(height), int(width),
    depth.ctypes.data)
  if rc != 0:
    raise MemoryError('raycast_render_depth failed')
  return depth


def render_depth(pts, faces, R, t, fx, fy, cx, cy, height, width, subpixel_bits=0):
  """subpixel_bits=0: exact window positions; 8: GL-style snapping to 1/256 px (what the CUDA
  rasteriser and OpenGL implementations do)."""
  pts = np.ascontiguousarray(pts, dtype=np.float64)
  faces = np.ascontiguousarray(faces, dtype=np.int32)
  R = np.ascontiguousarray(R, dtype=np.float64).reshape(9)
  t = np.ascontiguousarray(t, dtype=np.float64).reshape(3)
  depth = np.empty((height, width), dtype=np.float32)
  rc = _lib().oracle_render_depth_sub(
    pts.ctypes.data, pts.shape[0], faces.ctypes.data, faces.shape[0], R.ctypes.data,
    t.ctypes.data, float(fx), float(fy), float(cx), float(cy), int(height), int(width),
    int(subpixel_bits), depth.ctypes.data)
  if rc != 0:
    raise MemoryError('oracle_render_depth failed')
  return depth


class RendererOracle(object):
  """Renderer interface of bop_toolkit_lib/renderer.py:11-74 ('depth' mode only)."""

  def __init__(self, width, height, mode='depth', shading='phong', bg_color=(0, 0, 0, 0),
               subpixel_bits=0):
    self.width = width
    self.height = height
    self.models = {}
    self.subpixel_bits = subpixel_bits

  def add_object(self, obj_id, model_path=None, **kwargs):
    """`model_path` may be a path to a PLY (read with the package's loader) or kwargs
    pts=..., faces=... for in-memory synthetic meshes."""
    if 'pts' in kwargs:
      self.models[obj_id] = (np.asarray(kwargs['pts'], np.float64),
                             np.asarray(kwargs['faces'], np.int32))
    else:
      from bop_toolkit_b200 import inout
      m = inout.load_ply(model_path)
      self.models[obj_id] = (m['pts'], m['faces'].astype(np.int32))

  def remove_object(self, obj_id):
    del self.models[obj_id]

  def render_object(self, obj_id, R, t, fx, fy, cx, cy):
    pts, faces = self.models[obj_id]
    return {'depth': render_depth(pts, faces, R, t, fx, fy, cx, cy, self.height, self.width,
                                  self.subpixel_bits)}
