"""I/O of the pose-error path: the file formats either side of the kernels.

Contracts follow bop_toolkit_lib/inout.py (JSON :66-113, scene files :140-198, results CSV
:221-261, depth :39-46, PLY :401-583) so that files written here are read by the reference's
eval_calc_scores.py unchanged, and vice versa.  Implementations are vectorised numpy, not the
reference's per-element Python loops.
"""
import json
import os
import struct

import numpy as np


# ----------------------------------------------------------------------------- JSON
def load_json(path, keys_to_int=False):
  """JSON file -> object; with keys_to_int, digit-only dict keys become ints (inout.py:66-82)."""
  def conv(d):
    return {int(k) if k.lstrip('-').isdigit() else k: v for k, v in d.items()}
  with open(path, 'r') as f:
    return json.load(f, object_hook=conv) if keys_to_int else json.load(f)


def load_json_scene(path):
  """Per-scene BOP files (scene_camera / scene_gt / scene_gt_info: {im_id: ...}) with the image
  ids as ints.  Same result as load_json(path, keys_to_int=True) -- only the top level of these
  files has digit keys -- without an object hook on each of the ~1e4 inner dicts."""
  with open(path, 'r') as f:
    d = json.load(f)
  return {int(k) if k.lstrip('-').isdigit() else k: v for k, v in d.items()}


def dumps_errors_layout(content):
  """Text layout of inout.save_json (inout.py:85-113): one element per line, keys sorted,
  2-space indent, no trailing newline; inf serialises as `Infinity`."""
  if isinstance(content, dict):
    items = sorted(content.items(), key=lambda kv: kv[0])
    body = ',\n'.join('  "{}": {}'.format(k, json.dumps(v, sort_keys=True)) for k, v in items)
    return '{\n' + body + ('\n' if items else '') + '}'
  if isinstance(content, list):
    body = ',\n'.join('  {}'.format(json.dumps(e, sort_keys=True)) for e in content)
    return '[\n' + body + ('\n' if content else '') + ']'
  return json.dumps(content, sort_keys=True)


_INF = float('inf')


def _fmt_float(x):
  """The text json.dumps writes for a float: float.__repr__, `Infinity` / `-Infinity` / `NaN`."""
  if x != x:
    return 'NaN'
  if x == _INF:
    return 'Infinity'
  if x == -_INF:
    return '-Infinity'
  return float.__repr__(x)


def _fmt_scalar(v):
  t = type(v)
  if t is float:
    return _fmt_float(v)
  if t is int:
    return str(v)
  return json.dumps(v)


def error_record_tails(records):
  """Per record, the text after the `errors` dict in the sorted-key layout; the same for every
  tau file of a VSD run, so it is built once."""
  return ['}, "est_id": ' + _fmt_scalar(r['est_id']) + ', "im_id": ' + _fmt_scalar(r['im_id']) +
          ', "obj_id": ' + _fmt_scalar(r['obj_id']) + ', "score": ' + _fmt_scalar(r['score']) + '}'
          for r in records]


def dumps_error_records(records, tau_id=None, tails=None):
  """Exactly the text of dumps_errors_layout(records) -- or, with tau_id, of the per-tau view
  {'errors': {gt_id: [e[tau_id]]}} the reference deep-copies (eval_calc_errors.py:408-414) --
  built with string formatting instead of one json.dumps per record (the error files of a
  YCB-V-size VSD run hold 1.6 M records; json.dumps made the writer the slowest stage)."""
  if not records:
    return '[\n]'
  if tails is None:
    tails = error_record_tails(records)
  lines = []
  add = lines.append
  for r, tail in zip(records, tails):
    errs = r['errors']
    if len(errs) == 1:
      for g, e in errs.items():
        pass
      body = '"%d": [%s]' % (g, _fmt_float(e[tau_id]) if tau_id is not None else
                             ', '.join(map(_fmt_float, e)))
    else:
      body = ', '.join(
        '"%d": [%s]' % (g, _fmt_float(e[tau_id]) if tau_id is not None else ', '.join(map(_fmt_float, e)))
        for g, e in sorted(errs.items()))
    add('  {"errors": {' + body + tail)
  return '[\n' + ',\n'.join(lines) + '\n]'


def save_json(path, content):
  with open(path, 'w') as f:
    f.write(dumps_errors_layout(content))


# ---------------------------------------------------------------------- scene files
def load_scene_camera(path):
  """scene_camera.json -> {im_id: {'cam_K': 3x3, 'cam_R_w2c': 3x3, 'cam_t_w2c': 3x1, ...}}
  (inout.py:140-160)."""
  sc = load_json_scene(path)
  for cam in sc.values():
    if 'cam_K' in cam:
      cam['cam_K'] = np.array(cam['cam_K'], dtype=np.float64).reshape((3, 3))
    if 'cam_R_w2c' in cam:
      cam['cam_R_w2c'] = np.array(cam['cam_R_w2c'], dtype=np.float64).reshape((3, 3))
    if 'cam_t_w2c' in cam:
      cam['cam_t_w2c'] = np.array(cam['cam_t_w2c'], dtype=np.float64).reshape((3, 1))
  return sc


def load_scene_gt(path):
  """scene_gt.json -> {im_id: [{'obj_id', 'cam_R_m2c': 3x3, 'cam_t_m2c': 3x1}, ...]}
  (inout.py:182-198)."""
  sg = load_json_scene(path)
  for gts in sg.values():
    for gt in gts:
      if 'cam_R_m2c' in gt:
        gt['cam_R_m2c'] = np.array(gt['cam_R_m2c'], dtype=np.float64).reshape((3, 3))
      if 'cam_t_m2c' in gt:
        gt['cam_t_m2c'] = np.array(gt['cam_t_m2c'], dtype=np.float64).reshape((3, 1))
  return sg


def _load_bop_results_py(path):
  """Line-by-line parse exactly as the reference does it (inout.py:221-261)."""
  results = []
  header = 'scene_id,im_id,obj_id,score,R,t,time'
  with open(path, 'r') as f:
    for line_id, line in enumerate(f):
      if line_id == 0 and header in line:
        continue
      elems = line.split(',')
      if len(elems) != 7:
        raise ValueError('A line does not have 7 comma-sep. elements: {}'.format(line))
      results.append({
        'scene_id': int(elems[0]), 'im_id': int(elems[1]), 'obj_id': int(elems[2]),
        'score': float(elems[3]),
        'R': np.array(list(map(float, elems[4].split())), np.float64).reshape((3, 3)),
        't': np.array(list(map(float, elems[5].split())), np.float64).reshape((3, 1)),
        'time': float(elems[6])})
  return results


def load_bop_results(path, version='bop19'):
  """BOP19 results CSV -> list of dicts (inout.py:221-261): header line skipped, each line
  `scene_id,im_id,obj_id,score,R(9 space-sep),t(3 space-sep),time`.
  A plainly formatted file is parsed by the native reader (bop_parse_results_csv_host: the
  per-line numpy calls of the reference cost ~10 us per estimate); 'R' / 't' are then views into
  one array each -- read-only for the callers.  Anything irregular is parsed line by line the
  reference's way, which also raises the reference's errors."""
  if version != 'bop19':
    raise ValueError('Unknown version of BOP results.')
  try:
    from . import _lib
    L = _lib.lib()
  except (ImportError, OSError):
    return _load_bop_results_py(path)
  import ctypes
  with open(path, 'rb') as f:
    raw = f.read()
  cap = raw.count(b'\n') + 1
  ids = np.empty((cap, 3), np.int64)
  score, tm = np.empty(cap, np.float64), np.empty(cap, np.float64)
  R, t = np.empty((cap, 3, 3), np.float64), np.empty((cap, 3, 1), np.float64)
  n = ctypes.c_int64()
  rc = L.bop_parse_results_csv_host(raw, len(raw), cap, ids.ctypes.data, score.ctypes.data,
                                    R.ctypes.data, t.ctypes.data, tm.ctypes.data, ctypes.byref(n))
  if rc != 0:
    return _load_bop_results_py(path)
  n = n.value
  import gc
  was_on = gc.isenabled()
  gc.disable()        # 1e5..1e6 small containers: the generational collector would rescan them
  try:
    return [{'scene_id': a, 'im_id': b, 'obj_id': c, 'score': s, 'R': Ri, 't': ti, 'time': tt}
            for (a, b, c), s, Ri, ti, tt in zip(ids[:n].tolist(), score[:n].tolist(), list(R[:n]),
                                                list(t[:n]), tm[:n].tolist())]
  finally:
    if was_on:
      gc.enable()


def save_bop_results(path, results, version='bop19'):
  """Inverse of load_bop_results (inout.py:264-292)."""
  lines = ['scene_id,im_id,obj_id,score,R,t,time']
  for r in results:
    run_time = r['time'] if 'time' in r else -1
    lines.append('{scene_id},{im_id},{obj_id},{score},{R},{t},{time}'.format(
      scene_id=r['scene_id'], im_id=r['im_id'], obj_id=r['obj_id'], score=r['score'],
      R=' '.join(map(repr, np.asarray(r['R'], np.float64).flatten().tolist())),
      t=' '.join(map(repr, np.asarray(r['t'], np.float64).flatten().tolist())), time=run_time))
  with open(path, 'w') as f:
    f.write('\n'.join(lines))


# ---------------------------------------------------------------------------- depth
def load_depth_raw(path):
  """Depth image file -> HxW array in the file's own sample type (uint16 for BOP PNGs)."""
  import cv2
  d = cv2.imread(path, cv2.IMREAD_UNCHANGED)
  if d is None:
    raise IOError('cannot read depth image: ' + path)
  if d.ndim == 3:
    d = d[:, :, 0]
  return np.ascontiguousarray(d)


def load_depth_u16(path):
  """16-bit depth PNG -> HxW uint16 (raw samples; the kernels apply depth_scale in float32)."""
  return np.ascontiguousarray(load_depth_raw(path).astype(np.uint16))


def load_depth(path):
  """Depth PNG -> HxW float32 (inout.py:39-46)."""
  return load_depth_u16(path).astype(np.float32)


def save_depth(path, im):
  """HxW depth -> 16-bit PNG (inout.py:49-63: rounds to uint16)."""
  import cv2
  cv2.imwrite(path, np.round(im).astype(np.uint16))


# ------------------------------------------------------------------------------ PLY
_PLY_TYPES = {
  'char': 'i1', 'int8': 'i1', 'uchar': 'u1', 'uint8': 'u1', 'short': 'i2', 'int16': 'i2',
  'ushort': 'u2', 'uint16': 'u2', 'int': 'i4', 'int32': 'i4', 'uint': 'u4', 'uint32': 'u4',
  'float': 'f4', 'float32': 'f4', 'double': 'f8', 'float64': 'f8'}


def load_ply(path):
  """PLY (ascii or binary little/big endian, triangles) -> dict with 'pts' (Nx3 float64) and,
  when present, 'normals', 'colors', 'faces' (Fx3 float64, as the reference returns them,
  inout.py:474-476), 'texture_uv'.  Vertex values keep their float32 precision."""
  with open(path, 'rb') as f:
    data = f.read()
  end = data.find(b'end_header')
  if end < 0:
    raise ValueError('not a PLY file: ' + path)
  nl = data.find(b'\n', end)
  header = data[:nl].decode('utf8', 'replace').replace('\r', '').split('\n')
  body = data[nl + 1:]
  fmt, elements, texture_file = 'ascii', [], None
  for line in header:
    tok = line.split()
    if not tok:
      continue
    if line.startswith('comment TextureFile'):
      texture_file = tok[-1]
    elif tok[0] == 'format':
      fmt = tok[1]
    elif tok[0] == 'element':
      elements.append({'name': tok[1], 'count': int(tok[2]), 'props': []})
    elif tok[0] == 'property' and elements:
      if tok[1] == 'list':
        elements[-1]['props'].append(('list', tok[2], tok[3], tok[4]))
      else:
        elements[-1]['props'].append(('scalar', tok[1], tok[2]))
  model = {}
  if texture_file is not None:
    model['texture_file'] = texture_file
  if fmt == 'ascii':
    tokens = body.split()
    pos = 0
    for el in elements:
      n = el['count']
      if el['name'] == 'vertex':
        k = len(el['props'])
        arr = np.array(tokens[pos:pos + n * k], dtype=np.float64).reshape(n, k)
        pos += n * k
        _vertex_fields(model, arr, [p[2] for p in el['props']])
      elif el['name'] == 'face':
        faces, uvf = [], []
        for _ in range(n):
          for p in el['props']:
            cnt = int(tokens[pos]); pos += 1
            vals = tokens[pos:pos + cnt]; pos += cnt
            if p[3] in ('vertex_indices', 'vertex_index'):
              if cnt != 3:
                raise ValueError('Only triangular faces are supported.')
              faces.append(vals)
            elif p[3] == 'texcoord':
              uvf.append(vals)
        model['faces'] = np.array(faces, dtype=np.float64).reshape(n, 3)
        if uvf:
          model['texture_uv_face'] = np.array(uvf, dtype=np.float64)
      else:
        for _ in range(n):
          for p in el['props']:
            if p[0] == 'list':
              cnt = int(tokens[pos]); pos += 1 + cnt
            else:
              pos += 1
  else:
    bo = '<' if 'little' in fmt else '>'
    off = 0
    for el in elements:
      n = el['count']
      if all(p[0] == 'scalar' for p in el['props']):
        dt = np.dtype([(p[2], bo + _PLY_TYPES[p[1]]) for p in el['props']])
        rec = np.frombuffer(body, dtype=dt, count=n, offset=off)
        off += dt.itemsize * n
        if el['name'] == 'vertex':
          arr = np.stack([rec[name].astype(np.float64) for name in dt.names], axis=1)
          _vertex_fields(model, arr, list(dt.names))
      else:
        # Fixed-arity lists (triangles): count byte(s) + 3 indices [+ count + 6 uv].
        fields = []
        for j, p in enumerate(el['props']):
          if p[0] == 'scalar':
            fields.append(('s%d' % j, bo + _PLY_TYPES[p[1]]))
          else:
            arity = 6 if p[3] == 'texcoord' else 3
            fields.append(('c%d' % j, bo + _PLY_TYPES[p[1]]))
            fields.append(('v%d' % j, bo + _PLY_TYPES[p[2]], (arity,)))
        dt = np.dtype(fields)
        rec = np.frombuffer(body, dtype=dt, count=n, offset=off)
        off += dt.itemsize * n
        if el['name'] == 'face':
          for j, p in enumerate(el['props']):
            if p[0] == 'list' and p[3] in ('vertex_indices', 'vertex_index'):
              if n and not np.all(rec['c%d' % j] == 3):
                raise ValueError('Only triangular faces are supported.')
              model['faces'] = rec['v%d' % j].astype(np.float64)
            elif p[0] == 'list' and p[3] == 'texcoord':
              model['texture_uv_face'] = rec['v%d' % j].astype(np.float64)
  return model


def _vertex_fields(model, arr, names):
  idx = {n: i for i, n in enumerate(names)}
  model['pts'] = np.ascontiguousarray(arr[:, [idx['x'], idx['y'], idx['z']]])
  if all(k in idx for k in ('nx', 'ny', 'nz')):
    model['normals'] = arr[:, [idx['nx'], idx['ny'], idx['nz']]]
  if all(k in idx for k in ('red', 'green', 'blue')):
    model['colors'] = arr[:, [idx['red'], idx['green'], idx['blue']]]
  for u, v in (('texture_u', 'texture_v'), ('u', 'v')):
    if u in idx and v in idx:
      model['texture_uv'] = arr[:, [idx[u], idx[v]]]
      break


def save_ply(path, pts, faces=None, binary=True):
  """Minimal PLY writer (float32 vertices, int32 triangle indices) for synthetic datasets."""
  pts = np.asarray(pts, dtype=np.float32)
  faces = np.zeros((0, 3), np.int32) if faces is None else np.asarray(faces, dtype=np.int32)
  hdr = ['ply', 'format %s 1.0' % ('binary_little_endian' if binary else 'ascii'),
         'element vertex %d' % len(pts), 'property float x', 'property float y',
         'property float z', 'element face %d' % len(faces),
         'property list uchar int vertex_indices', 'end_header']
  with open(path, 'wb') as f:
    f.write(('\n'.join(hdr) + '\n').encode('utf8'))
    if binary:
      f.write(pts.astype('<f4').tobytes())
      rec = np.empty(len(faces), dtype=np.dtype([('c', 'u1'), ('v', '<i4', (3,))]))
      rec['c'] = 3
      rec['v'] = faces
      f.write(rec.tobytes())
    else:
      for p in pts:
        f.write(('%r %r %r\n' % (float(p[0]), float(p[1]), float(p[2]))).encode('utf8'))
      for t in faces:
        f.write(('3 %d %d %d\n' % (t[0], t[1], t[2])).encode('utf8'))
