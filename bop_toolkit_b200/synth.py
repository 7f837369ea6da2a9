"""Seeded synthetic inputs of BOP-eval shape: procedural closed meshes, LM-like
intrinsics, random GT poses with perturbed estimates, composite depth images.  Used by the
tests, the benchmark and smoke(); there is no dataset and no network on the GPU box."""
import math

import numpy as np

LM_K = np.array([[572.4114, 0.0, 325.2611], [0.0, 573.57043, 242.04899], [0.0, 0.0, 1.0]])
IM_W, IM_H = 640, 480


def make_mesh(n_lat=50, n_lon=100, diameter=200.0, seed=0, kind='blob'):
  """Closed genus-0 surface: a UV sphere with low-frequency radial noise (and superellipsoid
  squashing), scaled to `diameter` mm, origin at the bbox centre.  Vertices are float32-
  representable (like PLY `float` properties).  N = n_lon*(n_lat-1)+2, F = 2*n_lon*(n_lat-1).
  n_lat=50, n_lon=100 -> 4902 vertices / 9800 faces (BOP eval-model size)."""
  rng = np.random.default_rng(seed)
  th = np.linspace(0.0, math.pi, n_lat + 1)[1:-1]              # polar rings
  ph = np.linspace(0.0, 2.0 * math.pi, n_lon, endpoint=False)
  T, P = np.meshgrid(th, ph, indexing='ij')
  d = np.stack([np.sin(T) * np.cos(P), np.sin(T) * np.sin(P), np.cos(T)], axis=-1).reshape(-1, 3)
  d = np.vstack([[0.0, 0.0, 1.0], d, [0.0, 0.0, -1.0]])
  r = np.ones(d.shape[0])
  if kind == 'blob':
    for _ in range(6):
      k = rng.normal(size=3)
      k /= np.linalg.norm(k)
      r += 0.12 * rng.uniform(0.3, 1.0) * np.cos(rng.integers(1, 4) * np.arccos(np.clip(d @ k, -1, 1))
                                                 + rng.uniform(0, 2 * math.pi))
  scale = rng.uniform(0.6, 1.0, size=3) if kind == 'blob' else np.ones(3)
  pts = d * r[:, None] * scale[None, :]
  # faces
  faces = []
  def vid(i, j):
    return 1 + i * n_lon + (j % n_lon)
  n_rings = n_lat - 1
  for j in range(n_lon):
    faces.append((0, vid(0, j), vid(0, j + 1)))
  for i in range(n_rings - 1):
    for j in range(n_lon):
      a, b, c, e = vid(i, j), vid(i, j + 1), vid(i + 1, j), vid(i + 1, j + 1)
      faces.append((a, c, b))
      faces.append((b, c, e))
  last = 1 + n_rings * n_lon
  for j in range(n_lon):
    faces.append((last, vid(n_rings - 1, j + 1), vid(n_rings - 1, j)))
  faces = np.asarray(faces, dtype=np.int32)
  # centre on bbox and scale to the requested diameter (max pairwise distance ~ bbox diagonal
  # bound; use exact farthest-pair on the hull-ish point set by sampling).
  pts -= 0.5 * (pts.min(0) + pts.max(0))
  diam = _diameter(pts)
  pts *= diameter / diam
  pts = pts.astype(np.float32).astype(np.float64)
  return pts, faces


def _diameter(pts):
  sub = pts[np.linspace(0, len(pts) - 1, min(len(pts), 1500)).astype(int)]
  d2 = ((sub[:, None, :] - sub[None, :, :]) ** 2).sum(-1)
  return math.sqrt(float(d2.max()))


def make_dense_mesh(n_pts_target=100000, diameter=200.0, seed=0):
  n_lat = int(round(math.sqrt(n_pts_target / 2.0)))
  return make_mesh(n_lat, 2 * n_lat, diameter, seed)


def random_rotations(rng, n):
  q = rng.normal(size=(n, 4))
  q /= np.linalg.norm(q, axis=1, keepdims=True)
  w, x, y, z = q.T
  R = np.stack([
    1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
    2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
    2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], axis=1)
  return R.reshape(n, 3, 3)


def axis_angle(axis, ang):
  axis = axis / np.linalg.norm(axis)
  K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
  return np.eye(3) + math.sin(ang) * K + (1 - math.cos(ang)) * (K @ K)


def make_pose_pairs(n, seed=0, K=LM_K, width=IM_W, height=IM_H, z_range=(400.0, 1500.0),
                    max_rot_deg=15.0, t_sigma=(5.0, 5.0, 15.0), outlier_frac=0.05):
  """GT poses uniform on SO(3), t_z in z_range, origin projected into the central 80 % of the
  image; estimates = GT perturbed by <= max_rot_deg and N(0, t_sigma) mm; a fraction of gross
  outliers.  Returns (R_est, t_est, R_gt, t_gt) with shapes [n,3,3] / [n,3]."""
  rng = np.random.default_rng(seed)
  R_gt = random_rotations(rng, n)
  z = rng.uniform(z_range[0], z_range[1], size=n)
  u = rng.uniform(0.1 * width, 0.9 * width, size=n)
  v = rng.uniform(0.1 * height, 0.9 * height, size=n)
  t_gt = np.stack([(u - K[0, 2]) / K[0, 0] * z, (v - K[1, 2]) / K[1, 1] * z, z], axis=1)
  R_est = np.empty_like(R_gt)
  for i in range(n):
    ang = math.radians(rng.uniform(0.0, max_rot_deg))
    R_est[i] = R_gt[i] @ axis_angle(rng.normal(size=3), ang)
  t_est = t_gt + rng.normal(size=(n, 3)) * np.asarray(t_sigma)[None, :]
  n_out = int(round(outlier_frac * n))
  if n_out:
    idx = rng.choice(n, n_out, replace=False)
    R_est[idx] = random_rotations(rng, n_out)
    zo = rng.uniform(z_range[0], z_range[1], size=n_out)
    uo = rng.uniform(0.0, width, size=n_out)
    vo = rng.uniform(0.0, height, size=n_out)
    t_est[idx] = np.stack([(uo - K[0, 2]) / K[0, 0] * zo, (vo - K[1, 2]) / K[1, 1] * zo, zo], 1)
  return R_est, t_est, R_gt, t_gt


def degrade_depth(depth_gt_render, seed=0, bg_range=(1200.0, 2000.0), noise_mm=2.0,
                  missing_frac=0.03, occluder_prob=0.3, K=LM_K):
  """Turns a clean GT render [H,W] into a plausible sensor image: tilted background plane,
  Gaussian noise, zero 'missing-depth' pixels, optional occluder box in front of the object."""
  rng = np.random.default_rng(seed)
  H, W = depth_gt_render.shape
  ys, xs = np.mgrid[0:H, 0:W]
  z0 = rng.uniform(*bg_range)
  bg = z0 + rng.uniform(-0.3, 0.3) * (xs - W / 2) + rng.uniform(-0.3, 0.3) * (ys - H / 2)
  d = np.where(depth_gt_render > 0, depth_gt_render, bg).astype(np.float64)
  if rng.uniform() < occluder_prob and (depth_gt_render > 0).any():
    rr, cc = np.nonzero(depth_gt_render > 0)
    r0, r1, c0, c1 = rr.min(), rr.max(), cc.min(), cc.max()
    h, w = max(2, (r1 - r0) // 2), max(2, (c1 - c0) // 2)
    ro = rng.integers(r0, max(r0 + 1, r1 - h + 1))
    co = rng.integers(c0, max(c0 + 1, c1 - w + 1))
    zocc = 0.6 * depth_gt_render[depth_gt_render > 0].min()
    d[ro:ro + h, co:co + w] = zocc
  d += rng.normal(scale=noise_mm, size=d.shape)
  d[rng.uniform(size=d.shape) < missing_frac] = 0.0
  return np.maximum(d, 0.0).astype(np.float32)


def discrete_z_symmetries(n):
  """n-fold rotational symmetry about z as models_info 'symmetries_discrete' 4x4 lists
  (identity excluded; the toolkit adds it)."""
  out = []
  for i in range(1, n):
    a = 2.0 * math.pi * i / n
    c, s = math.cos(a), math.sin(a)
    out.append([c, -s, 0, 0, s, c, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1])
  return out


def write_bop_dataset(root, render_fn, dataset='lmo', split='test', scene_ids=(2, 3), n_images=3,
                      obj_ids=(1, 5, 6, 8), objs_per_image=3, ests_per_target=3, seed=0,
                      mesh_res=(12, 24), K=LM_K, width=IM_W, height=IM_H, method='synth',
                      depth_scales=(0.1, 1.0)):
  """Writes a small synthetic dataset in BOP layout under `root` (docs/bop_datasets_format.md):
    <root>/datasets/<dataset>/models_eval/{obj_XXXXXX.ply, models_info.json}
    <root>/datasets/<dataset>/<split>/<scene:06d>/{scene_camera.json, scene_gt.json, depth/*.png}
    <root>/datasets/<dataset>/test_targets_bop19.json
    <root>/results/<method>_<dataset>-<split>.csv
  render_fn(pts, faces, R, t, K) -> HxW float32 depth [mm] (the caller chooses the renderer).
  Exercises: several GT instances of one object in an image, targets without estimates,
  estimates for non-target objects, tied scores, gross outliers, depth_scale != 1.
  Returns a dict with the paths and the result file name."""
  import json
  import os
  import cv2
  from . import inout
  rng = np.random.default_rng(seed)
  ds = os.path.join(root, 'datasets', dataset)
  os.makedirs(os.path.join(ds, 'models_eval'), exist_ok=True)
  os.makedirs(os.path.join(ds, 'models'), exist_ok=True)     # calc_gt_info / calc_gt_masks use these
  os.makedirs(os.path.join(root, 'results'), exist_ok=True)
  models, info = {}, {}
  for k, oid in enumerate(obj_ids):
    diameter = 90.0 + 40.0 * k
    pts, faces = make_mesh(mesh_res[0], mesh_res[1], diameter, seed=seed * 100 + oid)
    models[oid] = (pts, faces)
    inout.save_ply(os.path.join(ds, 'models_eval', 'obj_%06d.ply' % oid), pts, faces)
    inout.save_ply(os.path.join(ds, 'models', 'obj_%06d.ply' % oid), pts, faces)
    mi = {'diameter': diameter, 'min_x': float(pts[:, 0].min()), 'min_y': float(pts[:, 1].min()),
          'min_z': float(pts[:, 2].min()), 'size_x': float(np.ptp(pts[:, 0])),
          'size_y': float(np.ptp(pts[:, 1])), 'size_z': float(np.ptp(pts[:, 2]))}
    if k % 4 == 1:
      mi['symmetries_discrete'] = discrete_z_symmetries(4)
    if k % 4 == 3:
      mi['symmetries_continuous'] = [{'axis': [0, 0, 1], 'offset': [0, 0, 0]}]
    info[str(oid)] = mi
  for folder in ('models_eval', 'models'):
    with open(os.path.join(ds, folder, 'models_info.json'), 'w') as f:
      json.dump(info, f)
  targets, lines = [], ['scene_id,im_id,obj_id,score,R,t,time']
  from concurrent.futures import ThreadPoolExecutor
  pool, writes = ThreadPoolExecutor(max_workers=8), []
  fmt = lambda a: ' '.join(repr(float(x)) for x in np.asarray(a).reshape(-1))
  for si, scene_id in enumerate(scene_ids):
    sdir = os.path.join(ds, split, '%06d' % scene_id)
    os.makedirs(os.path.join(sdir, 'depth'), exist_ok=True)
    scale = float(depth_scales[si % len(depth_scales)])
    cam, gt = {}, {}
    for im_id in range(n_images):
      Kim = K.copy()
      Kim[0, 2] += float(rng.uniform(-3, 3))     # per-image intrinsics differ slightly
      cam[str(im_id)] = {'cam_K': [float(x) for x in Kim.reshape(-1)], 'depth_scale': scale}
      chosen = list(rng.choice(obj_ids, size=min(objs_per_image, len(obj_ids)), replace=False))
      chosen.append(chosen[0])                    # a second instance of the first object
      _, _, R_g, t_g = make_pose_pairs(len(chosen), seed=int(rng.integers(1 << 30)), K=Kim,
                                       width=width, height=height, z_range=(500.0, 1100.0))
      gts, comp = [], np.zeros((height, width), np.float32)
      for j, oid in enumerate(chosen):
        gts.append({'cam_R_m2c': [float(x) for x in R_g[j].reshape(-1)],
                    'cam_t_m2c': [float(x) for x in t_g[j]], 'obj_id': int(oid)})
        d = render_fn(models[oid][0], models[oid][1], R_g[j], t_g[j], Kim)
        comp = np.where((d > 0) & ((comp == 0) | (d < comp)), d, comp)
      gt[str(im_id)] = gts
      depth = degrade_depth(comp, seed=int(rng.integers(1 << 30)), K=Kim)
      # PNG encoding runs on worker threads (cv2 releases the GIL); the files do not depend on it
      writes.append(pool.submit(cv2.imwrite, os.path.join(sdir, 'depth', '%06d.png' % im_id),
                                np.clip(np.round(depth / scale), 0, 65535).astype(np.uint16)))
      present = sorted(set(int(o) for o in chosen))
      for oid in present:
        targets.append({'im_id': im_id, 'inst_count': chosen.count(oid), 'obj_id': oid,
                        'scene_id': int(scene_id)})
      absent = [o for o in obj_ids if o not in present]
      if absent and im_id == 0:                   # a target without any GT instance
        targets.append({'im_id': im_id, 'inst_count': 1, 'obj_id': int(absent[0]),
                        'scene_id': int(scene_id)})
      for j, oid in enumerate(chosen):
        if im_id == 1 and j == 1:
          continue                                # a target without estimates
        for e in range(ests_per_target):
          ang = math.radians(rng.uniform(0.0, 12.0))
          R_e = R_g[j] @ axis_angle(rng.normal(size=3), ang)
          t_e = t_g[j] + rng.normal(size=3) * np.array([4.0, 4.0, 12.0])
          if e == 2 and j == 0:                   # gross outlier (fails the sphere gates)
            t_e = t_g[j] + np.array([300.0, -200.0, 400.0])
          score = 0.5 if e < 2 and j == 2 else float(np.round(rng.uniform(), 3))   # tied scores
          lines.append('%d,%d,%d,%r,%s,%s,%r' % (scene_id, im_id, oid, score, fmt(R_e), fmt(t_e), 0.25))
      lines.append('%d,%d,%d,%r,%s,%s,%r' % (scene_id, im_id, 99, 0.9, fmt(np.eye(3)),
                                             fmt([0, 0, 800.0]), 0.25))   # non-target object
    for w in writes:
      if not w.result():
        raise IOError('cannot write a depth image under ' + sdir)
    del writes[:]
    with open(os.path.join(sdir, 'scene_camera.json'), 'w') as f:
      json.dump(cam, f)
    with open(os.path.join(sdir, 'scene_gt.json'), 'w') as f:
      json.dump(gt, f)
  pool.shutdown(wait=True)
  with open(os.path.join(ds, 'test_targets_bop19.json'), 'w') as f:
    json.dump(targets, f)
  result_filename = '%s_%s-%s.csv' % (method, dataset, split)
  with open(os.path.join(root, 'results', result_filename), 'w') as f:
    f.write('\n'.join(lines))
  return {'datasets_path': os.path.join(root, 'datasets'), 'results_path': os.path.join(root, 'results'),
          'eval_path': os.path.join(root, 'eval'), 'result_filename': result_filename,
          'obj_ids': list(obj_ids), 'scene_ids': list(scene_ids)}
