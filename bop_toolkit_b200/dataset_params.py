"""Dataset layout tables consumed by the error-calculation driver.

Only the fields scripts/eval_calc_errors.py reads are provided: `get_model_params` ->
obj_ids, symmetric_obj_ids, model_tpath, models_info_path (bop_toolkit_lib/dataset_params.py:66-165)
and `get_split_params` -> base_path, split_path, im_size and the scene_camera / scene_gt /
scene_gt_info / depth / mask path templates (:315-646).  Depth/azimuth/elevation ranges, camera
files and the RGB side are not on the pose-error path.
"""
from os.path import join

_R = lambda a, b: list(range(a, b))

# dataset -> (object ids, ids evaluated with ADI under error_type='ad')
_MODELS = {
  'lm': (_R(1, 16), [3, 7, 10, 11]),
  'lmo': ([1, 5, 6, 8, 9, 10, 11, 12], [10, 11]),
  'tless': (_R(1, 31), _R(1, 31)),
  'tudl': (_R(1, 4), []),
  'tyol': (_R(1, 22), [3, 4, 5, 6, 7, 8, 10, 11, 12, 13, 15, 16, 17, 18, 19, 21]),
  'ruapc': (_R(1, 15), [8, 9, 12, 13]),
  'icmi': (_R(1, 7), [1, 2, 6]),
  'icbin': (_R(1, 3), [1]),
  'itodd': (_R(1, 29), [2, 3, 4, 5, 7, 8, 9, 11, 12, 14, 17, 18, 19, 23, 24, 25, 27, 28]),
  'hbs': ([1, 3, 4, 8, 9, 10, 12, 15, 17, 18, 19, 22, 23, 29, 32, 33], [10, 12, 18, 29]),
  'hb': (_R(1, 34), [6, 10, 11, 12, 13, 14, 18, 24, 29]),
  'ycbv': (_R(1, 22), [1, 13, 14, 16, 18, 19, 20, 21]),
  'hope': (_R(1, 29), None),
  'kit': (_R(1, 122), None),
  'bigbird': (_R(1, 111), None),
  'data2_real_source': (_R(1, 201), None),
  'data2_syn_source': (_R(1, 201), None),
  'data3_1_real_source': (_R(1, 201), None),
  'data3_1_syn_source': (_R(1, 201), None),
}

# dataset -> image size (width, height) of the default test sensor
_IM_SIZE = {
  'lm': (640, 480), 'lmo': (640, 480), 'tudl': (640, 480), 'tyol': (640, 480),
  'ruapc': (640, 480), 'icmi': (640, 480), 'icbin': (640, 480), 'itodd': (1280, 960),
  'hb': (640, 480), 'hbs': (640, 480), 'ycbv': (640, 480), 'hope': (1920, 1080),
  'kit': (640, 480), 'bigbird': (640, 480), 'gist': (1920, 1080), 'data2_source': (1920, 1080),
}
_TLESS_SIZE = {
  'train': {'primesense': (400, 400), 'kinect': (400, 400), 'canon': (1900, 1900),
            'render_reconst': (1280, 1024), 'pbr': (720, 540), 'synthetless': (400, 400)},
  'test': {'primesense': (720, 540), 'kinect': (720, 540), 'canon': (2560, 1920)},
}
_HB_SIZE = {'pbr': (640, 480), 'primesense': (640, 480), 'kinect': (1920, 1080)}


def _scene_ids(name, split, split_type):
  """Scene ids per dataset / split (dataset_params.py:336-584); `pbr` splits have 50 scenes."""
  if split_type == 'pbr':
    return list(range(50))
  by_split = lambda d: d[split]
  if name == 'lm':
    return _R(1, 16)
  if name == 'lmo':
    return by_split({'train': [1, 5, 6, 8, 9, 10, 11, 12], 'test': [2]})
  if name == 'tless':
    if split == 'train':
      return [1] if split_type == 'synthetless' else _R(1, 31)
    return _R(1, 21)
  if name == 'tudl':
    return _R(1, 4)
  if name == 'tyol':
    return _R(1, 22)
  if name == 'ruapc':
    return _R(1, 15)
  if name == 'icmi':
    return _R(1, 7)
  if name == 'icbin':
    return by_split({'train': _R(1, 3), 'test': _R(1, 4)})
  if name == 'itodd':
    return by_split({'train': [], 'val': [1], 'test': [1]})
  if name == 'hbs':
    return by_split({'train': [], 'val': [3, 5, 13], 'test': [3, 5, 13]})
  if name == 'hb':
    return by_split({'train': [], 'val': _R(1, 14), 'test': _R(1, 14)})
  if name == 'ycbv':
    if split == 'train':
      return {'real': _R(0, 48) + _R(60, 92), 'synt': _R(0, 80)}[split_type]
    return _R(48, 60)
  if name == 'hope':
    return by_split({'train': [], 'val': _R(1, 11), 'test': _R(1, 41)})
  if name == 'kit':
    return _R(1, 121)
  if name in ('bigbird', 'gist', 'data2_source'):
    return _R(1, 111)
  raise ValueError('Unknown BOP dataset ({}).'.format(name))


def get_model_params(datasets_path, dataset_name, model_type=None):
  """Object-model parameters of a dataset (dataset_params.py:66-165)."""
  if dataset_name not in _MODELS:
    raise KeyError(dataset_name)
  obj_ids, symmetric = _MODELS[dataset_name]
  if dataset_name == 'tless' and model_type is None:
    model_type = 'cad'
  if dataset_name == 'hbs':
    dataset_name = 'hb'
  folder = 'models' if model_type is None else 'models_' + model_type
  models_path = join(datasets_path, dataset_name, folder)
  return {
    'obj_ids': list(obj_ids),
    'symmetric_obj_ids': None if symmetric is None else list(symmetric),
    'model_tpath': join(models_path, 'obj_{obj_id:06d}.ply'),
    'models_info_path': join(models_path, 'models_info.json'),
  }


def get_split_params(datasets_path, dataset_name, split, split_type=None):
  """Split parameters (paths, image size) of a dataset (dataset_params.py:315-646).  As in the
  reference, 'name', 'split_type' and 'base_path' echo the arguments (so 'hbs' keeps its own
  base folder for the targets file) while the split folder uses the resolved names."""
  name_arg, type_arg = dataset_name, split_type
  depth_ext = '.tif' if (dataset_name == 'itodd' and split_type != 'pbr') else '.png'
  if dataset_name == 'tless':
    if split_type is None:
      split_type = 'primesense'
    im_size = _TLESS_SIZE[split][split_type]
  elif dataset_name in ('hb', 'hbs'):
    if split_type is None:
      split_type = 'primesense'
    im_size = _HB_SIZE[split_type]
  elif dataset_name in _IM_SIZE:
    im_size = _IM_SIZE[dataset_name]
    if dataset_name == 'tudl' and split == 'train' and split_type is None:
      split_type = 'render'
    if dataset_name == 'ycbv' and split == 'train' and split_type is None:
      split_type = 'real'
  else:
    raise ValueError('Unknown BOP dataset ({}).'.format(dataset_name))
  ids = _scene_ids(dataset_name, split, split_type)
  if dataset_name == 'hbs':
    dataset_name = 'hb'     # both versions share one directory
  base_path = join(datasets_path, name_arg)
  split_path = join(datasets_path, dataset_name, split)
  if split_type is not None:
    split_path += '_' + split_type
  scene = join(split_path, '{scene_id:06d}')
  return {
    'name': name_arg, 'split': split, 'split_type': type_arg,
    'base_path': base_path, 'split_path': split_path, 'im_size': im_size, 'scene_ids': ids,
    'scene_camera_tpath': join(scene, 'scene_camera.json'),
    'scene_gt_tpath': join(scene, 'scene_gt.json'),
    'scene_gt_info_tpath': join(scene, 'scene_gt_info.json'),
    'depth_tpath': join(scene, 'depth', '{im_id:06d}' + depth_ext),
    'mask_tpath': join(scene, 'mask', '{im_id:06d}_{gt_id:06d}.png'),
    'mask_visib_tpath': join(scene, 'mask_visib', '{im_id:06d}_{gt_id:06d}.png'),
  }
