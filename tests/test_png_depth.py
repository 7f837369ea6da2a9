"""Depth-PNG input path: the GPU inflate + unfilter kernels against OpenCV's libpng
decode (what `inout.load_depth` of the reference does through imageio, inout.py:39-46).

CPU part: the container walk / staging (host code of the product) and the device decoder's own
source compiled for the host (`bop_png_decode_selftest_host`) on PNGs from several encoders and
every deflate block type / PNG row filter.  GPU part: the kernels through `decode_depth_pngs` and
the driver's pipelined chunk loop.  The bar is bit-exact (integer samples)."""
import ctypes
import os
import struct
import zlib

import numpy as np
import pytest

from bop_toolkit_b200 import _lib

H, W = 480, 640


def _chunk(t, d):
  return struct.pack('>I', len(d)) + t + d + struct.pack('>I', zlib.crc32(t + d) & 0xffffffff)


def write_png16(path, im, filters, level=6, strategy=zlib.Z_DEFAULT_STRATEGY, idat=1 << 30,
                interlace=0, bit_depth=16):
  """Minimal PNG writer with a chosen row-filter cycle, zlib level/strategy and IDAT split."""
  h, w = im.shape
  be = im.astype('>u2').view(np.uint8).reshape(h, 2 * w).astype(np.int32)
  rows, prev = [], np.zeros(2 * w, np.int32)
  for r in range(h):
    ft, cur = filters[r % len(filters)], be[r]
    a = np.concatenate([[0, 0], cur[:-2]])
    b, c = prev, np.concatenate([[0, 0], prev[:-2]])
    if ft == 0:
      pred = 0
    elif ft == 1:
      pred = a
    elif ft == 2:
      pred = b
    elif ft == 3:
      pred = (a + b) >> 1
    else:
      p = a + b - c
      pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
      pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, b, c))
    rows.append(bytes([ft]) + ((cur - pred) & 255).astype(np.uint8).tobytes())
    prev = cur
  co = zlib.compressobj(level, zlib.DEFLATED, 15, 8, strategy)
  z = co.compress(b''.join(rows)) + co.flush()
  out = b'\x89PNG\r\n\x1a\n' + _chunk(b'IHDR', struct.pack('>IIBBBBB', w, h, bit_depth, 0, 0, 0, interlace))
  out += _chunk(b'tEXt', b'Software\0test')
  for i in range(0, len(z), idat):
    out += _chunk(b'IDAT', z[i:i + idat])
  out += _chunk(b'IEND', b'')
  with open(path, 'wb') as f:
    f.write(out)


def make_cases(d, seed=0):
  """[(path, HxW uint16)] covering OpenCV's encoder (Sub filter, RLE strategy, 8 KB IDATs), pypng
  style files (filter None, level 6, one IDAT -- how the BOP datasets were written,
  inout.py:49-63), stored / fixed / dynamic blocks, all five filters, long zero runs."""
  import cv2
  rng = np.random.default_rng(seed)
  yy, xx = np.mgrid[:H, :W]
  smooth = (900 + 0.3 * xx + 0.5 * yy + 40 * np.sin(xx / 37.0)).astype(np.uint16)
  noisy = (smooth + rng.normal(0, 2, (H, W))).astype(np.uint16)
  noisy[rng.random((H, W)) < 0.03] = 0
  holes = np.zeros((H, W), np.uint16)
  holes[100:200, 100:300] = noisy[100:200, 100:300]
  rand = rng.integers(0, 65536, (H, W)).astype(np.uint16)
  cases = []
  for name, im in (('smooth', smooth), ('noisy', noisy), ('holes', holes), ('rand', rand)):
    p = os.path.join(d, name + '_cv2.png')
    cv2.imwrite(p, im)
    cases.append((p, im))
    for lvl, strat, filt, idat in ((6, zlib.Z_DEFAULT_STRATEGY, [0], 1 << 30),
                                   (9, zlib.Z_DEFAULT_STRATEGY, [0, 1, 2, 3, 4], 8192),
                                   (1, zlib.Z_FIXED, [4], 1000),
                                   (0, zlib.Z_DEFAULT_STRATEGY, [3], 70000),
                                   (6, zlib.Z_HUFFMAN_ONLY, [2, 1], 1 << 30),
                                   (6, zlib.Z_RLE, [1], 8192)):
      p = os.path.join(d, '%s_l%d_s%d_f%s.png' % (name, lvl, strat, ''.join(map(str, filt))))
      write_png16(p, im, filt, lvl, strat, idat)
      cases.append((p, im))
  return cases


def make_periodic_cases(d):
  """Images whose rows repeat with a period of p pixels: zlib codes them as overlapping matches
  (distance 2p < length 258), the `src[i % dist]` path of the decoder's all-lane match copy --
  below, at and above the 32 lanes of a warp -- plus a period beyond the longest match."""
  rng = np.random.default_rng(9)
  cases = []
  for period in (1, 3, 5, 16, 17, 200):
    base = rng.integers(1, 60000, size=period).astype(np.uint16)
    im = np.tile(base, (H, W // period + 1))[:, :W].copy()
    im[H // 2:] = np.roll(im[H // 2:], 1, axis=1)          # a second phase, so rows are not all equal
    for lvl, filt in ((6, [0]), (9, [1])):
      p = os.path.join(d, 'periodic_%d_l%d_f%d.png' % (period, lvl, filt[0]))
      write_png16(p, im, filt, lvl)
      cases.append((p, im))
  return cases


def _scan(path):
  buf = bytearray(open(path, 'rb').read())
  cb = (ctypes.c_uint8 * len(buf)).from_buffer(buf)
  info, nb = _lib.PngInfo(), ctypes.c_int64()
  assert _lib.lib().bop_png_scan_host(cb, len(buf), ctypes.byref(info), ctypes.byref(nb)) == 0
  return cb, info, nb.value


def _host_twin(path, h, w):
  cb, info, nb = _scan(path)
  assert info.status == 0 and (info.width, info.height, info.bit_depth) == (w, h, 16)
  out = np.empty((h, w), np.uint16)
  rc = _lib.lib().bop_png_decode_selftest_host(cb, nb, h, w, out.ctypes.data)
  return rc, out


def test_decoder_source_on_host_matches_opencv(tmp_path):
  import cv2
  _lib.build()
  cases = make_cases(str(tmp_path))
  assert len(cases) == 28
  for path, im in cases:
    assert np.array_equal(cv2.imread(path, cv2.IMREAD_UNCHANGED), im), path   # the file is right
    rc, out = _host_twin(path, H, W)
    assert rc == 0 and np.array_equal(out, im), path
  odd = np.random.default_rng(1).integers(0, 3000, (37, 53)).astype(np.uint16)
  p = str(tmp_path / 'odd.png')
  cv2.imwrite(p, odd)
  rc, out = _host_twin(p, 37, 53)
  assert rc == 0 and np.array_equal(out, odd)
  for path, im in make_periodic_cases(str(tmp_path)):
    assert np.array_equal(cv2.imread(path, cv2.IMREAD_UNCHANGED), im), path
    rc, out = _host_twin(path, H, W)
    assert rc == 0 and np.array_equal(out, im), path


def test_staging_compacts_idat_and_reports_per_file_status(tmp_path):
  import cv2
  _lib.build()
  L = _lib.lib()
  cases = make_cases(str(tmp_path))[:8]
  rgb = str(tmp_path / 'rgb8.png')
  cv2.imwrite(rgb, np.zeros((H, W, 3), np.uint8))
  inter = str(tmp_path / 'interlaced.png')
  write_png16(inter, cases[0][1], [0], interlace=1)      # header only matters here
  junk = str(tmp_path / 'junk.png')
  open(junk, 'wb').write(b'not a png at all, but long enough to hold a header ....')
  paths = [c[0] for c in cases] + [rgb, inter, junk, str(tmp_path / 'missing.png')]
  n = len(paths)
  arr = (ctypes.c_char_p * n)(*[p.encode() for p in paths])
  offs, nbytes = np.zeros(n, np.int64), np.zeros(n, np.int32)
  infos, need = (_lib.PngInfo * n)(), ctypes.c_size_t()
  rc = L.bop_png_stage_files_host(arr, n, None, 0, offs.ctypes.data, nbytes.ctypes.data, infos, 4,
                                  ctypes.byref(need))
  assert rc == _lib.PNG_STAGING_TOO_SMALL and need.value > 0
  st = np.full(need.value, 0xAB, np.uint8)
  rc = L.bop_png_stage_files_host(arr, n, st.ctypes.data, st.size, offs.ctypes.data,
                                  nbytes.ctypes.data, infos, 4, ctypes.byref(need))
  assert rc == 0
  assert [infos[i].status for i in range(8, n)] == [-104, -104, -102, -101]
  assert (infos[8].bit_depth, infos[8].color_type) == (8, 2) and infos[9].interlace == 1
  for i, (path, im) in enumerate(cases):
    assert infos[i].status == 0 and offs[i] % 16 == 0 and nbytes[i] == infos[i].idat_bytes
    z = bytes(st[offs[i]:offs[i] + nbytes[i]])
    raw = zlib.decompress(z)                    # the compacted IDAT payloads ARE the zlib stream
    assert len(raw) == H * (2 * W + 1)
    assert not st[offs[i] + nbytes[i]:offs[i] + nbytes[i] + 8].any()     # zero tail for the bit reader
    out = np.empty((H, W), np.uint16)
    assert L.bop_png_decode_selftest_host(st.ctypes.data + int(offs[i]), int(nbytes[i]), H, W,
                                          out.ctypes.data) == 0
    assert np.array_equal(out, im)


def test_corrupt_streams_fail_cleanly(tmp_path):
  _lib.build()
  rng = np.random.default_rng(3)
  cases = make_cases(str(tmp_path))
  src = open(cases[2][0], 'rb').read()          # level 9, all filters, 8 KB IDATs
  detected = 0
  for k in range(300):
    b = bytearray(src)
    pos = int(rng.integers(60, len(b) - 20))
    b[pos] ^= 1 << int(rng.integers(0, 8))
    cb = (ctypes.c_uint8 * len(b)).from_buffer(b)
    info, nb = _lib.PngInfo(), ctypes.c_int64()
    _lib.lib().bop_png_scan_host(cb, len(b), ctypes.byref(info), ctypes.byref(nb))
    if info.status:
      detected += 1
      continue
    out = np.empty((H, W), np.uint16)
    rc = _lib.lib().bop_png_decode_selftest_host(cb, nb.value, H, W, out.ctypes.data)
    assert -10 <= rc <= 0
    detected += rc != 0
  assert detected > 30          # most single-bit flips desynchronise the Huffman stream
  # truncated stream
  cb, info, nb = _scan(cases[1][0])
  out = np.empty((H, W), np.uint16)
  assert _lib.lib().bop_png_decode_selftest_host(cb, nb // 2, H, W, out.ctypes.data) < 0


def _flip_stored_payload(src_path, dst_path):
  """A level-0 (stored blocks) file with one payload bit flipped: every block header stays valid,
  only the Adler-32 trailer can tell."""
  b = bytearray(open(src_path, 'rb').read())
  pos = b.index(b'IDAT') + 4 + 2 + 5 + 20000      # zlib header, first stored-block header, into the data
  b[pos] ^= 0x10
  open(dst_path, 'wb').write(bytes(b))


def test_adler32_mismatch_is_reported(tmp_path):
  _lib.build()
  cases = make_cases(str(tmp_path))
  stored = [c for c in cases if '_l0_' in c[0]][0]
  bad = str(tmp_path / 'bad_adler.png')
  _flip_stored_payload(stored[0], bad)
  cb, info, nb = _scan(bad)
  out = np.empty((H, W), np.uint16)
  assert _lib.lib().bop_png_decode_selftest_host(cb, nb, H, W, out.ctypes.data) == -10
  cb, info, nb = _scan(stored[0])
  assert _lib.lib().bop_png_decode_selftest_host(cb, nb, H, W, out.ctypes.data) == 0


# ------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_gpu_decode_matches_opencv(tmp_path):
  import cv2
  import torch
  from bop_toolkit_b200 import batch
  cases = make_cases(str(tmp_path))
  paths = [c[0] for c in cases]
  out = batch.decode_depth_pngs(paths, H, W, 'cuda:0', threads=4)
  assert out.shape == (len(cases), H, W) and out.dtype == torch.int16 and out.is_cuda
  got = out.cpu().numpy().view(np.uint16)
  for i, (path, im) in enumerate(cases):
    assert np.array_equal(got[i], im), path
  # overlapping matches (distance < length), copied by all lanes with src[i % dist]
  per = make_periodic_cases(str(tmp_path))
  got = batch.decode_depth_pngs([c[0] for c in per], H, W, 'cuda:0', threads=4).cpu().numpy().view(np.uint16)
  for i, (path, im) in enumerate(per):
    assert np.array_equal(got[i], im), path
  # many images: more than one wave of warps
  many = paths * 12
  got = batch.decode_depth_pngs(many, H, W, 'cuda:0').cpu().numpy().view(np.uint16)
  for i in range(len(many)):
    assert np.array_equal(got[i], cases[i % len(cases)][1])
  # odd width (16-bit stores) and a wide image (dynamic shared memory above the default)
  rng = np.random.default_rng(5)
  for h, w in ((37, 53), (8, 4001)):
    im = rng.integers(0, 4000, (h, w)).astype(np.uint16)
    p = str(tmp_path / ('s%dx%d.png' % (h, w)))
    cv2.imwrite(p, im)
    q = str(tmp_path / ('p%dx%d.png' % (h, w)))
    write_png16(q, im, [4, 3, 1, 2, 0], 6)
    got = batch.decode_depth_pngs([p, q], h, w, 'cuda:0').cpu().numpy().view(np.uint16)
    assert np.array_equal(got[0], im) and np.array_equal(got[1], im)


@pytest.mark.gpu
def test_gpu_decode_errors(tmp_path):
  import cv2
  from bop_toolkit_b200 import batch
  cases = make_cases(str(tmp_path))[:3]
  with pytest.raises(ValueError):           # wrong size
    batch.decode_depth_pngs([cases[0][0]], H, W + 2, 'cuda:0')
  with pytest.raises(IOError):
    batch.decode_depth_pngs([str(tmp_path / 'missing.png')], H, W, 'cuda:0')
  b = bytearray(open(cases[2][0], 'rb').read())
  cut = str(tmp_path / 'cut.png')
  # keep the container valid but zero the middle of the first IDAT payload
  pos = b.index(b'IDAT') + 4
  b[pos + 3000:pos + 3400] = bytes(400)
  open(cut, 'wb').write(bytes(b))
  with pytest.raises(IOError):
    batch.decode_depth_pngs([cases[0][0], cut], H, W, 'cuda:0')
  stored = [c for c in make_cases(str(tmp_path)) if '_l0_' in c[0]][0]
  bad = str(tmp_path / 'bad_adler.png')
  _flip_stored_payload(stored[0], bad)
  with pytest.raises(IOError, match='status -10'):      # valid structure, wrong Adler-32
    batch.decode_depth_pngs([stored[0], bad], H, W, 'cuda:0')
  # a 16-bit file the device path does not take (RGB): decoded by OpenCV, first channel
  rgb16 = str(tmp_path / 'rgb16.png')
  cv2.imwrite(rgb16, np.dstack([cases[0][1]] * 3))
  got = batch.decode_depth_pngs([rgb16, cases[1][0]], H, W, 'cuda:0').cpu().numpy().view(np.uint16)
  assert np.array_equal(got[0], cases[0][1]) and np.array_equal(got[1], cases[1][1])


@pytest.mark.gpu
def test_driver_vsd_same_errors_with_gpu_and_host_decode(tmp_path, monkeypatch, oracle_built):
  """eval_calc_errors VSD over a synthetic BOP dataset: depth decoded on the GPU (default) and by
  OpenCV on the host (BOP_B200_DEPTH_DECODE=host) give byte-identical error files."""
  from oracle import renderer_oracle
  from bop_toolkit_b200 import eval_calc_errors as ece, synth

  def render(pts, faces, R, t, K):
    return renderer_oracle.render_depth(pts, faces, R, t, K[0, 0], K[1, 1], K[0, 2], K[1, 2],
                                        synth.IM_H, synth.IM_W)
  info = synth.write_bop_dataset(str(tmp_path / 'ds'), render, dataset='lm', scene_ids=(1,),
                                 n_images=5, seed=11)
  outs = {}
  for mode in ('gpu', 'host'):
    monkeypatch.setenv('BOP_B200_DEPTH_DECODE', mode)
    ev = str(tmp_path / ('eval_' + mode))
    ece.calc_errors(ece.parse_args([
      '--error_type=vsd', '--n_top=-1', '--result_filenames=' + info['result_filename'],
      '--results_path=' + info['results_path'], '--eval_path=' + ev,
      '--datasets_path=' + info['datasets_path']]), log=lambda s: None)
    files = {}
    for dirpath, _, names in os.walk(ev):
      for name in names:
        full = os.path.join(dirpath, name)
        files[os.path.relpath(full, ev)] = open(full).read()
    assert len(files) >= 10
    outs[mode] = files
  assert outs['gpu'] == outs['host']
