// Depth-PNG input path on the device: 16-bit greyscale PNG files -> uint16 depth
// images in HBM, without libpng/zlib on the host cores.
//
// Replaces `inout.load_depth` (/root/reference bop_toolkit_lib/inout.py:39-46) as called once per
// image by scripts/eval_calc_errors.py:236-243.  Three stages:
//   1. bop_png_stage_files_host  host threads read the files straight into one pinned staging
//      buffer and compact each file's IDAT payloads in place (no decompression on the host);
//   2. png_inflate_kernel        one warp per image inflates the zlib stream (png_inflate.cuh):
//      the 32 lanes decode in lockstep on identical state with the tables in shared memory, lane 0
//      stores literals and every match is copied by all lanes;
//   3. png_unfilter_kernel       one warp per image undoes the PNG row filters (None / Sub / Up /
//      Average / Paeth, PNG spec 9.2) row by row in shared memory and writes little-endian uint16
//      samples, coalesced.
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <thread>
#include <vector>

#include "common.cuh"
#include "png_inflate.cuh"

using namespace bop_png;

namespace {

constexpr int kInflateWarps = 4;     // images per CTA of the inflate kernel
constexpr int kUnfilterWarps = 4;

inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
inline size_t image_stride(int height, int width) {       // inflated bytes of one image, padded
  return round_up((size_t)height * (2 * (size_t)width + 1) + 64, 256);
}

// Loads words [w0, w0 + count) of the stream into the ring (16-byte loads, zeros past the end).
__device__ __forceinline__ void ring_load(uint32_t* ring, const uint32_t* __restrict__ src,
                                          uint32_t nwords, uint32_t w0, uint32_t count, int lane) {
  for (uint32_t w = w0 + 4u * lane; w < w0 + count; w += 128u) {
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (w + 3u < nwords) {
      v = *reinterpret_cast<const uint4*>(src + w);
    } else {
      if (w < nwords) v.x = src[w];
      if (w + 1u < nwords) v.y = src[w + 1];
      if (w + 2u < nwords) v.z = src[w + 2];
    }
    *reinterpret_cast<uint4*>(ring + (w & (kRingWords - 1u))) = v;
  }
}

__global__ void __launch_bounds__(kInflateWarps * 32)
png_inflate_kernel(const uint8_t* __restrict__ comp, const int64_t* __restrict__ offsets,
                   const int32_t* __restrict__ comp_bytes, int n, uint8_t* raw, size_t stride,
                   uint32_t expect, int32_t* __restrict__ status,
                   uint32_t* __restrict__ adler_out) {
  __shared__ Tables tabs[kInflateWarps];
  __shared__ __align__(16) uint32_t rings[kInflateWarps][kRingWords];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int img = blockIdx.x * kInflateWarps + warp;
  if (img >= n) return;
  uint8_t* out = raw + (size_t)img * stride;
  const int32_t nb = comp_bytes[img];
  const uint32_t nwords = nb > 0 ? (uint32_t)((nb + 3) >> 2) : 0u;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(comp + offsets[img]);
  uint32_t* ring = rings[warp];
  ring_load(ring, src, nwords, 0u, kRingWords, lane);
  uint32_t ring_hi = kRingWords;           // words [ring_hi - kRingWords, ring_hi) are in the ring
  __syncwarp();
  // Warp-replicated decode (png_inflate.cuh): every lane holds the same State and takes the same
  // branches; lane 0 stores literals, all lanes copy matches.
  State s;
  int ev = nb < 6 ? (int)EV_TRUNCATED
                  : inflate_begin(s, ring, nwords, kRingWords - 1u,
                                  ring_hi >= nwords + 8u ? 0xffffffffu : ring_hi - kRingMargin, out, expect);
  if (ev == 0) {
    for (;;) {
      ev = inflate_run(s, tabs[warp], lane, 32);
      if (ev != EV_INPUT) break;
      // the reader is within kRingMargin words of the ring's end: everything more than that
      // behind it is consumed, so the oldest kRingRefill words can be replaced
      __syncwarp();
      ring_load(ring, src, nwords, ring_hi, kRingRefill, lane);
      ring_hi += kRingRefill;
      __syncwarp();
      s.br.limit = ring_hi >= nwords + 8u ? 0xffffffffu : ring_hi - kRingMargin;
    }
  }
  if (lane == 0) {
    if (ev == EV_DONE && s.o != expect) ev = EV_SHORT;
    status[img] = ev;
    adler_out[img] = ev == EV_DONE ? s.adler : 0u;
  }
}

__device__ __forceinline__ int paeth(int a, int b, int c) {
  const int p = a + b - c;
  const int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// Shared memory: per warp two rows of 2*width bytes (current and previous), 16-byte aligned.
__global__ void __launch_bounds__(kUnfilterWarps * 32)
png_unfilter_kernel(const uint8_t* __restrict__ raw, size_t stride, int n, int height, int width,
                    uint16_t* __restrict__ depth, int32_t* __restrict__ status,
                    const uint32_t* __restrict__ adler_expect) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int img = blockIdx.x * kUnfilterWarps + warp;
  if (img >= n) return;
  if (status[img] != 0) return;
  const int rowb = 2 * width;
  const int rowp = (rowb + 15) & ~15;
  uint8_t* cur = smem + (size_t)warp * 2 * rowp;
  uint8_t* prev = cur + rowp;
  for (int i = lane; i < rowp; i += 32) prev[i] = 0;
  const uint8_t* src = raw + (size_t)img * stride;
  uint16_t* dst = depth + (size_t)img * height * width;
  const int seg = (width + 31) >> 5;      // pixels per lane for the Sub scan
  int bad = 0;
  // Adler-32 (RFC 1950) of the inflated bytes, a row at a time: for a block of L bytes
  // A' = A + sum b_i, B' = B + L A + sum (L - i) b_i  (mod 65521); what libpng / zlib verify on the host.
  uint32_t adA = 1u, adB = 0u;
  const uint32_t L = (uint32_t)rowb + 1u;
  for (int r = 0; r < height; ++r, src += rowb + 1, dst += width) {
    const int ft = src[0];
    uint32_t sa = 0u, sw = 0u;
    if (lane == 0) { sa = (uint32_t)ft; sw = L * (uint32_t)ft; }
    for (int i = lane; i < rowb; i += 32) {
      const uint32_t b = src[1 + i];
      cur[i] = (uint8_t)b;
      sa += b;
      sw += (uint32_t)(rowb - i) * b;
    }
    sa = __reduce_add_sync(BOP_FULL_MASK, sa);
    sw = __reduce_add_sync(BOP_FULL_MASK, sw % 65521u);
    adB = (uint32_t)(((unsigned long long)adB + (unsigned long long)(L % 65521u) * adA + sw) % 65521ull);
    adA = (adA + sa) % 65521u;
    __syncwarp();
    if (ft == 1) {
      // Sub: per byte lane a prefix sum mod 256 along the row -> segment sums, warp scan, replay
      const int p0 = min(lane * seg, width), p1 = min(p0 + seg, width);
      uint32_t s0 = 0, s1 = 0;
      for (int p = p0; p < p1; ++p) { s0 += cur[2 * p]; s1 += cur[2 * p + 1]; }
      uint32_t packed = (s0 & 255u) | ((s1 & 255u) << 16);
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(BOP_FULL_MASK, packed, d);
        if (lane >= d) packed += up;
      }
      uint32_t excl = __shfl_up_sync(BOP_FULL_MASK, packed, 1);
      if (lane == 0) excl = 0;
      s0 = excl & 255u;
      s1 = (excl >> 16) & 255u;
      for (int p = p0; p < p1; ++p) {
        s0 = (s0 + cur[2 * p]) & 255u;
        s1 = (s1 + cur[2 * p + 1]) & 255u;
        cur[2 * p] = (uint8_t)s0;
        cur[2 * p + 1] = (uint8_t)s1;
      }
    } else if (ft == 2) {
      for (int i = lane; i < rowb; i += 32) cur[i] = (uint8_t)(cur[i] + prev[i]);
    } else if (ft == 3) {
      if (lane < 2) {
        int a = 0;
        for (int i = lane; i < rowb; i += 2) {
          a = (cur[i] + ((a + prev[i]) >> 1)) & 255;
          cur[i] = (uint8_t)a;
        }
      }
    } else if (ft == 4) {
      if (lane < 2) {
        int a = 0, c = 0;
        for (int i = lane; i < rowb; i += 2) {
          const int b = prev[i];
          a = (cur[i] + paeth(a, b, c)) & 255;
          cur[i] = (uint8_t)a;
          c = b;
        }
      }
    } else if (ft != 0) {
      bad = 1;
    }
    __syncwarp();
    // big-endian samples -> uint16, two pixels per lane where the row allows 32-bit stores
    if ((width & 1) == 0) {
      const uint32_t* c32 = reinterpret_cast<const uint32_t*>(cur);
      uint32_t* d32 = reinterpret_cast<uint32_t*>(dst);
      for (int i = lane; i < (width >> 1); i += 32) d32[i] = __byte_perm(c32[i], 0, 0x2301);
    } else {
      for (int p = lane; p < width; p += 32) dst[p] = (uint16_t)((cur[2 * p] << 8) | cur[2 * p + 1]);
    }
    __syncwarp();
    uint8_t* tmp = cur; cur = prev; prev = tmp;
  }
  if (lane == 0) {
    if (bad) status[img] = EV_BAD_FILTER;
    else if (((adB << 16) | adA) != adler_expect[img]) status[img] = EV_BAD_CHECKSUM;
  }
}

// ---- host side: PNG container ---------------------------------------------------------------
inline uint32_t be32(const uint8_t* p) {
  return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3];
}

// Walks the chunks of one PNG file held in `buf`; moves the IDAT payloads to the front of `buf`
// (dest <= src always) and fills `info`.  Returns the compacted byte count or a negative status.
int64_t compact_idat(uint8_t* buf, size_t nbytes, bop_png_info* info) {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  memset(info, 0, sizeof(*info));
  if (nbytes < 8 + 25 || memcmp(buf, sig, 8) != 0) return BOP_PNG_NOT_PNG;
  size_t pos = 8, w = 0;
  bool have_hdr = false, end = false;
  while (!end && pos + 12 <= nbytes) {
    const uint32_t len = be32(buf + pos);
    const uint8_t* type = buf + pos + 4;
    if ((size_t)len > nbytes - pos - 12) return BOP_PNG_TRUNCATED;
    const uint8_t* data = buf + pos + 8;
    if (!memcmp(type, "IHDR", 4)) {
      if (len != 13) return BOP_PNG_NOT_PNG;
      info->width = (int32_t)be32(data);
      info->height = (int32_t)be32(data + 4);
      info->bit_depth = data[8];
      info->color_type = data[9];
      info->interlace = data[12];
      have_hdr = true;
      if (data[10] != 0 || data[11] != 0) return BOP_PNG_UNSUPPORTED;
    } else if (!memcmp(type, "IDAT", 4)) {
      if (!have_hdr) return BOP_PNG_NOT_PNG;
      // the header (and possibly earlier IDAT data) is read before it can be overwritten: w <= pos
      memmove(buf + w, data, len);
      w += len;
    } else if (!memcmp(type, "IEND", 4)) {
      end = true;
    }
    pos += 12 + (size_t)len;
  }
  if (!have_hdr || w == 0) return BOP_PNG_TRUNCATED;
  info->idat_bytes = (int64_t)w;
  if (info->bit_depth != 16 || info->color_type != 0 || info->interlace != 0) return BOP_PNG_UNSUPPORTED;
  return (int64_t)w;
}

bool read_whole(const char* path, uint8_t* dst, size_t nbytes) {
  const int fd = open(path, O_RDONLY);
  if (fd < 0) return false;
  size_t got = 0;
  while (got < nbytes) {
    const ssize_t k = read(fd, dst + got, nbytes - got);
    if (k <= 0) break;
    got += (size_t)k;
  }
  close(fd);
  return got == nbytes;
}

template <class F>
void parallel_for(int n, int threads, F f) {
  threads = threads < 1 ? 1 : (threads > n ? (n > 0 ? n : 1) : threads);
  if (threads == 1) { for (int i = 0; i < n; ++i) f(i); return; }
  std::atomic<int> next{0};
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&] { for (int i; (i = next.fetch_add(1)) < n;) f(i); });
  for (auto& th : pool) th.join();
}

}  // namespace

extern "C" int bop_png_stage_files_host(const char* const* paths, int n, void* staging,
                                        size_t staging_bytes, int64_t* offsets, int32_t* comp_bytes,
                                        bop_png_info* infos, int threads, size_t* needed_bytes) {
  BOP_REQUIRE(n >= 0 && (n == 0 || (paths && offsets && comp_bytes && infos)) && needed_bytes,
              "bop_png_stage_files_host: bad arguments");
  std::vector<int64_t> sizes((size_t)n, -1);
  parallel_for(n, threads, [&](int i) {
    struct stat st;
    if (stat(paths[i], &st) == 0 && S_ISREG(st.st_mode)) sizes[i] = (int64_t)st.st_size;
  });
  size_t total = 0;
  for (int i = 0; i < n; ++i) {
    offsets[i] = (int64_t)total;
    comp_bytes[i] = 0;
    memset(&infos[i], 0, sizeof(bop_png_info));
    if (sizes[i] >= 0) total += round_up((size_t)sizes[i] + 8, 16);   // zero tail for the bit reader
    if (sizes[i] > 0x7fffff00ll) sizes[i] = -1;
  }
  *needed_bytes = total;
  if (total > staging_bytes || (total > 0 && !staging)) return BOP_PNG_STAGING_TOO_SMALL;
  uint8_t* base = static_cast<uint8_t*>(staging);
  parallel_for(n, threads, [&](int i) {
    bop_png_info& info = infos[i];
    if (sizes[i] < 0) { info.status = BOP_PNG_IO_ERROR; return; }
    uint8_t* dst = base + offsets[i];
    const size_t nb = (size_t)sizes[i];
    if (!read_whole(paths[i], dst, nb)) { info.status = BOP_PNG_IO_ERROR; return; }
    const int64_t w = compact_idat(dst, nb, &info);
    if (w < 0) { info.status = (int32_t)w; return; }
    memset(dst + w, 0, round_up(nb + 8, 16) - (size_t)w);
    comp_bytes[i] = (int32_t)w;
  });
  return 0;
}

extern "C" int bop_png_scan_host(void* file_bytes, size_t nbytes, bop_png_info* info,
                                 int64_t* idat_bytes) {
  BOP_REQUIRE(file_bytes && info && idat_bytes, "bop_png_scan_host: bad arguments");
  const int64_t w = compact_idat(static_cast<uint8_t*>(file_bytes), nbytes, info);
  info->status = w < 0 ? (int32_t)w : 0;
  *idat_bytes = w < 0 ? 0 : w;
  return 0;
}

extern "C" size_t bop_png_decode_workspace_bytes(int n, int height, int width) {
  if (n <= 0 || height <= 0 || width <= 0) return 256;
  return (size_t)n * image_stride(height, width) + 256 + round_up((size_t)n * 4, 256);
}

extern "C" int bop_png_decode_dev(const void* comp, const int64_t* offsets, const int32_t* comp_bytes,
                                  int n, int height, int width, uint16_t* depth_out,
                                  int32_t* status, void* workspace, size_t workspace_bytes,
                                  void* stream) {
  if (n == 0) return 0;
  BOP_REQUIRE(n > 0 && height > 0 && width > 0 && comp && offsets && comp_bytes && depth_out &&
              status && workspace, "bop_png_decode_dev: bad arguments");
  BOP_REQUIRE((size_t)height * (2 * (size_t)width + 1) < 0x7fffffffull, "bop_png_decode_dev: image too large");
  BOP_REQUIRE(workspace_bytes >= bop_png_decode_workspace_bytes(n, height, width),
              "bop_png_decode_dev: workspace too small (bop_png_decode_workspace_bytes)");
  BOP_REQUIRE(((uintptr_t)comp & 15) == 0, "bop_png_decode_dev: comp must be 16-byte aligned");
  const int rowp = (2 * width + 15) & ~15;
  const size_t smem = (size_t)kUnfilterWarps * 2 * rowp;
  BOP_REQUIRE(smem <= 200 * 1024, "bop_png_decode_dev: image too wide");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint8_t* raw = reinterpret_cast<uint8_t*>(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  const size_t stride = image_stride(height, width);
  uint32_t* adler = reinterpret_cast<uint32_t*>(raw + (size_t)n * stride);
  const uint32_t expect = (uint32_t)((size_t)height * (2 * (size_t)width + 1));
  bop_prof_begin(BOP_PROF_PNG, s);
  png_inflate_kernel<<<(n + kInflateWarps - 1) / kInflateWarps, kInflateWarps * 32, 0, s>>>(
    static_cast<const uint8_t*>(comp), offsets, comp_bytes, n, raw, stride, expect, status, adler);
  BOP_LAUNCH_CHECK();
  if (smem > 48 * 1024)
    BOP_CUDA(cudaFuncSetAttribute(png_unfilter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  png_unfilter_kernel<<<(n + kUnfilterWarps - 1) / kUnfilterWarps, kUnfilterWarps * 32, smem, s>>>(
    raw, stride, n, height, width, depth_out, status, adler);
  BOP_LAUNCH_CHECK();
  bop_prof_end(BOP_PROF_PNG, s);
  return 0;
}

// The device decoder's code compiled for the host: the CPU tests check it against zlib / OpenCV
// on the files the GPU path will see.  Not used by any product path.
extern "C" int bop_png_decode_selftest_host(const void* idat, int32_t comp_bytes, int height,
                                            int width, uint16_t* depth_out) {
  BOP_REQUIRE(idat && comp_bytes >= 0 && height > 0 && width > 0 && depth_out,
              "bop_png_decode_selftest_host: bad arguments");
  const size_t rowb = 2 * (size_t)width, expect = (size_t)height * (rowb + 1);
  std::vector<uint32_t> in(((size_t)comp_bytes + 3) / 4 + 4, 0u);
  memcpy(in.data(), idat, (size_t)comp_bytes);
  std::vector<uint8_t> raw(expect + 64, 0);
  Tables* t = new Tables;
  State s;
  uint32_t input_limit = 64;
  int ev = comp_bytes < 6 ? (int)EV_TRUNCATED
                          : inflate_begin(s, in.data(), (uint32_t)((comp_bytes + 3) >> 2), 0xffffffffu,
                                          input_limit, raw.data(), (uint32_t)expect);
  while (ev == 0 || ev == EV_COPY || ev == EV_INPUT) {
    ev = inflate_run(s, *t);
    if (ev == EV_INPUT) {        // the test hook hands the input out in slices, like the device ring
      input_limit += 64;
      s.br.limit = input_limit;
      continue;
    }
    if (ev != EV_COPY) break;
    uint8_t* o = raw.data() + s.o;
    for (uint32_t k = 0; k < s.m_len; ++k) o[k] = o[(ptrdiff_t)k - (ptrdiff_t)s.m_dist];
    s.o += s.m_len;
  }
  delete t;
  if (ev == EV_DONE && s.o != expect) ev = EV_SHORT;
  if (ev != EV_DONE) return ev;
  {
    uint32_t A = 1u, B = 0u;
    for (size_t i = 0; i < expect; ++i) {
      A = (A + raw[i]) % 65521u;
      B = (B + A) % 65521u;
    }
    if (((B << 16) | A) != s.adler) return EV_BAD_CHECKSUM;
  }
  std::vector<uint8_t> prev(rowb, 0), cur(rowb);
  for (int r = 0; r < height; ++r) {
    const uint8_t* src = raw.data() + (size_t)r * (rowb + 1);
    const int ft = src[0];
    if (ft > 4) return EV_BAD_FILTER;
    for (size_t i = 0; i < rowb; ++i) {
      const int a = i >= 2 ? cur[i - 2] : 0, b = prev[i], c = i >= 2 ? prev[i - 2] : 0;
      int pred = 0;
      if (ft == 1) pred = a;
      else if (ft == 2) pred = b;
      else if (ft == 3) pred = (a + b) >> 1;
      else if (ft == 4) {
        const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
        pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
      }
      cur[i] = (uint8_t)(src[1 + i] + pred);
    }
    for (int p = 0; p < width; ++p)
      depth_out[(size_t)r * width + p] = (uint16_t)((cur[2 * p] << 8) | cur[2 * p + 1]);
    prev.swap(cur);
  }
  return 0;
}
