// Device-side triangle rasteriser core (K1) for sm_100a.
//
// Semantics follow the reference's depth renderers (bop_toolkit_lib/renderer_vispy.py:123-146,
// 187-232, 433-466, 509-538; bop_toolkit_lib/renderer_py.py:121-144, 520-555), derived in
// SURVEY.md Appendix A:
//   * camera point (X,Y,Z) = R v + t with R, t and v in float32 (renderer_vispy.py:441-442,373-374);
//   * window position u = fx X/Z + cx, v = fy Y/Z + cy; image pixel (r, c) samples the ray through
//     (u, v) = (c + 0.5, r + 0.5);
//   * value = nearest-surface camera Z at the sample point (1/Z is affine in screen space), depth
//     test LESS, no culling, no blending; background = 0 (the tile holds max 1/Z, see z_merge_if);
//   * like every OpenGL implementation, window positions are snapped to a sub-pixel grid before
//     rasterisation (GL_SUBPIXEL_BITS; 8 bits here = 1/256 px), which makes the edge functions
//     exact integers: coverage is watertight by construction and the top-left rule resolves
//     samples that lie exactly on an edge, so a shared edge is drawn exactly once.
//
// Vertices are projected ONCE per (pose, vertex) by project_kernel (vsd.cu) into a 16-byte record;
// the per-triangle work here is three 128-bit gathers plus integer setup.
#pragma once
#include "common.cuh"

constexpr int kSubBits = 8;
constexpr int kSubOne = 1 << kSubBits;          // 256 sub-pixel units per pixel
constexpr int kFpClamp = 1 << 29;               // |x_fp| <= 2^29  (2^21 px)
constexpr int kSmallExtentFp = 1 << 15;         // 128 px: 32-bit edge functions are exact below

struct PoseCamF {
  float r[9];
  float t[3];
  float fx, fy, cxh, cyh;  // cxh = cx - 0.5, cyh = cy - 0.5: pixel centres land on integers
};

// Projected vertex record: snapped window position (sub-pixel units, pixel centres at multiples
// of 256), 1/Z, validity.
struct __align__(16) PVert {
  int x, y;
  float iz;
  int valid;
};

__device__ __forceinline__ PVert project_vertex(const PoseCamF& P, const float4 v) {
  PVert o;
  const float X = __fmaf_rn(P.r[0], v.x, __fmaf_rn(P.r[1], v.y, __fmaf_rn(P.r[2], v.z, P.t[0])));
  const float Y = __fmaf_rn(P.r[3], v.x, __fmaf_rn(P.r[4], v.y, __fmaf_rn(P.r[5], v.z, P.t[1])));
  const float Z = __fmaf_rn(P.r[6], v.x, __fmaf_rn(P.r[7], v.y, __fmaf_rn(P.r[8], v.z, P.t[2])));
  // Points on or behind the camera plane are outside the parity domain (the reference's near
  // plane would be <= 0 there, renderer_vispy.py:447-452): triangles touching them are dropped.
  o.valid = (Z > 1e-6f) ? 1 : 0;
  const float iz = __frcp_rn(o.valid ? Z : 1.f);
  o.iz = iz;
  const float px = __fmaf_rn(__fmul_rn(P.fx, X), iz, P.cxh);
  const float py = __fmaf_rn(__fmul_rn(P.fy, Y), iz, P.cyh);
  const float lim = (float)(kFpClamp >> kSubBits);
  o.x = __float2int_rn(fminf(fmaxf(px, -lim), lim) * (float)kSubOne);
  o.y = __float2int_rn(fminf(fmaxf(py, -lim), lim) * (float)kSubOne);
  return o;
}

// Per-triangle state of the 32-bit path.  Edge k is opposite vertex k:
//   E_k(x, y) = A_k * (x - x0) + B_k * (y - y0) + C_k     (pixel steps, relative to the bbox origin)
// oriented so that inside <=> E_k >= 0 for all k after the top-left bias (non-owning edges are
// biased by -1, which turns "E > 0" into "E - 1 >= 0" on integers).
struct TriSetup {
  int A[3], B[3], C[3];   // C already carries the top-left bias
  int bias[3];            // 0 or 1: added back to get the true edge value for interpolation
  // 1/Z is affine in screen space: izp(x, y) = iz0 + pa * (256 x - vx) + pb * (256 y - vy), with
  // (vx, vy) the snapped sub-pixel position of vertex 0.  The expression depends only on the
  // triangle and the pixel -- not on the tile, band or path that draws it -- so a pixel's depth
  // is the same bit pattern wherever it is rasterised.
  float pa, pb, iz0;
  int vx, vy;
  int bx0, by0, bx1, by1; // candidate pixel rectangle (inclusive), already clipped
};

// Depth plane of a triangle from its (oriented) edge steps.  A[k] = dE_k/dx, B[k] = dE_k/dy per
// PIXEL (256 sub-pixel units); sum_k A[k] = sum_k B[k] = 0, so the gradient only needs the
// differences of 1/Z, which are exact-ish in float (no cancellation against 1/Z itself).
__device__ __forceinline__ void tri_plane(float A1, float A2, float B1, float B2, float area,
                                          const PVert& v0, const PVert& v1, const PVert& v2,
                                          float& pa, float& pb) {
  const float d1 = __fsub_rn(v1.iz, v0.iz), d2 = __fsub_rn(v2.iz, v0.iz);
  const float k = __frcp_rn(area) * (1.0f / (float)kSubOne);
  pa = __fmul_rn(__fmaf_rn(A1, d1, __fmul_rn(A2, d2)), k);
  pb = __fmul_rn(__fmaf_rn(B1, d1, __fmul_rn(B2, d2)), k);
}

// Returns the number of candidate pixels (0: nothing to draw; < 0: -count, the triangle is too
// large for 32-bit edge functions and must take the 64-bit path).
// cull: 0 = none, +1 / -1 = drop triangles with negative / positive signed area (bop_model_dev.cull).
__device__ __forceinline__ int tri_setup(const PVert& v0, const PVert& v1, const PVert& v2, int cx0,
                                         int cy0, int cx1, int cy1, TriSetup& T, int cull = 0) {
  if (!(v0.valid & v1.valid & v2.valid)) return 0;
  const int minx = min(v0.x, min(v1.x, v2.x)), maxx = max(v0.x, max(v1.x, v2.x));
  const int miny = min(v0.y, min(v1.y, v2.y)), maxy = max(v0.y, max(v1.y, v2.y));
  // Pixel centre c is a candidate iff minx <= c*256 <= maxx.
  const int bx0 = max(cx0, (minx + kSubOne - 1) >> kSubBits), bx1 = min(cx1, maxx >> kSubBits);
  const int by0 = max(cy0, (miny + kSubOne - 1) >> kSubBits), by1 = min(cy1, maxy >> kSubBits);
  if (bx1 < bx0 || by1 < by0) return 0;
  const int n = (bx1 - bx0 + 1) * (by1 - by0 + 1);
  T.bx0 = bx0; T.bx1 = bx1; T.by0 = by0; T.by1 = by1;
  if (maxx - minx >= kSmallExtentFp || maxy - miny >= kSmallExtentFp) return -n;
  // Directed edges a->b: k=0: v1->v2, k=1: v2->v0, k=2: v0->v1.
  // E(p) = (bx-ax)*(py-ay) - (by-ay)*(px-ax); positive area = clockwise on screen (y down).
  const int ax[3] = {v1.x, v2.x, v0.x}, ay[3] = {v1.y, v2.y, v0.y};
  const int bx[3] = {v2.x, v0.x, v1.x}, by[3] = {v2.y, v0.y, v1.y};
  const int area2 = (v1.x - v0.x) * (v2.y - v0.y) - (v1.y - v0.y) * (v2.x - v0.x);
  if (area2 == 0) return 0;
  const int s = area2 > 0 ? 1 : -1;
  if (cull == -s) return 0;
  const int px0 = bx0 << kSubBits, py0 = by0 << kSubBits;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int dx = s * (bx[k] - ax[k]), dy = s * (by[k] - ay[k]);
    // value at the bbox origin pixel; per-pixel steps are dE/dx = -dy*256, dE/dy = dx*256
    const int e0 = dx * (py0 - ay[k]) - dy * (px0 - ax[k]);
    // top-left rule on the clockwise-normalised edge direction: left edge goes up (dy < 0), top
    // edge is horizontal and goes right (dy == 0, dx > 0)
    const int own = (dy < 0 || (dy == 0 && dx > 0)) ? 1 : 0;
    T.bias[k] = 1 - own;
    T.A[k] = -dy * kSubOne;
    T.B[k] = dx * kSubOne;
    T.C[k] = e0 - (1 - own);
  }
  tri_plane((float)T.A[1], (float)T.A[2], (float)T.B[1], (float)T.B[2], (float)(s * area2), v0, v1, v2,
            T.pa, T.pb);
  T.iz0 = v0.iz; T.vx = v0.x; T.vy = v0.y;
  return n;
}

// 1/x on the SFU (MUFU.RCP, <= 1 ulp): deterministic, so a pixel's depth does not depend on the
// path that drew it; far inside the 1e-3 relative depth tolerance.
__device__ __forceinline__ float rcp_fast(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// 1/Z at pixel (x, y) of a triangle's plane (see TriSetup).  256 x - vx is an exact float.
__device__ __forceinline__ float plane_row(float pb, float iz0, int y, int vy) {
  return __fmaf_rn(pb, (float)((y << kSubBits) - vy), iz0);
}
__device__ __forceinline__ float plane_iz(float pa, float rowbase, float fdx) {
  return __fmaf_rn(pa, fdx, rowbase);
}

// The z-tile keeps, per pixel, the LARGEST 1/Z drawn so far as float bits (0 = nothing drawn):
// depth test LESS on Z is GREATER on 1/Z, and because the reciprocal is monotone
// min_t rcp(iz_t) == rcp(max_t iz_t) bit for bit, so ONE reciprocal per pixel when the tile is
// read (tile_depth) replaces one per covered candidate in the raster loops.  Signed integer max:
// positive floats order like their bit patterns, and a non-positive 1/Z (rounding of a plane that
// crosses zero; outside the parity domain) is negative or zero as an integer and never wins.
// The store is a predicated shared-memory reduction (cover >= 0 <=> the three biased edge
// functions are all non-negative): no branch in the per-candidate loop.
#ifndef BOP_ZMERGE_SELECT
#define BOP_ZMERGE_SELECT 2
#endif
__device__ __forceinline__ void z_merge_if(unsigned saddr, float iz, int cover) {
#if BOP_ZMERGE_SELECT == 2
  // branch-free: an uncovered candidate (cover < 0) gets the sign bit and, merged as a signed
  // integer, never wins against a tile entry (>= 0); one LOP3 (a | (b & imm)) per candidate
  const int v = __float_as_int(iz) | (cover & (int)0x80000000);
  asm volatile("red.shared.max.s32 [%0], %1;" :: "r"(saddr), "r"(v) : "memory");
  return;
#elif BOP_ZMERGE_SELECT
  // branch-free variant: an uncovered candidate merges 0, which never wins
  const int v = __float_as_int(iz) & ~(cover >> 31);
  asm volatile("red.shared.max.s32 [%0], %1;" :: "r"(saddr), "r"(v) : "memory");
  return;
#endif
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ge.s32 p, %2, 0;\n\t@p red.shared.max.s32 [%0], %1;\n\t}"
      :: "r"(saddr), "r"(__float_as_int(iz)), "r"(cover) : "memory");
}
__device__ __forceinline__ unsigned smem_addr(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
// Depth (camera Z) of a tile entry; 0 where nothing was drawn (or the depth is not finite).
__device__ __forceinline__ float tile_depth(unsigned bits) {
  if ((int)bits <= 0) return 0.f;
  const float z = rcp_fast(__uint_as_float(bits));
  return __float_as_uint(z) < BOP_INF_BITS_F32 ? z : 0.f;
}

// Rasterises a small triangle (32-bit path) into the z-tile.  One thread; incremental edge
// stepping, 8 instructions per candidate pixel.  The triangle lists are sorted by candidate-
// rectangle size (assemble_kernel), so the 32 lanes of a warp run both loops with (nearly) the
// same trip counts.
__device__ __forceinline__ void tri_raster_small(const TriSetup& T, unsigned* __restrict__ z, int zx0,
                                                 int zy0, int zstride) {
  int r0 = T.C[0], r1 = T.C[1], r2 = T.C[2];
  const float fdx0 = (float)((T.bx0 << kSubBits) - T.vx);
  float fdy = (float)((T.by0 << kSubBits) - T.vy);
  unsigned zrow = smem_addr(z + (T.by0 - zy0) * zstride + (T.bx0 - zx0));
  const int w = T.bx1 - T.bx0 + 1;
  for (int h = T.by1 - T.by0 + 1; h > 0; --h) {
    int e0 = r0, e1 = r1, e2 = r2;
    const float rowbase = __fmaf_rn(T.pb, fdy, T.iz0);
    float fdx = fdx0;
    unsigned zp = zrow;
    for (int i = w; i > 0; --i) {
      z_merge_if(zp, plane_iz(T.pa, rowbase, fdx), e0 | e1 | e2);
      e0 += T.A[0]; e1 += T.A[1]; e2 += T.A[2];
      fdx += (float)kSubOne;
      zp += 4;
    }
    r0 += T.B[0]; r1 += T.B[1]; r2 += T.B[2];
    fdy += (float)kSubOne;
    zrow += 4 * zstride;
  }
}

// Warp-cooperative variant of the 32-bit path for triangles with many candidate pixels: the 32
// lanes tile the candidate rectangle in blocks of (lw x 32/lw) pixels, lw in {4, 8, 16, 32}.
__device__ __forceinline__ void tri_raster_warp(const TriSetup& T, int lane, unsigned* __restrict__ z,
                                                int zx0, int zy0, int zstride) {
  const int bw = T.bx1 - T.bx0 + 1;
  const int lw_log = bw > 16 ? 5 : (bw > 8 ? 4 : (bw > 4 ? 3 : 2));
  const int lx = lane & ((1 << lw_log) - 1), ly = lane >> lw_log;
  const int sx = 1 << lw_log, sy = 32 >> lw_log;
  const unsigned zb = smem_addr(z);
  for (int y = T.by0 + ly; y <= T.by1; y += sy) {
    const int dy = y - T.by0;
    const float rowbase = plane_row(T.pb, T.iz0, y, T.vy);
    for (int x = T.bx0 + lx; x <= T.bx1; x += sx) {
      const int dx = x - T.bx0;
      const int e0 = T.C[0] + T.A[0] * dx + T.B[0] * dy;
      const int e1 = T.C[1] + T.A[1] * dx + T.B[1] * dy;
      const int e2 = T.C[2] + T.A[2] * dx + T.B[2] * dy;
      z_merge_if(zb + 4u * (unsigned)((y - zy0) * zstride + (x - zx0)),
                 plane_iz(T.pa, rowbase, (float)((x << kSubBits) - T.vx)), e0 | e1 | e2);
    }
  }
}

// 64-bit path for large triangles; the candidate rectangle is strided over by `nlanes` lanes.
struct TriSetup64 {
  long long A[3], B[3], C[3];
  int bias[3];
  float pa, pb, iz0;
  int vx, vy;
  int bx0, by0, bx1, by1;
};

__device__ __forceinline__ bool tri_setup64(const PVert& v0, const PVert& v1, const PVert& v2,
                                            const TriSetup& Tb, TriSetup64& T, int cull = 0) {
  T.bx0 = Tb.bx0; T.bx1 = Tb.bx1; T.by0 = Tb.by0; T.by1 = Tb.by1;
  const long long ax[3] = {v1.x, v2.x, v0.x}, ay[3] = {v1.y, v2.y, v0.y};
  const long long bx[3] = {v2.x, v0.x, v1.x}, by[3] = {v2.y, v0.y, v1.y};
  const long long area2 = (long long)(v1.x - v0.x) * (v2.y - v0.y) - (long long)(v1.y - v0.y) * (v2.x - v0.x);
  if (area2 == 0) return false;
  const long long s = area2 > 0 ? 1 : -1;
  if (cull == -(int)s) return false;
  const long long px0 = (long long)T.bx0 << kSubBits, py0 = (long long)T.by0 << kSubBits;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const long long dx = s * (bx[k] - ax[k]), dy = s * (by[k] - ay[k]);
    const long long e0 = dx * (py0 - ay[k]) - dy * (px0 - ax[k]);
    const int own = (dy < 0 || (dy == 0 && dx > 0)) ? 1 : 0;
    T.bias[k] = 1 - own;
    T.A[k] = -dy * kSubOne;
    T.B[k] = dx * kSubOne;
    T.C[k] = e0 - (1 - own);
  }
  tri_plane((float)T.A[1], (float)T.A[2], (float)T.B[1], (float)T.B[2], (float)(s * area2), v0, v1, v2,
            T.pa, T.pb);
  T.iz0 = v0.iz; T.vx = v0.x; T.vy = v0.y;
  return true;
}

__device__ __forceinline__ void tri_raster_large(const TriSetup64& T, int lane, int nlanes,
                                                 unsigned* __restrict__ z, int zx0, int zy0,
                                                 int zstride) {
  const int bw = T.bx1 - T.bx0 + 1;
  const int n = bw * (T.by1 - T.by0 + 1);
  const unsigned zb = smem_addr(z);
  for (int c = lane; c < n; c += nlanes) {
    const int yy = c / bw, xx = c - yy * bw;
    const long long e0 = T.C[0] + T.A[0] * xx + T.B[0] * yy;
    const long long e1 = T.C[1] + T.A[1] * xx + T.B[1] * yy;
    const long long e2 = T.C[2] + T.A[2] * xx + T.B[2] * yy;
    {
      // Same plane expression as the 32-bit paths.
      const int x = T.bx0 + xx, y = T.by0 + yy;
      z_merge_if(zb + 4u * (unsigned)((y - zy0) * zstride + (x - zx0)),
                 plane_iz(T.pa, plane_row(T.pb, T.iz0, y, T.vy), (float)((x << kSubBits) - T.vx)),
                 ((e0 | e1 | e2) >= 0) ? 0 : -1);
    }
  }
}
