// Device-side triangle rasteriser core (K1) for sm_100a.
//
// Semantics follow the reference's depth renderers (bop_toolkit_lib/renderer_vispy.py:123-146,
// 187-232, 433-466, 509-538; bop_toolkit_lib/renderer_py.py:121-144, 520-555), derived in
// SURVEY.md Appendix A:
//   * camera point (X,Y,Z) = R v + t with R, t and v in float32 (renderer_vispy.py:441-442,373-374);
//   * window position u = fx X/Z + cx, v = fy Y/Z + cy; image pixel (r, c) samples the ray through
//     (u, v) = (c + 0.5, r + 0.5);
//   * value = nearest-surface camera Z at the sample point (1/Z is affine in screen space), depth
//     test LESS, no culling, no blending; background = 0;
//   * coverage ties (sample exactly on an edge) are resolved with the top-left rule, so that a
//     shared edge is drawn exactly once and a closed mesh is watertight.
//
// Watertightness in float32: the edge function of an edge is always evaluated from a CANONICAL
// endpoint order (lexicographic in the projected position) with a pinned operation sequence, so
// the two triangles sharing an edge see bit-identical values of opposite sign.
#pragma once
#include "common.cuh"

struct PoseCamF {
  float r[9];
  float t[3];
  float fx, fy, cxh, cyh;  // cxh = cx - 0.5, cyh = cy - 0.5: pixel centres land on integers
};

struct Vtx {
  float x, y, iz;  // shifted window position and 1/Z
};

// Projects one model vertex.  Returns false for points on or behind the camera plane (outside the
// parity domain: the reference's near plane would be <= 0 there, renderer_vispy.py:447-452).
__device__ __forceinline__ bool project_vertex(const PoseCamF& P, const float4 v, Vtx& o) {
  const float X = __fmaf_rn(P.r[0], v.x, __fmaf_rn(P.r[1], v.y, __fmaf_rn(P.r[2], v.z, P.t[0])));
  const float Y = __fmaf_rn(P.r[3], v.x, __fmaf_rn(P.r[4], v.y, __fmaf_rn(P.r[5], v.z, P.t[1])));
  const float Z = __fmaf_rn(P.r[6], v.x, __fmaf_rn(P.r[7], v.y, __fmaf_rn(P.r[8], v.z, P.t[2])));
  if (!(Z > 1e-6f)) return false;
  const float iz = __frcp_rn(Z);
  o.iz = iz;
  o.x = __fmaf_rn(__fmul_rn(P.fx, X), iz, P.cxh);
  o.y = __fmaf_rn(__fmul_rn(P.fy, Y), iz, P.cyh);
  return true;
}

struct TriSetup {
  float ax[3], ay[3], dx[3], dy[3];  // canonical edge k: origin (ax,ay), direction (dx,dy)
  float sg[3];                       // +1/-1: orientation x canonical flip
  unsigned tl;                       // bit k: edge k owns samples that lie exactly on it
  float iz[3];                       // 1/Z of the vertex OPPOSITE edge k
  int bx0, by0, bx1, by1;            // candidate pixel rectangle (inclusive), already clipped
};

__device__ __forceinline__ void make_edge(const Vtx& a, const Vtx& b, TriSetup& T, int k,
                                          float& flip) {
  const bool canon = (a.x < b.x) || (a.x == b.x && a.y < b.y);
  const float px = canon ? a.x : b.x, py = canon ? a.y : b.y;
  const float qx = canon ? b.x : a.x, qy = canon ? b.y : a.y;
  T.ax[k] = px;
  T.ay[k] = py;
  T.dx[k] = qx - px;
  T.dy[k] = qy - py;
  flip = canon ? 1.f : -1.f;
}

// Canonical edge function value at (x, y): dx*(y-ay) - dy*(x-ax), operation order pinned.
__device__ __forceinline__ float edge_eval(const TriSetup& T, int k, float x, float y) {
  const float t1 = __fmul_rn(T.dy[k], x - T.ax[k]);
  return __fmaf_rn(T.dx[k], y - T.ay[k], -t1);
}

// Builds the per-triangle state.  clip = inclusive pixel rectangle of the z-tile.
// Returns the number of candidate pixels (0 = nothing to draw).
__device__ __forceinline__ int tri_setup(const PoseCamF& P, const float4* __restrict__ verts,
                                         const int4 f, int cx0, int cy0, int cx1, int cy1,
                                         TriSetup& T) {
  Vtx v0, v1, v2;
  const bool ok0 = project_vertex(P, __ldg(verts + f.x), v0);
  const bool ok1 = project_vertex(P, __ldg(verts + f.y), v1);
  const bool ok2 = project_vertex(P, __ldg(verts + f.z), v2);
  if (!(ok0 && ok1 && ok2)) return 0;
  const float minx = fminf(v0.x, fminf(v1.x, v2.x)), maxx = fmaxf(v0.x, fmaxf(v1.x, v2.x));
  const float miny = fminf(v0.y, fminf(v1.y, v2.y)), maxy = fmaxf(v0.y, fmaxf(v1.y, v2.y));
  // Clamp in float before the int conversion (projections can be huge near Z -> 0).
  const int bx0 = max(cx0, (int)ceilf(fmaxf(minx, -1.0e6f)));
  const int bx1 = min(cx1, (int)floorf(fminf(maxx, 1.0e6f)));
  const int by0 = max(cy0, (int)ceilf(fmaxf(miny, -1.0e6f)));
  const int by1 = min(cy1, (int)floorf(fminf(maxy, 1.0e6f)));
  if (bx1 < bx0 || by1 < by0) return 0;
  T.bx0 = bx0; T.bx1 = bx1; T.by0 = by0; T.by1 = by1;
  // Directed edges: k=0: v1->v2 (opposite v0), k=1: v2->v0, k=2: v0->v1.
  float f0, f1, f2;
  make_edge(v1, v2, T, 0, f0);
  make_edge(v2, v0, T, 1, f1);
  make_edge(v0, v1, T, 2, f2);
  T.iz[0] = v0.iz; T.iz[1] = v1.iz; T.iz[2] = v2.iz;
  // Twice the signed area: directed edge v0->v1 evaluated at v2.  Positive = clockwise on screen
  // (y down).
  const float area2 = f2 * edge_eval(T, 2, v2.x, v2.y);
  if (!(area2 != 0.f)) return 0;  // degenerate (also rejects NaN)
  const float o = area2 > 0.f ? 1.f : -1.f;
  T.sg[0] = f0 * o; T.sg[1] = f1 * o; T.sg[2] = f2 * o;
  // Top-left ownership, on the edge direction as seen by the clockwise-normalised triangle:
  // left edge = going up (dy < 0), top edge = horizontal going right (dy == 0, dx > 0).
  unsigned tl = 0;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float ex = T.sg[k] * T.dx[k], ey = T.sg[k] * T.dy[k];
    if (ey < 0.f || (ey == 0.f && ex > 0.f)) tl |= 1u << k;
  }
  T.tl = tl;
  return (bx1 - bx0 + 1) * (by1 - by0 + 1);
}

// Tests one pixel centre and, if covered, merges its depth into the z-tile.
// zrow0/zx0: tile origin (pixel coords), zstride: tile row pitch in pixels.
__device__ __forceinline__ void tri_pixel(const TriSetup& T, int x, int y, unsigned* __restrict__ z,
                                          int zx0, int zy0, int zstride) {
  const float fx = (float)x, fy = (float)y;
  const float w0 = T.sg[0] * edge_eval(T, 0, fx, fy);
  const float w1 = T.sg[1] * edge_eval(T, 1, fx, fy);
  const float w2 = T.sg[2] * edge_eval(T, 2, fx, fy);
  const bool in0 = (w0 > 0.f) || (w0 == 0.f && (T.tl & 1u));
  const bool in1 = (w1 > 0.f) || (w1 == 0.f && (T.tl & 2u));
  const bool in2 = (w2 > 0.f) || (w2 == 0.f && (T.tl & 4u));
  if (in0 && in1 && in2) {
    const float num = w0 + w1 + w2;
    const float den = __fmaf_rn(w0, T.iz[0], __fmaf_rn(w1, T.iz[1], __fmul_rn(w2, T.iz[2])));
    const float Z = __fdiv_rn(num, den);
    if (Z > 0.f && Z < 3.0e38f) {
      const unsigned zb = __float_as_uint(Z);
      unsigned* p = z + (y - zy0) * zstride + (x - zx0);
      if (zb < *p) atomicMin(p, zb);
    }
  }
}
