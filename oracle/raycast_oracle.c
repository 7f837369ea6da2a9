/*
 * raycast_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A second, structurally different float64 restatement of the reference's depth render, used only
 * to cross-check raster_oracle.c (neither can be pinned against the reference's OpenGL output on
 * this box: PARITY UNPINNED).  raster_oracle.c works in the image plane (edge
 * functions of projected vertices, 1/Z interpolated affinely); this file works in 3-D: for every
 * pixel (r, c) the ray through the window position (c + 0.5, r + 0.5)
 *     dir = ((c + 0.5 - cx) / fx, (r + 0.5 - cy) / fy, 1)
 * (the sample point of renderer_vispy.py:187-232,529) is intersected with every
 * triangle in camera space (Moller-Trumbore); the value is the smallest positive Z of a hit
 * (depth test LESS, no culling: renderer_vispy.py:470; background 0: :531-536).  Because dir.z = 1
 * the ray parameter IS the camera Z.  A shared bug in the edge-function formulation (orientation,
 * top-left handling, 1/Z interpolation, pixel-centre offset) would show up as a disagreement on
 * interior pixels; exact-edge ties are the only place the two may legitimately differ.
 *
 * Pose and vertices are cast to float32 first, like the reference (renderer_vispy.py:373-374,
 * 441-442).  Triangles with a vertex at Z <= 1e-6 are skipped (outside the parity domain, as in
 * raster_oracle.c).  The per-triangle pixel rectangle (projected bounding box, one pixel of
 * margin) only bounds the search; it does not decide coverage.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

int raycast_render_depth(const double* pts, int n_pts, const int32_t* faces, int n_faces,
                         const double* R, const double* t, double fx, double fy, double cx,
                         double cy, int H, int W, float* depth) {
  double* zbuf = (double*)malloc(sizeof(double) * (size_t)H * W);
  double* cam = (double*)malloc(sizeof(double) * 3 * (size_t)n_pts);
  if (!zbuf || !cam) { free(zbuf); free(cam); return 1; }
  for (size_t i = 0; i < (size_t)H * W; ++i) zbuf[i] = INFINITY;
  double Rf[9], tf[3];
  for (int k = 0; k < 9; ++k) Rf[k] = (double)(float)R[k];
  for (int k = 0; k < 3; ++k) tf[k] = (double)(float)t[k];
  for (int i = 0; i < n_pts; ++i) {
    const double x = (double)(float)pts[3 * i + 0], y = (double)(float)pts[3 * i + 1],
                 z = (double)(float)pts[3 * i + 2];
    cam[3 * i + 0] = Rf[0] * x + Rf[1] * y + Rf[2] * z + tf[0];
    cam[3 * i + 1] = Rf[3] * x + Rf[4] * y + Rf[5] * z + tf[1];
    cam[3 * i + 2] = Rf[6] * x + Rf[7] * y + Rf[8] * z + tf[2];
  }
  for (int f = 0; f < n_faces; ++f) {
    const int i0 = faces[3 * f + 0], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
    if (i0 < 0 || i1 < 0 || i2 < 0 || i0 >= n_pts || i1 >= n_pts || i2 >= n_pts) continue;
    const double* a = cam + 3 * i0;
    const double* b = cam + 3 * i1;
    const double* c = cam + 3 * i2;
    if (!(a[2] > 1e-6 && b[2] > 1e-6 && c[2] > 1e-6)) continue;
    /* search rectangle */
    double umin = INFINITY, umax = -INFINITY, vmin = INFINITY, vmax = -INFINITY;
    const double* v[3] = {a, b, c};
    for (int k = 0; k < 3; ++k) {
      const double u = fx * v[k][0] / v[k][2] + cx - 0.5, w = fy * v[k][1] / v[k][2] + cy - 0.5;
      umin = fmin(umin, u); umax = fmax(umax, u); vmin = fmin(vmin, w); vmax = fmax(vmax, w);
    }
    int x0 = (int)floor(fmax(umin, -1.0e6)) - 1, x1 = (int)ceil(fmin(umax, 1.0e6)) + 1;
    int y0 = (int)floor(fmax(vmin, -1.0e6)) - 1, y1 = (int)ceil(fmin(vmax, 1.0e6)) + 1;
    if (x0 < 0) x0 = 0;
    if (y0 < 0) y0 = 0;
    if (x1 > W - 1) x1 = W - 1;
    if (y1 > H - 1) y1 = H - 1;
    const double e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]};
    const double e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
    for (int y = y0; y <= y1; ++y) {
      for (int x = x0; x <= x1; ++x) {
        const double d[3] = {((double)x + 0.5 - cx) / fx, ((double)y + 0.5 - cy) / fy, 1.0};
        /* Moller-Trumbore, ray origin = camera centre */
        const double p[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
        const double det = e1[0] * p[0] + e1[1] * p[1] + e1[2] * p[2];
        if (det == 0.0) continue;
        const double inv = 1.0 / det;
        const double s[3] = {-a[0], -a[1], -a[2]};
        const double uu = (s[0] * p[0] + s[1] * p[1] + s[2] * p[2]) * inv;
        if (uu < 0.0 || uu > 1.0) continue;
        const double q[3] = {s[1] * e1[2] - s[2] * e1[1], s[2] * e1[0] - s[0] * e1[2], s[0] * e1[1] - s[1] * e1[0]};
        const double vv = (d[0] * q[0] + d[1] * q[1] + d[2] * q[2]) * inv;
        if (vv < 0.0 || uu + vv > 1.0) continue;
        const double tt = (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]) * inv;
        double* pz = zbuf + (size_t)y * W + x;
        if (tt > 0.0 && tt < *pz) *pz = tt;
      }
    }
  }
  for (size_t i = 0; i < (size_t)H * W; ++i) depth[i] = isinf(zbuf[i]) ? 0.0f : (float)zbuf[i];
  free(zbuf);
  free(cam);
  return 0;
}
