/*
 * raster_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (float64) of what the reference's depth renderers compute.  The reference
 * renders through OpenGL (bop_toolkit_lib/renderer_vispy.py, renderer_py.py) or the un-vendored
 * bop_renderer (renderer_cpp.py); none of them can run on this box (no graphics stack), so rendered
 * pixels have no golden vectors: PARITY UNPINNED for the rasterisation step.
 * Everything below is derived from the reference sources:
 *
 *   - pose cast to float32 before use             renderer_vispy.py:441-442
 *   - vertices cast to float32                    renderer_vispy.py:373-374
 *   - eye depth = camera Z (v_eye_depth = -(u_mv*pos).z, CV->GL flip)  renderer_vispy.py:123-137,266-269
 *   - K -> GL projection, 'y_down', row flip  =>  image pixel (r,c) samples the ray through
 *     (u,v) = (c+0.5, r+0.5)                      renderer_vispy.py:187-232,529
 *   - depth test LESS, no culling, no blending    renderer_vispy.py:470
 *   - value = nearest-surface camera Z, perspective-correct (1/Z affine in screen space);
 *     background 0                                renderer_py.py:121-144,547-553; renderer_vispy.py:531-536
 *   - coverage ties: top-left rule (GL only guarantees "shared edges drawn once"; the tie-break
 *     is implementation-defined, we document ours).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file's shared object.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline double edge(double ax, double ay, double bx, double by, double px, double py) {
  /* Directed edge a->b evaluated at p; positive area = clockwise on screen (y down). */
  return (bx - ax) * (py - ay) - (by - ay) * (px - ax);
}

static inline int top_left(double dx, double dy) {
  /* Edge direction as seen by a clockwise-normalised triangle: left edge goes up (dy < 0), top
   * edge is horizontal and goes right (dy == 0, dx > 0). */
  return (dy < 0.0) || (dy == 0.0 && dx > 0.0);
}

/*
 * pts:   [n_pts][3] float64 (cast to float32 internally), faces: [n_faces][3] int32
 * R:     [9] row-major, t: [3]  (cast to float32 internally)
 * depth: [H][W] float32 output, 0 = background
 * Returns 0.
 */
int oracle_render_depth_sub(const double* pts, int n_pts, const int32_t* faces, int n_faces,
                            const double* R, const double* t, double fx, double fy, double cx,
                            double cy, int H, int W, int subpixel_bits, float* depth);

int oracle_render_depth(const double* pts, int n_pts, const int32_t* faces, int n_faces,
                        const double* R, const double* t, double fx, double fy, double cx,
                        double cy, int H, int W, float* depth) {
  return oracle_render_depth_sub(pts, n_pts, faces, n_faces, R, t, fx, fy, cx, cy, H, W, 0, depth);
}

/*
 * subpixel_bits > 0: window positions are snapped to a 2^-bits pixel grid before rasterisation,
 * as every OpenGL implementation does (GL_SUBPIXEL_BITS, typically 8); 0 = exact positions.
 */
int oracle_render_depth_sub(const double* pts, int n_pts, const int32_t* faces, int n_faces,
                            const double* R, const double* t, double fx, double fy, double cx,
                            double cy, int H, int W, int subpixel_bits, float* depth) {
  const double sub = subpixel_bits > 0 ? (double)(1 << subpixel_bits) : 0.0;
  double* zbuf = (double*)malloc(sizeof(double) * (size_t)H * W);
  double* proj = (double*)malloc(sizeof(double) * 3 * (size_t)n_pts); /* u', v', Z */
  if (!zbuf || !proj) { free(zbuf); free(proj); return 1; }
  for (size_t i = 0; i < (size_t)H * W; ++i) zbuf[i] = INFINITY;
  double Rf[9], tf[3];
  for (int k = 0; k < 9; ++k) Rf[k] = (double)(float)R[k];
  for (int k = 0; k < 3; ++k) tf[k] = (double)(float)t[k];
  for (int i = 0; i < n_pts; ++i) {
    const double x = (double)(float)pts[3 * i + 0], y = (double)(float)pts[3 * i + 1],
                 z = (double)(float)pts[3 * i + 2];
    const double X = Rf[0] * x + Rf[1] * y + Rf[2] * z + tf[0];
    const double Y = Rf[3] * x + Rf[4] * y + Rf[5] * z + tf[1];
    const double Z = Rf[6] * x + Rf[7] * y + Rf[8] * z + tf[2];
    proj[3 * i + 2] = Z;
    if (Z > 1e-6) {
      /* shifted so that pixel centres sit on integers */
      proj[3 * i + 0] = fx * X / Z + cx - 0.5;
      proj[3 * i + 1] = fy * Y / Z + cy - 0.5;
      if (sub > 0.0) {
        proj[3 * i + 0] = nearbyint(proj[3 * i + 0] * sub) / sub;
        proj[3 * i + 1] = nearbyint(proj[3 * i + 1] * sub) / sub;
      }
    } else {
      proj[3 * i + 0] = proj[3 * i + 1] = NAN;
    }
  }
  for (int f = 0; f < n_faces; ++f) {
    const int i0 = faces[3 * f + 0], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
    if (i0 < 0 || i1 < 0 || i2 < 0 || i0 >= n_pts || i1 >= n_pts || i2 >= n_pts) continue;
    const double z0 = proj[3 * i0 + 2], z1 = proj[3 * i1 + 2], z2 = proj[3 * i2 + 2];
    if (!(z0 > 1e-6 && z1 > 1e-6 && z2 > 1e-6)) continue; /* outside the parity domain */
    double x0 = proj[3 * i0], y0 = proj[3 * i0 + 1];
    double x1 = proj[3 * i1], y1 = proj[3 * i1 + 1];
    double x2 = proj[3 * i2], y2 = proj[3 * i2 + 1];
    double iz0 = 1.0 / z0, iz1 = 1.0 / z1, iz2 = 1.0 / z2;
    double area2 = edge(x0, y0, x1, y1, x2, y2);
    if (area2 == 0.0 || area2 != area2) continue;
    if (area2 < 0.0) { /* normalise to clockwise by swapping v1 and v2 */
      double tx = x1, ty = y1, tz = iz1;
      x1 = x2; y1 = y2; iz1 = iz2;
      x2 = tx; y2 = ty; iz2 = tz;
    }
    const double minx = fmin(x0, fmin(x1, x2)), maxx = fmax(x0, fmax(x1, x2));
    const double miny = fmin(y0, fmin(y1, y2)), maxy = fmax(y0, fmax(y1, y2));
    int bx0 = (int)ceil(fmax(minx, -1.0e6)), bx1 = (int)floor(fmin(maxx, 1.0e6));
    int by0 = (int)ceil(fmax(miny, -1.0e6)), by1 = (int)floor(fmin(maxy, 1.0e6));
    if (bx0 < 0) bx0 = 0;
    if (by0 < 0) by0 = 0;
    if (bx1 > W - 1) bx1 = W - 1;
    if (by1 > H - 1) by1 = H - 1;
    const int tl0 = top_left(x2 - x1, y2 - y1); /* edge v1->v2, opposite v0 */
    const int tl1 = top_left(x0 - x2, y0 - y2); /* edge v2->v0, opposite v1 */
    const int tl2 = top_left(x1 - x0, y1 - y0); /* edge v0->v1, opposite v2 */
    for (int y = by0; y <= by1; ++y) {
      for (int x = bx0; x <= bx1; ++x) {
        const double w0 = edge(x1, y1, x2, y2, (double)x, (double)y);
        const double w1 = edge(x2, y2, x0, y0, (double)x, (double)y);
        const double w2 = edge(x0, y0, x1, y1, (double)x, (double)y);
        const int in0 = (w0 > 0.0) || (w0 == 0.0 && tl0);
        const int in1 = (w1 > 0.0) || (w1 == 0.0 && tl1);
        const int in2 = (w2 > 0.0) || (w2 == 0.0 && tl2);
        if (in0 && in1 && in2) {
          const double Z = (w0 + w1 + w2) / (w0 * iz0 + w1 * iz1 + w2 * iz2);
          double* p = zbuf + (size_t)y * W + x;
          if (Z > 0.0 && Z < *p) *p = Z;
        }
      }
    }
  }
  for (size_t i = 0; i < (size_t)H * W; ++i) depth[i] = isinf(zbuf[i]) ? 0.0f : (float)zbuf[i];
  free(zbuf);
  free(proj);
  return 0;
}
