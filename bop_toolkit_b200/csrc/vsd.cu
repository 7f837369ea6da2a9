// K1 (batched depth rasteriser), K2 (fused depth->dist + visibility + tau sweep) and the fused
// K1+K2 VSD kernel for sm_100a.
//
// Reference path: pose_error.vsd (bop_toolkit_lib/pose_error.py:17-93) =
//   2x Renderer.render_object (renderer_vispy.py:433-466,509-538)
//   3x misc.depth_im_to_dist_im_fast (misc.py:143-163)
//   visibility.estimate_visib_mask_gt / _est (visibility.py:9-75)
//   intersection / union / per-tau cost (pose_error.py:61-91)
//
// Fused design: a pose pair never materialises its two rendered depth images in HBM.  A
// persistent CTA pulls (pair, row-band) work items from a device-side list, rasterises both poses
// of the band into shared-memory z-tiles (atomicMin on float-as-uint depth), then streams the
// matching rows of depth_test once (128-bit loads) and reduces (union, inter, per-tau cost) with
// warp shuffles + shared atomics + one global atomic per counter.  The pixel arithmetic mirrors
// the reference's dtype ladder operation by operation, so given equal rendered
// depths the integer counts -- and hence the errors -- are bit-identical to numpy's.
#include <limits.h>
#include <stddef.h>
#include <mutex>
#include <type_traits>
#include <vector>

#include "raster.cuh"

namespace {

constexpr int kMinTilePx = 4096;
constexpr int kMaxBandRows = 1024;   // rows of one work item (>= kMinTilePx / 4, the narrowest ROI)
constexpr int kMaxWidth = 4096;

struct Tuning {
  int tile_px = 20480;   // z-tile capacity per pose, pixels (2 poses x 4 B = 144 KB of smem; leaves L1 for gathers)
  int threads = 1024;
  int ctas_per_sm = 1;
  int small_tri_px = 256; // triangles with more candidate pixels go to the warp-cooperative path
};
Tuning g_tune;

// Pairs that share their GT pose (same object, image, GT pose: the k estimates eval_calc_errors.py
// :281-289 pairs with one GT) form a GROUP: the GT pose is projected, sorted and rasterised once
// per group and row band, its z-tile stays in shared memory while the estimates take turns.
constexpr int kMaxGroup = 32;

struct PairDesc {
  double kfx, kfy, kcx, kcy;
  double diameter;
  int x0, y0, w, h;       // ROI of the pair's group (GT pose + every active estimate); x0 and w are multiples of 4
  int obj, im;
  int n_poses;
  int nbs;                // row bands the triangle lists are sorted by (0: not sorted by band)
  int group_n;            // pairs in the group (meaningful in the group's first pair)
  int active;             // 0: the pair failed the sphere gate: no work, its counts stay 0
  int pad0, pad1;
};

struct WorkItem {
  int pair, row0, nrows, band;
};

struct PlanCounters {
  unsigned total;  // number of work items
  unsigned next;   // persistent-CTA ticket
  unsigned long long depth_groups;  // 4-pixel groups of depth_test actually fetched (statistics)
};

// overlapping_sphere_projections (bop_toolkit_lib/misc.py:309-331), float64.
__device__ __forceinline__ bool sphere_projections_overlap(double radius, const double* p1,
                                                           const double* p2) {
  if (p1[2] == 0.0 || p2[2] == 0.0) return false;
  const double ax = p1[0] / p1[2], ay = p1[1] / p1[2];
  const double bx = p2[0] / p2[2], by = p2[1] / p2[2];
  const double dx = ax - bx, dy = ay - by;
  const double dist = sqrt(dx * dx + dy * dy);
  const double thresh = radius * (1.0 / p1[2] + 1.0 / p2[2]);
  return dist < thresh;
}

__device__ __forceinline__ void make_posecam(const double* pose, double fx, double fy, double cx,
                                             double cy, PoseCamF& P) {
#pragma unroll
  for (int k = 0; k < 9; ++k) P.r[k] = (float)pose[k];
#pragma unroll
  for (int k = 0; k < 3; ++k) P.t[k] = (float)pose[9 + k];
  P.fx = (float)fx;
  P.fy = (float)fy;
  P.cxh = (float)(cx - 0.5);
  P.cyh = (float)(cy - 0.5);
}

// ---------------------------------------------------------------------------------
// Geometry pass: one CTA per item (pose pair, or single pose for render / GT info).
//   1. project every vertex of every pose (16-byte PVert: snapped window position + 1/Z) and reduce
//      the exact projected bounding boxes; the records of the pose being assembled live in shared
//      memory (global scratch for models above kSmemVerts);
//   2. plan: union ROI of the poses, split into row bands that fit the z-tile; one work item per
//      band is appended to the work list (this used to be a separate kernel);
//   3. per pose, assemble the triangles: gather the three records of every face, drop what can
//      never produce a pixel -- a vertex behind the camera plane, no pixel centre inside the image
//      under its bounding box, zero area, back-facing on a verified closed mesh (bop_model_dev.cull,
//      only while the camera is outside the model) -- and write the survivors as a list SORTED by
//      (row band, size class of the candidate-pixel rectangle).  A triangle that touches more than
//      one band goes to a "shared" segment behind the bands.  The raster pass of band b streams
//      segment b and the shared segment of both poses: it never sees a triangle of another band,
//      and 32 consecutive entries have (nearly) equal loop trip counts, so the lanes of a raster
//      warp stay converged.  Counting sort in shared memory: classify + histogram, prefix, scatter.
// cam_mode 0: K is [n_im][9] indexed through im_idx (VSD); 1: cam is [n][4] fx,fy,cx,cy (render).
// tri out:  [n][n_poses][tri_cap] records in three streams, tri_cnt[n][n_poses] valid;
// band_off: [n][n_poses][kBandOffN] start of every band segment, then of the shared segment.
// ---------------------------------------------------------------------------------
constexpr int kAsmThreads = 512;
constexpr int kSmemVerts = 12288;   // projected vertices the geometry kernel keeps in shared memory (192 KB)
constexpr int kClasses = 256;
constexpr int kMaxSortBands = 15;   // ROIs with more bands are not sorted by band (every band streams the whole list)
constexpr int kBandOffN = kMaxSortBands + 2;
constexpr int kMaxBins = (kMaxSortBands + 1) * kClasses;

// Sort class of a triangle that can draw something, 0 otherwise.  The class is the size of the
// candidate-pixel rectangle (rows, columns; each capped at 15), COMPLEMENTED so that ascending
// class order lists the most expensive triangles first: the cheap 1x1 batches fill the tail
// before the raster pass's barrier.  by0 / by1: first and last candidate row.
__device__ __forceinline__ int tri_class(const int4& r0, const int4& r1, const int4& r2, int W, int H,
                                         int cull, int& by0, int& by1) {
  if (!(r0.w & r1.w & r2.w)) return 0;
  const int minx = min(r0.x, min(r1.x, r2.x)), maxx = max(r0.x, max(r1.x, r2.x));
  const int miny = min(r0.y, min(r1.y, r2.y)), maxy = max(r0.y, max(r1.y, r2.y));
  const int w = min(W - 1, maxx >> kSubBits) - max(0, (minx + kSubOne - 1) >> kSubBits) + 1;
  by0 = max(0, (miny + kSubOne - 1) >> kSubBits);
  by1 = min(H - 1, maxy >> kSubBits);
  const int h = by1 - by0 + 1;
  if (w <= 0 || h <= 0) return 0;
  const long long area2 = (long long)(r1.x - r0.x) * (r2.y - r0.y) - (long long)(r1.y - r0.y) * (r2.x - r0.x);
  if (area2 == 0) return 0;
  if (cull == (area2 > 0 ? -1 : 1)) return 0;
  return kClasses - (min(h, 15) * 16 + min(w, 15));   // 1 (>= 15 x 15) .. 239 (1 x 1)
}

struct GeomArgs {
  const bop_model_dev* models;
  const int32_t* obj_idx;
  const double* pose_a;
  const double* pose_b;
  const int32_t* im_idx;
  const double* K;
  const double* diameters;
  const int32_t* group_off;   // [n_groups + 1] first pair of every group (pairs of a group are consecutive); NULL: every item is its own group
  int cam_mode, n_poses, H, W, gate, tile_px;
  int pv_stride, tri_cap;
  PVert* pv_global;
  int4* tri_a;
  int4* tri_b;
  float* tri_c;
  unsigned* stage;
  int* tri_cnt;
  int* band_off;
  PairDesc* desc;
  WorkItem* work;
  PlanCounters* ctr;
  unsigned max_items;
  int cx_off, cy_off;
};

__global__ void __launch_bounds__(kAsmThreads, 2)
geometry_kernel(const __grid_constant__ GeomArgs g) {
  extern __shared__ int4 s_dyn[];   // [n_pv] projected vertices (smem path), then [kMaxBins] ints
  __shared__ int s_box[4];          // union of the projected boxes of the group's active poses
  __shared__ int s_n;
  __shared__ int s_plan[4];         // y0, band rows, nb, nbs (bands the lists are sorted by; 0 = unsorted)
  __shared__ int s_wsum[kAsmThreads / 32];
  __shared__ int s_act[kMaxGroup];  // estimate q of the group passed the sphere gate
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n_poses = g.n_poses, W = g.W, H = g.H;
  // The group: pairs first .. first + cnt - 1 share object, image and GT pose.  Its poses: q < cnt is
  // the estimate of pair first + q (or, for single-pose work, THE pose); q == cnt is the GT pose.
  const int first = g.group_off ? g.group_off[blockIdx.x] : (int)blockIdx.x;
  const int cnt = g.group_off ? g.group_off[blockIdx.x + 1] - first : 1;
  if (cnt < 1 || cnt > kMaxGroup) __trap();
  const int Q = n_poses == 2 ? cnt + 1 : 1;
  const bop_model_dev M = g.models[g.obj_idx[first]];
  const bool smem_pv = M.n_pts <= kSmemVerts;
  int4* const s_pv = s_dyn;
  int* const s_bins = (int*)(s_dyn + (smem_pv ? min(g.pv_stride, kSmemVerts) : 0));
  auto pose_of = [&](int q) { return (n_poses == 2 && q == cnt) ? g.pose_b + 12 * (size_t)first : g.pose_a + 12 * (size_t)(first + q); };
  auto slot_of = [&](int q) { return n_poses == 2 ? (q == cnt ? first * 2 + 1 : (first + q) * 2) : first; };
  const double diam = g.diameters ? g.diameters[first] : M.diameter;
  double fx, fy, cx, cy;
  int im;
  if (g.cam_mode == 0) {
    im = g.im_idx[first];
    const double* Kp = g.K + 9 * (size_t)im;
    fx = Kp[0]; fy = Kp[4]; cx = Kp[2]; cy = Kp[5];
  } else {
    im = first;
    const double* c = g.K + 4 * (size_t)first;
    fx = c[0]; fy = c[1]; cx = c[2]; cy = c[3];
  }
  if (M.n_pts > g.pv_stride || M.n_faces > g.tri_cap) __trap();   // max_geom below the model's size: fail loudly
  if (n_poses == 2 && tid > 0 && tid < cnt) {
    // the pairs of a group must share object, image, GT pose and diameter (bop_vsd_grouped_dev's
    // contract; batch.group_pairs_by_gt builds such groups): anything else would silently score
    // the estimates of pair first + tid against the wrong GT render -- fail loudly instead
    bool same = g.obj_idx[first + tid] == g.obj_idx[first];
    if (g.cam_mode == 0) same = same && g.im_idx[first + tid] == g.im_idx[first];
    if (g.diameters) same = same && g.diameters[first + tid] == g.diameters[first];
    const long long* a = (const long long*)(g.pose_b + 12 * (size_t)first);
    const long long* b = (const long long*)(g.pose_b + 12 * (size_t)(first + tid));
#pragma unroll
    for (int k = 0; k < 12; ++k) same = same && a[k] == b[k];
    if (!same) __trap();
  }
  if (tid < cnt)
    s_act[tid] = (g.gate && n_poses == 2) ? (sphere_projections_overlap(0.5 * diam, pose_of(tid) + 9, pose_of(cnt) + 9) ? 1 : 0) : 1;
  if (tid < 4) s_box[tid] = (tid & 2) ? INT_MIN : INT_MAX;
  __syncthreads();
  bool any_active = false;
  for (int q = 0; q < cnt; ++q) any_active = any_active || s_act[q];
  const float4* __restrict__ verts = (const float4*)M.verts4;
  // The projected-vertex records are addressed through a compile-time switch so that the common
  // case compiles to shared-memory loads / stores (LDS / STS) instead of generic ones.
#define PV(i) (*(SMEM ? &s_pv[i] : &gpv[i]))
  auto body = [&](auto smem_tag) {
  constexpr bool SMEM = decltype(smem_tag)::value;
  int4* gpv = nullptr;   // global path: the records of the pose at hand (every list has its own scratch)
  // Projects every vertex of pose q; keep: store the records; want_box: reduce the projected box.
  auto project_pose = [&](int q, bool keep, bool want_box) {
    PoseCamF P;
    make_posecam(pose_of(q), fx, fy, cx + g.cx_off, cy + g.cy_off, P);
    gpv = (int4*)(g.pv_global + (size_t)slot_of(q) * g.pv_stride);
    int minx = INT_MAX, miny = INT_MAX, maxx = INT_MIN, maxy = INT_MIN;
    for (int i0 = tid; i0 < M.n_pts; i0 += 4 * kAsmThreads) {
      float4 vin[4];   // four loads in flight before the first store
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int i = i0 + k * kAsmThreads;
        vin[k] = i < M.n_pts ? __ldg(verts + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int i = i0 + k * kAsmThreads;
        if (i < M.n_pts) {
          const PVert v = project_vertex(P, vin[k]);
          if (keep) PV(i) = make_int4(v.x, v.y, __float_as_int(v.iz), v.valid);
          if (v.valid) {
            minx = min(minx, v.x); maxx = max(maxx, v.x);
            miny = min(miny, v.y); maxy = max(maxy, v.y);
          }
        }
      }
    }
    if (want_box) {
      minx = __reduce_min_sync(BOP_FULL_MASK, minx); miny = __reduce_min_sync(BOP_FULL_MASK, miny);
      maxx = __reduce_max_sync(BOP_FULL_MASK, maxx); maxy = __reduce_max_sync(BOP_FULL_MASK, maxy);
      if (lane == 0 && maxx >= minx) {
        atomicMin(&s_box[0], minx); atomicMin(&s_box[1], miny);
        atomicMax(&s_box[2], maxx); atomicMax(&s_box[3], maxy);
      }
    }
  };
  // Boxes of every active pose; the LAST pose (the GT pose, or the single pose) keeps its records
  // in shared memory for the assembly that follows (the global path keeps them all).
  if (any_active)
    for (int q = 0; q < Q; ++q)
      if (q == Q - 1 || s_act[q]) project_pose(q, q == Q - 1 || !SMEM, true);
  __syncthreads();
  // ---- plan (thread 0): ROI of the group, row bands, work items -------------------------------
  if (tid == 0) {
    int x0 = 0, y0 = 0, w = 0, h = 0, nb = 0, band = 1;
    if (s_box[2] >= s_box[0]) {
      const int bx0 = max(0, (s_box[0] + kSubOne - 1) >> kSubBits), by0 = max(0, (s_box[1] + kSubOne - 1) >> kSubBits);
      const int bx1 = min(W - 1, s_box[2] >> kSubBits), by1 = min(H - 1, s_box[3] >> kSubBits);
      if (bx1 >= bx0 && by1 >= by0) {
        x0 = bx0 & ~3;
        w = ((bx1 + 1 - x0) + 3) & ~3;
        y0 = by0;
        h = by1 + 1 - by0;
        const int cap_rows = min(kMaxBandRows, max(1, g.tile_px / w));
        nb = (h + cap_rows - 1) / cap_rows;
        band = (h + nb - 1) / nb;  // balanced bands
      }
    }
    const int nbs = (nb >= 2 && nb <= kMaxSortBands) ? nb : 0;
    s_box[0] = x0; s_box[1] = w; s_box[2] = h;   // (the box is consumed: reuse it to hand the ROI to the descriptor writers)
    s_plan[0] = y0; s_plan[1] = band; s_plan[2] = nb; s_plan[3] = nbs;
    if (nb > 0) {
      const unsigned start = atomicAdd(&g.ctr->total, (unsigned)nb);
      for (int b = 0; b < nb; ++b) {
        if (start + b >= g.max_items) break;  // cannot happen with the documented workspace size
        WorkItem wi;
        wi.pair = first;
        wi.row0 = b * band;
        wi.nrows = min(band, h - b * band);
        wi.band = b;
        g.work[start + b] = wi;
      }
    }
  }
  __syncthreads();
  const int roi_y0 = s_plan[0], band_rows = s_plan[1], nb = s_plan[2], nbs = s_plan[3];
  if (tid < cnt) {   // one descriptor per pair of the group
    PairDesc D;
    D.kfx = fx; D.kfy = fy; D.kcx = cx; D.kcy = cy;
    D.diameter = diam;
    D.x0 = s_box[0]; D.y0 = roi_y0; D.w = s_box[1]; D.h = s_box[2];
    D.obj = g.obj_idx[first]; D.im = im;
    D.n_poses = n_poses; D.nbs = nbs;
    D.group_n = tid == 0 ? cnt : 0;
    D.active = s_act[tid];
    D.pad0 = D.pad1 = 0;
    g.desc[first + tid] = D;
  }
  const float inv_band = 1.0f / (float)band_rows;
  const int nseg = nbs + 1;                 // band segments + the shared segment (unsorted: one segment)
  const int nbins = nseg * kClasses;
  const int4* __restrict__ faces = (const int4*)M.faces4;
  const int F = M.n_faces;
  const unsigned lt = (1u << lane) - 1u;
  // Assembly: the last pose first (its records are in shared memory already), then the estimates.
  bool have_records = true;
  for (int qq = 0; qq < Q; ++qq) {
    const int q = qq == 0 ? Q - 1 : qq - 1;
    const int gw = slot_of(q);
    int* const boff = g.band_off + (size_t)gw * kBandOffN;
    if (nb == 0 || (q < Q - 1 && !s_act[q])) {   // gated, or nothing inside the image
      if (tid == 0) g.tri_cnt[gw] = 0;
      if (tid < kBandOffN) boff[tid] = 0;
      continue;
    }
    __syncthreads();   // the scatter of the previous pose has finished with the records and the cursors
    if (!have_records) {
      if (SMEM) project_pose(q, true, false);
      else gpv = (int4*)(g.pv_global + (size_t)gw * g.pv_stride);
    }
    have_records = false;
    const double* const pose_q = pose_of(q);
    for (int i = tid; i < nbins; i += kAsmThreads) s_bins[i] = 0;
    if (tid == 0) s_n = 0;
    __syncthreads();
    // Culling is only output-preserving while the camera centre is outside the mesh: require it
    // outside the model's bounding sphere (|t| = distance of the camera from the model origin).
    int cull = M.cull;
    if (cull) {
      const double r = (double)M.cull_min_dist;
      const double* t = pose_q + 9;
      if (!(t[0] * t[0] + t[1] * t[1] + t[2] * t[2] > r * r)) cull = 0;
    }
    const size_t out0 = (size_t)gw * g.tri_cap;
    // Pass 1 -- classify: two faces per thread and iteration (both index loads and all six record
    // gathers are in flight before the first test).  The key of every face (0 = cannot draw) goes
    // to a per-list key array in face order (coalesced 16-bit stores) and is counted (one shared
    // atomic per distinct key and warp).
    unsigned short* const keys = (unsigned short*)(g.stage + out0);
    auto key_of = [&](int cls, int by0, int by1) {
      if (cls == 0 || nbs == 0) return cls;
      // row -> band without an integer division: (y + 0.5) / band_rows is at least 0.5 / 1024 away
      // from an integer and at most 16, far more than the float error of the product
      const int b0 = __float2int_rd(((float)(by0 - roi_y0) + 0.5f) * inv_band);
      const int b1 = __float2int_rd(((float)(by1 - roi_y0) + 0.5f) * inv_band);
      return (b0 == b1 ? b0 : nbs) * kClasses + cls;
    };
    int4 na4 = make_int4(0, 0, 0, 0), nb4 = na4;   // indices of the NEXT iteration (prefetched)
    if (tid < F) na4 = __ldg(faces + tid);
    if (tid + kAsmThreads < F) nb4 = __ldg(faces + tid + kAsmThreads);
    for (int f0 = 0; f0 < F; f0 += 2 * kAsmThreads) {
      const int fa = f0 + tid, fb = fa + kAsmThreads;
      const int4 ia = na4, ib = nb4;
      if (fa + 2 * kAsmThreads < F) na4 = __ldg(faces + fa + 2 * kAsmThreads);
      if (fb + 2 * kAsmThreads < F) nb4 = __ldg(faces + fb + 2 * kAsmThreads);
      int ka = 0, kb = 0;
      int4 a0, a1, a2, b0, b1, b2;
      if (fa < F) { a0 = PV(ia.x); a1 = PV(ia.y); a2 = PV(ia.z); }
      if (fb < F) { b0 = PV(ib.x); b1 = PV(ib.y); b2 = PV(ib.z); }
      int y0a = 0, y1a = 0, y0b = 0, y1b = 0;
      if (fa < F) {
        BOP_DCHECK((unsigned)ia.x < (unsigned)M.n_pts && (unsigned)ia.y < (unsigned)M.n_pts && (unsigned)ia.z < (unsigned)M.n_pts);
        ka = key_of(tri_class(a0, a1, a2, W, H, cull, y0a, y1a), y0a, y1a); keys[fa] = (unsigned short)ka;
      }
      if (fb < F) {
        BOP_DCHECK((unsigned)ib.x < (unsigned)M.n_pts && (unsigned)ib.y < (unsigned)M.n_pts && (unsigned)ib.z < (unsigned)M.n_pts);
        kb = key_of(tri_class(b0, b1, b2, W, H, cull, y0b, y1b), y0b, y1b); keys[fb] = (unsigned short)kb;
      }
      BOP_DCHECK(ka >= 0 && ka < nbins && kb >= 0 && kb < nbins);
      const unsigned pa = __match_any_sync(BOP_FULL_MASK, ka), pb = __match_any_sync(BOP_FULL_MASK, kb);
      if (ka && (pa & lt) == 0) atomicAdd(&s_bins[ka], __popc(pa));
      if (kb && (pb & lt) == 0) atomicAdd(&s_bins[kb], __popc(pb));
    }
    __syncthreads();
    // Exclusive prefix over the bins (each thread owns ceil(nbins / 512) consecutive bins; warp scan
    // + one pass over the warp totals): s_bins becomes the write cursor of every key.
    {
      const int per = (nbins + kAsmThreads - 1) / kAsmThreads;
      const int first = tid * per;
      int sum = 0;
      for (int k = 0; k < per; ++k)
        if (first + k < nbins) sum += s_bins[first + k];
      int incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(BOP_FULL_MASK, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) s_wsum[warp] = incl;
      __syncthreads();
      int wbase = 0;
      for (int k = 0; k < warp; ++k) wbase += s_wsum[k];
      int run = wbase + incl - sum;
      for (int k = 0; k < per; ++k)
        if (first + k < nbins) { const int c = s_bins[first + k]; s_bins[first + k] = run; run += c; }
      if (tid == kAsmThreads - 1) s_n = run;   // the last thread's running total = number of survivors
    }
    __syncthreads();
    const int n_surv = s_n;
    if (n_surv > g.tri_cap) __trap();   // caller passed a max_geom below (n_faces + 1) / 2: fail loudly
    if (tid < kBandOffN) boff[tid] = tid < nseg ? s_bins[tid * kClasses] : n_surv;   // read before the scatter moves the cursors
    if (tid == 0) g.tri_cnt[gw] = n_surv;
    __syncthreads();
    // Pass 2 -- scatter, again in face order: indices and keys are coalesced loads fetched one
    // iteration ahead, so nothing in the loop waits on a dependent global load; a surviving face
    // takes a slot under its key (one shared atomic per distinct key and warp), gathers its three
    // records from shared memory once more and writes its 36 bytes.
    na4 = nb4 = make_int4(0, 0, 0, 0);
    int nka = 0, nkb = 0;
    if (tid < F) { na4 = __ldg(faces + tid); nka = keys[tid]; }
    if (tid + kAsmThreads < F) { nb4 = __ldg(faces + tid + kAsmThreads); nkb = keys[tid + kAsmThreads]; }
    for (int f0 = 0; f0 < F; f0 += 2 * kAsmThreads) {
      const int fa = f0 + tid, fb = fa + kAsmThreads;
      const int4 ia = na4, ib = nb4;
      const int ka = nka, kb = nkb;
      nka = nkb = 0;
      if (fa + 2 * kAsmThreads < F) { na4 = __ldg(faces + fa + 2 * kAsmThreads); nka = keys[fa + 2 * kAsmThreads]; }
      if (fb + 2 * kAsmThreads < F) { nb4 = __ldg(faces + fb + 2 * kAsmThreads); nkb = keys[fb + 2 * kAsmThreads]; }
      const unsigned pa = __match_any_sync(BOP_FULL_MASK, ka), pb = __match_any_sync(BOP_FULL_MASK, kb);
      const int la = __ffs(pa) - 1, lb = __ffs(pb) - 1;
      int sa = 0, sb = 0;
      if (ka && lane == la) sa = atomicAdd(&s_bins[ka], __popc(pa));
      if (kb && lane == lb) sb = atomicAdd(&s_bins[kb], __popc(pb));
      sa = __shfl_sync(BOP_FULL_MASK, sa, la) + __popc(pa & lt);
      sb = __shfl_sync(BOP_FULL_MASK, sb, lb) + __popc(pb & lt);
      BOP_DCHECK(!ka || (sa >= 0 && sa < n_surv));
      BOP_DCHECK(!kb || (sb >= 0 && sb < n_surv));
      if (ka) {
        const int4 r0 = PV(ia.x), r1 = PV(ia.y), r2 = PV(ia.z);
        g.tri_a[out0 + sa] = make_int4(r0.x, r0.y, r1.x, r1.y);
        g.tri_b[out0 + sa] = make_int4(r2.x, r2.y, r0.z, r1.z);
        g.tri_c[out0 + sa] = __int_as_float(r2.z);
      }
      if (kb) {
        const int4 r0 = PV(ib.x), r1 = PV(ib.y), r2 = PV(ib.z);
        g.tri_a[out0 + sb] = make_int4(r0.x, r0.y, r1.x, r1.y);
        g.tri_b[out0 + sb] = make_int4(r2.x, r2.y, r0.z, r1.z);
        g.tri_c[out0 + sb] = __int_as_float(r2.z);
      }
    }
  }
  };
  if (smem_pv) body(std::true_type{}); else body(std::false_type{});
#undef PV
}

}  // namespace

// =====================================================================================
// Pixel arithmetic shared by the fused kernel and the stand-alone K2 pass.
// =====================================================================================
namespace {

struct PixParams {
  double taus[BOP_MAX_TAUS];
  float delta_f32;   // `d_diff <= delta` is a float32 comparison (visibility.py:35-37, NEP 50)
  int n_taus;
  int normalized;
  int visib_mode;
  int H, W;
  int depth_dtype;
  int vec_ok;        // rows of depth_test are 16-byte (f32) / 8-byte (u16) aligned
  // float32 decision path (step cost): thresholds and their undecidable bands (see vsd_pixel_fast)
  float tau_f[BOP_MAX_TAUS], tau_eps[BOP_MAX_TAUS];
  int fast;          // 1: try the float32 decision path first
  // Histogram form of the step-cost sweep (taus ascending): a pixel adds 1 to the bin
  // cnt = #(tau_k <= d) instead of 1 to every counter k < cnt.  tau_tab = [-3e38, taus..., 3e38...]
  // so that tau_tab[cnt] <= d < tau_tab[cnt + 1]; cnt is guessed from the spacing of the first two
  // taus (exact for the evenly spaced BOP sweep) and verified against the table.
  float tau_tab[BOP_MAX_TAUS + 2];
  float guess_scale, guess_off;
  int hist;          // 1: the HIST kernel variant may be used
};

constexpr int MODE_STEP = 0, MODE_TLINEAR = 1, MODE_RENDER = 2, MODE_MASK = 3, MODE_GTINFO = 4;

constexpr int kHistBits = 5;   // per-thread bin width: a thread sees <= 31 pixels per work item (checked on the host)

template <int MODE, int NT, bool HIST = false>
struct Acc {
  int uni = 0, inter = 0;
  int c[NT];
  __device__ Acc() {
#pragma unroll
    for (int k = 0; k < NT; ++k) c[k] = 0;
  }
};
template <int NT>
struct Acc<MODE_TLINEAR, NT, false> {
  int uni = 0, inter = 0;
  double c[NT];
  __device__ Acc() {
#pragma unroll
    for (int k = 0; k < NT; ++k) c[k] = 0.0;
  }
};
// Step cost, ascending taus: NT + 1 bins of kHistBits bits in one 64-bit register.
template <int NT>
struct Acc<MODE_STEP, NT, true> {
  static_assert((NT + 1) * kHistBits <= 64, "histogram does not fit 64 bits");
  int uni = 0, inter = 0;
  unsigned long long hist = 0ull;
};

// misc.depth_im_to_dist_im_fast (misc.py:158-161): sqrt((X*d)^2 + (Y*d)^2 + d^2) in float64,
// every product and sum rounded separately (numpy does not contract to FMA).
__device__ __forceinline__ double dist_of(double X, double Y, float d) {
  const double dd = (double)d;
  const double a = __dmul_rn(X, dd), b = __dmul_rn(Y, dd);
  return __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(a, a), __dmul_rn(b, b)), __dmul_rn(dd, dd)));
}

// One pixel of pose_error.vsd lines 48-91.  dt/de/dg: depth_test, rendered est, rendered gt.
// NT >= pp.n_taus is the compile-time length of the tau sweep; pp.taus[k >= n_taus] = +inf adds nothing.
template <int MODE, int NT, bool HIST = false>
__device__ __forceinline__ void vsd_pixel(const PixParams& pp, double X, double Y, float dt,
                                          float de, float dg, double diameter,
                                          Acc<MODE, NT, HIST>& acc) {
  const bool he = de != 0.f, hg = dg != 0.f;   // d_model > 0 on the distance image
  if (!he && !hg) return;
  const double Dt = dist_of(X, Y, dt);
  const float Dt32 = (float)Dt;
  const bool t_zero = Dt == 0.0, t_pos = Dt > 0.0;
  double Dg = 0.0, De = 0.0;
  bool vg = false, ve = false;
  if (hg) {
    Dg = dist_of(X, Y, dg);
    const bool c = __fsub_rn((float)Dg, Dt32) <= pp.delta_f32;
    vg = (pp.visib_mode == BOP_VISIB_BOP19) ? (c || t_zero) : (c && t_pos);
  }
  if (he) {
    De = dist_of(X, Y, de);
    const bool c = __fsub_rn((float)De, Dt32) <= pp.delta_f32;
    ve = (pp.visib_mode == BOP_VISIB_BOP19) ? (c || t_zero) : (c && t_pos);
    ve = ve || vg;   // visibility.py:74: OR (visib_gt AND d_est > 0)
  }
  if (!(vg || ve)) return;
  acc.uni += 1;
  if (vg && ve) {
    acc.inter += 1;
    double d = fabs(__dsub_rn(Dg, De));
    if (pp.normalized) d = __ddiv_rn(d, diameter);
    if constexpr (HIST) {
      int cnt = 0;
#pragma unroll
      for (int k = 0; k < NT; ++k) cnt += (d >= pp.taus[k]) ? 1 : 0;
      acc.hist += 1ull << (kHistBits * cnt);
    } else {
#pragma unroll
      for (int k = 0; k < NT; ++k) {
        if constexpr (MODE == MODE_TLINEAR) {
          double c = __ddiv_rn(d, pp.taus[k]);
          acc.c[k] += (c > 1.0) ? 1.0 : c;
        } else {
          acc.c[k] += (d >= pp.taus[k]) ? 1 : 0;
        }
      }
    }
  }
}

// Float32 decision path of vsd_pixel for the step cost.  Every quantity the reference compares is
// |depth difference| * s with s = sqrt(1 + X^2 + Y^2) (dist = depth * s for all three images), so
// the float64 distance images are only needed when a comparison is too close to call:
//   visibility  fl32(D_model) - fl32(D_test) <= delta : a = (d_model - d_test) * s, undecidable
//               within 2.5e-7 * (d_model + d_test) * s + 3e-6 * |a| of delta (float32 roundings of
//               the two distances <= 6e-8 relative each; a itself is good to < 4e-7 relative);
//   cost        |D_gt - D_est| (/ diameter) >= tau_k : dn = |d_gt - d_est| * s (* 1/diameter),
//               undecidable within 3e-6 * tau_k of tau_k (dn is good to < 6e-7 relative).
// Returns false without touching acc when anything is undecidable; the caller then runs
// vsd_pixel.  Decided pixels get exactly the counts vsd_pixel would give them.
// s2x = fl32(1 + X^2) of the column, yy = fl32(Y^2) of the row.
// HIST: s_tau = PixParams::tau_tab in shared memory (indexed per lane).
template <int NT, bool HIST>
__device__ __forceinline__ bool vsd_pixel_fast(const PixParams& pp, const float* __restrict__ s_tau,
                                               float s2x, float yy, float dt, float de, float dg,
                                               float inv_diam, Acc<MODE_STEP, NT, HIST>& acc) {
  const bool he = de != 0.f, hg = dg != 0.f;
  if (!he && !hg) return true;
  float s;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(__fadd_rn(s2x, yy)));
  const float adt = fabsf(dt);                 // the distance image of a negative depth is positive
  const bool t_zero = adt == 0.f, t_pos = adt > 0.f;
  const bool bop19 = pp.visib_mode == BOP_VISIB_BOP19;
  const bool c_matters = bop19 ? !t_zero : t_pos;
  const float ks = 2.5e-7f * s;
  bool vg = false, ve = false;
  if (hg) {
    const float a = (dg - adt) * s;
    if (c_matters && !(fabsf(a - pp.delta_f32) > fmaf(ks, dg + adt, 3e-6f * fabsf(a)))) return false;
    const bool c = a <= pp.delta_f32;
    vg = bop19 ? (c || t_zero) : (c && t_pos);
  }
  if (he) {
    const float a = (de - adt) * s;
    if (c_matters && !(fabsf(a - pp.delta_f32) > fmaf(ks, de + adt, 3e-6f * fabsf(a)))) return false;
    const bool c = a <= pp.delta_f32;
    ve = bop19 ? (c || t_zero) : (c && t_pos);
    ve = ve || vg;
  }
  if (!(vg || ve)) return true;
  if (vg && ve) {
    const float dn = fabsf(dg - de) * s * inv_diam;
    if constexpr (HIST) {
      // cnt = #(tau_k <= dn).  The taus ascend, so only the neighbours tau_tab[cnt] <= dn <
      // tau_tab[cnt + 1] can lie within their band of dn.
      int cnt = min(max(__float2int_rd(fmaf(dn, pp.guess_scale, pp.guess_off)), 0), pp.n_taus);
      float lo = s_tau[cnt], hi = s_tau[cnt + 1];
      if (!(lo <= dn && dn < hi)) {   // unevenly spaced taus, or a rounding of the guess at a threshold
        cnt = 0;
#pragma unroll
        for (int k = 0; k < NT; ++k) cnt += (dn >= pp.tau_f[k]) ? 1 : 0;
        lo = s_tau[cnt]; hi = s_tau[cnt + 1];
      }
      if (!(dn - lo > fmaf(3e-6f, fabsf(lo), 1e-30f)) || !(hi - dn > fmaf(3e-6f, fabsf(hi), 1e-30f))) return false;
      acc.inter += 1;
      acc.hist += 1ull << (kHistBits * cnt);
    } else {
      bool near = false;
#pragma unroll
      for (int k = 0; k < NT; ++k) near = near || !(fabsf(dn - pp.tau_f[k]) > pp.tau_eps[k]);
      if (near) return false;
      acc.inter += 1;
#pragma unroll
      for (int k = 0; k < NT; ++k) acc.c[k] += (dn >= pp.tau_f[k]) ? 1 : 0;
    }
  }
  acc.uni += 1;
  return true;
}

// Loads 4 horizontally adjacent depth_test samples as float32 millimetres.
__device__ __forceinline__ void load_depth4(const PixParams& pp, const void* __restrict__ depth,
                                            float scale, int im, int row, int col, float out[4]) {
  const size_t base = ((size_t)im * pp.H + row) * (size_t)pp.W + col;
  if (pp.depth_dtype == BOP_DEPTH_F32) {
    const float* p = (const float*)depth + base;
    if (pp.vec_ok && col + 3 < pp.W) {
      const float4 v = ld_stream_f4(p);
      out[0] = v.x; out[1] = v.y; out[2] = v.z; out[3] = v.w;
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) out[k] = (col + k < pp.W) ? __ldg(p + k) : 0.f;
    }
  } else {
    const unsigned short* p = (const unsigned short*)depth + base;
    unsigned short u[4];
    if (pp.vec_ok && col + 3 < pp.W) {
      const uint2 v = ld_stream_u2(p);
      u[0] = v.x & 0xffff; u[1] = v.x >> 16; u[2] = v.y & 0xffff; u[3] = v.y >> 16;
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) u[k] = (col + k < pp.W) ? __ldg(p + k) : 0;
    }
    // depth_im (float32) *= depth_scale  (scripts/eval_calc_errors.py:243): float32 product.
#pragma unroll
    for (int k = 0; k < 4; ++k) out[k] = __fmul_rn((float)u[k], scale);
  }
}

}  // namespace

// =====================================================================================
// Fused persistent kernel: rasterise (K1) + pixel pass (K2) per (pair, row band).
// MODE_RENDER writes the z-tile to depth_out instead of running the pixel pass.
// =====================================================================================
namespace {

constexpr int kQueueCap = 2048;   // deferred (large) triangles per pose and work item
constexpr int kPrefetchWarpBytes = 1152;     // one batch of 32 records (36 B each) in flight per warp

struct FusedArgs {
  const bop_model_dev* models;
  const PairDesc* desc;
  const WorkItem* work;
  PlanCounters* ctr;
  // Assembled triangles (assemble_kernel): [n][n_poses][tri_cap] records in three streams,
  // tri_cnt[n][n_poses] of them valid.  ta = (x0,y0,x1,y1), tb = (x2,y2,iz0,iz1), tc = iz2.
  const int4* tri_a;
  const int4* tri_b;
  const float* tri_c;
  const int* tri_cnt;
  const int* band_off;          // [n][n_poses][kBandOffN] segment starts of every list (geometry_kernel)
  int tri_cap;
  const void* depth_test;       // VSD
  const float* depth_scale;     // VSD, u16 input only
  unsigned long long* counts;   // VSD step: [n][2+BOP_MAX_TAUS] u64
  double* tl_sums;              // VSD tlinear: [n][BOP_MAX_TAUS] f64
  float* depth_out;             // RENDER: [n][H][W]
  int* mask_bbox;               // MASK: [n][8] = xmin,ymin,xmax,ymax of the est mask, then of the gt mask
  unsigned char* mask_full;     // GTINFO (optional): [n][H][W] 255 where the object projects
  unsigned char* mask_visib;    // GTINFO (optional): [n][H][W] 255 where it is visible
  int canvas_w;                 // clip width of the raster canvas (= pp.W except GTINFO)
  int mx, my;                   // GTINFO: margin of the canvas around the image (canvas = image + 2 margins)
  unsigned max_items;
  int tile_px;
  int small_tri_px;
  PixParams pp;
};

// VARIANT 0: <= 512 threads, 1 CTA/SM (up to 128 registers); VARIANT 1: up to 1024 resident
// threads per SM (64 registers): 1024-thread CTAs or two 512-thread CTAs per SM.
// NT: compile-time length of the tau sweep (>= pp.n_taus; only the VSD modes use it).
// HIST: step cost accumulated as a histogram over ascending taus (see PixParams).
template <int MODE, int VARIANT, int NT, bool HIST>
__global__ void __launch_bounds__(VARIANT == 0 ? 512 : 1024, 1)
fused_kernel(const __grid_constant__ FusedArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int wpad = (a.pp.W + 3) & ~3;
  unsigned* const zt = (unsigned*)smem_raw;                       // [2][tile_px]
  double* const xh = (double*)(smem_raw + (size_t)2 * a.tile_px * 4);  // [wpad]       X = (x - cx) / fx
  double* const yh = xh + wpad;                                   // [kMaxBandRows]  Y = (y - cy) / fy
  float* const s2x = (float*)(yh + kMaxBandRows);                 // [wpad]          fl32(1 + X^2)
  float* const yyf = s2x + wpad;                                  // [kMaxBandRows]  fl32(Y^2)
  int* const queue = (int*)(yyf + kMaxBandRows);                  // [kQueueCap]
  int4* const prefetch = (int4*)(queue + kQueueCap);              // [nwarps] x 1152 B of triangle records in flight

  __shared__ PairDesc sD;
  __shared__ WorkItem sW;
  __shared__ unsigned s_ticket;
  __shared__ int s_q;
  __shared__ int s_next;
  // The triangles of a work item: the band's segment and the shared segment of pose 0, then of
  // pose 1 (geometry_kernel), addressed as one virtual index range.
  __shared__ int s_seg0[4];   // first list entry of each segment
  __shared__ int s_segc[5];   // virtual index where each segment starts; [4] = total
  __shared__ int s_slot[4];   // list (pair * n_poses + pose) each segment belongs to
  __shared__ long long s_base[4];
  __shared__ unsigned long long s_cnt[2 + BOP_MAX_TAUS + 1];
  __shared__ double s_sum[BOP_MAX_TAUS];
  __shared__ float s_tau[BOP_MAX_TAUS + 2];

  const int tid = threadIdx.x, nthr = blockDim.x;
  const int lane = tid & 31, warp = tid >> 5, nwarps = nthr >> 5;
  const unsigned total = min(a.ctr->total, a.max_items);
  if (HIST && tid < BOP_MAX_TAUS + 2) s_tau[tid] = a.pp.tau_tab[tid];

  for (;;) {
    __syncthreads();
    if (tid == 0) {
      const unsigned t = atomicAdd(&a.ctr->next, 1u);
      s_ticket = t;
      if (t < total) sW = a.work[t];
      s_q = 0;
      s_next = 0;
    }
    if (tid < 2 + BOP_MAX_TAUS + 1) s_cnt[tid] = 0ull;
    if (tid < BOP_MAX_TAUS) s_sum[tid] = 0.0;
    __syncthreads();
    if (s_ticket >= total) break;
    // Segment table of a raster phase: the band's segment and the shared segment of the estimate's
    // list of pair `pe` (tile 0), then of the GT list of pair `pg` (tile 1); -1 = not in this phase.
    auto set_segments = [&](int pe, int pg, int np, int nbs) {
      int c = 0;
      for (int sg = 0; sg < 4; ++sg) {
        const int p = sg >> 1;
        const int pair = p == 0 ? pe : pg;
        int lo = 0, hi = 0, slot = 0;
        if (pair >= 0 && p < np) {
          slot = pair * np + p;
          const int* off = a.band_off + (size_t)slot * kBandOffN;
          if (nbs == 0) { if (!(sg & 1)) { lo = off[0]; hi = off[1]; } }
          else if (sg & 1) { lo = off[nbs]; hi = off[nbs + 1]; }
          else { lo = off[sW.band]; hi = off[sW.band + 1]; }
        }
        s_seg0[sg] = lo; s_segc[sg] = c; s_slot[sg] = slot;
        s_base[sg] = (long long)slot * a.tri_cap + lo - c;   // list entry of virtual index g = s_base + g
        c += hi - lo;
      }
      s_segc[4] = c;
    };
    {
      // Cooperative copy of the descriptor of the group's first pair (word-wise).
      const int* src = (const int*)(a.desc + sW.pair);
      int* dst = (int*)&sD;
      for (int i = tid; i < (int)(sizeof(PairDesc) / 4); i += nthr) dst[i] = src[i];
      if (tid == nthr - 1) {
        const int np = src[offsetof(PairDesc, n_poses) / 4], nbs = src[offsetof(PairDesc, nbs) / 4];
        const int gn = src[offsetof(PairDesc, group_n) / 4];
        set_segments(gn > 1 ? -1 : sW.pair, sW.pair, np, nbs);   // a group rasterises its GT pose first, alone
      }
    }
    __syncthreads();
    const int w = sD.w, nrows = sW.nrows;
    const int cx0 = sD.x0, cx1 = min(sD.x0 + w - 1, a.canvas_w - 1);
    const int cy0 = sD.y0 + sW.row0, cy1 = cy0 + nrows - 1;
    const int npx = nrows * w;
    const int n_poses = sD.n_poses;
    const int n_grp = sD.group_n > 1 ? sD.group_n : 1;
    BOP_DCHECK(npx <= a.tile_px && nrows <= kMaxBandRows && (w & 3) == 0 && cx0 >= 0 && cy0 >= 0);
    // Clear the z-tiles (0 = nothing drawn; they hold max 1/Z, raster.cuh) and build the x-ray
    // table (misc.py:137: (xs - cx) / fx).
    {
      const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
      for (int p = (n_grp > 1 ? 1 : 0); p < n_poses; ++p) {   // a group clears the estimate's tile per estimate
        uint4* z4 = (uint4*)(zt + (size_t)p * a.tile_px);
        for (int i = tid; i < (npx >> 2); i += nthr) z4[i] = zero4;
      }
      if (MODE == MODE_STEP || MODE == MODE_TLINEAR) {
        for (int c = tid; c < w; c += nthr) {
          const double X = __ddiv_rn(__dsub_rn((double)(cx0 + c), sD.kcx), sD.kfx);
          xh[c] = X;
          s2x[c] = (float)(1.0 + X * X);
        }
        for (int r = tid; r < nrows; r += nthr) {
          const double Y = __ddiv_rn(__dsub_rn((double)(cy0 + r), sD.kcy), sD.kfy);   // misc.py:138
          yh[r] = Y;
          yyf[r] = (float)(Y * Y);
        }
      }
    }
    __syncthreads();

    // A raster phase: every triangle of the current segment table into the z-tiles.  Ends with all
    // warps past the barrier behind the deferred queue.
    auto raster_phase = [&]() {
      // One loop over the assembled triangles of BOTH poses (they draw into different z-tiles, so
      // no barrier is needed between them).  Triangles with many candidate pixels are deferred to
      // a queue and drawn warp-cooperatively afterwards.
      const int c1 = s_segc[1], c2 = s_segc[2], c3 = s_segc[3], nT = s_segc[4];
      // virtual index -> (pose, list entry)
      auto locate = [&](int g, int& p) -> size_t {
        const int sg = (g >= c1 ? 1 : 0) + (g >= c2 ? 1 : 0) + (g >= c3 ? 1 : 0);
        p = sg >> 1;
        BOP_DCHECK(g >= 0 && g < nT && s_seg0[sg] + (g - s_segc[sg]) < a.tri_cnt[s_slot[sg]]);
        return (size_t)(s_base[sg] + g);
      };
      // Warps claim 32 consecutive triangles at a time from a shared ticket (triangle cost varies
      // by class; static striding would leave warps idle at the barrier).  The records of the NEXT
      // batch are fetched by cp.async into a per-warp staging area (every lane copies and later
      // reads its own 36 bytes) while the current batch is rasterised, so the L2 / HBM latency of
      // the list is hidden without spending registers on it.
      auto ticket = [&]() {
        int b = 0;
        if (lane == 0) b = atomicAdd(&s_next, 32);
        return __shfl_sync(BOP_FULL_MASK, b, 0);
      };
      int4* const pre_a = prefetch + (size_t)warp * 64 + lane;   // [warp][2][32] int4, then [warp][32] float
      int4* const pre_b = pre_a + 32;
      float* const pre_c = (float*)(prefetch + (size_t)nwarps * 64) + warp * 32 + lane;
      const unsigned sa = smem_addr(pre_a), sb = smem_addr(pre_b), sc = smem_addr(pre_c);
      int pose_next = 0;
      auto fetch = [&](int base) {
        const int g = base + lane;
        if (g < nT) {
          const size_t ti = locate(g, pose_next);
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(sa), "l"(a.tri_a + ti) : "memory");
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(sb), "l"(a.tri_b + ti) : "memory");
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(sc), "l"(a.tri_c + ti) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
      };
      int base = ticket();
      if (base < nT) fetch(base);
      while (base < nT) {
        const int next = ticket();
        const int g = base + lane;
        const int p = pose_next;
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        const int4 ra = *pre_a, rb = *pre_b;
        const float rc = *pre_c;
        if (next < nT) fetch(next);   // the staging area is free again: this lane has its record in registers
        base = next;
        if (g < nT) {
          const PVert v0 = {ra.x, ra.y, __int_as_float(rb.z), 1};
          const PVert v1 = {ra.z, ra.w, __int_as_float(rb.w), 1};
          const PVert v2 = {rb.x, rb.y, rc, 1};
          TriSetup T;
          const int n = tri_setup(v0, v1, v2, cx0, cy0, cx1, cy1, T);
          if (n != 0) {
            unsigned* const z = zt + (size_t)p * a.tile_px;
            // the clipped candidate rectangle lies inside this work item's tile
            BOP_DCHECK(T.bx0 >= cx0 && T.bx1 <= cx1 && T.by0 >= cy0 && T.by1 <= cy1 && (T.by1 - cy0) * w + (T.bx1 - cx0) < a.tile_px);
            if (n > 0 && n <= a.small_tri_px) {
              tri_raster_small(T, z, cx0, cy0, w);
            } else {
              const int slot = atomicAdd(&s_q, 1);
              if (slot < kQueueCap) {
                queue[slot] = g;
              } else if (n > 0) {   // queue full: draw it here (slow but correct)
                tri_raster_small(T, z, cx0, cy0, w);
              } else {
                TriSetup64 T64;
                if (tri_setup64(v0, v1, v2, T, T64)) tri_raster_large(T64, 0, 1, z, cx0, cy0, w);
              }
            }
          }
        }
      }
      __syncthreads();
      {
        const int nq = min(s_q, kQueueCap);
        for (int qi = warp; qi < nq; qi += nwarps) {
          const int g = queue[qi];
          int p;
          const size_t ti = locate(g, p);
          const int4 ra = __ldg(a.tri_a + ti), rb = __ldg(a.tri_b + ti);
          const float rc = __ldg(a.tri_c + ti);
          const PVert v0 = {ra.x, ra.y, __int_as_float(rb.z), 1};
          const PVert v1 = {ra.z, ra.w, __int_as_float(rb.w), 1};
          const PVert v2 = {rb.x, rb.y, rc, 1};
          unsigned* const z = zt + (size_t)p * a.tile_px;
          TriSetup T;
          const int n = tri_setup(v0, v1, v2, cx0, cy0, cx1, cy1, T);
          if (n > 0) {
            tri_raster_warp(T, lane, z, cx0, cy0, w);
          } else if (n < 0) {
            TriSetup64 T64;
            if (tri_setup64(v0, v1, v2, T, T64)) tri_raster_large(T64, lane, 32, z, cx0, cy0, w);
          }
        }
      }
      __syncthreads();

    };
    // Phases of a work item.  A single pair: one raster phase (both poses) and its pixel pass.
    // A group: phase -1 rasterises the GT pose into tile 1, then every active estimate in turn is
    // rasterised into tile 0 and scored against it.
    for (int j = (n_grp == 1 ? 0 : -1); j < n_grp; ++j) {
    const int pair = sW.pair + (j > 0 ? j : 0);
    if (n_grp > 1 && j >= 0) {
      if (!a.desc[pair].active) continue;   // uniform over the CTA
      __syncthreads();                      // the previous estimate's pixel pass is done with tile 0 and the counters
      if (tid == 0) { s_q = 0; s_next = 0; set_segments(pair, -1, n_poses, sD.nbs); }
      if (tid < 2 + BOP_MAX_TAUS + 1) s_cnt[tid] = 0ull;
      if (tid < BOP_MAX_TAUS) s_sum[tid] = 0.0;
      const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
      uint4* z4 = (uint4*)zt;
      for (int i = tid; i < (npx >> 2); i += nthr) z4[i] = zero4;
      __syncthreads();
    }
    raster_phase();
    if (j < 0) continue;

    if (MODE == MODE_RENDER) {
      float* out = a.depth_out + (size_t)sD.im * a.pp.H * a.pp.W;
      for (int i = tid; i < npx; i += nthr) {
        const float zv = tile_depth(zt[i]);
        if (zv != 0.f) {
          const int r = i / w, c = i - r * w;
          out[(size_t)(cy0 + r) * a.pp.W + (cx0 + c)] = zv;
        }
      }
      continue;
    }

    if (MODE == MODE_MASK) {
      // pose_error.cus / cou_bb_proj (pose_error.py:256-286,300-330): mask = rendered depth > 0;
      // counts of the est / gt / intersection / union masks and the pixel bbox of each mask.
      int ne = 0, ng = 0, ni = 0, nu = 0;
      int bb[8] = {INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MAX, INT_MAX, INT_MIN, INT_MIN};
      for (int i = tid; i < npx; i += nthr) {
        const bool e = tile_depth(zt[i]) != 0.f, g = tile_depth(zt[a.tile_px + i]) != 0.f;
        if (!(e || g)) continue;
        const int r = i / w, x = cx0 + (i - r * w), y = cy0 + r;
        nu += 1; ni += (e && g) ? 1 : 0;
        if (e) { ne += 1; bb[0] = min(bb[0], x); bb[1] = min(bb[1], y); bb[2] = max(bb[2], x); bb[3] = max(bb[3], y); }
        if (g) { ng += 1; bb[4] = min(bb[4], x); bb[5] = min(bb[5], y); bb[6] = max(bb[6], x); bb[7] = max(bb[7], y); }
      }
      const int cnt[4] = {warp_sum(nu), warp_sum(ni), warp_sum(ne), warp_sum(ng)};
#pragma unroll
      for (int k = 0; k < 8; ++k)
        bb[k] = (k & 2) ? __reduce_max_sync(BOP_FULL_MASK, bb[k]) : __reduce_min_sync(BOP_FULL_MASK, bb[k]);
      if (lane == 0) {
        unsigned long long* c = a.counts + (size_t)sW.pair * (2 + BOP_MAX_TAUS);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (cnt[k]) atomicAdd(c + k, (unsigned long long)cnt[k]);
        int* o = a.mask_bbox + (size_t)sW.pair * 8;
        if (cnt[2]) { atomicMin(o + 0, bb[0]); atomicMin(o + 1, bb[1]); atomicMax(o + 2, bb[2]); atomicMax(o + 3, bb[3]); }
        if (cnt[3]) { atomicMin(o + 4, bb[4]); atomicMin(o + 5, bb[5]); atomicMax(o + 6, bb[6]); atomicMax(o + 7, bb[7]); }
      }
      continue;
    }

    if (MODE == MODE_GTINFO) {
      // scripts/calc_gt_info.py:113-175 / calc_gt_masks.py:110-133 for one GT instance: the pose
      // is rendered on a canvas larger than the image (margins mx, my); px_count_all counts the
      // whole silhouette, the visibility test (visibility.estimate_visib_mask_gt, bop19) runs on
      // the pixels inside the image.  Boxes are in IMAGE coordinates (may leave the image).
      int n_all = 0, n_valid = 0, n_vis = 0;
      int bb[8] = {INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MAX, INT_MAX, INT_MIN, INT_MIN};
      const float scale = (a.pp.depth_dtype == BOP_DEPTH_U16 && a.depth_scale) ? a.depth_scale[sD.im] : 1.f;
      const size_t im_base = (size_t)sD.im * a.pp.H * (size_t)a.pp.W;
      const size_t out_base = (size_t)sW.pair * a.pp.H * (size_t)a.pp.W;
      for (int i = tid; i < npx; i += nthr) {
        const float zv = tile_depth(zt[i]);
        if (zv == 0.f) continue;
        const int r = i / w, x = cx0 + (i - r * w) - a.mx, y = cy0 + r - a.my;
        n_all += 1;
        bb[0] = min(bb[0], x); bb[1] = min(bb[1], y); bb[2] = max(bb[2], x); bb[3] = max(bb[3], y);
        if (x < 0 || y < 0 || x >= a.pp.W || y >= a.pp.H) continue;
        const size_t off = (size_t)y * a.pp.W + x;
        float dt;
        if (a.pp.depth_dtype == BOP_DEPTH_F32) dt = __ldg((const float*)a.depth_test + im_base + off);
        else dt = __fmul_rn((float)__ldg((const unsigned short*)a.depth_test + im_base + off), scale);
        const double X = __ddiv_rn(__dsub_rn((double)x, sD.kcx), sD.kfx);
        const double Y = __ddiv_rn(__dsub_rn((double)y, sD.kcy), sD.kfy);
        const double Dt = dist_of(X, Y, dt), Dg = dist_of(X, Y, zv);
        if (Dt > 0.0) n_valid += 1;
        const bool vis = (__fsub_rn((float)Dg, (float)Dt) <= a.pp.delta_f32) || Dt == 0.0;
        if (a.mask_full) a.mask_full[out_base + off] = 255;
        if (vis) {
          n_vis += 1;
          bb[4] = min(bb[4], x); bb[5] = min(bb[5], y); bb[6] = max(bb[6], x); bb[7] = max(bb[7], y);
          if (a.mask_visib) a.mask_visib[out_base + off] = 255;
        }
      }
      const int cnt[3] = {warp_sum(n_all), warp_sum(n_valid), warp_sum(n_vis)};
#pragma unroll
      for (int k = 0; k < 8; ++k)
        bb[k] = (k & 2) ? __reduce_max_sync(BOP_FULL_MASK, bb[k]) : __reduce_min_sync(BOP_FULL_MASK, bb[k]);
      if (lane == 0) {
        unsigned long long* c = a.counts + (size_t)sW.pair * (2 + BOP_MAX_TAUS);
#pragma unroll
        for (int k = 0; k < 3; ++k)
          if (cnt[k]) atomicAdd(c + k, (unsigned long long)cnt[k]);
        int* o = a.mask_bbox + (size_t)sW.pair * 8;
        if (cnt[0]) { atomicMin(o + 0, bb[0]); atomicMin(o + 1, bb[1]); atomicMax(o + 2, bb[2]); atomicMax(o + 3, bb[3]); }
        if (cnt[2]) { atomicMin(o + 4, bb[4]); atomicMin(o + 5, bb[5]); atomicMax(o + 6, bb[6]); atomicMax(o + 7, bb[7]); }
      }
      continue;
    }

    // ---- pixel pass ------------------------------------------------------------------
    Acc<MODE, NT, HIST> acc;
    int n_loaded = 0;
    const int w4 = w >> 2;
    const int ngroups = nrows * w4;
    const double diameter = sD.diameter;
    const float inv_diam = a.pp.normalized ? (float)(1.0 / diameter) : 1.f;
    const bool fast = MODE == MODE_STEP && a.pp.fast != 0;
    const float scale = (a.pp.depth_dtype == BOP_DEPTH_U16 && a.depth_scale) ? a.depth_scale[sD.im] : 1.f;
    for (int g = tid; g < ngroups; g += nthr) {
      const int r = g / w4, c4 = g - r * w4;
      const uint4 ze = *(const uint4*)(zt + r * w + 4 * c4);
      const uint4 zg = *(const uint4*)(zt + a.tile_px + r * w + 4 * c4);
      if (((ze.x | ze.y | ze.z | ze.w) | (zg.x | zg.y | zg.z | zg.w)) == 0u)
        continue;   // no model pixel in this group: depth_test is not even loaded
      const int row = cy0 + r, col = cx0 + 4 * c4;
      BOP_DCHECK(row >= 0 && row < a.pp.H && col >= 0 && col < a.pp.W);
      float dt[4];
      load_depth4(a.pp, a.depth_test, scale, sD.im, row, col, dt);
      ++n_loaded;
      const float yy = yyf[r];
      const unsigned zev[4] = {ze.x, ze.y, ze.z, ze.w};
      const unsigned zgv[4] = {zg.x, zg.y, zg.z, zg.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float de = tile_depth(zev[k]), dg = tile_depth(zgv[k]);
        if constexpr (MODE == MODE_STEP) {
          if (fast && vsd_pixel_fast<NT, HIST>(a.pp, s_tau, s2x[4 * c4 + k], yy, dt[k], de, dg, inv_diam, acc))
            continue;
        }
        vsd_pixel<MODE, NT, HIST>(a.pp, xh[4 * c4 + k], yh[r], dt[k], de, dg, diameter, acc);
      }
    }
    // ---- reduce: warp shuffle -> shared atomics -> one global atomic per counter ---------
    {
      const int u = warp_sum(acc.uni), in = warp_sum(acc.inter), nl = warp_sum(n_loaded);
      if (lane == 0) {
        if (u) atomicAdd(&s_cnt[0], (unsigned long long)u);
        if (in) atomicAdd(&s_cnt[1], (unsigned long long)in);
        if (nl) atomicAdd(&a.ctr->depth_groups, (unsigned long long)nl);
      }
      if constexpr (HIST) {
        // bins 1 .. n_taus (bin 0 = below every tau adds to no counter)
#pragma unroll
        for (int c = 1; c <= NT; ++c) {
          if (c <= a.pp.n_taus) {
            const int s = warp_sum((int)((acc.hist >> (kHistBits * c)) & ((1u << kHistBits) - 1u)));
            if (lane == 0 && s) atomicAdd(&s_cnt[2 + c], (unsigned long long)s);
          }
        }
      } else {
#pragma unroll
        for (int k = 0; k < NT; ++k) {
          if (k < a.pp.n_taus) {
            if constexpr (MODE == MODE_TLINEAR) {
              const double s = warp_sum((double)acc.c[k]);
              if (lane == 0 && s != 0.0) atomicAdd(&s_sum[k], s);
            } else {
              const int s = warp_sum((int)acc.c[k]);
              if (lane == 0 && s) atomicAdd(&s_cnt[2 + k], (unsigned long long)s);
            }
          }
        }
      }
    }
    __syncthreads();
    if constexpr (HIST) {
      // counter k = #(d >= tau_k) = bins k + 1 .. n_taus
      if (tid < 2 + a.pp.n_taus) {
        unsigned long long v = 0ull;
        if (tid < 2) v = s_cnt[tid];
        else for (int c = tid - 1; c <= a.pp.n_taus; ++c) v += s_cnt[2 + c];
        if (v) atomicAdd(a.counts + (size_t)pair * (2 + BOP_MAX_TAUS) + tid, v);
      }
    } else if (tid < 2 + BOP_MAX_TAUS && s_cnt[tid])
      atomicAdd(a.counts + (size_t)pair * (2 + BOP_MAX_TAUS) + tid, s_cnt[tid]);
    if (MODE == MODE_TLINEAR && tid < a.pp.n_taus && s_sum[tid] != 0.0)
      atomicAdd(a.tl_sums + (size_t)pair * BOP_MAX_TAUS + tid, s_sum[tid]);
    }   // phases
  }
}

// errors[pair][k] = (cost_k + (union - inter)) / union, or 1.0 when union == 0
// (pose_error.py:61-65,75-76,90).
__global__ void vsd_finalize_kernel(const unsigned long long* __restrict__ counts,
                                    const double* __restrict__ tl_sums, int n, int n_taus,
                                    int cost_type, double* __restrict__ errors,
                                    long long* __restrict__ counts_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long* c = counts + (size_t)i * (2 + BOP_MAX_TAUS);
  const unsigned long long uni = c[0], inter = c[1];
  if (counts_out) {
    counts_out[(size_t)i * (2 + n_taus) + 0] = (long long)uni;
    counts_out[(size_t)i * (2 + n_taus) + 1] = (long long)inter;
  }
  for (int k = 0; k < n_taus; ++k) {
    double e;
    if (uni == 0ull) {
      e = 1.0;
    } else if (cost_type == BOP_COST_TLINEAR) {
      e = (tl_sums[(size_t)i * BOP_MAX_TAUS + k] + (double)(uni - inter)) / (double)uni;
    } else {
      e = (double)(c[2 + k] + (uni - inter)) / (double)uni;
    }
    errors[(size_t)i * n_taus + k] = e;
    if (counts_out) counts_out[(size_t)i * (2 + n_taus) + 2 + k] = (long long)c[2 + k];
  }
}

// pose_error.cus (pose_error.py:279-285): 1 - inter / union, 1.0 when the union is empty;
// pose_error.cou_bb_proj (:318-330) with misc.calc_2d_bbox (misc.py:202-218; w = xmax - xmin,
// h = ymax - ymin) and misc.iou (misc.py:236-263) on integer pixel coordinates.  A pair with an
// empty mask has no bbox (the reference raises on it): cou_bb is NaN there.
__global__ void mask_finalize_kernel(const unsigned long long* __restrict__ counts,
                                     const int* __restrict__ bbox, int n, double* __restrict__ cus,
                                     double* __restrict__ cou_bb, long long* __restrict__ counts_out,
                                     int* __restrict__ bbox_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long* c = counts + (size_t)i * (2 + BOP_MAX_TAUS);
  const unsigned long long uni = c[0], inter = c[1];
  if (cus) cus[i] = uni > 0 ? 1.0 - (double)inter / (double)uni : 1.0;
  if (counts_out)
    for (int k = 0; k < 4; ++k) counts_out[(size_t)i * 4 + k] = (long long)c[k];
  const int* b = bbox + (size_t)i * 8;
  if (bbox_out) {
    for (int p = 0; p < 2; ++p) {
      const bool has = c[2 + p] > 0;
      bbox_out[i * 8 + 4 * p + 0] = has ? b[4 * p + 0] : 0;
      bbox_out[i * 8 + 4 * p + 1] = has ? b[4 * p + 1] : 0;
      bbox_out[i * 8 + 4 * p + 2] = has ? b[4 * p + 2] - b[4 * p + 0] : -1;
      bbox_out[i * 8 + 4 * p + 3] = has ? b[4 * p + 3] - b[4 * p + 1] : -1;
    }
  }
  if (cou_bb) {
    if (c[2] == 0 || c[3] == 0) {
      cou_bb[i] = __longlong_as_double(0x7ff8000000000000ll);
    } else {
      const long long ax0 = b[0], ay0 = b[1], ax1 = b[2], ay1 = b[3];
      const long long bx0 = b[4], by0 = b[5], bx1 = b[6], by1 = b[7];
      const long long wi = min(ax1, bx1) - max(ax0, bx0), hi = min(ay1, by1) - max(ay0, by0);
      double iou = 0.0;
      if (wi > 0 && hi > 0) {
        const long long ai = wi * hi, aa = (ax1 - ax0) * (ay1 - ay0), ab = (bx1 - bx0) * (by1 - by0);
        iou = (double)ai / (double)(aa + ab - ai);
      }
      cou_bb[i] = 1.0 - iou;
    }
  }
}

// Record fields of scripts/calc_gt_info.py:146-175: px_count_all / valid / visib, visib_fract,
// bbox_obj and bbox_visib as (x, y, w, h) = misc.calc_2d_bbox; both boxes are [-1,-1,-1,-1]
// unless px_count_visib > 0 (:158-169).
__global__ void gt_info_finalize_kernel(const unsigned long long* __restrict__ counts,
                                        const int* __restrict__ bbox, int n,
                                        long long* __restrict__ counts_out,
                                        double* __restrict__ visib_fract, int* __restrict__ bbox_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long* c = counts + (size_t)i * (2 + BOP_MAX_TAUS);
  for (int k = 0; k < 3; ++k) counts_out[(size_t)i * 3 + k] = (long long)c[k];
  visib_fract[i] = c[0] > 0 ? (double)c[2] / (double)c[0] : 0.0;
  const int* b = bbox + (size_t)i * 8;
  for (int p = 0; p < 2; ++p) {
    const bool has = c[2] > 0;
    bbox_out[i * 8 + 4 * p + 0] = has ? b[4 * p + 0] : -1;
    bbox_out[i * 8 + 4 * p + 1] = has ? b[4 * p + 1] : -1;
    bbox_out[i * 8 + 4 * p + 2] = has ? b[4 * p + 2] - b[4 * p + 0] : -1;
    bbox_out[i * 8 + 4 * p + 3] = has ? b[4 * p + 3] - b[4 * p + 1] : -1;
  }
}

__global__ void fill_i32_pattern_kernel(int* __restrict__ p, int n8) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n8 * 8) p[i] = (i & 2) ? INT_MIN : INT_MAX;
}

}  // namespace

// =====================================================================================
// Stand-alone K2: pre-rendered depth images in HBM (e.g. from a foreign Renderer).
// grid = (n_pairs, chunks); each thread owns groups of 4 horizontally adjacent pixels.
// Pure streaming: 3 x 4 B per pixel, 128-bit loads; depth_test is only fetched where a model
// pixel exists.
// =====================================================================================
namespace {

constexpr int kK2Threads = 256;
constexpr int kK2GroupsPerThread = 8;

template <int MODE>
__global__ void __launch_bounds__(kK2Threads)
pixel_pass_kernel(const void* __restrict__ depth_test, const float* __restrict__ depth_scale,
                  const int32_t* __restrict__ im_idx, const double* __restrict__ K,
                  const float* __restrict__ depth_est, const float* __restrict__ depth_gt,
                  const double* __restrict__ diameters, unsigned long long* __restrict__ counts,
                  double* __restrict__ tl_sums, const __grid_constant__ PixParams pp) {
  const int pair = blockIdx.x;
  const int W = pp.W, H = pp.H;
  const int w4 = (W + 3) >> 2;
  const int ngroups = H * w4;
  const int g0 = blockIdx.y * (kK2Threads * kK2GroupsPerThread);
  if (g0 >= ngroups) return;
  const int im = im_idx[pair];
  const double* Kp = K + 9 * (size_t)im;
  const double fx = Kp[0], fy = Kp[4], cx = Kp[2], cy = Kp[5];
  const double diameter = diameters[pair];
  const float scale = (pp.depth_dtype == BOP_DEPTH_U16 && depth_scale) ? depth_scale[im] : 1.f;
  const float* pe = depth_est + (size_t)pair * H * W;
  const float* pg = depth_gt + (size_t)pair * H * W;
  const bool vec = pp.vec_ok != 0;

  Acc<MODE, BOP_MAX_TAUS, false> acc;
#pragma unroll 2
  for (int j = 0; j < kK2GroupsPerThread; ++j) {
    const int g = g0 + j * kK2Threads + threadIdx.x;
    if (g >= ngroups) break;
    const int row = g / w4, col = (g - row * w4) << 2;
    float de[4], dg[4];
    if (vec) {
      const float4 a = ld_stream_f4(pe + (size_t)row * W + col);
      const float4 b = ld_stream_f4(pg + (size_t)row * W + col);
      de[0] = a.x; de[1] = a.y; de[2] = a.z; de[3] = a.w;
      dg[0] = b.x; dg[1] = b.y; dg[2] = b.z; dg[3] = b.w;
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const bool in = col + k < W;
        de[k] = in ? __ldg(pe + (size_t)row * W + col + k) : 0.f;
        dg[k] = in ? __ldg(pg + (size_t)row * W + col + k) : 0.f;
      }
    }
    bool any = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) any = any || de[k] != 0.f || dg[k] != 0.f;
    if (!any) continue;
    float dt[4];
    load_depth4(pp, depth_test, scale, im, row, col, dt);
    const double Y = __ddiv_rn(__dsub_rn((double)row, cy), fy);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (de[k] != 0.f || dg[k] != 0.f) {
        const double X = __ddiv_rn(__dsub_rn((double)(col + k), cx), fx);
        vsd_pixel<MODE, BOP_MAX_TAUS, false>(pp, X, Y, dt[k], de[k], dg[k], diameter, acc);
      }
    }
  }
  __shared__ unsigned long long s_cnt[2 + BOP_MAX_TAUS];
  __shared__ double s_sum[BOP_MAX_TAUS];
  if (threadIdx.x < 2 + BOP_MAX_TAUS) s_cnt[threadIdx.x] = 0ull;
  if (threadIdx.x < BOP_MAX_TAUS) s_sum[threadIdx.x] = 0.0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int u = warp_sum(acc.uni), in = warp_sum(acc.inter);
  if (lane == 0) {
    if (u) atomicAdd(&s_cnt[0], (unsigned long long)u);
    if (in) atomicAdd(&s_cnt[1], (unsigned long long)in);
  }
#pragma unroll
  for (int k = 0; k < BOP_MAX_TAUS; ++k) {
    if (k < pp.n_taus) {
      if (MODE == MODE_TLINEAR) {
        const double s = warp_sum((double)acc.c[k]);
        if (lane == 0 && s != 0.0) atomicAdd(&s_sum[k], s);
      } else {
        const int s = warp_sum((int)acc.c[k]);
        if (lane == 0 && s) atomicAdd(&s_cnt[2 + k], (unsigned long long)s);
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < 2 + BOP_MAX_TAUS && s_cnt[threadIdx.x])
    atomicAdd(counts + (size_t)pair * (2 + BOP_MAX_TAUS) + threadIdx.x, s_cnt[threadIdx.x]);
  if (MODE == MODE_TLINEAR && threadIdx.x < pp.n_taus && s_sum[threadIdx.x] != 0.0)
    atomicAdd(tl_sums + (size_t)pair * BOP_MAX_TAUS + threadIdx.x, s_sum[threadIdx.x]);
}

}  // namespace

// =====================================================================================
// Host launchers / C ABI
// =====================================================================================
namespace {

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

constexpr size_t kGeomBudgetBytes = (size_t)8 << 30;   // assembled-triangle scratch per chunk

struct WsLayout {
  size_t ctr, counts, tl, bbox, mbox, desc, work, tri_cnt, band_off, goff, tri_a, tri_b, tri_c, stage, pv, total;
  unsigned max_items;
  int chunk;       // items processed per pass (the batch is split so that the scratch fits the budget)
  int pv_stride;   // PVert records per (item, pose) (global fallback for models above kSmemVerts)
  int tri_cap;     // assembled triangles per (item, pose)
};

// max_geom = max over the models used of max(n_pts, (n_faces + 1) / 2).
WsLayout ws_layout(int n, int H, int W, int n_poses, int max_geom) {
  WsLayout L;
  L.pv_stride = (max_geom + 7) & ~7;
  L.tri_cap = 2 * L.pv_stride;
  const bool pv_global = max_geom > kSmemVerts;
  const size_t per_item_pv = pv_global ? (size_t)n_poses * L.pv_stride * sizeof(PVert) : 0;
  const size_t per_item = per_item_pv + (size_t)n_poses * L.tri_cap * 40;
  size_t chunk = kGeomBudgetBytes / per_item;
  if (chunk < 1) chunk = 1;
  if (chunk > (size_t)(n > 0 ? n : 1)) chunk = (size_t)(n > 0 ? n : 1);
  L.chunk = (int)chunk;
  const int c = L.chunk;
  const int wp = (W + 3) & ~3;
  const int rows_per_band = kMinTilePx / (wp > 0 ? wp : 1);
  const size_t items_per = (size_t)H / (size_t)(rows_per_band > 0 ? rows_per_band : 1) + 2;
  size_t mi = items_per * (size_t)c;
  if (mi > 0x7fffffffull) mi = 0x7fffffffull;
  L.max_items = (unsigned)mi;
  const size_t ntri = (size_t)c * n_poses * L.tri_cap;
  size_t off = 0;
  L.ctr = off;     off = align_up(off + sizeof(PlanCounters), 256);
  L.counts = off;  off = align_up(off + (size_t)c * (2 + BOP_MAX_TAUS) * 8, 256);
  L.tl = off;      off = align_up(off + (size_t)c * BOP_MAX_TAUS * 8, 256);
  L.bbox = off;    off = align_up(off + (size_t)c * 2 * sizeof(int4), 256);
  L.mbox = off;    off = align_up(off + (size_t)c * 8 * sizeof(int), 256);
  L.desc = off;    off = align_up(off + (size_t)c * sizeof(PairDesc), 256);
  L.work = off;    off = align_up(off + (size_t)L.max_items * sizeof(WorkItem), 256);
  L.tri_cnt = off; off = align_up(off + (size_t)c * 2 * sizeof(int), 256);
  L.band_off = off; off = align_up(off + (size_t)c * 2 * kBandOffN * sizeof(int), 256);
  L.goff = off;    off = align_up(off + ((size_t)c + 1) * sizeof(int), 256);
  L.tri_a = off;   off = align_up(off + ntri * sizeof(int4), 256);
  L.tri_b = off;   off = align_up(off + ntri * sizeof(int4), 256);
  L.tri_c = off;   off = align_up(off + ntri * sizeof(float), 256);
  L.stage = off;   off = align_up(off + ntri * sizeof(unsigned), 256);
  L.pv = off;      off = align_up(off + (size_t)c * per_item_pv, 256);
  L.total = off;
  return L;
}

bool g_exact_pixels = false;   // bop_vsd_set_exact: disable the float32 decision path (tests)
bool g_no_hist = false;        // bop_vsd_set_exact(2): float32 decisions with per-tau counters instead of the histogram (tests)

int fill_pix_params(const bop_vsd_params* prm, const void* depth_test, PixParams& pp) {
  BOP_REQUIRE(prm, "vsd: null params");
  BOP_REQUIRE(prm->n_taus >= 1 && prm->n_taus <= BOP_MAX_TAUS, "vsd: n_taus must be in [1, BOP_MAX_TAUS]");
  BOP_REQUIRE(prm->height > 0 && prm->width > 0 && prm->width <= kMaxWidth, "vsd: bad image size");
  BOP_REQUIRE(prm->cost_type == BOP_COST_STEP || prm->cost_type == BOP_COST_TLINEAR,
              "Unknown pixel matching cost.");
  BOP_REQUIRE(prm->visib_mode == BOP_VISIB_BOP19 || prm->visib_mode == BOP_VISIB_BOP18,
              "Unknown visibility mode.");
  BOP_REQUIRE(prm->depth_dtype == BOP_DEPTH_F32 || prm->depth_dtype == BOP_DEPTH_U16,
              "vsd: unknown depth dtype");
  // Unused slots hold +inf: "d >= inf" and min(d / inf, 1) add nothing, so the sweep needs no length check.
  bool fast = prm->cost_type == BOP_COST_STEP && !g_exact_pixels && prm->delta == prm->delta;
  for (int k = 0; k < BOP_MAX_TAUS; ++k) {
    pp.taus[k] = k < prm->n_taus ? prm->taus[k] : (double)INFINITY;
    pp.tau_f[k] = (float)pp.taus[k];
    pp.tau_eps[k] = k < prm->n_taus ? 3e-6f * fabsf(pp.tau_f[k]) + 1e-30f : 0.f;
    if (k < prm->n_taus && !(pp.taus[k] == pp.taus[k])) fast = false;   // NaN tau: exact path only
  }
  pp.fast = fast ? 1 : 0;
  bool ascending = fast;
  for (int k = 1; k < prm->n_taus; ++k) ascending = ascending && pp.tau_f[k - 1] <= pp.tau_f[k];
  for (int k = 0; k < prm->n_taus; ++k) ascending = ascending && fabsf(pp.tau_f[k]) < 1e37f;
  pp.hist = ascending && !g_no_hist ? 1 : 0;
  pp.tau_tab[0] = -3e38f;
  for (int k = 0; k <= BOP_MAX_TAUS; ++k) pp.tau_tab[1 + k] = k < prm->n_taus ? pp.tau_f[k] : 3e38f;
  const float step = prm->n_taus >= 2 ? pp.tau_f[1] - pp.tau_f[0] : 0.f;
  pp.guess_scale = step > 0.f ? 1.f / step : 0.f;
  pp.guess_off = step > 0.f ? 1.f - pp.tau_f[0] / step : 0.f;
  pp.delta_f32 = (float)prm->delta;
  pp.n_taus = prm->n_taus;
  pp.normalized = prm->normalized_by_diameter;
  pp.visib_mode = prm->visib_mode;
  pp.H = prm->height;
  pp.W = prm->width;
  pp.depth_dtype = prm->depth_dtype;
  const size_t al = prm->depth_dtype == BOP_DEPTH_F32 ? 16 : 8;
  pp.vec_ok = (prm->width % 4 == 0) && (((uintptr_t)depth_test) % al == 0);
  return 0;
}

int g_sm_count = 0;
int sm_count() {
  if (g_sm_count == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  return g_sm_count;
}

void set_geometry(FusedArgs& a, unsigned char* ws, const WsLayout& L) {
  a.tri_a = (const int4*)(ws + L.tri_a);
  a.tri_b = (const int4*)(ws + L.tri_b);
  a.tri_c = (const float*)(ws + L.tri_c);
  a.tri_cnt = (const int*)(ws + L.tri_cnt);
  a.band_off = (const int*)(ws + L.band_off);
  a.tri_cap = L.tri_cap;
}

// z-tile capacity actually used for images of width W: the tuned value, lowered when the per-column
// ray tables of a wide image would not leave room for it in shared memory.
// Dynamic shared memory one CTA of the fused kernel may use: the SM's 227 KB split over the resident
// CTAs, minus the kernel's static shared memory and the 1 KB the driver reserves per CTA.
long long fused_smem_budget() {
  return (long long)227 * 1024 / g_tune.ctas_per_sm - 3072;
}
long long fused_smem_fixed(int W) {   // everything but the z-tiles
  const int wp = (W + 3) & ~3;
  return (long long)wp * 12 + (long long)kMaxBandRows * 12 + (long long)kQueueCap * 4 +
         (long long)(g_tune.threads / 32) * kPrefetchWarpBytes;
}
int effective_tile_px(int W) {
  const long long room = fused_smem_budget() - fused_smem_fixed(W);
  long long fit = (room / 8) & ~3ll;
  if (fit < kMinTilePx) fit = kMinTilePx;
  return g_tune.tile_px < fit ? g_tune.tile_px : (int)fit;
}

template <int MODE, int NT, bool HIST>
int launch_fused_nt(const FusedArgs& a, int W, cudaStream_t s) {
  const size_t smem = (size_t)2 * a.tile_px * 4 + (size_t)fused_smem_fixed(W);
  // (a tuning that asks for more CTAs per SM than its tile allows simply runs with fewer resident)
  BOP_REQUIRE(smem <= (size_t)225 * 1024, "fused kernel: z-tile + tables exceed shared memory (lower tile_px)");
  const int grid = sm_count() * g_tune.ctas_per_sm;
  const int slot = MODE == MODE_RENDER ? BOP_PROF_RENDER : ((MODE == MODE_MASK || MODE == MODE_GTINFO) ? BOP_PROF_MASK : BOP_PROF_VSD_FUSED);
  const bool dense = g_tune.threads * g_tune.ctas_per_sm > 512;
  if (dense)
    BOP_CUDA(cudaFuncSetAttribute(fused_kernel<MODE, 1, NT, HIST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  else
    BOP_CUDA(cudaFuncSetAttribute(fused_kernel<MODE, 0, NT, HIST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  bop_prof_begin(slot, s);
  if (dense)
    fused_kernel<MODE, 1, NT, HIST><<<grid, g_tune.threads, smem, s>>>(a);
  else
    fused_kernel<MODE, 0, NT, HIST><<<grid, g_tune.threads, smem, s>>>(a);
  bop_prof_end(slot, s);
  BOP_LAUNCH_CHECK();
  return 0;
}

template <int MODE>
int launch_fused(const FusedArgs& a, int W, cudaStream_t s) {
  if constexpr (MODE == MODE_STEP) {   // the step-cost sweep is unrolled to 4, 10 (the BOP19 default) or 16 taus
    // A thread sees at most 4 * ceil(tile_px / 4 / threads) pixels of a work item: the histogram
    // bins (kHistBits bits per thread) must hold that.
    const bool hist = a.pp.hist && 4 * ((a.tile_px / 4 + g_tune.threads - 1) / g_tune.threads) < (1 << kHistBits);
    if (a.pp.n_taus <= 4) return hist ? launch_fused_nt<MODE, 4, true>(a, W, s) : launch_fused_nt<MODE, 4, false>(a, W, s);
    if (a.pp.n_taus <= 10) return hist ? launch_fused_nt<MODE, 10, true>(a, W, s) : launch_fused_nt<MODE, 10, false>(a, W, s);
  }
  return launch_fused_nt<MODE, BOP_MAX_TAUS, false>(a, W, s);
}

// bbox + plan; shared by VSD and render.
int launch_plan(const bop_model_dev* models, const int32_t* obj_idx, const double* pose_a,
                const double* pose_b, const int32_t* im_idx, const double* K,
                const double* diameters, int cam_mode, int n,
                int n_poses, int H, int W, int gate, unsigned char* ws, const WsLayout& L,
                cudaStream_t s, int cx_off = 0, int cy_off = 0, const int32_t* group_off = nullptr,
                int n_groups = 0) {
  // counters + counts + tl sums are contiguous at the head of the workspace.
  BOP_CUDA(cudaMemsetAsync(ws, 0, L.bbox, s));
  GeomArgs g;
  g.models = models; g.obj_idx = obj_idx; g.pose_a = pose_a; g.pose_b = pose_b; g.im_idx = im_idx;
  g.K = K; g.diameters = diameters; g.cam_mode = cam_mode; g.n_poses = n_poses; g.H = H; g.W = W;
  g.group_off = group_off;
  g.gate = gate; g.tile_px = effective_tile_px(W); g.pv_stride = L.pv_stride; g.tri_cap = L.tri_cap;
  g.pv_global = (PVert*)(ws + L.pv); g.tri_a = (int4*)(ws + L.tri_a); g.tri_b = (int4*)(ws + L.tri_b);
  g.tri_c = (float*)(ws + L.tri_c); g.stage = (unsigned*)(ws + L.stage); g.tri_cnt = (int*)(ws + L.tri_cnt);
  g.band_off = (int*)(ws + L.band_off); g.desc = (PairDesc*)(ws + L.desc); g.work = (WorkItem*)(ws + L.work);
  g.ctr = (PlanCounters*)(ws + L.ctr); g.max_items = L.max_items; g.cx_off = cx_off; g.cy_off = cy_off;
  bop_prof_begin(BOP_PROF_VSD_PLAN, s);
  const size_t geom_smem = (size_t)(L.pv_stride < kSmemVerts ? L.pv_stride : kSmemVerts) * sizeof(PVert) +
                           (size_t)kMaxBins * sizeof(int);
  BOP_CUDA(cudaFuncSetAttribute(geometry_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)geom_smem));
  geometry_kernel<<<group_off ? n_groups : n, kAsmThreads, geom_smem, s>>>(g);
  bop_prof_end(BOP_PROF_VSD_PLAN, s);
  BOP_LAUNCH_CHECK();
  return 0;
}

}  // namespace

extern "C" int bop_vsd_set_tuning(int tile_px, int threads, int ctas_per_sm, int small_tri_px) {
  if (tile_px) {
    BOP_REQUIRE(tile_px >= kMinTilePx && tile_px % 4 == 0 && tile_px <= 26624, "tuning: tile_px out of range");
    g_tune.tile_px = tile_px;
  }
  if (threads) {
    BOP_REQUIRE(threads >= 64 && threads <= 1024 && threads % 32 == 0, "tuning: threads out of range");
    g_tune.threads = threads;
  }
  if (ctas_per_sm) {
    BOP_REQUIRE(ctas_per_sm >= 1 && ctas_per_sm <= 8, "tuning: ctas_per_sm out of range");
    g_tune.ctas_per_sm = ctas_per_sm;
  }
  if (small_tri_px) {
    BOP_REQUIRE(small_tri_px >= 1, "tuning: small_tri_px out of range");
    g_tune.small_tri_px = small_tri_px;
  }
  return 0;
}

// Test hook: exact != 0 makes every pixel take the float64 path (the float32 decision path must
// give bit-identical counts; tests compare the two).
extern "C" int bop_vsd_set_exact(int exact) {
  g_exact_pixels = exact == 1;
  g_no_hist = exact == 2;
  return 0;
}

// Bytes of depth_test the fused kernel fetched in the most recent pass that used `workspace`
// (synchronises the device).  With host-mapped (pinned) depth images this is the PCIe traffic.
extern "C" int bop_vsd_depth_bytes_fetched(const void* workspace, int depth_dtype, uint64_t* bytes) {
  BOP_REQUIRE(workspace && bytes, "bop_vsd_depth_bytes_fetched: null pointer");
  PlanCounters c;
  BOP_CUDA(cudaMemcpy(&c, workspace, sizeof(c), cudaMemcpyDeviceToHost));
  *bytes = c.depth_groups * (depth_dtype == BOP_DEPTH_U16 ? 8ull : 16ull);
  return 0;
}

// Test hook: copies the assembled-triangle list of (pair, pose) left in `workspace` by the most
// recent bop_vsd_dev call with these sizes to the host: xy[k][6] = snapped window positions
// (1/256 px) of the three vertices of list entry k.  Synchronises the device.
extern "C" int bop_vsd_debug_triangle_list(const void* workspace, int n_pairs, int height, int width,
                                           int max_geom, int pair, int pose, int32_t* xy, int cap,
                                           int* count) {
  BOP_REQUIRE(workspace && xy && count && pair >= 0 && pair < n_pairs && (pose == 0 || pose == 1),
              "bop_vsd_debug_triangle_list: bad arguments");
  const WsLayout L = ws_layout(n_pairs, height, width, 2, max_geom);
  BOP_REQUIRE(pair < L.chunk, "bop_vsd_debug_triangle_list: pair outside the last chunk");
  const unsigned char* ws = (const unsigned char*)workspace;
  BOP_CUDA(cudaDeviceSynchronize());
  int n = 0;
  BOP_CUDA(cudaMemcpy(&n, ws + L.tri_cnt + sizeof(int) * ((size_t)pair * 2 + pose), sizeof(int), cudaMemcpyDeviceToHost));
  *count = n;
  if (n > cap) n = cap;
  if (n <= 0) return 0;
  int4* ha = (int4*)malloc(sizeof(int4) * (size_t)n * 2);
  BOP_REQUIRE(ha, "bop_vsd_debug_triangle_list: out of memory");
  int4* hb = ha + n;
  const size_t off = ((size_t)pair * 2 + pose) * L.tri_cap;
  cudaError_t e = cudaMemcpy(ha, ws + L.tri_a + sizeof(int4) * off, sizeof(int4) * n, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(hb, ws + L.tri_b + sizeof(int4) * off, sizeof(int4) * n, cudaMemcpyDeviceToHost);
  for (int k = 0; e == cudaSuccess && k < n; ++k) {
    xy[6 * k + 0] = ha[k].x; xy[6 * k + 1] = ha[k].y; xy[6 * k + 2] = ha[k].z; xy[6 * k + 3] = ha[k].w;
    xy[6 * k + 4] = hb[k].x; xy[6 * k + 5] = hb[k].y;
  }
  free(ha);
  BOP_CUDA(e);
  return 0;
}

extern "C" size_t bop_vsd_workspace_bytes(int n_pairs, int height, int width, int max_geom) {
  return ws_layout(n_pairs, height, width, 2, max_geom).total;
}
extern "C" size_t bop_render_workspace_bytes(int n, int height, int width, int max_geom) {
  return ws_layout(n, height, width, 1, max_geom).total;
}

// pose_error.vsd for a batch of pairs.  group_off_host (host array, [n_groups + 1], NULL = every
// pair on its own): consecutive pairs [off[g], off[g + 1]) share object, image, GT pose and
// diameter -- the k estimates scripts/eval_calc_errors.py:281-289 pairs with one GT.  The geometry
// and raster work of the GT pose is then done once per group instead of once per pair.
static int vsd_impl(const bop_model_dev* models, const int32_t* obj_idx, const double* pose_est,
                    const double* pose_gt, const void* depth_test, const float* depth_scale,
                    const int32_t* im_idx, const double* K, const double* diameters, int n_pairs,
                    const int32_t* group_off_host, int n_groups, int max_geom,
                    const bop_vsd_params* prm, double* errors_out, int64_t* counts_out,
                    void* workspace, size_t workspace_bytes, void* stream, void* geometry_stream = nullptr) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(models && obj_idx && pose_est && pose_gt && depth_test && im_idx && K && errors_out,
              "bop_vsd_dev: null pointer");
  BOP_REQUIRE(max_geom >= 1, "bop_vsd_dev: max_geom must be >= 1");
  PixParams pp;
  if (int rc = fill_pix_params(prm, depth_test, pp)) return rc;
  const WsLayout L = ws_layout(n_pairs, pp.H, pp.W, 2, max_geom);
  BOP_REQUIRE(workspace && workspace_bytes >= L.total, "bop_vsd_dev: workspace too small");
  BOP_REQUIRE(((uintptr_t)workspace) % 256 == 0, "bop_vsd_dev: workspace must be 256-byte aligned");
  if (group_off_host) {
    BOP_REQUIRE(n_groups >= 1 && group_off_host[0] == 0 && group_off_host[n_groups] == n_pairs,
                "bop_vsd_grouped_dev: group offsets must start at 0 and end at n_pairs");
    for (int gI = 0; gI < n_groups; ++gI) {
      const int k = group_off_host[gI + 1] - group_off_host[gI];
      BOP_REQUIRE(k >= 1 && k <= kMaxGroup, "bop_vsd_grouped_dev: a group holds 1 .. 32 pairs");
    }
  }
  cudaStream_t s = (cudaStream_t)stream;
  // Two-stream form (bop_vsd_split_dev): the geometry pass runs on `sg`, the raster + pixel pass on
  // `s`; an event orders each chunk's raster pass behind its geometry pass and the next chunk's
  // geometry pass behind the raster pass that still reads the workspace.
  cudaStream_t sg = geometry_stream ? (cudaStream_t)geometry_stream : s;
  const bool split = sg != s;
  auto order = [](cudaStream_t first, cudaStream_t then) -> int {
    cudaEvent_t ev;
    BOP_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    cudaError_t e = cudaEventRecord(ev, first);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(then, ev, 0);
    cudaEventDestroy(ev);   // released once the recorded work has completed
    BOP_CUDA(e);
    return 0;
  };
  unsigned char* ws = (unsigned char*)workspace;
  std::vector<int32_t> rebased;
  int g0 = 0;   // first group of the chunk
  for (int c0 = 0; c0 < n_pairs;) {
    if (split && c0 > 0) {
      if (int rc = order(s, sg)) return rc;
    }
    int nc, ng = 0;
    if (group_off_host) {   // whole groups, at most L.chunk pairs (a group never exceeds a chunk: chunk >= kMaxGroup or n_pairs)
      int g1 = g0;
      while (g1 < n_groups && group_off_host[g1 + 1] - c0 <= (L.chunk > kMaxGroup ? L.chunk : kMaxGroup)) ++g1;
      BOP_REQUIRE(g1 > g0, "bop_vsd_grouped_dev: workspace chunk smaller than a group");
      nc = group_off_host[g1] - c0;
      ng = g1 - g0;
      rebased.resize((size_t)ng + 1);
      for (int k = 0; k <= ng; ++k) rebased[k] = group_off_host[g0 + k] - c0;
      BOP_CUDA(cudaMemcpyAsync(ws + L.goff, rebased.data(), sizeof(int32_t) * ((size_t)ng + 1), cudaMemcpyHostToDevice, sg));
      BOP_CUDA(cudaStreamSynchronize(sg));   // `rebased` is reused by the next chunk
      g0 = g1;
    } else {
      nc = n_pairs - c0 < L.chunk ? n_pairs - c0 : L.chunk;
    }
    if (int rc = launch_plan(models, obj_idx + c0, pose_est + 12 * (size_t)c0, pose_gt + 12 * (size_t)c0,
                             im_idx + c0, K, diameters ? diameters + c0 : nullptr, 0, nc, 2, pp.H,
                             pp.W, prm->gate, ws, L, sg, 0, 0,
                             group_off_host ? (const int32_t*)(ws + L.goff) : nullptr, ng))
      return rc;
    if (split) {
      if (int rc = order(sg, s)) return rc;
    }
    FusedArgs a;
    a.models = models;
    a.desc = (const PairDesc*)(ws + L.desc);
    a.work = (const WorkItem*)(ws + L.work);
    a.ctr = (PlanCounters*)(ws + L.ctr);
    set_geometry(a, ws, L);
    a.depth_test = depth_test;
    a.depth_scale = depth_scale;
    a.counts = (unsigned long long*)(ws + L.counts);
    a.tl_sums = (double*)(ws + L.tl);
    a.depth_out = nullptr;
    a.max_items = L.max_items;
    a.tile_px = effective_tile_px(pp.W);
    a.small_tri_px = g_tune.small_tri_px;
    a.pp = pp;
    a.canvas_w = pp.W;
    a.mx = a.my = 0;
    a.mask_full = a.mask_visib = nullptr;
    a.mask_bbox = nullptr;
    int rc = prm->cost_type == BOP_COST_TLINEAR ? launch_fused<MODE_TLINEAR>(a, pp.W, s)
                                                : launch_fused<MODE_STEP>(a, pp.W, s);
    if (rc) return rc;
    vsd_finalize_kernel<<<(nc + 127) / 128, 128, 0, s>>>(
        a.counts, a.tl_sums, nc, pp.n_taus, prm->cost_type, errors_out + (size_t)c0 * pp.n_taus,
        counts_out ? (long long*)counts_out + (size_t)c0 * (2 + pp.n_taus) : nullptr);
    BOP_LAUNCH_CHECK();
    c0 += nc;
  }
  return 0;
}

extern "C" int bop_vsd_dev(const bop_model_dev* models, const int32_t* obj_idx,
                           const double* pose_est, const double* pose_gt, const void* depth_test,
                           const float* depth_scale, const int32_t* im_idx, const double* K,
                           const double* diameters, int n_pairs, int max_geom,
                           const bop_vsd_params* prm, double* errors_out, int64_t* counts_out,
                           void* workspace, size_t workspace_bytes, void* stream) {
  return vsd_impl(models, obj_idx, pose_est, pose_gt, depth_test, depth_scale, im_idx, K, diameters, n_pairs,
                  nullptr, 0, max_geom, prm, errors_out, counts_out, workspace, workspace_bytes, stream);
}

extern "C" int bop_vsd_split_dev(const bop_model_dev* models, const int32_t* obj_idx,
                                 const double* pose_est, const double* pose_gt, const void* depth_test,
                                 const float* depth_scale, const int32_t* im_idx, const double* K,
                                 const double* diameters, int n_pairs, int max_geom,
                                 const bop_vsd_params* prm, double* errors_out, int64_t* counts_out,
                                 void* workspace, size_t workspace_bytes, void* geometry_stream,
                                 void* raster_stream) {
  return vsd_impl(models, obj_idx, pose_est, pose_gt, depth_test, depth_scale, im_idx, K, diameters, n_pairs,
                  nullptr, 0, max_geom, prm, errors_out, counts_out, workspace, workspace_bytes, raster_stream,
                  geometry_stream);
}

extern "C" int bop_vsd_grouped_dev(const bop_model_dev* models, const int32_t* obj_idx,
                                   const double* pose_est, const double* pose_gt, const void* depth_test,
                                   const float* depth_scale, const int32_t* im_idx, const double* K,
                                   const double* diameters, int n_pairs, const int32_t* group_off_host,
                                   int n_groups, int max_geom, const bop_vsd_params* prm,
                                   double* errors_out, int64_t* counts_out, void* workspace,
                                   size_t workspace_bytes, void* stream) {
  BOP_REQUIRE(group_off_host, "bop_vsd_grouped_dev: null group offsets");
  return vsd_impl(models, obj_idx, pose_est, pose_gt, depth_test, depth_scale, im_idx, K, diameters, n_pairs,
                  group_off_host, n_groups, max_geom, prm, errors_out, counts_out, workspace, workspace_bytes,
                  stream);
}

extern "C" int bop_render_depth_dev(const bop_model_dev* models, const int32_t* obj_idx,
                                    const double* pose, const double* cam, int n, int max_geom,
                                    int height, int width, float* depth_out, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  if (n <= 0) return 0;
  BOP_REQUIRE(models && obj_idx && pose && cam && depth_out, "bop_render_depth_dev: null pointer");
  BOP_REQUIRE(height > 0 && width > 0 && width <= kMaxWidth, "bop_render_depth_dev: bad image size");
  BOP_REQUIRE(max_geom >= 1, "bop_render_depth_dev: max_geom must be >= 1");
  const WsLayout L = ws_layout(n, height, width, 1, max_geom);
  BOP_REQUIRE(workspace && workspace_bytes >= L.total, "bop_render_depth_dev: workspace too small");
  BOP_REQUIRE(((uintptr_t)workspace) % 256 == 0, "bop_render_depth_dev: workspace must be 256-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  unsigned char* ws = (unsigned char*)workspace;
  BOP_CUDA(cudaMemsetAsync(depth_out, 0, sizeof(float) * (size_t)n * height * width, s));
  for (int c0 = 0; c0 < n; c0 += L.chunk) {
    const int nc = n - c0 < L.chunk ? n - c0 : L.chunk;
    if (int rc = launch_plan(models, obj_idx + c0, pose + 12 * (size_t)c0, pose + 12 * (size_t)c0, nullptr,
                             cam + 4 * (size_t)c0, nullptr, 1, nc, 1, height, width, 0, ws, L, s))
      return rc;
    FusedArgs a;
    memset(&a, 0, sizeof(a));
    a.models = models;
    a.desc = (const PairDesc*)(ws + L.desc);
    a.work = (const WorkItem*)(ws + L.work);
    a.ctr = (PlanCounters*)(ws + L.ctr);
    set_geometry(a, ws, L);
    a.depth_out = depth_out + (size_t)c0 * height * width;
    a.max_items = L.max_items;
    a.tile_px = effective_tile_px(width);
    a.small_tri_px = g_tune.small_tri_px;
    a.pp.H = height;
    a.pp.W = width;
    a.canvas_w = width;
    a.pp.n_taus = 1;
    if (int rc = launch_fused<MODE_RENDER>(a, width, s)) return rc;
  }
  return 0;
}

extern "C" int bop_vsd_from_depth_dev(const void* depth_test, const float* depth_scale,
                                      const int32_t* im_idx, const double* K,
                                      const float* depth_est, const float* depth_gt,
                                      const double* diameters, int n_pairs,
                                      const bop_vsd_params* prm, double* errors_out,
                                      int64_t* counts_out, void* workspace,
                                      size_t workspace_bytes, void* stream) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(depth_test && im_idx && K && depth_est && depth_gt && diameters && errors_out,
              "bop_vsd_from_depth_dev: null pointer");
  PixParams pp;
  if (int rc = fill_pix_params(prm, depth_test, pp)) return rc;
  pp.vec_ok = pp.vec_ok && ((uintptr_t)depth_est % 16 == 0) && ((uintptr_t)depth_gt % 16 == 0);
  // Only the counters at the head of the layout are used here (no projected vertices).
  WsLayout L = ws_layout(n_pairs, pp.H, pp.W, 2, 1);
  {
    size_t off = 0;
    L.ctr = off;    off = align_up(off + sizeof(PlanCounters), 256);
    L.counts = off; off = align_up(off + (size_t)n_pairs * (2 + BOP_MAX_TAUS) * 8, 256);
    L.tl = off;     off = align_up(off + (size_t)n_pairs * BOP_MAX_TAUS * 8, 256);
    L.bbox = off;
  }
  BOP_REQUIRE(workspace && workspace_bytes >= L.bbox, "bop_vsd_from_depth_dev: workspace too small");
  cudaStream_t s = (cudaStream_t)stream;
  unsigned char* ws = (unsigned char*)workspace;
  BOP_CUDA(cudaMemsetAsync(ws, 0, L.bbox, s));
  unsigned long long* counts = (unsigned long long*)(ws + L.counts);
  double* tl = (double*)(ws + L.tl);
  const int ngroups = pp.H * ((pp.W + 3) >> 2);
  const int chunk = kK2Threads * kK2GroupsPerThread;
  dim3 grid(n_pairs, (ngroups + chunk - 1) / chunk);
  bop_prof_begin(BOP_PROF_K2, s);
  if (prm->cost_type == BOP_COST_TLINEAR)
    pixel_pass_kernel<MODE_TLINEAR><<<grid, kK2Threads, 0, s>>>(depth_test, depth_scale, im_idx, K, depth_est,
                                                               depth_gt, diameters, counts, tl, pp);
  else
    pixel_pass_kernel<MODE_STEP><<<grid, kK2Threads, 0, s>>>(depth_test, depth_scale, im_idx, K, depth_est,
                                                            depth_gt, diameters, counts, tl, pp);
  bop_prof_end(BOP_PROF_K2, s);
  BOP_LAUNCH_CHECK();
  vsd_finalize_kernel<<<(n_pairs + 127) / 128, 128, 0, s>>>(counts, tl, n_pairs, pp.n_taus,
                                                            prm->cost_type, errors_out,
                                                            (long long*)counts_out);
  BOP_LAUNCH_CHECK();
  return 0;
}

// =====================================================================================
// bop_vsd_from_depth_static_dev: bop_vsd_from_depth_dev for callers that must not allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// No counterpart in the reference.
//
// The scratch of the K2 pass -- plan counters, per-pair integer counts [n][2 + BOP_MAX_TAUS] u64
// and tlinear sums [n][BOP_MAX_TAUS] f64: 272 B per pair -- lives in a file-scope __device__
// array sized for BOP_STATIC_MAX_PAIRS pairs (1,114,368 B = 1.06 MiB; every GPU that loads the
// library holds it until the process exits).  In HOST code `g_k2_static_scratch` is only a
// stand-in for the symbol, not a device address: the address is looked up once per device, with
// that device current, by cudaGetSymbolAddress and that pointer is what the kernels receive.
// The function calls no cudaMalloc / cudaFree / allocating library routine (it also leaves the
// optional per-kernel profiling events alone, which are created on first use).  Every call on a
// device shares the one array, which keeps the previous call's contents: the part a call uses is
// cleared at the start of the call (a memset node on the caller's stream), and two calls that
// OVERLAP on the GPU (different streams, or graphs replayed concurrently) race on it -- keep
// calls of this entry point on one stream or order them with events.
// =====================================================================================
namespace {

constexpr int kStaticMaxPairs = BOP_STATIC_MAX_PAIRS;
constexpr size_t kStaticCtrBytes = 256;
constexpr size_t kStaticCountsBytes = (size_t)kStaticMaxPairs * (2 + BOP_MAX_TAUS) * 8;   // multiple of 256
constexpr size_t kStaticTlBytes = (size_t)kStaticMaxPairs * BOP_MAX_TAUS * 8;             // multiple of 256
constexpr size_t kStaticScratchBytes = kStaticCtrBytes + kStaticCountsBytes + kStaticTlBytes;
static_assert(kStaticCountsBytes % 256 == 0 && kStaticTlBytes % 256 == 0, "static scratch layout must stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_k2_static_scratch[kStaticScratchBytes];

constexpr int kStaticMaxDevices = 64;
unsigned char* g_k2_static_ptr[kStaticMaxDevices];   // device address of the array, per device (0 = not looked up yet)
std::mutex g_k2_static_mutex;

// Device address of g_k2_static_scratch on the CURRENT device (looked up on the first call per device).
int k2_static_scratch(unsigned char** out) {
  int dev = -1;
  BOP_CUDA(cudaGetDevice(&dev));
  BOP_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "bop_vsd_from_depth_static_dev: device index out of range");
  std::lock_guard<std::mutex> lock(g_k2_static_mutex);
  if (!g_k2_static_ptr[dev]) {
    void* ptr = nullptr;
    BOP_CUDA(cudaGetSymbolAddress(&ptr, g_k2_static_scratch));
    BOP_REQUIRE(ptr && ((uintptr_t)ptr) % 256 == 0, "bop_vsd_from_depth_static_dev: bad symbol address");
    g_k2_static_ptr[dev] = (unsigned char*)ptr;
  }
  *out = g_k2_static_ptr[dev];
  return 0;
}

}  // namespace

extern "C" int bop_vsd_from_depth_static_dev(const void* depth_test, const float* depth_scale,
                                             const int32_t* im_idx, const double* K,
                                             const float* depth_est, const float* depth_gt,
                                             const double* diameters, int n_pairs,
                                             const bop_vsd_params* prm, double* errors_out,
                                             int64_t* counts_out, void* stream) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(n_pairs <= kStaticMaxPairs,
              "bop_vsd_from_depth_static_dev: n_pairs exceeds BOP_STATIC_MAX_PAIRS (4096); split the batch or use "
              "bop_vsd_from_depth_dev with a caller-provided workspace");
  BOP_REQUIRE(depth_test && im_idx && K && depth_est && depth_gt && diameters && errors_out,
              "bop_vsd_from_depth_static_dev: null pointer");
  PixParams pp;
  if (int rc = fill_pix_params(prm, depth_test, pp)) return rc;
  pp.vec_ok = pp.vec_ok && ((uintptr_t)depth_est % 16 == 0) && ((uintptr_t)depth_gt % 16 == 0);
  unsigned char* ws = nullptr;
  if (int rc = k2_static_scratch(&ws)) return rc;
  unsigned long long* counts = (unsigned long long*)(ws + kStaticCtrBytes);
  double* tl = (double*)(ws + kStaticCtrBytes + kStaticCountsBytes);
  cudaStream_t s = (cudaStream_t)stream;
  // the array keeps the previous call's contents: clear the counters and this call's rows
  BOP_CUDA(cudaMemsetAsync(ws, 0, kStaticCtrBytes + (size_t)n_pairs * (2 + BOP_MAX_TAUS) * 8, s));
  BOP_CUDA(cudaMemsetAsync(tl, 0, (size_t)n_pairs * BOP_MAX_TAUS * 8, s));
  const int ngroups = pp.H * ((pp.W + 3) >> 2);
  const int chunk = kK2Threads * kK2GroupsPerThread;
  dim3 grid(n_pairs, (ngroups + chunk - 1) / chunk);
  if (prm->cost_type == BOP_COST_TLINEAR)
    pixel_pass_kernel<MODE_TLINEAR><<<grid, kK2Threads, 0, s>>>(depth_test, depth_scale, im_idx, K, depth_est,
                                                               depth_gt, diameters, counts, tl, pp);
  else
    pixel_pass_kernel<MODE_STEP><<<grid, kK2Threads, 0, s>>>(depth_test, depth_scale, im_idx, K, depth_est,
                                                            depth_gt, diameters, counts, tl, pp);
  BOP_LAUNCH_CHECK();
  vsd_finalize_kernel<<<(n_pairs + 127) / 128, 128, 0, s>>>(counts, tl, n_pairs, pp.n_taus, prm->cost_type,
                                                            errors_out, (long long*)counts_out);
  BOP_LAUNCH_CHECK();
  return 0;
}

// pose_error.cus (pose_error.py:256-286) and pose_error.cou_bb_proj (:300-330) for a batch of
// pose pairs: both poses are rasterised into shared-memory z-tiles and only mask statistics
// leave the SM.  cam: [n][4] fx,fy,cx,cy per pair.
extern "C" int bop_mask_overlap_dev(const bop_model_dev* models, const int32_t* obj_idx,
                                    const double* pose_est, const double* pose_gt,
                                    const double* cam, int n_pairs, int max_geom, int height,
                                    int width, int gate, double* cus_out, double* cou_bb_out,
                                    int64_t* counts_out, int32_t* bbox_out, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(models && obj_idx && pose_est && pose_gt && cam, "bop_mask_overlap_dev: null pointer");
  BOP_REQUIRE(height > 0 && width > 0 && width <= kMaxWidth, "bop_mask_overlap_dev: bad image size");
  BOP_REQUIRE(max_geom >= 1, "bop_mask_overlap_dev: max_geom must be >= 1");
  const WsLayout L = ws_layout(n_pairs, height, width, 2, max_geom);
  BOP_REQUIRE(workspace && workspace_bytes >= L.total, "bop_mask_overlap_dev: workspace too small");
  BOP_REQUIRE(((uintptr_t)workspace) % 256 == 0, "bop_mask_overlap_dev: workspace must be 256-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  unsigned char* ws = (unsigned char*)workspace;
  for (int c0 = 0; c0 < n_pairs; c0 += L.chunk) {
    const int nc = n_pairs - c0 < L.chunk ? n_pairs - c0 : L.chunk;
    if (int rc = launch_plan(models, obj_idx + c0, pose_est + 12 * (size_t)c0, pose_gt + 12 * (size_t)c0,
                             nullptr, cam + 4 * (size_t)c0, nullptr, 1, nc, 2, height, width, gate, ws, L, s))
      return rc;
    fill_i32_pattern_kernel<<<(nc * 8 + 255) / 256, 256, 0, s>>>((int*)(ws + L.mbox), nc);
    BOP_LAUNCH_CHECK();
    FusedArgs a;
    memset(&a, 0, sizeof(a));
    a.models = models;
    a.desc = (const PairDesc*)(ws + L.desc);
    a.work = (const WorkItem*)(ws + L.work);
    a.ctr = (PlanCounters*)(ws + L.ctr);
    set_geometry(a, ws, L);
    a.counts = (unsigned long long*)(ws + L.counts);
    a.mask_bbox = (int*)(ws + L.mbox);
    a.max_items = L.max_items;
    a.tile_px = effective_tile_px(width);
    a.small_tri_px = g_tune.small_tri_px;
    a.pp.H = height;
    a.pp.W = width;
    a.canvas_w = width;
    a.pp.n_taus = 1;
    if (int rc = launch_fused<MODE_MASK>(a, width, s)) return rc;
    mask_finalize_kernel<<<(nc + 127) / 128, 128, 0, s>>>(
        a.counts, a.mask_bbox, nc, cus_out ? cus_out + c0 : nullptr, cou_bb_out ? cou_bb_out + c0 : nullptr,
        counts_out ? (long long*)counts_out + (size_t)c0 * 4 : nullptr, bbox_out ? bbox_out + (size_t)c0 * 8 : nullptr);
    BOP_LAUNCH_CHECK();
  }
  return 0;
}

// Mask statistics of pre-rendered depth images / binary masks (non-zero = inside): the part of
// pose_error.cus / cou_bb_proj after the two render calls, and pose_error.cou_mask
// (pose_error.py:235-253).  grid = (n_pairs, chunks); 128-bit loads when rows allow it.
namespace {
__global__ void __launch_bounds__(256)
mask_stats_kernel(const float* __restrict__ de, const float* __restrict__ dg, int H, int W,
                  unsigned long long* __restrict__ counts, int* __restrict__ mbox) {
  const int pair = blockIdx.x;
  const size_t base = (size_t)pair * H * W;
  const int npx = H * W;
  int ne = 0, ng = 0, ni = 0, nu = 0;
  int bb[8] = {INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MAX, INT_MAX, INT_MIN, INT_MIN};
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < npx; i += gridDim.y * blockDim.x) {
    const bool e = __ldg(de + base + i) != 0.f, g = __ldg(dg + base + i) != 0.f;
    if (!(e || g)) continue;
    const int y = i / W, x = i - y * W;
    nu += 1; ni += (e && g) ? 1 : 0;
    if (e) { ne += 1; bb[0] = min(bb[0], x); bb[1] = min(bb[1], y); bb[2] = max(bb[2], x); bb[3] = max(bb[3], y); }
    if (g) { ng += 1; bb[4] = min(bb[4], x); bb[5] = min(bb[5], y); bb[6] = max(bb[6], x); bb[7] = max(bb[7], y); }
  }
  const int cnt[4] = {warp_sum(nu), warp_sum(ni), warp_sum(ne), warp_sum(ng)};
#pragma unroll
  for (int k = 0; k < 8; ++k)
    bb[k] = (k & 2) ? __reduce_max_sync(BOP_FULL_MASK, bb[k]) : __reduce_min_sync(BOP_FULL_MASK, bb[k]);
  if ((threadIdx.x & 31) == 0) {
    unsigned long long* c = counts + (size_t)pair * (2 + BOP_MAX_TAUS);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (cnt[k]) atomicAdd(c + k, (unsigned long long)cnt[k]);
    int* o = mbox + (size_t)pair * 8;
    if (cnt[2]) { atomicMin(o + 0, bb[0]); atomicMin(o + 1, bb[1]); atomicMax(o + 2, bb[2]); atomicMax(o + 3, bb[3]); }
    if (cnt[3]) { atomicMin(o + 4, bb[4]); atomicMin(o + 5, bb[5]); atomicMax(o + 6, bb[6]); atomicMax(o + 7, bb[7]); }
  }
}
}  // namespace

extern "C" size_t bop_mask_stats_workspace_bytes(int n_pairs) {
  return align_up((size_t)n_pairs * (2 + BOP_MAX_TAUS) * 8, 256) + align_up((size_t)n_pairs * 8 * sizeof(int), 256);
}

extern "C" int bop_mask_overlap_from_depth_dev(const float* depth_est, const float* depth_gt,
                                               int n_pairs, int height, int width, double* cus_out,
                                               double* cou_bb_out, int64_t* counts_out,
                                               int32_t* bbox_out, void* workspace,
                                               size_t workspace_bytes, void* stream) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(depth_est && depth_gt, "bop_mask_overlap_from_depth_dev: null pointer");
  BOP_REQUIRE(height > 0 && width > 0, "bop_mask_overlap_from_depth_dev: bad image size");
  BOP_REQUIRE(workspace && workspace_bytes >= bop_mask_stats_workspace_bytes(n_pairs),
              "bop_mask_overlap_from_depth_dev: workspace too small");
  cudaStream_t s = (cudaStream_t)stream;
  unsigned char* ws = (unsigned char*)workspace;
  const size_t cbytes = align_up((size_t)n_pairs * (2 + BOP_MAX_TAUS) * 8, 256);
  unsigned long long* counts = (unsigned long long*)ws;
  int* mbox = (int*)(ws + cbytes);
  BOP_CUDA(cudaMemsetAsync(ws, 0, cbytes, s));
  fill_i32_pattern_kernel<<<(n_pairs * 8 + 255) / 256, 256, 0, s>>>(mbox, n_pairs);
  BOP_LAUNCH_CHECK();
  const int chunks = max(1, min(64, (height * width + 256 * 16 - 1) / (256 * 16)));
  bop_prof_begin(BOP_PROF_MASK, s);
  mask_stats_kernel<<<dim3(n_pairs, chunks), 256, 0, s>>>(depth_est, depth_gt, height, width, counts, mbox);
  bop_prof_end(BOP_PROF_MASK, s);
  BOP_LAUNCH_CHECK();
  mask_finalize_kernel<<<(n_pairs + 127) / 128, 128, 0, s>>>(counts, mbox, n_pairs, cus_out, cou_bb_out,
                                                             (long long*)counts_out, bbox_out);
  BOP_LAUNCH_CHECK();
  return 0;
}

// scripts/calc_gt_info.py:113-175 and scripts/calc_gt_masks.py:110-133 for a batch of GT
// instances.  margin_x / margin_y: the reference renders calc_gt_info on a 3x canvas (margins =
// image size, :76-79,121-128) so that px_count_all and bbox_obj include the part of the object
// outside the image; calc_gt_masks uses no margin.
extern "C" size_t bop_gt_info_workspace_bytes(int n, int height, int width, int margin_x,
                                              int margin_y, int max_geom) {
  return ws_layout(n, height + 2 * margin_y, width + 2 * margin_x, 1, max_geom).total;
}

extern "C" int bop_gt_info_dev(const bop_model_dev* models, const int32_t* obj_idx,
                               const double* pose_gt, const void* depth_test,
                               const float* depth_scale, const int32_t* im_idx, const double* K,
                               int n, int max_geom, int height, int width, int depth_dtype,
                               double delta, int margin_x, int margin_y, int64_t* counts_out,
                               double* visib_fract_out, int32_t* bbox_out,
                               unsigned char* mask_out, unsigned char* mask_visib_out,
                               void* workspace, size_t workspace_bytes, void* stream) {
  if (n <= 0) return 0;
  BOP_REQUIRE(models && obj_idx && pose_gt && depth_test && im_idx && K && counts_out &&
              visib_fract_out && bbox_out, "bop_gt_info_dev: null pointer");
  BOP_REQUIRE(height > 0 && width > 0 && width <= kMaxWidth, "bop_gt_info_dev: bad image size");
  BOP_REQUIRE(margin_x >= 0 && margin_y >= 0, "bop_gt_info_dev: negative margin");
  BOP_REQUIRE(depth_dtype == BOP_DEPTH_F32 || depth_dtype == BOP_DEPTH_U16, "bop_gt_info_dev: unknown depth dtype");
  const int cw = width + 2 * margin_x, ch = height + 2 * margin_y;
  BOP_REQUIRE(((cw + 3) & ~3) <= effective_tile_px(cw), "bop_gt_info_dev: canvas wider than the z-tile");
  const WsLayout L = ws_layout(n, ch, cw, 1, max_geom);
  BOP_REQUIRE(workspace && workspace_bytes >= L.total, "bop_gt_info_dev: workspace too small");
  BOP_REQUIRE(((uintptr_t)workspace) % 256 == 0, "bop_gt_info_dev: workspace must be 256-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  unsigned char* ws = (unsigned char*)workspace;
  if (mask_out) BOP_CUDA(cudaMemsetAsync(mask_out, 0, (size_t)n * height * width, s));
  if (mask_visib_out) BOP_CUDA(cudaMemsetAsync(mask_visib_out, 0, (size_t)n * height * width, s));
  for (int c0 = 0; c0 < n; c0 += L.chunk) {
    const int nc = n - c0 < L.chunk ? n - c0 : L.chunk;
    if (int rc = launch_plan(models, obj_idx + c0, pose_gt + 12 * (size_t)c0, pose_gt + 12 * (size_t)c0,
                             im_idx + c0, K, nullptr, 0, nc, 1, ch, cw, 0, ws, L, s, margin_x, margin_y))
      return rc;
    fill_i32_pattern_kernel<<<(nc * 8 + 255) / 256, 256, 0, s>>>((int*)(ws + L.mbox), nc);
    BOP_LAUNCH_CHECK();
    FusedArgs a;
    memset(&a, 0, sizeof(a));
    a.models = models;
    a.desc = (const PairDesc*)(ws + L.desc);
    a.work = (const WorkItem*)(ws + L.work);
    a.ctr = (PlanCounters*)(ws + L.ctr);
    set_geometry(a, ws, L);
    a.depth_test = depth_test;
    a.depth_scale = depth_scale;
    a.counts = (unsigned long long*)(ws + L.counts);
    a.mask_bbox = (int*)(ws + L.mbox);
    a.mask_full = mask_out ? mask_out + (size_t)c0 * height * width : nullptr;
    a.mask_visib = mask_visib_out ? mask_visib_out + (size_t)c0 * height * width : nullptr;
    a.max_items = L.max_items;
    a.tile_px = effective_tile_px(cw);
    a.small_tri_px = g_tune.small_tri_px;
    a.pp.H = height;
    a.pp.W = width;
    a.pp.n_taus = 1;
    a.pp.delta_f32 = (float)delta;
    a.pp.depth_dtype = depth_dtype;
    a.canvas_w = cw;
    a.mx = margin_x;
    a.my = margin_y;
    if (int rc = launch_fused<MODE_GTINFO>(a, width, s)) return rc;
    gt_info_finalize_kernel<<<(nc + 127) / 128, 128, 0, s>>>(
        a.counts, a.mask_bbox, nc, (long long*)counts_out + (size_t)c0 * 3, visib_fract_out + c0,
        bbox_out + (size_t)c0 * 8);
    BOP_LAUNCH_CHECK();
  }
  return 0;
}
