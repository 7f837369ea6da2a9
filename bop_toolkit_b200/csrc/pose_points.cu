// K3 (MSSD / MSPD / ADD / PROJ), K4 (ADI) and K5 (re / te) for sm_100a.
//
// Reference semantics (gist-ailab/bop_toolkit):
//   mssd  bop_toolkit_lib/pose_error.py:96-118   min_S max_i ||(R_e p_i + t_e) - (R_g S_R p_i + R_g S_t + t_g)||
//   mspd  bop_toolkit_lib/pose_error.py:121-144  same on pinhole projections K[R|t][p;1] / w  (misc.py:93-107)
//   add   bop_toolkit_lib/pose_error.py:147-161  mean_i ||(R_e p_i + t_e) - (R_g p_i + t_g)||
//   proj  bop_toolkit_lib/pose_error.py:217-232  mean_i of the 2-D projection distance
//   adi   bop_toolkit_lib/pose_error.py:164-184  mean_j min_i ||(R_g p_j + t_g) - (R_e p_i + t_e)||
//   re    bop_toolkit_lib/pose_error.py:187-202, te :205-214
//
// The rigid part is folded in float64 (the reference's dtype) before the per-vertex loop:
// (R_e p + t_e) - (R_gs p + t_gs) = dR p + dt with dR, dt formed per (pair, symmetry) -- this
// removes the |t| ~ 1 m cancellation from the inner loop and halves its work.  What is left per
// vertex has no cancellation against large terms, so for float32-exact model points (every PLY
// model) MSSD evaluates |dR p + dt|^2 in float32 and MSPD evaluates the difference of the two
// projections as (D.x n.w - n.x D.w) / (n.w (n.w + D.w)) with n = K [R_gs | t_gs] [p; 1] and
// D = (K [R_e | t_e] - K [R_gs | t_gs]) [p; 1] in float32: every factor is good to ~1e-7 relative
// and the error of the result stays below 1e-6 relative (bar: 1e-5).  ADD / PROJ use the same two
// per-vertex forms and add the norms up in float64.  Models whose points are not float32-
// representable, ill-conditioned projections (a vertex within 10 % of the camera plane) and the
// final square roots / means stay in float64.  ADI runs in
// the estimated model frame (distances are invariant under the rigid map R_e^T(. - t_e)), so the
// search set is the static model.
#include "common.cuh"

namespace {

#ifndef BOP_SYM_CTAS
#define BOP_SYM_CTAS 3
#endif
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

template <typename PT>
__device__ __forceinline__ void load_pt(const PT* __restrict__ pts, int i, double& x, double& y,
                                        double& z) {
  x = (double)__ldg(pts + 3 * (size_t)i + 0);
  y = (double)__ldg(pts + 3 * (size_t)i + 1);
  z = (double)__ldg(pts + 3 * (size_t)i + 2);
}

__device__ __forceinline__ bool spheres_overlap(const double* pe, const double* pg, double diameter) {
  // scripts/eval_calc_errors.py:305-309: np.linalg.norm(t_e - t_g) < diameter
  const double dx = pe[9] - pg[9], dy = pe[10] - pg[10], dz = pe[11] - pg[11];
  return sqrt(dx * dx + dy * dy + dz * dz) < diameter;
}

// ---------------------------------------------------------------------------------
// K5: re / te
// ---------------------------------------------------------------------------------
__global__ void re_te_kernel(const double* __restrict__ pe, const double* __restrict__ pg, int n,
                             double* __restrict__ re, double* __restrict__ te) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* e = pe + 12 * (size_t)i;
  const double* g = pg + 12 * (size_t)i;
  if (re) {
    // trace(R_e . inv(R_g)) with a true inverse (pose_error.py:195 uses np.linalg.inv,
    // not the transpose): inv = adj / det.
    const double a = g[0], b = g[1], c = g[2], d = g[3], f = g[4], h = g[5], k = g[6], l = g[7],
                 m = g[8];
    const double A = f * m - h * l, B = -(d * m - h * k), C = d * l - f * k;
    const double D = -(b * m - c * l), E = a * m - c * k, F = -(a * l - b * k);
    const double G = b * h - c * f, H = -(a * h - c * d), I = a * f - b * d;
    const double det = a * A + b * B + c * C;
    // inv = 1/det * [[A D G],[B E H],[C F I]];  trace(Re*inv) = sum_ij Re[i][j]*inv[j][i]
    const double tr = (e[0] * A + e[1] * B + e[2] * C + e[3] * D + e[4] * E + e[5] * F + e[6] * G +
                       e[7] * H + e[8] * I) /
                      det;
    double cosv = 0.5 * (tr - 1.0);
    cosv = fmin(1.0, fmax(-1.0, cosv));
    re[i] = 180.0 * acos(cosv) / 3.141592653589793;
  }
  if (te) {
    const double dx = g[9] - e[9], dy = g[10] - e[10], dz = g[11] - e[11];
    te[i] = sqrt(dx * dx + dy * dy + dz * dz);
  }
}

// ---------------------------------------------------------------------------------
// K3: symmetric max errors (MSSD / MSPD)
// grid = (n_pairs, sym_groups); each CTA owns up to BOP_MAX_SYMS_PER_CTA symmetries of
// one pair and all points.  A warp keeps 128 points (4 per lane) in registers and sweeps the
// CTA's symmetries over them, so points are loaded once.  Per-symmetry maxima are
// kept in shared memory as uint64 bit patterns (non-negative doubles order like
// integers); the CTA's min over symmetries goes to out[] with one atomicMin.
// ---------------------------------------------------------------------------------
struct SymMaxSmem {
  double xf[BOP_MAX_SYMS_PER_CTA][12];
  double Pe[12];
  unsigned long long symmax[BOP_MAX_SYMS_PER_CTA];
  // float64 paths: running maximum of every lane per symmetry (32 KB): the cross-lane reduction
  // happens once per CTA and symmetry instead of once per 128 points and symmetry.  (The float32
  // paths keep theirs in registers, four symmetries at a time.)
  double run[kWarps][BOP_MAX_SYMS_PER_CTA][32];
  // float32 copies for the float32 per-vertex paths: xg = dR|dt (MSSD) or P_gs (MSPD); xd = P_e - P_gs (MSPD)
  alignas(16) float xg[BOP_MAX_SYMS_PER_CTA][12];   // rows read as three float4
  alignas(16) float xd[BOP_MAX_SYMS_PER_CTA][12];
};

// 1/x for the projections: the hardware's 20-bit float64 seed (MUFU.RCP64H) + one Newton step,
// relative error ~1e-12 -- far inside the 1e-5 bar and a third of the instructions of the
// compiler's IEEE division.  Zero, denormal, huge and non-finite values take the IEEE path, so
// x = 0 still yields +-inf like numpy.
__device__ __forceinline__ double rcp_f64(double x) {
  const double ax = fabs(x);
  if (!(ax > 1e-300 && ax < 1e300)) return 1.0 / x;
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  return fma(r, fma(-x, r, 1.0), r);
}

template <typename PT, int MODE, bool F32>  // MODE 0 = MSSD, 1 = MSPD; F32: float32 per-vertex arithmetic
__device__ __forceinline__ void
sym_max_body(const bop_model_dev& M, int pair, const double* __restrict__ pose_est,
             const double* __restrict__ pose_gt, const double* __restrict__ Kmat,
             int syms_per_cta, int gate, unsigned long long* __restrict__ out_bits,
             SymMaxSmem& sm) {
  const int S = M.n_syms;
  const int s0 = blockIdx.y * syms_per_cta;
  if (s0 >= S) return;
  const int ns = min(syms_per_cta, S - s0);
  const double* pe = pose_est + 12 * (size_t)pair;
  const double* pg = pose_gt + 12 * (size_t)pair;
  if (MODE == 0 && gate && !spheres_overlap(pe, pg, M.diameter)) return;  // stays +inf

  // Per-symmetry folded transforms: MSSD 12 doubles (dR, dt); MSPD 12 (P_gs); P_e shared.
  double (*xf)[12] = sm.xf;
  double* Pe = sm.Pe;
  unsigned long long* symmax = sm.symmax;

  const int tid = threadIdx.x;
  if (tid < ns) {
    const double* sym = M.syms + 12 * (size_t)(s0 + tid);
    double Rgs[9], tgs[3];
    mat3_mul(pg, sym, Rgs);            // R_gt . sym.R          (pose_error.py:113)
    mat3_vec(pg, sym + 9, tgs);        // R_gt . sym.t + t_gt   (pose_error.py:114)
    tgs[0] += pg[9]; tgs[1] += pg[10]; tgs[2] += pg[11];
    if (MODE == 0) {
#pragma unroll
      for (int k = 0; k < 9; ++k) xf[tid][k] = pe[k] - Rgs[k];
#pragma unroll
      for (int k = 0; k < 3; ++k) xf[tid][9 + k] = pe[9 + k] - tgs[k];
    } else {
      // P = K.[R|t] (misc.py:103), stored row-major 3x4.
      const double* Kp = Kmat + 9 * (size_t)pair;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c)
          xf[tid][r * 4 + c] = Kp[r * 3 + 0] * Rgs[0 * 3 + c] + Kp[r * 3 + 1] * Rgs[1 * 3 + c] +
                               Kp[r * 3 + 2] * Rgs[2 * 3 + c];
        xf[tid][r * 4 + 3] = Kp[r * 3 + 0] * tgs[0] + Kp[r * 3 + 1] * tgs[1] + Kp[r * 3 + 2] * tgs[2];
      }
    }
    symmax[tid] = 0ull;
    if (F32) {
      if (MODE == 0) {
#pragma unroll
        for (int k = 0; k < 12; ++k) sm.xg[tid][k] = (float)xf[tid][k];
      } else {
        // P_e - P_gs = K [R_e - R_gs | t_e - t_gs], folded in float64 (no cancellation is left for float32)
        const double* Kp = Kmat + 9 * (size_t)pair;
        double dRt[12];
#pragma unroll
        for (int k = 0; k < 9; ++k) dRt[k] = pe[k] - Rgs[k];
#pragma unroll
        for (int k = 0; k < 3; ++k) dRt[9 + k] = pe[9 + k] - tgs[k];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
#pragma unroll
          for (int c = 0; c < 3; ++c)
            sm.xd[tid][r * 4 + c] = (float)(Kp[r * 3 + 0] * dRt[0 * 3 + c] + Kp[r * 3 + 1] * dRt[1 * 3 + c] +
                                            Kp[r * 3 + 2] * dRt[2 * 3 + c]);
          sm.xd[tid][r * 4 + 3] = (float)(Kp[r * 3 + 0] * dRt[9] + Kp[r * 3 + 1] * dRt[10] + Kp[r * 3 + 2] * dRt[11]);
        }
#pragma unroll
        for (int k = 0; k < 12; ++k) sm.xg[tid][k] = (float)xf[tid][k];
      }
    }
  }
  if (!F32) {
    for (int i = tid; i < kWarps * ns * 32; i += kThreads) {   // only the symmetries this CTA owns
      const int wi = i / (ns * 32);
      (&sm.run[wi][0][0])[i - wi * ns * 32] = 0.0;
    }
  }
  if (MODE == 1 && tid == 32) {
    const double* Kp = Kmat + 9 * (size_t)pair;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int c = 0; c < 3; ++c)
        Pe[r * 4 + c] = Kp[r * 3 + 0] * pe[0 * 3 + c] + Kp[r * 3 + 1] * pe[1 * 3 + c] +
                        Kp[r * 3 + 2] * pe[2 * 3 + c];
      Pe[r * 4 + 3] = Kp[r * 3 + 0] * pe[9] + Kp[r * 3 + 1] * pe[10] + Kp[r * 3 + 2] * pe[11];
    }
  }
  __syncthreads();

  const PT* __restrict__ pts = (const PT*)M.pts;
  const int N = M.n_pts;
  const int warp = tid >> 5, lane = tid & 31;
  if (F32) {
    // float32 sweeps.  A warp takes 128 points (4 per lane; the last pass repeats point N-1, which
    // cannot change a maximum) and visits the CTA's symmetries four at a time with their running
    // maxima in registers: one warp reduction per (warp, symmetry) at the end, nothing per pass.
    // MSSD: |dR p + dt|^2 with dR, dt folded in float64 and rounded once (9 FFMA + 3 for the norm).
    // MSPD: per (point, symmetry) n = P_gs [p; 1] and D = (P_e - P_gs) [p; 1] (18 FFMA), then
    // du = (D.x n.w - n.x D.w) / (n.w (n.w + D.w)) (likewise dv): the projections themselves
    // (hundreds of pixels) never meet, only their difference is formed, from factors that are
    // each good to ~1e-7 relative.  One MUFU reciprocal per (point, symmetry).
    constexpr int kPerLane = 4, kSymChunk = 4;
    const int npass = (N + 32 * kPerLane - 1) / (32 * kPerLane);
    for (int c0 = 0; c0 < ns; c0 += kSymChunk) {
      const int nc = min(kSymChunk, ns - c0);
      float runmax[kSymChunk];
#pragma unroll
      for (int j = 0; j < kSymChunk; ++j) runmax[j] = 0.f;
      for (int pass = warp; pass < npass; pass += kWarps) {
        float px[kPerLane], py[kPerLane], pz[kPerLane];
#pragma unroll
        for (int k = 0; k < kPerLane; ++k) {
          const int i = min((pass * kPerLane + k) * 32 + lane, N - 1);
          px[k] = (float)__ldg(pts + 3 * (size_t)i + 0);
          py[k] = (float)__ldg(pts + 3 * (size_t)i + 1);
          pz[k] = (float)__ldg(pts + 3 * (size_t)i + 2);
        }
#pragma unroll
        for (int j = 0; j < kSymChunk; ++j) {
          if (j < nc) {
            const float4* g4 = reinterpret_cast<const float4*>(sm.xg[c0 + j]);
            const float4 g0 = g4[0], g1 = g4[1], g2 = g4[2];
            float m2 = runmax[j];
            if (MODE == 0) {
              // T = dR row-major (g0.x .. g2.x), dt = (g2.y, g2.z, g2.w)
#pragma unroll
              for (int k = 0; k < kPerLane; ++k) {
                const float dx = fmaf(g0.x, px[k], fmaf(g0.y, py[k], fmaf(g0.z, pz[k], g2.y)));
                const float dy = fmaf(g0.w, px[k], fmaf(g1.x, py[k], fmaf(g1.y, pz[k], g2.z)));
                const float dz = fmaf(g1.z, px[k], fmaf(g1.w, py[k], fmaf(g2.x, pz[k], g2.w)));
                m2 = fmaxf(m2, fmaf(dx, dx, fmaf(dy, dy, dz * dz)));   // a NaN never wins, like fmax
              }
            } else {
              const float4* d4 = reinterpret_cast<const float4*>(sm.xd[c0 + j]);
              const float4 d0 = d4[0], d1 = d4[1], d2r = d4[2];
#pragma unroll
              for (int k = 0; k < kPerLane; ++k) {
                const float nx = fmaf(g0.x, px[k], fmaf(g0.y, py[k], fmaf(g0.z, pz[k], g0.w)));
                const float ny = fmaf(g1.x, px[k], fmaf(g1.y, py[k], fmaf(g1.z, pz[k], g1.w)));
                const float nw = fmaf(g2.x, px[k], fmaf(g2.y, py[k], fmaf(g2.z, pz[k], g2.w)));
                const float dx = fmaf(d0.x, px[k], fmaf(d0.y, py[k], fmaf(d0.z, pz[k], d0.w)));
                const float dy = fmaf(d1.x, px[k], fmaf(d1.y, py[k], fmaf(d1.z, pz[k], d1.w)));
                const float dw = fmaf(d2r.x, px[k], fmaf(d2r.y, py[k], fmaf(d2r.z, pz[k], d2r.w)));
                float inv;
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(nw * (nw + dw)));
                const float du = fmaf(dx, nw, -(nx * dw)) * inv, dv = fmaf(dy, nw, -(ny * dw)) * inv;
                m2 = fmaxf(m2, fmaf(du, du, dv * dv));
              }
            }
            runmax[j] = m2;
          }
        }
      }
#pragma unroll
      for (int j = 0; j < kSymChunk; ++j) {
        if (j < nc) {
          float m = runmax[j];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(BOP_FULL_MASK, m, o));
          if (lane == 0) atomicMax(&symmax[c0 + j], (unsigned long long)__double_as_longlong((double)m));
        }
      }
    }
  }
  if (MODE == 1 && !F32) {
    // MSPD: the projection of the ESTIMATED pose does not depend on the symmetry, and float64
    // division is the expensive instruction here.  A warp takes 128 points (4 per lane), projects
    // them with P_e once, then sweeps the symmetries: 9 DFMA + ONE reciprocal per (point, symmetry).
    constexpr int kPerLane = 4;
    const int npass = (N + 32 * kPerLane - 1) / (32 * kPerLane);
    for (int pass = warp; pass < npass; pass += kWarps) {
      double px[kPerLane], py[kPerLane], pz[kPerLane], eu[kPerLane], ev[kPerLane];
      bool ok[kPerLane];
#pragma unroll
      for (int k = 0; k < kPerLane; ++k) {
        const int i = (pass * kPerLane + k) * 32 + lane;
        ok[k] = i < N;
        px[k] = py[k] = pz[k] = eu[k] = ev[k] = 0.0;
        if (ok[k]) {
          load_pt(pts, i, px[k], py[k], pz[k]);
          const double ex = Pe[0] * px[k] + Pe[1] * py[k] + Pe[2] * pz[k] + Pe[3];
          const double ey = Pe[4] * px[k] + Pe[5] * py[k] + Pe[6] * pz[k] + Pe[7];
          const double ew = Pe[8] * px[k] + Pe[9] * py[k] + Pe[10] * pz[k] + Pe[11];
          const double iw = rcp_f64(ew);     // pts_im /= pts_im[2, :]  (misc.py:106)
          eu[k] = ex * iw;
          ev[k] = ey * iw;
        }
      }
      for (int sy = 0; sy < ns; ++sy) {
        double T[12];
#pragma unroll
        for (int k = 0; k < 12; ++k) T[k] = xf[sy][k];
        double m2 = 0.0;
#pragma unroll
        for (int k = 0; k < kPerLane; ++k) {
          if (ok[k]) {
            const double gx = T[0] * px[k] + T[1] * py[k] + T[2] * pz[k] + T[3];
            const double gy = T[4] * px[k] + T[5] * py[k] + T[6] * pz[k] + T[7];
            const double gw = T[8] * px[k] + T[9] * py[k] + T[10] * pz[k] + T[11];
            const double inv = rcp_f64(gw);
            const double du = eu[k] - gx * inv, dv = ev[k] - gy * inv;
            const double d2 = du * du + dv * dv;
            m2 = d2 > m2 ? d2 : m2;   // NaN never wins, like fmax
          }
        }
        sm.run[warp][sy][lane] = fmax(sm.run[warp][sy][lane], m2);
      }
    }
  }
  if (MODE == 0 && !F32) {
    // MSSD: same sweep -- a warp keeps 128 points in registers and visits every symmetry, so the
    // points are loaded (and converted to float64) once instead of once per symmetry.
    constexpr int kPerLane = 4;
    const int npass = (N + 32 * kPerLane - 1) / (32 * kPerLane);
    for (int pass = warp; pass < npass; pass += kWarps) {
      double px[kPerLane], py[kPerLane], pz[kPerLane];
      bool ok[kPerLane];
#pragma unroll
      for (int k = 0; k < kPerLane; ++k) {
        const int i = (pass * kPerLane + k) * 32 + lane;
        ok[k] = i < N;
        px[k] = py[k] = pz[k] = 0.0;
        if (ok[k]) load_pt(pts, i, px[k], py[k], pz[k]);
      }
      for (int sy = 0; sy < ns; ++sy) {
        double T[12];
#pragma unroll
        for (int k = 0; k < 12; ++k) T[k] = xf[sy][k];
        double m2 = 0.0;
#pragma unroll
        for (int k = 0; k < kPerLane; ++k) {
          if (ok[k]) {
            const double dx = T[0] * px[k] + T[1] * py[k] + T[2] * pz[k] + T[9];
            const double dy = T[3] * px[k] + T[4] * py[k] + T[5] * pz[k] + T[10];
            const double dz = T[6] * px[k] + T[7] * py[k] + T[8] * pz[k] + T[11];
            const double d2 = dx * dx + dy * dy + dz * dz;
            m2 = d2 > m2 ? d2 : m2;   // NaN never wins, like fmax
          }
        }
        sm.run[warp][sy][lane] = fmax(sm.run[warp][sy][lane], m2);
      }
    }
  }
  __syncthreads();
  if (!F32) {
    for (int sy = 0; sy < ns; ++sy) {   // 256 running maxima per symmetry -> one value
      const double m = warp_max(sm.run[warp][sy][lane]);
      if (lane == 0) atomicMax(&symmax[sy], (unsigned long long)__double_as_longlong(m));
    }
    __syncthreads();
  }
  if (tid == 0) {
    unsigned long long best = symmax[0];
    for (int s = 1; s < ns; ++s) best = min(best, symmax[s]);
    const double e = sqrt(__longlong_as_double((long long)best));
    atomicMin(out_bits + pair, (unsigned long long)__double_as_longlong(e));
  }
}

template <int MODE>
__global__ void __launch_bounds__(kThreads, BOP_SYM_CTAS)
sym_max_kernel(const bop_model_dev* __restrict__ models, const int32_t* __restrict__ obj_idx,
               const double* __restrict__ pose_est, const double* __restrict__ pose_gt,
               const double* __restrict__ Kmat, int syms_per_cta, int gate, int exact64,
               unsigned long long* __restrict__ out_bits) {
  __shared__ SymMaxSmem sm;
  const int pair = blockIdx.x;
  const bop_model_dev M = models[obj_idx[pair]];
  bool f32 = !M.pts_f64 && !exact64;
  if (MODE == 1 && f32) {
    // The float32 projection difference is good to ~1e-7 x (w at the model centre / smallest w):
    // keep it for poses whose every vertex has w >= 10 % of the centre's (any sensible pose); a
    // model that comes close to or crosses the camera plane is ill-conditioned and takes the
    // float64 path.  |p| <= cull_min_dist for every model point, so w >= k2.t - |k2| r.
    const double* Kp = Kmat + 9 * (size_t)pair;
    const double* pe = pose_est + 12 * (size_t)pair;
    const double* pg = pose_gt + 12 * (size_t)pair;
    const double kr = sqrt(Kp[6] * Kp[6] + Kp[7] * Kp[7] + Kp[8] * Kp[8]) * (double)M.cull_min_dist;
    const double we = Kp[6] * pe[9] + Kp[7] * pe[10] + Kp[8] * pe[11];
    f32 = we - kr >= 0.1 * we;
    const int s0 = blockIdx.y * syms_per_cta, s1 = min(M.n_syms, s0 + syms_per_cta);
    for (int sy = s0; sy < s1 && f32; ++sy) {
      const double* st = M.syms + 12 * (size_t)sy + 9;
      double tg[3];
      mat3_vec(pg, st, tg);
      const double wg = Kp[6] * (tg[0] + pg[9]) + Kp[7] * (tg[1] + pg[10]) + Kp[8] * (tg[2] + pg[11]);
      f32 = wg - kr >= 0.1 * wg;
    }
  }
  if (M.pts_f64)
    sym_max_body<double, MODE, false>(M, pair, pose_est, pose_gt, Kmat, syms_per_cta, gate, out_bits, sm);
  else if (!f32)
    sym_max_body<float, MODE, false>(M, pair, pose_est, pose_gt, Kmat, syms_per_cta, gate, out_bits, sm);
  else
    sym_max_body<float, MODE, true>(M, pair, pose_est, pose_gt, Kmat, syms_per_cta, gate, out_bits, sm);
}

// ---------------------------------------------------------------------------------
// K3: mean errors (ADD / PROJ).  grid = (n_pairs, splits); partial sums are combined with
// one double atomicAdd per CTA into sum[]; a finalize kernel divides by N / applies the gate.
// ---------------------------------------------------------------------------------
template <typename PT, int MODE, bool F32>  // MODE 0 = ADD, 1 = PROJ; F32: float32 per-vertex arithmetic
__device__ __forceinline__ void
mean_err_body(const bop_model_dev& M, int pair, const double* __restrict__ pose_est,
              const double* __restrict__ pose_gt, const double* __restrict__ Kmat, int gate,
              double* __restrict__ sum, double* wsum) {
  const double* pe = pose_est + 12 * (size_t)pair;
  const double* pg = pose_gt + 12 * (size_t)pair;
  if (MODE == 0 && gate && !spheres_overlap(pe, pg, M.diameter)) return;
  const int N = M.n_pts;
  const int per = (N + gridDim.y - 1) / gridDim.y;
  const int i0 = blockIdx.y * per, i1 = min(N, i0 + per);
  if (i0 >= i1) return;

  double T[12], Pe[12];
  if (MODE == 0) {
#pragma unroll
    for (int k = 0; k < 12; ++k) T[k] = pe[k] - pg[k];
  } else {
    const double* Kp = Kmat + 9 * (size_t)pair;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        Pe[r * 4 + c] = Kp[r * 3 + 0] * pe[c] + Kp[r * 3 + 1] * pe[3 + c] + Kp[r * 3 + 2] * pe[6 + c];
        T[r * 4 + c] = Kp[r * 3 + 0] * pg[c] + Kp[r * 3 + 1] * pg[3 + c] + Kp[r * 3 + 2] * pg[6 + c];
      }
      Pe[r * 4 + 3] = Kp[r * 3 + 0] * pe[9] + Kp[r * 3 + 1] * pe[10] + Kp[r * 3 + 2] * pe[11];
      T[r * 4 + 3] = Kp[r * 3 + 0] * pg[9] + Kp[r * 3 + 1] * pg[10] + Kp[r * 3 + 2] * pg[11];
    }
  }
  const PT* __restrict__ pts = (const PT*)M.pts;
  double acc = 0.0;
  if (F32) {
    // float32 per vertex as in the MSSD / MSPD sweeps (dR, dt and P_e - P_g are folded in float64
    // above and rounded once); the norms are summed in float64.
    float G[12], D[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) { G[k] = (float)T[k]; D[k] = MODE == 1 ? (float)(Pe[k] - T[k]) : 0.f; }
    for (int i = i0 + threadIdx.x; i < i1; i += kThreads) {
      const float x = (float)__ldg(pts + 3 * (size_t)i + 0), y = (float)__ldg(pts + 3 * (size_t)i + 1),
                  z = (float)__ldg(pts + 3 * (size_t)i + 2);
      if (MODE == 0) {
        const float dx = fmaf(G[0], x, fmaf(G[1], y, fmaf(G[2], z, G[9])));
        const float dy = fmaf(G[3], x, fmaf(G[4], y, fmaf(G[5], z, G[10])));
        const float dz = fmaf(G[6], x, fmaf(G[7], y, fmaf(G[8], z, G[11])));
        acc += (double)sqrtf(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
      } else {
        const float nx = fmaf(G[0], x, fmaf(G[1], y, fmaf(G[2], z, G[3])));
        const float ny = fmaf(G[4], x, fmaf(G[5], y, fmaf(G[6], z, G[7])));
        const float nw = fmaf(G[8], x, fmaf(G[9], y, fmaf(G[10], z, G[11])));
        const float dx = fmaf(D[0], x, fmaf(D[1], y, fmaf(D[2], z, D[3])));
        const float dy = fmaf(D[4], x, fmaf(D[5], y, fmaf(D[6], z, D[7])));
        const float dw = fmaf(D[8], x, fmaf(D[9], y, fmaf(D[10], z, D[11])));
        float inv;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(nw * (nw + dw)));
        const float du = fmaf(dx, nw, -(nx * dw)) * inv, dv = fmaf(dy, nw, -(ny * dw)) * inv;
        acc += (double)sqrtf(fmaf(du, du, dv * dv));
      }
    }
  }
  for (int i = i0 + threadIdx.x; !F32 && i < i1; i += kThreads) {
    double x, y, z;
    load_pt(pts, i, x, y, z);
    if (MODE == 0) {
      const double dx = T[0] * x + T[1] * y + T[2] * z + T[9];
      const double dy = T[3] * x + T[4] * y + T[5] * z + T[10];
      const double dz = T[6] * x + T[7] * y + T[8] * z + T[11];
      acc += sqrt(dx * dx + dy * dy + dz * dz);
    } else {
      const double ex = Pe[0] * x + Pe[1] * y + Pe[2] * z + Pe[3];
      const double ey = Pe[4] * x + Pe[5] * y + Pe[6] * z + Pe[7];
      const double ew = Pe[8] * x + Pe[9] * y + Pe[10] * z + Pe[11];
      const double gx = T[0] * x + T[1] * y + T[2] * z + T[3];
      const double gy = T[4] * x + T[5] * y + T[6] * z + T[7];
      const double gw = T[8] * x + T[9] * y + T[10] * z + T[11];
      const double du = ex / ew - gx / gw, dv = ey / ew - gy / gw;
      acc += sqrt(du * du + dv * dv);
    }
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) s += wsum[w];
    atomicAdd(sum + pair, s);
  }
}

template <int MODE>
__global__ void __launch_bounds__(kThreads)
mean_err_kernel(const bop_model_dev* __restrict__ models, const int32_t* __restrict__ obj_idx,
                const double* __restrict__ pose_est, const double* __restrict__ pose_gt,
                const double* __restrict__ Kmat, int gate, int exact64, double* __restrict__ sum) {
  __shared__ double wsum[kWarps];
  const int pair = blockIdx.x;
  const bop_model_dev M = models[obj_idx[pair]];
  bool f32 = !M.pts_f64 && !exact64;
  if (MODE == 1 && f32) {
    // same conditioning test as MSPD: every vertex of both poses at >= 10 % of the centre's depth
    const double* Kp = Kmat + 9 * (size_t)pair;
    const double kr = sqrt(Kp[6] * Kp[6] + Kp[7] * Kp[7] + Kp[8] * Kp[8]) * (double)M.cull_min_dist;
    const double* te = pose_est + 12 * (size_t)pair + 9;
    const double* tg = pose_gt + 12 * (size_t)pair + 9;
    const double we = Kp[6] * te[0] + Kp[7] * te[1] + Kp[8] * te[2];
    const double wg = Kp[6] * tg[0] + Kp[7] * tg[1] + Kp[8] * tg[2];
    f32 = we - kr >= 0.1 * we && wg - kr >= 0.1 * wg;
  }
  if (M.pts_f64)
    mean_err_body<double, MODE, false>(M, pair, pose_est, pose_gt, Kmat, gate, sum, wsum);
  else if (!f32)
    mean_err_body<float, MODE, false>(M, pair, pose_est, pose_gt, Kmat, gate, sum, wsum);
  else
    mean_err_body<float, MODE, true>(M, pair, pose_est, pose_gt, Kmat, gate, sum, wsum);
}

// out[i] = gated ? +inf : out[i] / n_pts   (in place on the sum buffer)
__global__ void mean_finalize_kernel(const bop_model_dev* __restrict__ models,
                                     const int32_t* __restrict__ obj_idx,
                                     const double* __restrict__ pose_est,
                                     const double* __restrict__ pose_gt, int n, int gate,
                                     double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const bop_model_dev M = models[obj_idx[i]];
  if (gate && !spheres_overlap(pose_est + 12 * (size_t)i, pose_gt + 12 * (size_t)i, M.diameter)) {
    out[i] = __longlong_as_double((long long)BOP_INF_BITS_F64);
    return;
  }
  out[i] = out[i] / (double)M.n_pts;
}

__global__ void fill_u64_kernel(unsigned long long* p, unsigned long long v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// ---------------------------------------------------------------------------------
// K4: ADI brute-force nearest neighbour.
// Queries q_j = A p_j + b with A = R_e^T R_g, b = R_e^T (t_g - t_e) (float64, rounded once
// to float32); search set = the model's own float32 vertices (verts4), staged through
// shared memory in 1024-point tiles and broadcast to all threads; each thread owns Q
// queries in registers.  Inner step per (query, target): 3 FADD + FMUL + 2 FFMA + FMNMX.
// The contraction length is 3, so tensor cores have nothing to offer.
// grid = (n_pairs, query_blocks).
// ---------------------------------------------------------------------------------
constexpr int kAdiQ = 4;
constexpr int kAdiTile = 1024;

struct AdiSmem {
  double Ab[12];
  float4 tile[kAdiTile];
  double wsum[kWarps];
};

template <typename PT>
__device__ __forceinline__ void
adi_body(const bop_model_dev& M, int pair, const double* __restrict__ pose_est,
         const double* __restrict__ pose_gt, int gate, double* __restrict__ sum, AdiSmem& sm) {
  const int N = M.n_pts;
  const int q0 = blockIdx.y * (kThreads * kAdiQ);
  if (q0 >= N) return;
  const double* pe = pose_est + 12 * (size_t)pair;
  const double* pg = pose_gt + 12 * (size_t)pair;
  if (gate && !spheres_overlap(pe, pg, M.diameter)) return;

  double* Ab = sm.Ab;
  float4* tile = sm.tile;
  double* wsum = sm.wsum;
  const int tid = threadIdx.x;
  if (tid < 9) {
    const int r = tid / 3, c = tid % 3;  // A = R_e^T R_g
    Ab[tid] = pe[0 * 3 + r] * pg[0 * 3 + c] + pe[1 * 3 + r] * pg[1 * 3 + c] + pe[2 * 3 + r] * pg[2 * 3 + c];
  } else if (tid < 12) {
    const int r = tid - 9;  // b = R_e^T (t_g - t_e)
    Ab[tid] = pe[0 * 3 + r] * (pg[9] - pe[9]) + pe[1 * 3 + r] * (pg[10] - pe[10]) +
              pe[2 * 3 + r] * (pg[11] - pe[11]);
  }
  __syncthreads();

  const PT* __restrict__ pts = (const PT*)M.pts;
  float qx[kAdiQ], qy[kAdiQ], qz[kAdiQ], best[kAdiQ];
#pragma unroll
  for (int q = 0; q < kAdiQ; ++q) {
    const int j = q0 + q * kThreads + tid;
    double x = 0, y = 0, z = 0;
    if (j < N) load_pt(pts, j, x, y, z);
    qx[q] = (float)(Ab[0] * x + Ab[1] * y + Ab[2] * z + Ab[9]);
    qy[q] = (float)(Ab[3] * x + Ab[4] * y + Ab[5] * z + Ab[10]);
    qz[q] = (float)(Ab[6] * x + Ab[7] * y + Ab[8] * z + Ab[11]);
    best[q] = __int_as_float(0x7f7fffff);
  }

  const float4* __restrict__ tgt = (const float4*)M.verts4;
  for (int t0 = 0; t0 < N; t0 += kAdiTile) {
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kAdiTile / kThreads; ++k) {
      const int i = t0 + k * kThreads + tid;
      tile[k * kThreads + tid] = (i < N) ? __ldg(tgt + i) : make_float4(1e18f, 1e18f, 1e18f, 0.f);
    }
    __syncthreads();
#pragma unroll 8
    for (int i = 0; i < kAdiTile; ++i) {
      const float4 p = tile[i];
#pragma unroll
      for (int q = 0; q < kAdiQ; ++q) {
        const float dx = qx[q] - p.x, dy = qy[q] - p.y, dz = qz[q] - p.z;
        best[q] = fminf(best[q], fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
      }
    }
  }
  double acc = 0.0;
#pragma unroll
  for (int q = 0; q < kAdiQ; ++q) {
    const int j = q0 + q * kThreads + tid;
    if (j < N) acc += sqrt((double)best[q]);
  }
  acc = warp_sum(acc);
  if ((tid & 31) == 0) wsum[tid >> 5] = acc;
  __syncthreads();
  if (tid == 0) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) s += wsum[w];
    atomicAdd(sum + pair, s);
  }
}

__global__ void __launch_bounds__(kThreads)
adi_kernel(const bop_model_dev* __restrict__ models, const int32_t* __restrict__ obj_idx,
           const double* __restrict__ pose_est, const double* __restrict__ pose_gt, int gate,
           double* __restrict__ sum) {
  __shared__ AdiSmem sm;
  const int pair = blockIdx.x;
  const bop_model_dev M = models[obj_idx[pair]];
  if (M.pts_f64)
    adi_body<double>(M, pair, pose_est, pose_gt, gate, sum, sm);
  else
    adi_body<float>(M, pair, pose_est, pose_gt, gate, sum, sm);
}

}  // namespace

// ---------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------
extern "C" int bop_re_te_dev(const double* pose_est, const double* pose_gt, int n_pairs,
                             double* re_out, double* te_out, void* stream) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(pose_est && pose_gt, "bop_re_te_dev: null pose pointer");
  cudaStream_t s = (cudaStream_t)stream;
  re_te_kernel<<<(n_pairs + 127) / 128, 128, 0, s>>>(pose_est, pose_gt, n_pairs, re_out, te_out);
  BOP_LAUNCH_CHECK();
  return 0;
}

static int g_points_exact = 0;   // bop_points_set_exact: float64 per-vertex arithmetic for MSSD / MSPD / ADD / PROJ (tests)

// Test hook: exact != 0 makes MSSD / MSPD evaluate every vertex in float64 (the reference's dtype)
// instead of the float32 forms described at the top of this file; tests compare the two.
extern "C" int bop_points_set_exact(int exact) {
  g_points_exact = exact != 0;
  return 0;
}

static int launch_sym_max(int mode, const bop_model_dev* models, const int32_t* obj_idx,
                          const double* pose_est, const double* pose_gt, const double* K,
                          int n_pairs, int max_syms, int gate, double* out, cudaStream_t s) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(models && obj_idx && pose_est && pose_gt && out, "mssd/mspd: null pointer");
  BOP_REQUIRE(max_syms >= 1, "mssd/mspd: max_syms must be >= 1");
  // Symmetries per CTA: enough CTAs to fill the machine, at most 32 per CTA.
  int groups = 1;
  const int want_ctas = 148 * 8;
  while (groups < max_syms && n_pairs * groups < want_ctas) ++groups;
  int spc = (max_syms + groups - 1) / groups;
  if (spc > BOP_MAX_SYMS_PER_CTA) spc = BOP_MAX_SYMS_PER_CTA;
  groups = (max_syms + spc - 1) / spc;
  unsigned long long* bits = (unsigned long long*)out;
  fill_u64_kernel<<<(n_pairs + 255) / 256, 256, 0, s>>>(bits, BOP_INF_BITS_F64, n_pairs);
  BOP_LAUNCH_CHECK();
  dim3 grid(n_pairs, groups);
  bop_prof_begin(mode == 0 ? BOP_PROF_MSSD : BOP_PROF_MSPD, s);
  if (mode == 0)
    sym_max_kernel<0><<<grid, kThreads, 0, s>>>(models, obj_idx, pose_est, pose_gt, K, spc, gate, g_points_exact, bits);
  else
    sym_max_kernel<1><<<grid, kThreads, 0, s>>>(models, obj_idx, pose_est, pose_gt, K, spc, gate, g_points_exact, bits);
  bop_prof_end(mode == 0 ? BOP_PROF_MSSD : BOP_PROF_MSPD, s);
  BOP_LAUNCH_CHECK();
  return 0;
}

extern "C" int bop_mssd_dev(const bop_model_dev* models, const int32_t* obj_idx,
                            const double* pose_est, const double* pose_gt, int n_pairs,
                            int max_syms, int gate, double* out, void* stream) {
  return launch_sym_max(0, models, obj_idx, pose_est, pose_gt, nullptr, n_pairs, max_syms, gate,
                        out, (cudaStream_t)stream);
}

extern "C" int bop_mspd_dev(const bop_model_dev* models, const int32_t* obj_idx,
                            const double* pose_est, const double* pose_gt, const double* K,
                            int n_pairs, int max_syms, double* out, void* stream) {
  BOP_REQUIRE(K || n_pairs <= 0, "bop_mspd_dev: null K");
  return launch_sym_max(1, models, obj_idx, pose_est, pose_gt, K, n_pairs, max_syms, 0, out,
                        (cudaStream_t)stream);
}

static int launch_mean(int mode, const bop_model_dev* models, const int32_t* obj_idx,
                       const double* pose_est, const double* pose_gt, const double* K, int n_pairs,
                       int max_pts, int gate, double* out, cudaStream_t s) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(models && obj_idx && pose_est && pose_gt && out, "add/proj: null pointer");
  BOP_REQUIRE(max_pts >= 1, "add/proj: max_pts must be >= 1");
  BOP_CUDA(cudaMemsetAsync(out, 0, sizeof(double) * (size_t)n_pairs, s));
  int splits = 1;
  const int want_ctas = 148 * 8;
  while (n_pairs * splits < want_ctas && max_pts / (splits + 1) >= 4 * kThreads) ++splits;
  dim3 grid(n_pairs, splits);
  if (mode == 0)
    mean_err_kernel<0><<<grid, kThreads, 0, s>>>(models, obj_idx, pose_est, pose_gt, K, gate, g_points_exact, out);
  else
    mean_err_kernel<1><<<grid, kThreads, 0, s>>>(models, obj_idx, pose_est, pose_gt, K, 0, g_points_exact, out);
  BOP_LAUNCH_CHECK();
  mean_finalize_kernel<<<(n_pairs + 255) / 256, 256, 0, s>>>(models, obj_idx, pose_est, pose_gt,
                                                            n_pairs, mode == 0 ? gate : 0, out);
  BOP_LAUNCH_CHECK();
  return 0;
}

extern "C" int bop_add_dev(const bop_model_dev* models, const int32_t* obj_idx,
                           const double* pose_est, const double* pose_gt, int n_pairs, int max_pts,
                           int gate, double* out, void* stream) {
  return launch_mean(0, models, obj_idx, pose_est, pose_gt, nullptr, n_pairs, max_pts, gate, out,
                     (cudaStream_t)stream);
}

extern "C" int bop_proj_dev(const bop_model_dev* models, const int32_t* obj_idx,
                            const double* pose_est, const double* pose_gt, const double* K,
                            int n_pairs, int max_pts, double* out, void* stream) {
  BOP_REQUIRE(K || n_pairs <= 0, "bop_proj_dev: null K");
  return launch_mean(1, models, obj_idx, pose_est, pose_gt, K, n_pairs, max_pts, 0, out,
                     (cudaStream_t)stream);
}

extern "C" int bop_adi_dev(const bop_model_dev* models, const int32_t* obj_idx,
                           const double* pose_est, const double* pose_gt, int n_pairs, int max_pts,
                           int gate, double* out, void* stream) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(models && obj_idx && pose_est && pose_gt && out, "bop_adi_dev: null pointer");
  BOP_REQUIRE(max_pts >= 1, "bop_adi_dev: max_pts must be >= 1");
  cudaStream_t s = (cudaStream_t)stream;
  BOP_CUDA(cudaMemsetAsync(out, 0, sizeof(double) * (size_t)n_pairs, s));
  const int qblocks = (max_pts + kThreads * kAdiQ - 1) / (kThreads * kAdiQ);
  dim3 grid(n_pairs, qblocks);
  bop_prof_begin(BOP_PROF_ADI, s);
  adi_kernel<<<grid, kThreads, 0, s>>>(models, obj_idx, pose_est, pose_gt, gate, out);
  bop_prof_end(BOP_PROF_ADI, s);
  BOP_LAUNCH_CHECK();
  mean_finalize_kernel<<<(n_pairs + 255) / 256, 256, 0, s>>>(models, obj_idx, pose_est, pose_gt,
                                                            n_pairs, gate, out);
  BOP_LAUNCH_CHECK();
  return 0;
}
