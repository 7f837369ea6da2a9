// Library-level plumbing of libbop_b200: error reporting, launch counter, model upload
// and the host-buffer (`*_host`) convenience layer declared in include/bop_b200.h.
#include <atomic>
#include <string>
#include <math.h>
#include <vector>

#include "common.cuh"

namespace {
thread_local std::string g_err;
std::atomic<uint64_t> g_launches{0};
}  // namespace

int bop_set_error_cuda(cudaError_t e, const char* what, const char* file, int line) {
  char buf[1024];
  snprintf(buf, sizeof(buf), "CUDA error %d (%s) in `%s` at %s:%d", (int)e, cudaGetErrorString(e),
           what, file, line);
  g_err = buf;
  return (int)e ? (int)e : -1;
}
int bop_set_error_msg(int code, const char* msg) {
  g_err = msg;
  return code ? code : -1;
}
void bop_count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

// ---------------------------------------------------------------------------------
// Per-kernel timing: a ring of CUDA event pairs per slot, recorded on the launching stream.
// ---------------------------------------------------------------------------------
namespace {
constexpr int kProfRing = 256;
struct ProfSlot {
  cudaEvent_t beg[kProfRing], end[kProfRing];
  bool made = false;
  uint64_t n = 0;   // completed begin/end pairs
};
ProfSlot g_prof[BOP_PROF_SLOTS];
bool g_prof_on = false;
}  // namespace

void bop_prof_begin(int slot, cudaStream_t s) {
  if (!g_prof_on) return;
  ProfSlot& p = g_prof[slot];
  if (!p.made) {
    for (int i = 0; i < kProfRing; ++i) { cudaEventCreate(&p.beg[i]); cudaEventCreate(&p.end[i]); }
    p.made = true;
  }
  cudaEventRecord(p.beg[p.n % kProfRing], s);
}
void bop_prof_end(int slot, cudaStream_t s) {
  if (!g_prof_on) return;
  ProfSlot& p = g_prof[slot];
  cudaEventRecord(p.end[p.n % kProfRing], s);
  p.n++;
}

extern "C" int bop_profile_enable(int on) {
  g_prof_on = on != 0;
  if (!g_prof_on) for (int i = 0; i < BOP_PROF_SLOTS; ++i) g_prof[i].n = 0;
  return 0;
}
extern "C" int bop_profile_reset(void) {
  for (int i = 0; i < BOP_PROF_SLOTS; ++i) g_prof[i].n = 0;
  return 0;
}
// Sum of the device durations (ms) of the last `last_n` timed launches of `slot` (all if 0).
// Synchronises on the recorded events.
extern "C" int bop_profile_sum_ms(int slot, int last_n, double* sum_ms, int* count) {
  BOP_REQUIRE(slot >= 0 && slot < BOP_PROF_SLOTS && sum_ms && count, "bop_profile_sum_ms: bad arguments");
  ProfSlot& p = g_prof[slot];
  uint64_t n = p.n < (uint64_t)kProfRing ? p.n : (uint64_t)kProfRing;
  if (last_n > 0 && (uint64_t)last_n < n) n = (uint64_t)last_n;
  double sum = 0.0;
  for (uint64_t k = 0; k < n; ++k) {
    const uint64_t i = (p.n - 1 - k) % kProfRing;
    BOP_CUDA(cudaEventSynchronize(p.end[i]));
    float ms = 0.f;
    BOP_CUDA(cudaEventElapsedTime(&ms, p.beg[i], p.end[i]));
    sum += ms;
  }
  *sum_ms = sum;
  *count = (int)n;
  return 0;
}

extern "C" int bop_abi_version(void) { return BOP_B200_ABI_VERSION; }
extern "C" const char* bop_last_error(void) { return g_err.c_str(); }
extern "C" int bop_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}
extern "C" uint64_t bop_kernel_launch_count(void) { return g_launches.load(); }

// ---------------------------------------------------------------------------------
// Model upload
// ---------------------------------------------------------------------------------
extern "C" int bop_model_upload(const double* pts, int n_pts, const int32_t* faces, int n_faces,
                                const double* syms, int n_syms, double diameter,
                                bop_model_dev* desc_host_out, bop_model_dev** desc_dev_out) {
  BOP_REQUIRE(pts && n_pts > 0 && desc_host_out, "bop_model_upload: bad arguments");
  BOP_REQUIRE(bop_device_count() > 0, "bop_model_upload: no CUDA device (there is no CPU fallback)");
  bop_model_dev d;
  memset(&d, 0, sizeof(d));
  // Points stay float64 unless every coordinate is float32-representable (PLY floats are).
  bool f32_exact = true;
  for (size_t i = 0; i < (size_t)n_pts * 3; ++i)
    if ((double)(float)pts[i] != pts[i]) { f32_exact = false; break; }
  std::vector<float> v4((size_t)n_pts * 4);
  for (int i = 0; i < n_pts; ++i) {
    v4[4 * (size_t)i + 0] = (float)pts[3 * (size_t)i + 0];
    v4[4 * (size_t)i + 1] = (float)pts[3 * (size_t)i + 1];
    v4[4 * (size_t)i + 2] = (float)pts[3 * (size_t)i + 2];
    v4[4 * (size_t)i + 3] = 0.f;
  }
  void* p = nullptr;
  if (f32_exact) {
    std::vector<float> p3((size_t)n_pts * 3);
    for (size_t i = 0; i < p3.size(); ++i) p3[i] = (float)pts[i];
    BOP_CUDA(cudaMalloc(&p, p3.size() * sizeof(float)));
    BOP_CUDA(cudaMemcpy(p, p3.data(), p3.size() * sizeof(float), cudaMemcpyHostToDevice));
  } else {
    BOP_CUDA(cudaMalloc(&p, (size_t)n_pts * 3 * sizeof(double)));
    BOP_CUDA(cudaMemcpy(p, pts, (size_t)n_pts * 3 * sizeof(double), cudaMemcpyHostToDevice));
  }
  d.pts = p;
  d.pts_f64 = f32_exact ? 0 : 1;
  d.cull = 0;       /* the convenience layer never culls; see batch.ModelSet for the verified path */
  double r2 = 0.0;   // bounding sphere about the model origin (also bounds w in the MSPD float32 path)
  for (int i = 0; i < n_pts; ++i) {
    const double x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
    const double q = x * x + y * y + z * z;
    if (q > r2) r2 = q;
  }
  d.cull_min_dist = (float)(sqrt(r2) * 1.001);
  BOP_CUDA(cudaMalloc(&p, v4.size() * sizeof(float)));
  BOP_CUDA(cudaMemcpy(p, v4.data(), v4.size() * sizeof(float), cudaMemcpyHostToDevice));
  d.verts4 = (const float*)p;
  if (n_faces > 0 && faces) {
    std::vector<int32_t> f4((size_t)n_faces * 4);
    for (int i = 0; i < n_faces; ++i) {
      for (int k = 0; k < 3; ++k) {
        const int32_t v = faces[3 * (size_t)i + k];
        BOP_REQUIRE(v >= 0 && v < n_pts, "bop_model_upload: face index out of range");
        f4[4 * (size_t)i + k] = v;
      }
      f4[4 * (size_t)i + 3] = 0;
    }
    BOP_CUDA(cudaMalloc(&p, f4.size() * sizeof(int32_t)));
    BOP_CUDA(cudaMemcpy(p, f4.data(), f4.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    d.faces4 = (const int32_t*)p;
    d.n_faces = n_faces;
  }
  if (n_syms > 0 && syms) {
    BOP_CUDA(cudaMalloc(&p, (size_t)n_syms * 12 * sizeof(double)));
    BOP_CUDA(cudaMemcpy(p, syms, (size_t)n_syms * 12 * sizeof(double), cudaMemcpyHostToDevice));
    d.syms = (const double*)p;
    d.n_syms = n_syms;
  }
  d.diameter = diameter;
  d.n_pts = n_pts;
  *desc_host_out = d;
  if (desc_dev_out) {
    BOP_CUDA(cudaMalloc(&p, sizeof(bop_model_dev)));
    BOP_CUDA(cudaMemcpy(p, &d, sizeof(d), cudaMemcpyHostToDevice));
    *desc_dev_out = (bop_model_dev*)p;
  }
  return 0;
}

extern "C" int bop_model_free(bop_model_dev* desc_host, bop_model_dev* desc_dev) {
  if (desc_host) {
    cudaFree((void*)desc_host->pts);
    cudaFree((void*)desc_host->verts4);
    cudaFree((void*)desc_host->faces4);
    cudaFree((void*)desc_host->syms);
    memset(desc_host, 0, sizeof(*desc_host));
  }
  if (desc_dev) cudaFree(desc_dev);
  return 0;
}

// ---------------------------------------------------------------------------------
// Host-buffer layer.  Simple synchronous staging: cudaMalloc + cudaMemcpy around the `_dev`
// entry points, on the legacy default stream.  (The Python package stages through torch's
// caching allocator and pinned memory instead; this layer is for bindings without torch.)
// ---------------------------------------------------------------------------------
namespace {

struct DevBuf {
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  int alloc(size_t bytes) {
    BOP_CUDA(cudaMalloc(&p, bytes ? bytes : 1));
    return 0;
  }
  int upload(const void* h, size_t bytes) {
    if (int rc = alloc(bytes)) return rc;
    if (bytes) BOP_CUDA(cudaMemcpy(p, h, bytes, cudaMemcpyHostToDevice));
    return 0;
  }
  template <typename T> T* as() { return (T*)p; }
};

int download(void* h, const void* d, size_t bytes) {
  BOP_CUDA(cudaMemcpy(h, d, bytes, cudaMemcpyDeviceToHost));
  return 0;
}

int zeros_i32(DevBuf& b, int n) {
  if (int rc = b.alloc(sizeof(int32_t) * (size_t)n)) return rc;
  BOP_CUDA(cudaMemset(b.p, 0, sizeof(int32_t) * (size_t)n));
  return 0;
}

#define RC(x) do { if (int _rc = (x)) return _rc; } while (0)

}  // namespace

extern "C" int bop_re_te_host(const double* pose_est, const double* pose_gt, int n_pairs,
                              double* re_out, double* te_out) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(bop_device_count() > 0, "no CUDA device (there is no CPU fallback)");
  const size_t pb = sizeof(double) * 12 * (size_t)n_pairs, ob = sizeof(double) * (size_t)n_pairs;
  DevBuf pe, pg, re, te;
  RC(pe.upload(pose_est, pb)); RC(pg.upload(pose_gt, pb));
  RC(re.alloc(ob)); RC(te.alloc(ob));
  RC(bop_re_te_dev(pe.as<double>(), pg.as<double>(), n_pairs, re_out ? re.as<double>() : nullptr,
                   te_out ? te.as<double>() : nullptr, nullptr));
  if (re_out) RC(download(re_out, re.p, ob));
  if (te_out) RC(download(te_out, te.p, ob));
  return 0;
}

namespace {
// kind: 0 mssd, 1 mspd, 2 add, 3 adi
int points_host(int kind, const bop_model_dev* model_dev, int max_n, const double* pose_est,
                const double* pose_gt, const double* K, int n_pairs, int gate, double* out) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(bop_device_count() > 0, "no CUDA device (there is no CPU fallback)");
  BOP_REQUIRE(model_dev && pose_est && pose_gt && out, "null pointer");
  const size_t pb = sizeof(double) * 12 * (size_t)n_pairs, ob = sizeof(double) * (size_t)n_pairs;
  DevBuf pe, pg, kk, oi, o;
  RC(pe.upload(pose_est, pb)); RC(pg.upload(pose_gt, pb));
  RC(zeros_i32(oi, n_pairs)); RC(o.alloc(ob));
  if (kind == 1) {
    BOP_REQUIRE(K, "null K");
    RC(kk.upload(K, sizeof(double) * 9 * (size_t)n_pairs));
  }
  switch (kind) {
    case 0: RC(bop_mssd_dev(model_dev, oi.as<int32_t>(), pe.as<double>(), pg.as<double>(), n_pairs, max_n, gate, o.as<double>(), nullptr)); break;
    case 1: RC(bop_mspd_dev(model_dev, oi.as<int32_t>(), pe.as<double>(), pg.as<double>(), kk.as<double>(), n_pairs, max_n, o.as<double>(), nullptr)); break;
    case 2: RC(bop_add_dev(model_dev, oi.as<int32_t>(), pe.as<double>(), pg.as<double>(), n_pairs, max_n, gate, o.as<double>(), nullptr)); break;
    default: RC(bop_adi_dev(model_dev, oi.as<int32_t>(), pe.as<double>(), pg.as<double>(), n_pairs, max_n, gate, o.as<double>(), nullptr)); break;
  }
  return download(out, o.p, ob);
}
}  // namespace

extern "C" int bop_mssd_host(const bop_model_dev* model_dev, int n_syms, const double* pose_est,
                             const double* pose_gt, int n_pairs, int gate, double* out) {
  return points_host(0, model_dev, n_syms, pose_est, pose_gt, nullptr, n_pairs, gate, out);
}
extern "C" int bop_mspd_host(const bop_model_dev* model_dev, int n_syms, const double* pose_est,
                             const double* pose_gt, const double* K, int n_pairs, double* out) {
  return points_host(1, model_dev, n_syms, pose_est, pose_gt, K, n_pairs, 0, out);
}
extern "C" int bop_add_host(const bop_model_dev* model_dev, int n_pts, const double* pose_est,
                            const double* pose_gt, int n_pairs, int gate, double* out) {
  return points_host(2, model_dev, n_pts, pose_est, pose_gt, nullptr, n_pairs, gate, out);
}
extern "C" int bop_adi_host(const bop_model_dev* model_dev, int n_pts, const double* pose_est,
                            const double* pose_gt, int n_pairs, int gate, double* out) {
  return points_host(3, model_dev, n_pts, pose_est, pose_gt, nullptr, n_pairs, gate, out);
}

extern "C" int bop_render_depth_host(const bop_model_dev* model_dev, const double* pose,
                                     const double* cam, int n, int height, int width,
                                     float* depth_out) {
  if (n <= 0) return 0;
  BOP_REQUIRE(bop_device_count() > 0, "no CUDA device (there is no CPU fallback)");
  BOP_REQUIRE(model_dev && pose && cam && depth_out, "null pointer");
  DevBuf p, c, oi, o, ws;
  const size_t ob = sizeof(float) * (size_t)n * height * width;
  RC(p.upload(pose, sizeof(double) * 12 * (size_t)n));
  RC(c.upload(cam, sizeof(double) * 4 * (size_t)n));
  RC(zeros_i32(oi, n)); RC(o.alloc(ob));
  bop_model_dev mh;
  BOP_CUDA(cudaMemcpy(&mh, model_dev, sizeof(mh), cudaMemcpyDeviceToHost));
  const int max_geom = mh.n_pts > (mh.n_faces + 1) / 2 ? mh.n_pts : (mh.n_faces + 1) / 2;
  const size_t wsb = bop_render_workspace_bytes(n, height, width, max_geom);
  RC(ws.alloc(wsb));
  RC(bop_render_depth_dev(model_dev, oi.as<int32_t>(), p.as<double>(), c.as<double>(), n, max_geom,
                          height, width, o.as<float>(), ws.p, wsb, nullptr));
  return download(depth_out, o.p, ob);
}

extern "C" int bop_vsd_host(const bop_model_dev* model_dev, const double* pose_est,
                            const double* pose_gt, const void* depth_test,
                            const float* depth_scale, const int32_t* im_idx, const double* K,
                            int n_im, int n_pairs, const bop_vsd_params* prm, double* errors_out) {
  if (n_pairs <= 0) return 0;
  BOP_REQUIRE(bop_device_count() > 0, "no CUDA device (there is no CPU fallback)");
  BOP_REQUIRE(model_dev && pose_est && pose_gt && depth_test && im_idx && K && prm && errors_out && n_im > 0,
              "null pointer");
  const size_t px = (size_t)prm->height * prm->width;
  const size_t db = px * (size_t)n_im * (prm->depth_dtype == BOP_DEPTH_U16 ? 2 : 4);
  const size_t pb = sizeof(double) * 12 * (size_t)n_pairs;
  const size_t eb = sizeof(double) * (size_t)n_pairs * prm->n_taus;
  DevBuf pe, pg, d, sc, ii, kk, oi, e, ws;
  RC(pe.upload(pose_est, pb)); RC(pg.upload(pose_gt, pb));
  RC(d.upload(depth_test, db));
  if (depth_scale) RC(sc.upload(depth_scale, sizeof(float) * (size_t)n_im));
  RC(ii.upload(im_idx, sizeof(int32_t) * (size_t)n_pairs));
  RC(kk.upload(K, sizeof(double) * 9 * (size_t)n_im));
  RC(zeros_i32(oi, n_pairs)); RC(e.alloc(eb));
  bop_model_dev mh;
  BOP_CUDA(cudaMemcpy(&mh, model_dev, sizeof(mh), cudaMemcpyDeviceToHost));
  const int max_geom = mh.n_pts > (mh.n_faces + 1) / 2 ? mh.n_pts : (mh.n_faces + 1) / 2;
  const size_t wsb = bop_vsd_workspace_bytes(n_pairs, prm->height, prm->width, max_geom);
  RC(ws.alloc(wsb));
  RC(bop_vsd_dev(model_dev, oi.as<int32_t>(), pe.as<double>(), pg.as<double>(), d.p,
                 depth_scale ? sc.as<float>() : nullptr, ii.as<int32_t>(), kk.as<double>(), nullptr, n_pairs,
                 max_geom, prm, e.as<double>(), nullptr, ws.p, wsb, nullptr));
  return download(errors_out, e.p, eb);
}
