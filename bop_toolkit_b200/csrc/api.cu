// Library-level plumbing of libbop_b200: error reporting, launch counter, model upload
// and the host-buffer (`*_host`) convenience layer declared in include/bop_b200.h.
#include <atomic>
#include <string>
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
