// Library state, error reporting and device checks for libnerf_b200.
#include <atomic>
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace nb {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static thread_local int g_limit_mlp_clusters = 0, g_limit_wgrad_ctas = 0;
int limit_mlp_clusters() { return g_limit_mlp_clusters; }
int limit_wgrad_ctas() { return g_limit_wgrad_ctas; }

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  count_launch();
  return 0;
}

int init_mlp_tc(int device);      // mlp_tc.cu
int init_mlp_simt(int device);    // mlp_simt.cu
int init_mlp_bwd(int device);     // mlp_bwd.cu
int init_mlp_ingrad(int device);  // mlp_ingrad.cu

}  // namespace nb

extern "C" {

int nb_version(void) { return NB_VERSION; }

const char* nb_last_error(void) { return nb::g_err; }

int64_t nb_launch_count(void) { return nb::g_launches.load(std::memory_order_relaxed); }

int nb_init(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    nb::set_error("nb_init: no CUDA device (%s); libnerf_b200 has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "count=0");
    return NB_E_DEVICE;
  }
  if (device < 0 || device >= n) {
    nb::set_error("nb_init: device %d out of range (%d devices)", device, n);
    return NB_E_ARG;
  }
  int major = 0, minor = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device);
  if (major != 10) {
    nb::set_error("nb_init: device %d is sm_%d%d; this library contains sm_100a code only", device, major, minor);
    return NB_E_DEVICE;
  }
  int cur = 0;
  cudaGetDevice(&cur);
  cudaSetDevice(device);
  int rc = nb::init_mlp_simt(device);
  if (rc == 0) rc = nb::init_mlp_tc(device);
  if (rc == 0) rc = nb::init_mlp_bwd(device);
  if (rc == 0) rc = nb::init_mlp_ingrad(device);
  cudaSetDevice(cur);
  return rc;
}

int nb_shutdown(void) { return 0; }

int nb_set_launch_limits(int mlp_clusters, int wgrad_ctas) {
  if (mlp_clusters < 0 || wgrad_ctas < 0) {
    nb::set_error("nb_set_launch_limits: negative limit");
    return NB_E_ARG;
  }
  nb::g_limit_mlp_clusters = mlp_clusters;
  nb::g_limit_wgrad_ctas = wgrad_ctas;
  return 0;
}

}  // extern "C"
