// Library-level entry points: version, thread-local error text, device check.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace rg {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* where) {
  set_error("CUDA error at %s: %s", where, cudaGetErrorString(e));
  return RG_E_CUDA;
}

}  // namespace rg

extern "C" {

int rg_version(void) { return RG_VERSION; }

const char* rg_last_error(void) { return rg::g_err; }

int rg_device_check(int device) {
  cudaDeviceProp p;
  cudaError_t e = cudaGetDeviceProperties(&p, device);
  if (e != cudaSuccess) return rg::cuda_fail(e, "cudaGetDeviceProperties");
  if (p.major != 10) {
    rg::set_error("libraygun_b200 is built for sm_100a only; device %d is sm_%d%d", device, p.major, p.minor);
    return RG_E_ARCH;
  }
  return RG_OK;
}

}  // extern "C"
