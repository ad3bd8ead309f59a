// api.cu -- error string, ABI/version probes.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace mdi {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace mdi

extern "C" const char* mdi_last_error(void) { return mdi::g_err; }
extern "C" int mdi_abi_version(void) { return MDI_ABI_VERSION; }
extern "C" int mdi_sizeof_view(void) { return (int)sizeof(MdiView); }

extern "C" int mdi_check_device(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) {
    cudaGetLastError();
    mdi::set_error("mdi_check_device: no CUDA device %d (count %d)", device, n);
    return MDI_ENOGPU;
  }
  cudaDeviceProp p;
  MDI_CUDA(cudaGetDeviceProperties(&p, device), "mdi_check_device");
  if (p.major != 10) {
    mdi::set_error("mdi_check_device: device %d is sm_%d%d, this library is sm_100a only", device,
                   p.major, p.minor);
    return MDI_ENOGPU;
  }
  return MDI_OK;
}
