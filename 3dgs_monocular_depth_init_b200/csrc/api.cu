// api.cu -- error string, ABI/version probes.
#include <stdarg.h>
#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace mdi {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ---- optional per-kernel device timing (CUDA events around every launch) -----------------------
bool g_prof = false;
namespace {
struct ProfRange { const char* name; cudaEvent_t e0, e1; cudaStream_t st; bool closed; };
std::mutex g_prof_mutex;
std::vector<ProfRange> g_ranges;
}  // namespace
void prof_mark(const char* name, cudaStream_t st) {
  if (!g_prof) return;
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  if (!g_ranges.empty() && !g_ranges.back().closed) {
    cudaEventRecord(g_ranges.back().e1, g_ranges.back().st);
    g_ranges.back().closed = true;
  }
  if (name) {
    ProfRange r{name, nullptr, nullptr, st, false};
    cudaEventCreate(&r.e0); cudaEventCreate(&r.e1);
    cudaEventRecord(r.e0, st);
    g_ranges.push_back(r);
  }
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

extern "C" void mdi_profile_enable(int on) { mdi::g_prof = on != 0; }

// Synchronises, sums the recorded ranges by kernel name, clears the record.  names: '\n'-separated
// list written into names_buf; ms / calls: per name.  Returns the number of names (<= cap).
extern "C" int mdi_profile_collect(char* names_buf, int buf_len, float* ms, int32_t* calls, int cap) {
  std::lock_guard<std::mutex> lock(mdi::g_prof_mutex);
  std::vector<std::string> names;
  std::vector<float> tot;
  std::vector<int> cnt;
  for (auto& r : mdi::g_ranges) {
    if (r.closed && cudaEventSynchronize(r.e1) == cudaSuccess) {
      float t = 0.f;
      cudaEventElapsedTime(&t, r.e0, r.e1);
      size_t k = 0;
      for (; k < names.size(); ++k) if (names[k] == r.name) break;
      if (k == names.size()) { names.push_back(r.name); tot.push_back(0.f); cnt.push_back(0); }
      tot[k] += t; cnt[k] += 1;
    }
    cudaEventDestroy(r.e0); cudaEventDestroy(r.e1);
  }
  mdi::g_ranges.clear();
  cudaGetLastError();
  std::string joined;
  int n = 0;
  for (size_t k = 0; k < names.size() && n < cap; ++k, ++n) {
    if (k) joined += "\n";
    joined += names[k];
    ms[n] = tot[k]; calls[n] = cnt[k];
  }
  if (names_buf && buf_len > 0) {
    strncpy(names_buf, joined.c_str(), (size_t)buf_len - 1);
    names_buf[buf_len - 1] = 0;
  }
  return n;
}
