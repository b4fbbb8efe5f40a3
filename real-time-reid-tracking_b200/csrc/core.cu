// Error reporting, ABI version and device check.
#include "common.cuh"

namespace reidb200 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace reidb200

extern "C" const char* reidb200_last_error(void) { return reidb200::g_err; }

extern "C" int reidb200_abi_version(void) { return 8; }

extern "C" int reidb200_check_device(void) {
  int dev = 0;
  RB_CUDA(cudaGetDevice(&dev));
  int major = 0, minor = 0;
  RB_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  RB_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (major != 10) {
    reidb200::set_error("device %d is sm_%d%d; libreidb200 is built for sm_100a only", dev, major, minor);
    return REIDB200_ECUDA;
  }
  return REIDB200_OK;
}

// ---- optional per-kernel timing with CUDA events recorded on the launch stream (bench.py roofline numbers)
#include <map>
#include <string>
#include <vector>
#include <mutex>

namespace reidb200 {
namespace {
struct ProfSpan { cudaEvent_t a, b; };
std::mutex g_prof_mu;
bool g_prof_on = false;
std::map<std::string, std::vector<ProfSpan>> g_prof;
}  // namespace

ProfScope::ProfScope(const char* name, cudaStream_t st) : name_(name), st_(st), on_(false) {
  if (!g_prof_on) return;
  on_ = true;
  cudaEventCreate(&a_);
  cudaEventCreate(&b_);
  cudaEventRecord(a_, st_);
}
ProfScope::~ProfScope() {
  if (!on_) return;
  cudaEventRecord(b_, st_);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof[name_].push_back(ProfSpan{a_, b_});
}
}  // namespace reidb200

extern "C" int reidb200_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(reidb200::g_prof_mu);
  reidb200::g_prof_on = on != 0;
  for (auto& kv : reidb200::g_prof)
    for (auto& s : kv.second) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
  reidb200::g_prof.clear();
  return REIDB200_OK;
}

// Sum of the recorded spans of `name` (after synchronising them): launches and total milliseconds.
extern "C" int reidb200_profile_read(const char* name, int* launches, double* total_ms) {
  RB_REQUIRE(name && launches && total_ms, "profile_read: bad argument");
  std::lock_guard<std::mutex> lk(reidb200::g_prof_mu);
  *launches = 0;
  *total_ms = 0.0;
  auto it = reidb200::g_prof.find(name);
  if (it == reidb200::g_prof.end()) return REIDB200_OK;
  for (auto& s : it->second) {
    RB_CUDA(cudaEventSynchronize(s.b));
    float ms = 0.f;
    RB_CUDA(cudaEventElapsedTime(&ms, s.a, s.b));
    *total_ms += ms;
    *launches += 1;
  }
  return REIDB200_OK;
}

// Names seen so far, '\n'-separated, into buf (truncated to cap).
extern "C" int reidb200_profile_names(char* buf, int cap) {
  RB_REQUIRE(buf && cap > 0, "profile_names: bad argument");
  std::lock_guard<std::mutex> lk(reidb200::g_prof_mu);
  std::string all;
  for (auto& kv : reidb200::g_prof) { all += kv.first; all += '\n'; }
  snprintf(buf, cap, "%s", all.c_str());
  return REIDB200_OK;
}
