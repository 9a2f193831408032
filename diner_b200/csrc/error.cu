// Error reporting + ABI version for the diner_b200 C ABI.
#include "common.cuh"
#include <string.h>

namespace dnr {
static thread_local char g_last_error[512] = "";
void set_last_error(const char* where, cudaError_t e) {
  snprintf(g_last_error, sizeof(g_last_error), "%s: %s (%s)", where, cudaGetErrorName(e), cudaGetErrorString(e));
}
void set_last_error_msg(const char* msg) { snprintf(g_last_error, sizeof(g_last_error), "%s", msg); }

constexpr int kMaxDevices = 64;
static DeviceInfo g_devices[kMaxDevices];      // zero-initialised: num_sms == 0 means "not looked up yet"
DeviceInfo* device_info() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
  DeviceInfo* d = &g_devices[dev];
  if (d->num_sms == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return nullptr;
    d->num_sms = n;
  }
  return d;
}
}  // namespace dnr

DNR_EXPORT const char* dnr_last_error() { return dnr::g_last_error; }
DNR_EXPORT int dnr_abi_version() { return 19; }
