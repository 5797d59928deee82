// kvm_abi.cu - error reporting and device queries of the C ABI.
#include <stdarg.h>
#include <string.h>

#include "kvm_common.cuh"

static thread_local char g_err[512] = "";

void kvm_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int kvm_num_sms() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
      sms = 148;  // B200
    cached = sms;
    cached_dev = dev;
  }
  return cached;
}

extern "C" const char* kvm_last_error(void) { return g_err; }
extern "C" int kvm_abi_version(void) { return KVM_ABI_VERSION; }
extern "C" int kvm_sm_count(void) {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
    kvm_set_error("kvm_sm_count: no CUDA device");
    return KVM_ERR_CUDA;
  }
  return sms;
}
