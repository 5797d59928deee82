// kvm_abi.cu - error reporting and device queries of the C ABI.
#include <stdarg.h>
#include <stdlib.h>
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

// CTAs the streaming kernels leave out of a machine-filling grid (kvm_stream_grid); settable for
// experiments: KVM_TAIL_SLOTS=0 restores full grids.
static int g_tail_slots = -1;
int kvm_tail_slots() {
  if (g_tail_slots < 0) {
    const char* e = getenv("KVM_TAIL_SLOTS");
    int v = e ? atoi(e) : 16;
    g_tail_slots = v < 0 ? 0 : v;
  }
  return g_tail_slots;
}
extern "C" int kvm_set_tail_slots(int32_t slots) {
  g_tail_slots = slots < 0 ? 0 : slots;
  return KVM_OK;
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
