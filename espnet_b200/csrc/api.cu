// api.cu -- error string, ABI version and per-device attribute cache.
#include "common.cuh"

#include <string.h>

namespace b200 {

static thread_local char g_err[512] = {0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return B200_ERR_CUDA;
}

// Device attributes are immutable, so a racy fill of this cache is benign.
static int g_sms[64] = {0};
static int g_smem[64] = {0};

int num_sms() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (g_sms[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    g_sms[dev] = n;
  }
  return g_sms[dev];
}

int max_smem_optin() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 227 * 1024;
  if (g_smem[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess || n <= 0)
      n = 227 * 1024;
    g_smem[dev] = n;
  }
  return g_smem[dev];
}

}  // namespace b200

extern "C" int b200_ctc_abi_version(void) { return B200_CTC_ABI_VERSION; }
extern "C" const char* b200_last_error(void) { return b200::g_err; }
