// api.cu -- error string, ABI version and per-device attribute cache.
#include "common.cuh"

#include <string.h>

#include <mutex>
#include <vector>

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

namespace {
struct AttrKey {
  const void* kern;
  int dev, a, b;
  bool operator==(const AttrKey& o) const { return kern == o.kern && dev == o.dev && a == o.a && b == o.b; }
};
struct AttrEntry {
  AttrKey key;
  int value;
};
// a handful of kernels x devices: a linear scan under a mutex is cheaper than the driver call it saves
std::mutex g_attr_mu;
std::vector<AttrEntry> g_smem_set, g_occ;
}  // namespace

cudaError_t set_max_smem_cached(const void* kern, int bytes) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  {
    std::lock_guard<std::mutex> lk(g_attr_mu);
    for (const AttrEntry& x : g_smem_set)
      if (x.key.kern == kern && x.key.dev == dev && x.value >= bytes) return cudaSuccess;
  }
  e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lk(g_attr_mu);
  for (AttrEntry& x : g_smem_set)
    if (x.key.kern == kern && x.key.dev == dev) { x.value = bytes; return cudaSuccess; }
  g_smem_set.push_back(AttrEntry{AttrKey{kern, dev, 0, 0}, bytes});
  return cudaSuccess;
}

cudaError_t max_blocks_per_sm_cached(const void* kern, int threads, size_t smem, int* out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const AttrKey key{kern, dev, threads, (int)smem};
  {
    std::lock_guard<std::mutex> lk(g_attr_mu);
    for (const AttrEntry& x : g_occ)
      if (x.key == key) { *out = x.value; return cudaSuccess; }
  }
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(out, kern, threads, smem);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lk(g_attr_mu);
  g_occ.push_back(AttrEntry{key, *out});
  return cudaSuccess;
}

}  // namespace b200

extern "C" int b200_ctc_abi_version(void) { return B200_CTC_ABI_VERSION; }
extern "C" const char* b200_last_error(void) { return b200::g_err; }
