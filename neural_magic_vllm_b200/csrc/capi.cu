// C-ABI plumbing of libb200vllm: error reporting, enum parsing, device attribute queries.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    // clear the sticky-less launch error so that the next call starts clean
    cudaGetLastError();
    set_error("%s: CUDA error: %s", what, cudaGetErrorString(e));
    return B200_ECUDA;
  }
  return B200_OK;
}

// read per launch (a getenv is noise next to a launch) so that a test can flip it in-process
bool pdl_enabled() {
  const char* e = getenv("B200_PDL");
  return !(e != nullptr && e[0] == '0');
}

int num_sms() {
  static int cached[kMaxDevices] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    cudaGetLastError();
    return 148;
  }
  if (dev >= 0 && dev < kMaxDevices && cached[dev] > 0) return cached[dev];
  int v = 0;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) {
    cudaGetLastError();
    v = 148;
  }
  if (dev >= 0 && dev < kMaxDevices) cached[dev] = v;
  return v;
}

int env_int_once(const char* name, int* cache, int unset) {
  // cache[0]: 0 = not read yet; cache[1]: the value (`unset` when the variable is absent)
  if (cache[0] == 0) {
    const char* e = getenv(name);
    cache[1] = e != nullptr ? atoi(e) : unset;
    cache[0] = 1;
  }
  return cache[1];
}

}  // namespace b200

extern "C" {

int b200_abi_version(void) { return B200_VLLM_ABI_VERSION; }

const char* b200_last_error(void) { return b200::g_err; }

int b200_parse_kv_cache_dtype(const char* s) {
  if (s == nullptr) return B200_EINVAL;
  if (!strcmp(s, "auto")) return B200_KV_AUTO;
  if (!strcmp(s, "fp8") || !strcmp(s, "fp8_e4m3")) return B200_KV_FP8_E4M3;
  if (!strcmp(s, "fp8_e5m2")) return B200_KV_FP8_E5M2;
  b200::set_error("Unsupported data type of kv cache: %s", s);
  return B200_EINVAL;
}

// csrc/cuda_utils_kernels.cu:12-21
int b200_get_device_attribute(int32_t attribute, int32_t device_id, int32_t* value) {
  int device = device_id;
  if (device_id < 0) {
    if (cudaGetDevice(&device) != cudaSuccess) {
      b200::set_error("cudaGetDevice failed");
      return B200_ECUDA;
    }
  }
  int v = 0;
  cudaError_t e = cudaDeviceGetAttribute(&v, static_cast<cudaDeviceAttr>(attribute), device);
  if (e != cudaSuccess) {
    cudaGetLastError();
    b200::set_error("cudaDeviceGetAttribute(%d, %d): %s", attribute, device, cudaGetErrorString(e));
    return B200_ECUDA;
  }
  *value = v;
  return B200_OK;
}

// csrc/cuda_utils_kernels.cu:23-29
int b200_get_max_shared_memory_per_block_device_attribute(int32_t device_id, int32_t* value) {
  return b200_get_device_attribute((int32_t)cudaDevAttrMaxSharedMemoryPerBlockOptin, device_id,
                                   value);
}

}  // extern "C"
