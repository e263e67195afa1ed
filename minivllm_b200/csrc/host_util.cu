// host_util.cu — error plumbing and TMA descriptor encoding for libmvb200.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace mvb200 {

static thread_local char g_err[512] = "";
static thread_local const char* g_kernel = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

// distinct kernel families launched by this process (names are string literals; a handful at most)
static const char* g_seen[64];
static int g_seen_n = 0;

void note_kernel(const char* name) {
  if (g_kernel == name) return;
  g_kernel = name;
  int n = g_seen_n;
  for (int i = 0; i < n; ++i)
    if (g_seen[i] == name || strcmp(g_seen[i], name) == 0) return;
  if (n < 64) {
    g_seen[n] = name;
    g_seen_n = n + 1;  // racing threads may drop or duplicate an entry: diagnostic only
  }
}

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "%s: %s", what, cudaGetErrorString(e));
  return MVB200_OK;
}

// Per-device caches: cudaFuncSetAttribute and the SM count are properties of the CURRENT device, and one
// thread may drive several GPUs (torch.cuda.device switches in a single-process multi-GPU host).
static constexpr int kMaxDevices = 64;

static int current_device_slot() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= kMaxDevices) d = 0;
  return d;
}

int device_sm_count() {
  static int count[kMaxDevices];
  const int d = current_device_slot();
  if (count[d] <= 0) {
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, d);
    count[d] = n > 0 ? n : 148;
  }
  return count[d];
}

int optin_dynamic_smem(const void* kernel, int bytes, SmemOptIn* cache, const char* what) {
  const int d = current_device_slot();
  if (cache->bytes[d] >= bytes) return MVB200_OK;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess)
    return fail(MVB200_ECUDA, "%s: smem opt-in %d B failed on device %d: %s", what, bytes, d,
                cudaGetErrorString(e));
  cache->bytes[d] = bytes;
  return MVB200_OK;
}

int resident_ctas_per_sm(const void* kernel, int threads, int smem_bytes, ResidentCtas* cache, const char* what,
                         int* out) {
  const int d = current_device_slot();
  if (cache->smem[d] != smem_bytes || cache->ctas[d] <= 0) {
    int n = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, (size_t)smem_bytes);
    if (e != cudaSuccess)
      return fail(MVB200_ECUDA, "%s: occupancy query failed on device %d: %s", what, d, cudaGetErrorString(e));
    if (n < 1)
      return fail(MVB200_EUNSUPPORTED, "%s: not one CTA (%d threads, %d B smem) fits an SM of device %d", what,
                  threads, smem_bytes, d);
    cache->smem[d] = smem_bytes;
    cache->ctas[d] = n;
  }
  *out = cache->ctas[d];
  return MVB200_OK;
}

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("MVB200_NO_PDL");
    return !(e && e[0] && e[0] != '0');
  }();
  return on;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn resolve_encode() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

int make_tensor_map(CUtensorMap* map, const void* base, int elem_bytes, int rank,
                    const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                    bool swizzle128) {
  EncodeTiledFn enc = resolve_encode();
  if (!enc) return fail(MVB200_ECUDA, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t gdims[5];
  cuuint64_t gstr[5];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i];
  }
  CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16
                                           : (elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_UINT32
                                                              : CU_TENSOR_MAP_DATA_TYPE_UINT8);
  CUresult r = enc(map, dt, (cuuint32_t)rank, const_cast<void*>(base), gdims, gstr, gbox, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(MVB200_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return MVB200_OK;
}

}  // namespace mvb200

extern "C" {
int mvb200_version(void) { return MVB200_VERSION; }
const char* mvb200_last_error(void) { return mvb200::g_err; }
const char* mvb200_last_kernel(void) { return mvb200::g_kernel; }
int mvb200_kernels_seen(char* buf, int cap) {
  int used = 0;
  if (cap > 0) buf[0] = 0;
  for (int i = 0; i < mvb200::g_seen_n; ++i) {
    int w = snprintf(buf + used, cap > used ? (size_t)(cap - used) : 0, "%s%s", i ? "," : "",
                     mvb200::g_seen[i]);
    if (w < 0 || used + w >= cap) break;
    used += w;
  }
  return mvb200::g_seen_n;
}
}
