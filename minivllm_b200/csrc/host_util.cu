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

void note_kernel(const char* name) { g_kernel = name; }

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "%s: %s", what, cudaGetErrorString(e));
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
}
