// Host-side plumbing of libdualhyp.so: last-error string, device init, TMA descriptor encoding through a
// driver entry point resolved at run time (the library links against cudart only, so it loads -- and its
// symbols can be inspected -- on machines without a GPU driver).
#include <stdarg.h>
#include <string.h>

#include <stdlib.h>

#include "common.cuh"
#include "../../include/dualhyp.h"

namespace dh {

static thread_local char g_err[1024] = "";
static int g_sm_count = 0;
static int g_device = -1;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static encode_tiled_fn g_encode = nullptr;

static int resolve_encode() {
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
    set_error("cuTensorMapEncodeTiled not available from the driver (%s)", cudaGetErrorString(e));
    return -1;
  }
  g_encode = reinterpret_cast<encode_tiled_fn>(fn);
  return 0;
}

int make_tensor_map(CUtensorMap* map, const void* base, CUtensorMapDataType dtype, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swizzle) {
  if (resolve_encode()) return -1;
  // The encoder is a driver-API call and needs a current context.  PyTorch's autograd worker threads (where
  // backward kernels are launched from) may never have bound the primary context: bind it once per thread.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    cudaFree(0);
    ctx_bound = true;
  }
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t gbox[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
  }
  CUresult r = g_encode(map, dtype, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, gbox, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): base=%p rank=%d dims=[%llu,%llu,%llu] stride1=%llu box=[%u,%u,%u]", (int)r,
              base, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
              (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 1 ? strides_bytes[0] : 0), box[0],
              rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0);
    return -1;
  }
  return 0;
}

int sm_count() { return g_sm_count > 0 ? g_sm_count : 148; }

int pdl_level() {
  static int v = -1;
  if (v < 0) {
    // 0: off.  1: eager launches only (neutral on the power-capped training step, 35.8 vs 35.8 ms).  2 (default): also under
    // stream capture -- programmatic edges in the decode graph, where the GEMMs fetch frozen weight tiles ahead of the
    // wait (DH_B_FROZEN): 2.07 -> 1.97 ms per token step of 128 sequences.
    const char* e = getenv("DH_PDL");
    v = (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 2;
  }
  return v;
}
bool pdl_enabled() { return pdl_level() >= 1; }

}  // namespace dh

extern "C" int dh_abi_version(void) { return DH_ABI_VERSION; }
extern "C" const char* dh_last_error(void) { return dh::g_err; }
extern "C" int dh_sm_count(void) { return dh::sm_count(); }

extern "C" int dh_init(int device) {
  using namespace dh;
  if (g_device == device && g_sm_count > 0) return 0;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    set_error("dh_init: no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    return -1;
  }
  DH_CHECK(device >= 0 && device < n, "dh_init: device %d out of range (%d devices)", device, n);
  cudaDeviceProp prop;
  DH_CUDA(cudaGetDeviceProperties(&prop, device));
  DH_CHECK(prop.major == 10, "dh_init: device %d is sm_%d%d; libdualhyp is built for sm_100a only", device, prop.major,
           prop.minor);
  g_sm_count = prop.multiProcessorCount;
  g_device = device;
  if (resolve_encode()) return -1;
  return 0;
}
