// libb2seg.so — ABI bookkeeping: version, thread-local error text, device query,
// TMA descriptor encoding via the driver entry point.
#include <stdlib.h>
#include <cudaTypedefs.h>
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace {
thread_local char g_err[512] = "";
}

void b2seg_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" int b2seg_abi_version(void) { return B2SEG_ABI_VERSION; }

extern "C" const char* b2seg_last_error(void) { return g_err; }

extern "C" int b2seg_device_info(char* name, int len, int* sms, int* cc) {
  int dev = 0;
  B2_CHECK_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  B2_CHECK_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (name && len > 0) {
    strncpy(name, prop.name, (size_t)len - 1);
    name[len - 1] = 0;
  }
  if (sms) *sms = prop.multiProcessorCount;
  if (cc) *cc = prop.major * 10 + prop.minor;
  return B2SEG_OK;
}

bool b2_pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B2SEG_PDL");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}

int b2_encode_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                        const uint64_t* strides_bytes, const uint32_t* box, int swizzle) {
  static PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
  if (!encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
      b2seg_set_error("cuTensorMapEncodeTiled entry point unavailable (%s)",
                      cudaGetErrorString(e));
      return B2SEG_ECUDA;
    }
    encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
  }
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bdim[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) != 0) {
    b2seg_set_error("TMA base address %p not 16-byte aligned", base);
    return B2SEG_EINVAL;
  }
  for (int i = 0; i + 1 < rank; ++i) {
    if (gstr[i] % 16 != 0) {
      b2seg_set_error("TMA stride[%d]=%llu not a multiple of 16 bytes", i,
                      (unsigned long long)gstr[i]);
      return B2SEG_EINVAL;
    }
  }
  CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank,
                      const_cast<void*>(base), gdim, gstr, bdim, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE,
                      swizzle == 1 ? CU_TENSOR_MAP_SWIZZLE_128B : swizzle == 2 ? CU_TENSOR_MAP_SWIZZLE_64B
                                   : swizzle == 3 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE,
                      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b2seg_set_error("cuTensorMapEncodeTiled failed: CUresult %d (rank %d, dims %llu,%llu,%llu,%llu "
                    "box %u,%u,%u,%u)",
                    (int)r, rank, (unsigned long long)dims[0],
                    (unsigned long long)(rank > 1 ? dims[1] : 0),
                    (unsigned long long)(rank > 2 ? dims[2] : 0),
                    (unsigned long long)(rank > 3 ? dims[3] : 0), box[0], rank > 1 ? box[1] : 0,
                    rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    return B2SEG_ECUDA;
  }
  return B2SEG_OK;
}
