// Library-wide plumbing of the C-ABI: error string, device check, TMA descriptor encoding (driver entry point is
// resolved at run time so the library links without libcuda and loads on a CPU-only build box).
#include "common.cuh"
#include "../../include/avion_b200.h"
#include <stdarg.h>
#include <string.h>

static thread_local char g_err[512] = "";

void avn_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* avn_last_error(void) { return g_err; }
extern "C" int avn_version(void) { return 101; }
// binding check: a host-language mirror of avn_gemm_desc (ctypes, cgo, JNI ...) compares its own size with this one at load time
extern "C" int avn_sizeof_gemm_desc(void) { return (int)sizeof(avn_gemm_desc); }

int avn_num_sms() {
    static int sms[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (sms[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        sms[dev] = n;
    }
    return sms[dev];
}

extern "C" int avn_device_check(void) {
    int dev = 0, major = 0;
    AVN_CHECK_CUDA(cudaGetDevice(&dev));
    AVN_CHECK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    AVN_REQUIRE(major == 10, "avion_b200 kernels are built for sm_100a only; device has compute capability %d.x", major);
    return AVN_OK;
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
    static PFN_encodeTiled fn = nullptr;
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiled)p;
    }
    return fn;
}

int avn_make_tmap(CUtensorMap* out, const void* base, CUtensorMapDataType dtype, int elem_bytes, int rank,
                  const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box, bool swizzle128) {
    PFN_encodeTiled enc = get_encode();
    if (enc == nullptr) {
        avn_set_error("cuTensorMapEncodeTiled not available from the driver");
        return AVN_ERR_DRIVER;
    }
    cuuint64_t gdim[5];
    cuuint64_t gstr[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
    for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
    (void)elem_bytes;
    CUresult r = enc(out, dtype, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        avn_set_error("cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu,%llu,%llu] strides [%llu,%llu] box [%u,%u,%u] base %p",
                      (int)r, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                      (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 1 ? strides_bytes[0] : 0),
                      (unsigned long long)(rank > 2 ? strides_bytes[1] : 0), box[0], rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0, base);
        return AVN_ERR_DRIVER;
    }
    return AVN_OK;
}
