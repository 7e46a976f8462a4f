// runtime.cu -- error reporting, device checks, version.
#include "common.cuh"
#include <atomic>
#include <mutex>

namespace b2d {

char* err_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

int set_err(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace b2d

static std::atomic<long long> g_launches{0};
namespace b2d { void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); } }

extern "C" int b2d_version(void) { return 100; }  // 0.1.0
extern "C" long long b2d_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" const char* b2d_last_error(void) { return b2d::err_buf(); }

extern "C" int b2d_check_device(void) {
  int dev = 0, major = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return b2d::set_err(B2D_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess) return b2d::set_err(B2D_ERR_CUDA, "cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
  if (major != 10)
    return b2d::set_err(B2D_ERR_ARCH, "libb200diff is built for sm_100a only; device %d is sm_%d*", dev, major);
  return B2D_OK;
}

// ---------------------------------------------------------------------------
// TMA tensor maps
// ---------------------------------------------------------------------------
#include "tmap.cuh"

namespace b2d {

PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, []() {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  });
  return fn;
}

static int make_tmap_typed(CUtensorMap* m, CUtensorMapDataType dtype, const void* base, int rank,
                           const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                           CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B);

int make_tmap_16b(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box) {
  return make_tmap_typed(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, base, rank, dims, strides_bytes, box);
}

int make_tmap_f32(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                  const uint64_t* strides_bytes, const uint32_t* box) {
  return make_tmap_typed(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, base, rank, dims, strides_bytes, box);
}

int make_tmap_f32_linear(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                         const uint64_t* strides_bytes, const uint32_t* box) {
  return make_tmap_typed(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, base, rank, dims, strides_bytes, box,
                         CU_TENSOR_MAP_SWIZZLE_NONE);
}

static int make_tmap_typed(CUtensorMap* m, CUtensorMapDataType dtype, const void* base, int rank,
                           const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                           CUtensorMapSwizzle swz) {
  PFN_encodeTiled enc = get_encode_tiled();
  B2D_REQUIRE(enc != nullptr, B2D_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  B2D_REQUIRE(aligned16(base), B2D_ERR_ALIGN, "tensor map base %p is not 16-byte aligned", base);
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) {
    gstr[i] = strides_bytes[i];
    B2D_REQUIRE(gstr[i] % 16 == 0, B2D_ERR_ALIGN, "tensor map stride %d (%llu B) not a multiple of 16", i,
                (unsigned long long)gstr[i]);
  }
  CUresult r = enc(m, dtype, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx,
                   es, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B2D_REQUIRE(r == CUDA_SUCCESS, B2D_ERR_CUDA,
              "cuTensorMapEncodeTiled failed (%d): rank=%d dims=[%llu,%llu,%llu,%llu] box=[%u,%u,%u,%u]", (int)r,
              rank, (unsigned long long)gdim[0], (unsigned long long)(rank > 1 ? gdim[1] : 0),
              (unsigned long long)(rank > 2 ? gdim[2] : 0), (unsigned long long)(rank > 3 ? gdim[3] : 0), bx[0],
              rank > 1 ? bx[1] : 0, rank > 2 ? bx[2] : 0, rank > 3 ? bx[3] : 0);
  return B2D_OK;
}

}  // namespace b2d
