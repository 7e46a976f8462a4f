// tmap.cuh -- host-side TMA tensor-map construction (driver entry point resolved
// through the runtime so libcuda need not be linked).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include "common.cuh"

namespace b2d {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);

PFN_encodeTiled get_encode_tiled();

// bf16 tensor map, SWIZZLE_128B, zero fill out of bounds.
// dims[0] is the innermost (contiguous) dimension; strides_bytes[i] is the pitch of dims[i+1].
int make_tmap_16b(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box);
// same for fp32 elements (box[0] * 4 bytes <= 128)
int make_tmap_f32(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                  const uint64_t* strides_bytes, const uint32_t* box);
// fp32, no swizzle (plain row-major box in shared memory)
int make_tmap_f32_linear(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                         const uint64_t* strides_bytes, const uint32_t* box);

}  // namespace b2d
