// Host helper shared by the TMA-fed kernels: encode an fp16 tiled tensor map with the 128-byte swizzle
// (cuTensorMapEncodeTiled fetched through the runtime's driver entry point; no -lcuda link dependency).
#pragma once
#include <cuda.h>
#include <cstdint>

namespace ace {

// rank <= 5; dims innermost first; strides_bytes has rank-1 entries (dims 1..rank-1); box innermost first.
// OOB elements read as zero.  Returns false when the driver entry point is unavailable or the encode fails.
bool make_tmap_f16_sw128(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                         const uint32_t* box);

// same for an fp32 tensor (used for the TMA reduce-add epilogue)
bool make_tmap_f32_sw128(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                         const uint32_t* box);

}  // namespace ace
