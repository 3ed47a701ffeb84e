// Host-side TMA descriptor encoding (defined in runtime.cu).
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace pb200 {

// bf16 tiled tensor map, dims[0] innermost (contiguous); strides_bytes has rank-1 entries
// (dims 1..rank-1), each a multiple of 16.  Out-of-bounds box elements read as zero.
int encode_tma_bf16(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);

}  // namespace pb200
