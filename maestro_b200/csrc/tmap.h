// Cached CUtensorMap construction (bf16 tiled maps, rank 2 or 3) shared by the TMA-fed kernels.
#pragma once
#include <cuda.h>
#include <stdint.h>

// dims[i] in elements (dims[0] innermost / contiguous), strides[i] in ELEMENTS for dims 1..rank-1,
// box[i] in elements, swizzle_bytes in {32, 64, 128}.  Returns 0 or a negative code.
int mb_get_tmap(CUtensorMap* out, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides,
                const uint32_t* box, int swizzle_bytes);
int mb_num_sms();
