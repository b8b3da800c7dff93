// Host-side TMA tensor-map encoding (cuTensorMapEncodeTiled through the runtime's driver entry
// point, so the library does not link libcuda directly).
#pragma once
#include <cuda.h>
#include <cstdint>

namespace fm {

// dims / box: innermost first; strides_bytes: rank - 1 entries (dimension 1 upwards).
// Out-of-bounds elements are zero-filled.  Returns false on failure.
bool encode_tiled(CUtensorMap* m, CUtensorMapDataType dt, int rank, const void* ptr, const uint64_t* dims,
                  const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swizzle);

}  // namespace fm
