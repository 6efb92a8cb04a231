// tmap.h -- opaque 128-byte TMA descriptor passed to kernels as a __grid_constant__ parameter.
#pragma once
#include <stdint.h>

namespace olb {

struct alignas(64) TensorMap {
    uint64_t opaque[16];
};

// Descriptor over a frame-interleaved fp32 block [frames][units] (unit index fastest) with a
// box of box_units x box_frames elements.  Out-of-bounds parts of a box read as zero and are
// clipped on store.
int make_tmap_frames(TensorMap *out, const float *base, uint64_t frames, uint64_t units, uint32_t box_units, uint32_t box_frames);

}  // namespace olb
