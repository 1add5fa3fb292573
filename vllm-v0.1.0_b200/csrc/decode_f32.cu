// Instantiates the paged decode kernels for one element type (kept in its own
// translation unit so the three types compile in parallel).
#define PA_DECODE_IMPL
#include "decode_persistent.cuh"

namespace pa {
int launch_decode_f32(const DecodeLaunch& l) { return l.persistent ? launch_persistent_dtype<F32>(l) : launch_dtype<F32>(l); }
}  // namespace pa
