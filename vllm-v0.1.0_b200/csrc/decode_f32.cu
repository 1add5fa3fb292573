// Instantiates the paged decode kernels for one element type (kept in its own
// translation unit so the three types compile in parallel).
#define PA_DECODE_IMPL
#include "decode_attention.cuh"

namespace pa {
int launch_decode_f32(const DecodeLaunch& l) { return launch_dtype<F32>(l); }
}  // namespace pa
