// NeRF rendering (Testbed::render_nerf / NerfTracer, src/testbed_nerf.cu:1556-1987). Placeholder
// body until the ray-tile kernel lands later in this round.
#include "testbed.h"

namespace ngpb {

void Testbed::render_nerf(uint32_t, uint32_t, const float*, float, float, float, float, uint32_t, uint32_t, uint32_t, float*, float*) {
	throw std::runtime_error{"ngp-b200: render_nerf is not built yet"};
}

} // namespace ngpb
