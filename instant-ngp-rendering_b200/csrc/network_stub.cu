// Placeholder until mlp.cu lands (replaced in the same round).
#include "module.h"
namespace ngpb {
Module* create_network_with_input_encoding_module(uint32_t, uint32_t, const Json&, const Json&) {
	throw std::runtime_error{"ngp-b200: networks are not built yet"};
}
Module* create_nerf_network_module(uint32_t, uint32_t, uint32_t, uint32_t, const Json&, const Json&, const Json&, const Json&) {
	throw std::runtime_error{"ngp-b200: NerfNetwork is not built yet"};
}
}
