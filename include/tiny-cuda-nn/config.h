/* create_from_config (include/tiny-cuda-nn/config.h:46-63): {"loss", "optimizer", "encoding", "network"} JSON ->
 * TrainableModel.  The JSON is passed as text; any JSON object type with a dump() method (e.g. nlohmann::json, which the
 * reference's signature uses) is accepted through the template overload. */
#pragma once
#include "trainer.h"

namespace tcnn {

struct TrainableModel {
	std::shared_ptr<Network> network;
	std::shared_ptr<Trainer> trainer;
};

inline TrainableModel create_from_config(uint32_t n_input_dims, uint32_t n_output_dims, const std::string& config_json, uint32_t seed = 1337) {
	auto network = std::make_shared<Network>(config_json, n_input_dims, n_output_dims);
	auto trainer = std::make_shared<Trainer>(network, seed);
	return {network, trainer};
}
inline TrainableModel create_from_config(uint32_t n_input_dims, uint32_t n_output_dims, const char* config_json, uint32_t seed = 1337) {
	return create_from_config(n_input_dims, n_output_dims, std::string(config_json), seed);
}
template <typename Json, typename = decltype(std::declval<const Json&>().dump())>
TrainableModel create_from_config(uint32_t n_input_dims, uint32_t n_output_dims, const Json& config, uint32_t seed = 1337) {
	return create_from_config(n_input_dims, n_output_dims, std::string(config.dump()), seed);
}

} // namespace tcnn
