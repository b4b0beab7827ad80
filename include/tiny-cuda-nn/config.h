/* create_from_config (include/tiny-cuda-nn/config.h:46-63): {"loss", "optimizer", "encoding", "network"} ->
 * TrainableModel{loss, optimizer, network, trainer}.  Takes tcnn::json like the reference, JSON text, or any JSON object type with
 * a dump() method (e.g. a host's own nlohmann::json). */
#pragma once
#include "random.h"
#include "trainer.h"

namespace tcnn {

struct TrainableModel {
	std::shared_ptr<Loss<network_precision_t>> loss;
	std::shared_ptr<Optimizer<network_precision_t>> optimizer;
	std::shared_ptr<NetworkWithInputEncoding<network_precision_t>> network;
	std::shared_ptr<Trainer<float, network_precision_t, network_precision_t>> trainer;
};

inline TrainableModel create_from_config(uint32_t n_input_dims, uint32_t n_output_dims, json config) {
	std::shared_ptr<Loss<network_precision_t>> loss{create_loss<network_precision_t>(config.value("loss", json::object()))};
	std::shared_ptr<Optimizer<network_precision_t>> optimizer{create_optimizer<network_precision_t>(config.value("optimizer", json::object()))};
	auto network = std::make_shared<NetworkWithInputEncoding<network_precision_t>>(n_input_dims, n_output_dims, config.value("encoding", json::object()), config.value("network", json::object()));
	auto trainer = std::make_shared<Trainer<float, network_precision_t, network_precision_t>>(network, optimizer, loss);
	return {loss, optimizer, network, trainer};
}
inline TrainableModel create_from_config(uint32_t n_input_dims, uint32_t n_output_dims, const std::string& config_json) { return create_from_config(n_input_dims, n_output_dims, json::parse(config_json)); }
inline TrainableModel create_from_config(uint32_t n_input_dims, uint32_t n_output_dims, const char* config_json) { return create_from_config(n_input_dims, n_output_dims, json::parse(std::string(config_json))); }
template <typename Json, typename = decltype(std::declval<const Json&>().dump()), typename = typename std::enable_if<!std::is_same<Json, json>::value>::type>
TrainableModel create_from_config(uint32_t n_input_dims, uint32_t n_output_dims, const Json& config) { return create_from_config(n_input_dims, n_output_dims, json::parse(std::string(config.dump()))); }

} // namespace tcnn
