/* NetworkWithInputEncoding<T> (include/tiny-cuda-nn/network_with_input_encoding.h:41-190) with the inference interface of
 * DifferentiableObject (object.h:126-179), over a C-ABI plan.  T must be __half (TCNN_HALF_PRECISION, common.h:99). */
#pragma once
#include <memory>

#include "gpu_matrix.h"

namespace tcnn {

template <typename T>
class NetworkWithInputEncoding {
	static_assert(std::is_same<T, __half>::value, "the fused path computes in fp16 (network_precision_t)");
public:
	/* network_with_input_encoding.h:55: the two JSON objects of create_from_config */
	NetworkWithInputEncoding(uint32_t n_dims_to_encode, uint32_t n_output_dims, const json& encoding, const json& network)
	: m_encoding_config(encoding), m_network_config(network) { create(json{{"encoding", encoding}, {"network", network}}.dump(), n_dims_to_encode, n_output_dims); }
	/* whole config as text ({"encoding": ..., "network": ..., optionally "loss" / "optimizer"}) */
	NetworkWithInputEncoding(const std::string& config_json, uint32_t n_input_dims, uint32_t n_output_dims) {
		const json c = json::parse(config_json);
		m_encoding_config = c.value("encoding", json::object()); m_network_config = c.value("network", json::object());
		create(config_json, n_input_dims, n_output_dims);
	}
	NetworkWithInputEncoding(const NetworkWithInputEncoding&) = delete;
	NetworkWithInputEncoding& operator=(const NetworkWithInputEncoding&) = delete;
	~NetworkWithInputEncoding() { if (m_plan) tcnn_b200_plan_destroy(m_plan); }

	/* object.h:147-176: input [D x B] CM fp32, output [O x B] CM fp32 */
	void inference(cudaStream_t stream, const GPUMatrixDynamic<float>& input, GPUMatrixDynamic<float>& output, bool use_inference_params = true) {
		(void)use_inference_params; /* Adam keeps one parameter copy: inference parameters == training parameters (trainer.h:329-333) */
		check(input, m_info.n_input_dims, "input");
		if (output.m() != m_info.n_output_dims || output.n() != input.n() || output.layout() != CM)
			throw std::runtime_error("Network::inference: output must be a column-major [n_output_dims x batch] matrix");
		if (!m_params) throw std::runtime_error("Network::inference: no parameters set (set_params)");
		TCNN_B200_CHECK_THROW(tcnn_b200_inference(m_plan, stream, input.n(), input.data(), m_params, nullptr, output.data(), workspace(input.n())));
	}
	void inference(const GPUMatrixDynamic<float>& input, GPUMatrixDynamic<float>& output, bool use_inference_params = true) { inference(nullptr, input, output, use_inference_params); }
	/* object.h:126-145: fp16 output padded to 16 rows */
	void inference_mixed_precision(cudaStream_t stream, const GPUMatrixDynamic<float>& input, GPUMatrixDynamic<T>& output, bool use_inference_params = true) {
		(void)use_inference_params;
		check(input, m_info.n_input_dims, "input");
		if (output.m() != m_info.padded_output_width || output.n() != input.n() || output.layout() != CM)
			throw std::runtime_error("Network::inference_mixed_precision: output must be a column-major [padded_output_width x batch] matrix");
		if (!m_params) throw std::runtime_error("Network::inference_mixed_precision: no parameters set (set_params)");
		TCNN_B200_CHECK_THROW(tcnn_b200_inference(m_plan, stream, input.n(), input.data(), m_params, output.data(), nullptr, workspace(input.n())));
	}

	void set_params(T* params, T* inference_params, T* gradients) { m_params = inference_params ? inference_params : params; (void)gradients; }
	T* params() const { return m_params; }
	size_t n_params() const { return m_info.n_params; }
	uint32_t input_width() const { return m_info.n_input_dims; }
	uint32_t output_width() const { return m_info.n_output_dims; }
	uint32_t padded_output_width() const { return m_info.padded_output_width; }
	uint32_t required_input_alignment() const { return 1; }
	json hyperparams() const { return {{"otype", "NetworkWithInputEncoding"}, {"encoding", m_encoding_config}, {"network", m_network_config}}; }
	std::string name() const { return "NetworkWithInputEncoding"; }
	const json& encoding_config() const { return m_encoding_config; }
	const json& network_config() const { return m_network_config; }
	const tcnn_b200_info& info() const { return m_info; }
	tcnn_b200_plan* plan() const { return m_plan; }

	void* workspace(uint32_t batch_size) {
		const size_t need = tcnn_b200_workspace_bytes(m_plan, batch_size);
		if (need > m_workspace->size()) m_workspace->resize(need);
		return m_workspace->data();
	}
	static void check(const GPUMatrixDynamic<float>& m, uint32_t rows, const char* what) {
		if (m.layout() != CM || m.m() != rows) throw std::runtime_error(std::string("tcnn_b200: ") + what + " must be a column-major [" + std::to_string(rows) + " x batch] matrix");
		check_batch_size(m.n());
	}

private:
	void create(const std::string& config_json, uint32_t n_in, uint32_t n_out) {
		TCNN_B200_CHECK_THROW(tcnn_b200_plan_create(config_json.c_str(), n_in, n_out, &m_plan));
		TCNN_B200_CHECK_THROW(tcnn_b200_plan_info(m_plan, &m_info));
	}
	json m_encoding_config, m_network_config;
	tcnn_b200_plan* m_plan = nullptr;
	tcnn_b200_info m_info{};
	T* m_params = nullptr;
	std::shared_ptr<GPUMemory<char>> m_workspace = make_scratch();
};

using Network = NetworkWithInputEncoding<network_precision_t>;

} // namespace tcnn
