/* tcnn::cpp::Module — the raw-pointer API the PyTorch extension of the reference is written against
 * (include/tiny-cuda-nn/cpp_api.h:86-115, src/cpp_api.cu:62-165), over the tcnn_b200 C ABI.  All pointers are device
 * pointers owned by the caller: input [B][D] fp32, output [B][16] fp16, params / dL_dparams [n_params] fp16. */
#pragma once
#include <functional>

#include "trainer.h"

namespace tcnn { namespace cpp {

enum class Precision { Fp32, Fp16 };
inline uint32_t batch_size_granularity() { return tcnn_b200_batch_size_granularity(); }
inline float default_loss_scale(Precision p) { return p == Precision::Fp16 ? 128.0f : 1.0f; }
inline Precision preferred_precision() { return Precision::Fp16; }
inline bool has_networks() { return true; }
inline void free_temporary_memory() { free_all_gpu_memory_arenas(); }
/* set_log_callback (cpp_api.h:58-66, src/common_host.cu:50-73): the library itself logs nothing - every problem is an exception -
 * so the callback is only kept for hosts that install one */
enum class LogSeverity { Info, Debug, Warning, Error, Success };
inline std::function<void(LogSeverity, const std::string&)>& log_callback() { static std::function<void(LogSeverity, const std::string&)> cb; return cb; }
inline void set_log_callback(const std::function<void(LogSeverity, const std::string&)>& callback) { log_callback() = callback; }

struct Context { GPUMemory<char> saved; }; /* what backward needs: the encoded batch */

class Module {
public:
	Module(uint32_t n_input_dims, uint32_t n_output_dims, const std::string& encoding_json, const std::string& network_json)
	: m_network("{\"encoding\": " + encoding_json + ", \"network\": " + network_json + "}", n_input_dims, n_output_dims) {}
	/* one of the two objects alone: {"network": ...} = Identity encoding + network, {"encoding": ...} = grid encoding */
	Module(uint32_t n_input_dims, uint32_t n_output_dims, const std::string& config_json, Precision precision = Precision::Fp16)
	: m_network(config_json, n_input_dims, n_output_dims), m_precision(precision) {}

	void inference(cudaStream_t stream, uint32_t n_elements, const float* input, void* output, void* params) {
		TCNN_B200_CHECK_THROW(tcnn_b200_inference(m_network.plan(), stream, n_elements, input, params, output, nullptr, m_network.workspace(n_elements)));
	}
	Context forward(cudaStream_t stream, uint32_t n_elements, const float* input, void* output, void* params, bool prepare_input_gradients) {
		(void)prepare_input_gradients; /* nothing to prepare: dy/dx is recomputed in backward */
		Context ctx;
		ctx.saved.resize(tcnn_b200_workspace_bytes(m_network.plan(), n_elements));
		TCNN_B200_CHECK_THROW(tcnn_b200_forward(m_network.plan(), stream, n_elements, input, params, output, ctx.saved.data()));
		return ctx;
	}
	void backward(cudaStream_t stream, const Context& ctx, uint32_t n_elements, float* dL_dinput, const void* dL_doutput, void* dL_dparams,
		const float* input, const void* /*output*/, const void* params) {
		TCNN_B200_CHECK_THROW(tcnn_b200_backward(m_network.plan(), stream, n_elements, input, params, dL_doutput, dL_dparams, dL_dinput, ctx.saved.data()));
	}
	/* src/cpp_api.cu:111-127: second order of a grid encoding (dL_dparams == nullptr: GradientMode::Ignore) */
	void backward_backward_input(cudaStream_t stream, const Context& /*ctx*/, uint32_t n_elements, const float* dL_ddLdinput, const float* input,
		const void* dL_doutput, void* dL_dparams, void* dL_ddLdoutput, float* dL_dinput, const void* params) {
		TCNN_B200_CHECK_THROW(tcnn_b200_backward_backward_input(m_network.plan(), stream, n_elements, input, params, dL_ddLdinput, dL_doutput,
			dL_ddLdoutput, dL_dparams, dL_dinput, nullptr));
	}
	/* GridEncoding::set_max_level (encodings/grid_interface.h:105-107) of the encoding inside this module */
	void set_max_level(float value) { TCNN_B200_CHECK_THROW(tcnn_b200_plan_set_max_level(m_network.plan(), value)); }
	void initialize_params(size_t seed, float* params_full_precision) {
		TCNN_B200_CHECK_THROW(tcnn_b200_init_params(m_network.plan(), nullptr, seed, 0, params_full_precision));
	}
	uint32_t n_input_dims() const { return m_network.input_width(); }
	uint32_t n_output_dims() const { return m_network.padded_output_width(); } /* padded, as in the reference */
	size_t n_params() const { return m_network.n_params(); }
	Precision param_precision() const { return m_precision; }
	Precision output_precision() const { return m_precision; }
	std::string name() const { return "NetworkWithInputEncoding"; }
	json hyperparams() const { return m_network.hyperparams(); } /* cpp_api.h:112 */

private:
	Network m_network;
	Precision m_precision = Precision::Fp16;
};

inline Module* create_network_with_input_encoding(uint32_t n_input_dims, uint32_t n_output_dims, const std::string& encoding_json, const std::string& network_json) {
	return new Module(n_input_dims, n_output_dims, encoding_json, network_json);
}

/* src/cpp_api.cu:151-153 */
inline Module* create_network(uint32_t n_input_dims, uint32_t n_output_dims, const std::string& network_json) {
	return new Module(n_input_dims, n_output_dims, "{\"network\": " + network_json + "}");
}
/* src/cpp_api.cu:156-165; n_output_dims() is the row width of the output (next multiple of 8 of L * F).  Precision::Fp32:
 * params, output, dL_doutput and dL_dparams of the module are fp32 buffers. */
inline Module* create_encoding(uint32_t n_input_dims, const std::string& encoding_json, Precision requested_precision = Precision::Fp16) {
	const bool f32 = requested_precision == Precision::Fp32;
	return new Module(n_input_dims, 0, "{\"encoding\": " + encoding_json + (f32 ? ", \"precision\": \"fp32\"}" : "}"), requested_precision);
}

}} // namespace tcnn::cpp
