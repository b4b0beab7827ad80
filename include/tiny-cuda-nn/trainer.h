/* Network, Trainer and TrainableModel over the tcnn_b200 C ABI.
 *
 * Mirrors, for the HashGrid + FullyFusedMLP path:
 *   Network / DifferentiableObject::inference     include/tiny-cuda-nn/object.h:147-179, network.h
 *   Trainer<float, half, half>                    include/tiny-cuda-nn/trainer.h:47-361 (training_step :163-190, loss :205-207,
 *                                                 optimizer_step :155-157, initialize_params :68-87, buffer layout :75,322-336)
 * Differences by construction: loss and optimizer are part of the plan (no separate Loss / Optimizer objects); the forward
 * context only carries the loss scalar (activations never leave the chip); there is no CUDA-graph re-capture per step. */
#pragma once
#include <memory>

#include "gpu_matrix.h"

namespace tcnn {

class Network {
public:
	Network(const std::string& config_json, uint32_t n_input_dims, uint32_t n_output_dims) {
		TCNN_B200_CHECK_THROW(tcnn_b200_plan_create(config_json.c_str(), n_input_dims, n_output_dims, &m_plan));
		TCNN_B200_CHECK_THROW(tcnn_b200_plan_info(m_plan, &m_info));
	}
	Network(const Network&) = delete;
	Network& operator=(const Network&) = delete;
	~Network() { if (m_plan) tcnn_b200_plan_destroy(m_plan); }

	/* object.h:147-176: input [D x B] CM fp32, output [O x B] CM fp32 */
	void inference(cudaStream_t stream, const GPUMatrixDynamic<float>& input, GPUMatrixDynamic<float>& output) {
		check(input, m_info.n_input_dims, "input");
		if (output.m() != m_info.n_output_dims || output.n() != input.n() || output.layout() != CM)
			throw std::runtime_error("Network::inference: output must be a column-major [n_output_dims x batch] matrix");
		if (!m_params) throw std::runtime_error("Network::inference: no parameters set (set_params)");
		TCNN_B200_CHECK_THROW(tcnn_b200_inference(m_plan, stream, input.n(), input.data(), m_params, nullptr, output.data(), workspace(input.n())));
	}
	void inference(const GPUMatrixDynamic<float>& input, GPUMatrixDynamic<float>& output) { inference(nullptr, input, output); }
	/* object.h:126-145: fp16 output padded to 16 rows */
	void inference_mixed_precision(cudaStream_t stream, const GPUMatrixDynamic<float>& input, GPUMatrixDynamic<__half>& output) {
		check(input, m_info.n_input_dims, "input");
		if (output.m() != m_info.padded_output_width || output.n() != input.n() || output.layout() != CM)
			throw std::runtime_error("Network::inference_mixed_precision: output must be a column-major [padded_output_width x batch] matrix");
		TCNN_B200_CHECK_THROW(tcnn_b200_inference(m_plan, stream, input.n(), input.data(), m_params, output.data(), nullptr, workspace(input.n())));
	}

	void set_params(__half* params, __half* inference_params, __half* gradients) { m_params = inference_params ? inference_params : params; (void)gradients; }
	size_t n_params() const { return m_info.n_params; }
	uint32_t input_width() const { return m_info.n_input_dims; }
	uint32_t output_width() const { return m_info.n_output_dims; }
	uint32_t padded_output_width() const { return m_info.padded_output_width; }
	const tcnn_b200_info& info() const { return m_info; }
	tcnn_b200_plan* plan() const { return m_plan; }

	void* workspace(uint32_t batch_size) {
		const size_t need = tcnn_b200_workspace_bytes(m_plan, batch_size);
		if (need > m_workspace.size()) m_workspace.resize(need);
		return m_workspace.data();
	}
	static void check(const GPUMatrixDynamic<float>& m, uint32_t rows, const char* what) {
		if (m.layout() != CM || m.m() != rows) throw std::runtime_error(std::string("tcnn_b200: ") + what + " must be a column-major [" + std::to_string(rows) + " x batch] matrix");
		check_batch_size(m.n());
	}

private:
	tcnn_b200_plan* m_plan = nullptr;
	tcnn_b200_info m_info{};
	__half* m_params = nullptr;
	GPUMemory<char> m_workspace;
};

class Trainer {
public:
	struct ForwardContext { float* loss_device; };

	explicit Trainer(std::shared_ptr<Network> network, uint32_t seed = 1337) : m_network(std::move(network)) {
		const size_t P = m_network->n_params();
		m_params_buffer.resize(8 * P); /* trainer.h:75: fp32 master | fp16 params | fp16 gradients */
		m_params_full_precision = reinterpret_cast<float*>(m_params_buffer.data());
		m_params = reinterpret_cast<__half*>(m_params_buffer.data() + 4 * P);
		m_param_gradients = reinterpret_cast<__half*>(m_params_buffer.data() + 6 * P);
		m_first_moments.resize(P); m_second_moments.resize(P); m_param_steps.resize(P); m_loss.resize(1);
		m_network->set_params(m_params, m_params, m_param_gradients);
		initialize_params(seed);
	}

	void initialize_params(uint32_t seed = 1337) { /* trainer.h:68-87 */
		TCNN_B200_CHECK_THROW(tcnn_b200_init_params(m_network->plan(), nullptr, seed, 1, m_params_full_precision));
		TCNN_B200_CHECK_THROW(tcnn_b200_cast_params(nullptr, m_network->n_params(), m_params_full_precision, m_params));
		CUDA_CHECK_THROW(cudaMemset(m_param_gradients, 0, 2 * m_network->n_params()));
		m_first_moments.memset(0); m_second_moments.memset(0); m_param_steps.memset(0); m_loss.memset(0);
		m_step = 0; m_grads_clean = true;
		CUDA_CHECK_THROW(cudaDeviceSynchronize());
	}

	/* trainer.h:163-190.  input [D x B] CM, target [O x B] CM. */
	std::unique_ptr<ForwardContext> training_step(cudaStream_t stream, const GPUMatrixDynamic<float>& input, const GPUMatrix<float>& target, bool run_optimizer = true) {
		Network::check(input, m_network->input_width(), "input");
		if (target.m() != m_network->output_width() || target.n() != input.n()) throw std::runtime_error("Trainer::training_step: target must be [n_output_dims x batch]");
		const uint32_t B = input.n();
		void* ws = m_network->workspace(B);
		if (run_optimizer && m_grads_clean) {
			++m_step;
			TCNN_B200_CHECK_THROW(tcnn_b200_training_step(m_network->plan(), stream, B, input.data(), target.data(), loss_scale(), m_step, m_params_full_precision, m_params,
				m_param_gradients, m_first_moments.data(), m_second_moments.data(), m_param_steps.data(), m_loss.data(), nullptr, ws));
		} else {
			TCNN_B200_CHECK_THROW(tcnn_b200_training_fwd_bwd(m_network->plan(), stream, B, input.data(), target.data(), m_params, m_param_gradients, loss_scale(),
				m_grads_clean ? 0 : 1, m_loss.data(), nullptr, ws));
			m_grads_clean = false;
			if (run_optimizer) optimizer_step(stream, loss_scale());
		}
		return std::unique_ptr<ForwardContext>(new ForwardContext{m_loss.data()});
	}
	std::unique_ptr<ForwardContext> training_step(const GPUMatrixDynamic<float>& input, const GPUMatrix<float>& target) { return training_step(nullptr, input, target); }

	void optimizer_step(cudaStream_t stream, float loss_scale) { /* trainer.h:155-157 */
		++m_step;
		TCNN_B200_CHECK_THROW(tcnn_b200_adam_step(m_network->plan(), stream, loss_scale, m_step, m_params_full_precision, m_params, m_param_gradients,
			m_first_moments.data(), m_second_moments.data(), m_param_steps.data(), 1));
		m_grads_clean = true;
	}

	/* trainer.h:205-207: synchronises and returns the mean loss of the step */
	float loss(cudaStream_t stream, const ForwardContext& ctx) const {
		float v = 0.f;
		CUDA_CHECK_THROW(cudaMemcpyAsync(&v, ctx.loss_device, sizeof(float), cudaMemcpyDeviceToHost, stream));
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		return v;
	}

	void update_hyperparams(const std::string& optimizer_json) { TCNN_B200_CHECK_THROW(tcnn_b200_plan_update_optimizer(m_network->plan(), optimizer_json.c_str())); }

	float loss_scale() const { return default_loss_scale<network_precision_t>(); }
	float* params_full_precision() const { return m_params_full_precision; }
	__half* params() const { return m_params; }
	__half* param_gradients() const { return m_param_gradients; }
	size_t n_params() const { return m_network->n_params(); }
	uint32_t step() const { return m_step; }
	std::shared_ptr<Network> model() const { return m_network; }

private:
	std::shared_ptr<Network> m_network;
	GPUMemory<char> m_params_buffer;
	float* m_params_full_precision = nullptr;
	__half* m_params = nullptr;
	__half* m_param_gradients = nullptr;
	GPUMemory<float> m_first_moments, m_second_moments, m_loss;
	GPUMemory<uint32_t> m_param_steps;
	uint32_t m_step = 0;
	bool m_grads_clean = true;
};

} // namespace tcnn
