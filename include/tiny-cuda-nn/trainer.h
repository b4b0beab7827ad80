/* Trainer<T, PARAMS_T, COMPUTE_T> over the tcnn_b200 C ABI (include/tiny-cuda-nn/trainer.h:47-361): the parameter buffer in the
 * reference's layout (:75, 322-336), initialize_params (:68-87), training_step with the reference's full signature (:163-190),
 * optimizer_step (:155-157), loss (:205-207), update_hyperparams (:213-216), serialize / deserialize (:275-315).
 * By construction: forward, loss, backward (and on one GPU the optimizer) of a step are ONE call into the library - no
 * activations, loss values or dL/d(output) exist in device memory, and nothing is re-captured per step. */
#pragma once
#include "loss.h"
#include "network_with_input_encoding.h"
#include "optimizer.h"

namespace tcnn {

template <typename T, typename PARAMS_T, typename COMPUTE_T = PARAMS_T>
class Trainer {
	static_assert(std::is_same<T, float>::value && std::is_same<PARAMS_T, __half>::value && std::is_same<COMPUTE_T, __half>::value,
		"the fused path trains fp32 inputs with fp16 parameters and fp16 compute: Trainer<float, __half, __half>");
public:
	using Model = NetworkWithInputEncoding<PARAMS_T>;
	struct ForwardContext : public Context { /* trainer.h:89-95; the fused step keeps only what hosts read back */
		float* loss_device = nullptr;
		GPUMatrix<COMPUTE_T> output;   /* [padded_output_width x B], filled only if the trainer was asked to (keep_output) */
	};

	Trainer(std::shared_ptr<Model> model, std::shared_ptr<Optimizer<PARAMS_T>> optimizer, std::shared_ptr<Loss<PARAMS_T>> loss, uint32_t seed = 1337)
	: m_model(std::move(model)), m_optimizer(std::move(optimizer)), m_loss(std::move(loss)), m_seed(seed) {
		const json config = {{"encoding", m_model->encoding_config()}, {"network", m_model->network_config()}, {"loss", m_loss->config()}, {"optimizer", m_optimizer->plan_json()}};
		TCNN_B200_CHECK_THROW(tcnn_b200_plan_create(config.dump().c_str(), m_model->input_width(), m_model->output_width(), &m_plan));
		m_optimizer->attach(m_plan);
		initialize_params();
	}
	Trainer(const Trainer&) = delete;
	Trainer& operator=(const Trainer&) = delete;
	virtual ~Trainer() { if (m_optimizer) m_optimizer->attach(nullptr); if (m_plan) tcnn_b200_plan_destroy(m_plan); }

	void initialize_params() { /* trainer.h:68-87 */
		const size_t P = m_model->n_params();
		m_optimizer->allocate((uint32_t)P);
		m_params_buffer.resize(8 * P); /* fp32 master | fp16 params | fp16 gradients */
		m_params_buffer.memset(0);
		reset_param_pointers();
		TCNN_B200_CHECK_THROW(tcnn_b200_init_params(m_plan, nullptr, m_seed, 1, m_params_full_precision));
		TCNN_B200_CHECK_THROW(tcnn_b200_cast_params(nullptr, P, m_params_full_precision, m_params));
		m_loss_value.resize(1); m_loss_value.memset(0);
		m_grads_clean = true;
		CUDA_CHECK_THROW(cudaDeviceSynchronize());
	}

	/* trainer.h:163-190.  input [D x B] CM, target [O x B] CM. */
	std::unique_ptr<ForwardContext> training_step(
		cudaStream_t stream,
		const GPUMatrixDynamic<T>& input,
		const GPUMatrix<float>& target,
		const GPUMatrix<float>* data_pdf = nullptr,
		bool run_optimizer = true,
		GPUMatrixDynamic<T>* dL_dinput = nullptr,
		bool use_inference_params = false,
		GradientMode param_gradients_mode = GradientMode::Overwrite,
		const GPUMatrix<COMPUTE_T>* external_dL_dy = nullptr
	) {
		if (data_pdf) throw std::runtime_error("Trainer::training_step: data_pdf is not implemented by the fused training step");
		if (dL_dinput) throw std::runtime_error("Trainer::training_step: dL_dinput is not implemented by the fused training step (cpp::Module::backward computes it)");
		if (external_dL_dy) throw std::runtime_error("Trainer::training_step: external_dL_dy is not implemented by the fused training step (cpp::Module::backward takes dL/d(output))");
		if (param_gradients_mode == GradientMode::Accumulate) throw std::runtime_error("Trainer::training_step: GradientMode::Accumulate is not implemented");
		(void)use_inference_params; /* one parameter copy (trainer.h:329-333) */
		Model::check(input, m_model->input_width(), "input");
		if (target.m() != m_model->output_width() || target.n() != input.n()) throw std::runtime_error("Trainer::training_step: target must be [n_output_dims x batch]");
		const uint32_t B = input.n();
		void* ws = m_model->workspace(B);
		auto ctx = std::unique_ptr<ForwardContext>(new ForwardContext());
		ctx->loss_device = m_loss_value.data();
		void* out = nullptr;
		if (m_keep_output) { ctx->output = GPUMatrix<COMPUTE_T>(m_model->padded_output_width(), B); out = ctx->output.data(); }
		if (param_gradients_mode == GradientMode::Ignore) run_optimizer = false;
		if (run_optimizer && m_grads_clean) {
			m_optimizer->advance_step();
			TCNN_B200_CHECK_THROW(tcnn_b200_training_step(m_plan, stream, B, input.data(), target.data(), loss_scale(), m_optimizer->step(), m_params_full_precision, m_params,
				m_param_gradients, m_optimizer->first_moments(), m_optimizer->second_moments(), m_optimizer->param_steps(), m_loss_value.data(), out, ws));
		} else {
			TCNN_B200_CHECK_THROW(tcnn_b200_training_fwd_bwd(m_plan, stream, B, input.data(), target.data(), m_params, m_param_gradients, loss_scale(),
				m_grads_clean ? 0 : 1, m_loss_value.data(), out, ws));
			m_grads_clean = false;
			if (run_optimizer) optimizer_step(stream, loss_scale());
		}
		return ctx;
	}
	std::unique_ptr<ForwardContext> training_step(
		const GPUMatrixDynamic<T>& input, const GPUMatrix<float>& target, const GPUMatrix<float>* data_pdf = nullptr, bool run_optimizer = true,
		GPUMatrixDynamic<T>* dL_dinput = nullptr, bool use_inference_params = false, GradientMode param_gradients_mode = GradientMode::Overwrite,
		const GPUMatrix<COMPUTE_T>* external_dL_dy = nullptr
	) { return training_step(nullptr, input, target, data_pdf, run_optimizer, dL_dinput, use_inference_params, param_gradients_mode, external_dL_dy); }

	void optimizer_step(cudaStream_t stream, float loss_scale) { /* trainer.h:155-157 */
		m_optimizer->step(stream, loss_scale, m_params_full_precision, m_params, m_param_gradients);
		m_grads_clean = true;
	}
	void optimizer_step(float loss_scale) { optimizer_step(nullptr, loss_scale); }

	/* trainer.h:205-207: synchronises and returns the mean loss of the step */
	float loss(cudaStream_t stream, const ForwardContext& ctx) const {
		float v = 0.f;
		CUDA_CHECK_THROW(cudaMemcpyAsync(&v, ctx.loss_device, sizeof(float), cudaMemcpyDeviceToHost, stream));
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		return v;
	}
	float loss(const ForwardContext& ctx) const { return loss(nullptr, ctx); }

	/* trainer.h:213-216: {"optimizer": {...}, "loss": {...}} */
	void update_hyperparams(const json& params) {
		m_optimizer->update_hyperparams(params.value("optimizer", json::object()));
		m_loss->update_hyperparams(params.value("loss", json::object()));
	}
	json hyperparams() const { return {{"otype", "Trainer"}, {"optimizer", m_optimizer->hyperparams()}, {"loss", m_loss->hyperparams()}}; }

	/* trainer.h:243-273 */
	void set_params_full_precision(const float* params, size_t n_params, bool device_ptr = false) {
		if (n_params != m_model->n_params()) throw std::runtime_error("Can't set fp params because buffer has the wrong size.");
		CUDA_CHECK_THROW(cudaMemcpy(m_params_full_precision, params, n_params * sizeof(float), device_ptr ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice));
		TCNN_B200_CHECK_THROW(tcnn_b200_cast_params(nullptr, n_params, m_params_full_precision, m_params));
		CUDA_CHECK_THROW(cudaDeviceSynchronize());
	}
	void set_params(const PARAMS_T* params, size_t n_params, bool device_ptr = false) {
		if (n_params != m_model->n_params()) throw std::runtime_error("Can't set params because buffer has the wrong size.");
		CUDA_CHECK_THROW(cudaMemcpy(m_params, params, n_params * sizeof(PARAMS_T), device_ptr ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice));
		std::vector<PARAMS_T> h(n_params); std::vector<float> f(n_params); /* master weights follow the fp16 values */
		CUDA_CHECK_THROW(cudaMemcpy(h.data(), m_params, n_params * sizeof(PARAMS_T), cudaMemcpyDeviceToHost));
		for (size_t i = 0; i < n_params; ++i) f[i] = __half2float(h[i]);
		CUDA_CHECK_THROW(cudaMemcpy(m_params_full_precision, f.data(), n_params * sizeof(float), cudaMemcpyHostToDevice));
	}

	/* trainer.h:275-288 (+ adam.h:278-286) */
	json serialize(bool serialize_optimizer = false) {
		const size_t n = m_model->n_params();
		json data;
		data["n_params"] = n;
		data["params_type"] = "__half";
		data["params_binary"] = Optimizer<PARAMS_T>::to_binary(m_params, sizeof(PARAMS_T) * n);
		if (serialize_optimizer) data["optimizer"] = m_optimizer->serialize();
		return data;
	}
	/* trainer.h:290-315 */
	void deserialize(const json& data) {
		const std::string type = data.value("params_type", "__half");
		if (type == "float") {
			GPUMemory<float> p; Optimizer<PARAMS_T>::from_binary(data["params_binary"], p);
			set_params_full_precision(p.data(), p.size(), true);
		} else if (type == "__half") {
			GPUMemory<PARAMS_T> p; Optimizer<PARAMS_T>::from_binary(data["params_binary"], p);
			set_params(p.data(), p.size(), true);
		} else {
			throw std::runtime_error{"Trainer: snapshot parameters must be of type float of __half"};
		}
		if (data.contains("optimizer")) m_optimizer->deserialize(data["optimizer"]);
		CUDA_CHECK_THROW(cudaMemset(m_param_gradients, 0, sizeof(PARAMS_T) * m_model->n_params()));
		m_grads_clean = true;
		CUDA_CHECK_THROW(cudaDeviceSynchronize());
	}

	float loss_scale() const { return default_loss_scale<PARAMS_T>(); }
	float* params_full_precision() const { return m_params_full_precision; }
	PARAMS_T* params() const { return m_params; }
	PARAMS_T* params_inference() const { return m_params; }
	PARAMS_T* param_gradients() const { return m_param_gradients; }
	size_t n_params() const { return m_model->n_params(); }
	uint32_t step() const { return m_optimizer->step(); }
	std::shared_ptr<Model> model() { return m_model; }
	std::shared_ptr<Optimizer<PARAMS_T>> optimizer() { return m_optimizer; }
	/* facade extension: also write the network output [padded_output_width x B] of every training step into ForwardContext::output */
	void keep_output(bool keep) { m_keep_output = keep; }

private:
	void reset_param_pointers() { /* trainer.h:322-336 */
		const size_t P = m_model->n_params();
		m_params_full_precision = reinterpret_cast<float*>(m_params_buffer.data());
		m_params = reinterpret_cast<PARAMS_T*>(m_params_buffer.data() + 4 * P);
		m_param_gradients = reinterpret_cast<PARAMS_T*>(m_params_buffer.data() + 6 * P);
		m_model->set_params(m_params, m_params, m_param_gradients);
	}
	std::shared_ptr<Model> m_model;
	std::shared_ptr<Optimizer<PARAMS_T>> m_optimizer;
	std::shared_ptr<Loss<PARAMS_T>> m_loss;
	tcnn_b200_plan* m_plan = nullptr;
	uint32_t m_seed;
	GPUMemory<char> m_params_buffer;
	float* m_params_full_precision = nullptr;
	PARAMS_T* m_params = nullptr;
	PARAMS_T* m_param_gradients = nullptr;
	GPUMemory<float> m_loss_value;
	bool m_grads_clean = true, m_keep_output = false;
};

} // namespace tcnn
