/* Optimizer<T> and create_optimizer (include/tiny-cuda-nn/optimizer.h:36-68, src/optimizer.cu:50-80) for Adam
 * (optimizers/adam.h:121-327): hyper-parameters, the optimizer state (first / second moments, per-parameter step counters) and
 * its snapshot format (adam.h:278-299).  step() runs the fused Adam kernel of the C ABI on the plan the trainer bound it to. */
#pragma once
#include "gpu_memory.h"

namespace tcnn {

template <typename T>
class Optimizer {
public:
	explicit Optimizer(const json& params) : m_params(json::object()) {
		std::string otype = (params.is_null() ? json::object() : params).value("otype", "Adam");
		for (auto& c : otype) c = (char)std::tolower((unsigned char)c);
		if (otype != "adam") throw std::runtime_error("Invalid optimizer type: " + otype + " (this library implements Adam)");
		update_hyperparams(params.is_null() ? json::object() : params);
	}
	virtual ~Optimizer() {}

	/* AdamOptimizer::allocate (adam.h:128-148) */
	void allocate(uint32_t n_weights, const std::vector<std::pair<uint32_t, uint32_t>>& = {}) {
		if (n_weights <= m_first_moments.size()) return;
		m_first_moments.resize(n_weights); m_second_moments.resize(n_weights); m_param_steps.resize(n_weights);
		m_first_moments.memset(0); m_second_moments.memset(0); m_param_steps.memset(0);
		m_n_weights = n_weights;
	}
	/* AdamOptimizer::step (adam.h:150-188) */
	void step(cudaStream_t stream, float loss_scale, float* weights_full_precision, T* weights, const T* gradients) {
		if (!m_plan) throw std::runtime_error("Optimizer::step: the optimizer is not attached to a trainer");
		++m_current_step;
		TCNN_B200_CHECK_THROW(tcnn_b200_adam_step(m_plan, stream, loss_scale, m_current_step, weights_full_precision, weights, const_cast<T*>(gradients),
			m_first_moments.data(), m_second_moments.data(), m_param_steps.data(), 1));
	}
	float learning_rate() const { return m_params.value("learning_rate", 1e-3f); }
	void set_learning_rate(float val) { update_hyperparams(json{{"learning_rate", val}}); }
	uint32_t step() const { return m_current_step; }
	uint32_t n_weights() const { return m_n_weights; }
	T* custom_weights() const { return nullptr; }
	size_t n_nested() const { return 0; }

	/* adam.h:210-258: keys given replace the stored ones; the plan (if attached) re-reads them */
	void update_hyperparams(const json& params) {
		for (const auto& kv : params.items()) m_params[kv.first] = kv.second;
		if (m_plan) TCNN_B200_CHECK_THROW(tcnn_b200_plan_update_optimizer(m_plan, plan_json().dump().c_str()));
	}
	json hyperparams() const { json h = m_params; h["otype"] = "Adam"; return h; }
	const json& config() const { return m_params; }
	/* the keys the C ABI's Adam knows (it rejects unknown ones); the reference's samples carry keys of other optimizers
	 * alongside (samples/mlp_learning_an_image.cu:125-131: "only used when the optimizer is Shampoo"), which Adam ignores */
	json plan_json() const {
		static const char* known[] = {"otype", "learning_rate", "beta1", "beta2", "epsilon", "l2_reg", "relative_decay", "absolute_decay", "adabound",
			"clipping_magnitude", "non_matrix_learning_rate_factor", "optimize_matrix_params", "optimize_non_matrix_params"};
		json out = json::object();
		for (const char* k : known) if (m_params.contains(k)) out[k] = m_params[k];
		return out;
	}

	/* adam.h:278-299 */
	json serialize() const {
		json data;
		data["current_step"] = m_current_step;
		data["base_learning_rate"] = learning_rate();
		data["first_moments_binary"] = to_binary(m_first_moments.data(), m_n_weights * sizeof(float));
		data["second_moments_binary"] = to_binary(m_second_moments.data(), m_n_weights * sizeof(float));
		data["param_steps_binary"] = to_binary(m_param_steps.data(), m_n_weights * sizeof(uint32_t));
		return data;
	}
	void deserialize(const json& data) {
		from_binary(data["first_moments_binary"], m_first_moments);
		from_binary(data["second_moments_binary"], m_second_moments);
		if (data.contains("param_steps_binary")) from_binary(data["param_steps_binary"], m_param_steps);
		else { m_param_steps.resize(m_second_moments.size()); m_param_steps.memset(0); }
		m_n_weights = (uint32_t)m_first_moments.size();
		m_current_step = data["current_step"].get<uint32_t>();
		set_learning_rate(data["base_learning_rate"].get<float>());
	}

	/* facade plumbing: the trainer attaches the plan whose Adam kernel this object drives, and the fused step advances the counter */
	void attach(tcnn_b200_plan* plan) { m_plan = plan; }
	void advance_step() { ++m_current_step; }
	float* first_moments() const { return m_first_moments.data(); }
	float* second_moments() const { return m_second_moments.data(); }
	uint32_t* param_steps() const { return m_param_steps.data(); }

	static json::binary_t to_binary(const void* device, size_t n_bytes) { /* gpu_memory_json.h:36-41 */
		json::binary_t b;
		b.resize(n_bytes);
		if (n_bytes) CUDA_CHECK_THROW(cudaMemcpy(b.data(), device, n_bytes, cudaMemcpyDeviceToHost));
		return b;
	}
	template <typename U> static void from_binary(const json& j, GPUMemory<U>& dst) { /* gpu_memory_json.h:48-72: binary value or {"bytes": [...]} */
		json::binary_t tmp;
		const json::binary_t* b = &tmp;
		if (j.is_binary()) b = &j.get_binary();
		else if (j.is_object()) { for (const json& e : j["bytes"]) tmp.push_back((uint8_t)e.get<unsigned>()); }
		else throw std::runtime_error("Invalid json type: must be either binary or object");
		dst.resize(b->size() / sizeof(U));
		if (!b->empty()) CUDA_CHECK_THROW(cudaMemcpy(dst.data(), b->data(), dst.size() * sizeof(U), cudaMemcpyHostToDevice));
	}

private:
	json m_params;
	tcnn_b200_plan* m_plan = nullptr;
	GPUMemory<float> m_first_moments, m_second_moments;
	GPUMemory<uint32_t> m_param_steps;
	uint32_t m_n_weights = 0, m_current_step = 0;
};

template <typename T>
Optimizer<T>* create_optimizer(const json& params) { return new Optimizer<T>(params); }

} // namespace tcnn
