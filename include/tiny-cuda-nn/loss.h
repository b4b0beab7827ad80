/* Loss<T> and create_loss (include/tiny-cuda-nn/loss.h:40-76, src/loss.cu:53-93).  On this path the loss is evaluated inside the
 * fused MLP kernel's epilogue, so the object is a handle on its configuration: the trainer hands it to the plan. */
#pragma once
#include "common.h"

namespace tcnn {

template <typename T>
class Loss {
public:
	explicit Loss(const json& params) : m_params(params.is_null() ? json::object() : params) {
		std::string otype = m_params.value("otype", "RelativeL2"); /* src/loss.cu:86 */
		for (auto& c : otype) c = (char)std::tolower((unsigned char)c);
		if (otype != "l2" && otype != "relativel2") throw std::runtime_error("Invalid loss type: " + m_params.value("otype", "RelativeL2") + " (this library evaluates L2 and RelativeL2)");
	}
	virtual ~Loss() {}
	virtual void update_hyperparams(const json& params) { for (const auto& kv : params.items()) m_params[kv.first] = kv.second; }
	virtual json hyperparams() const { return m_params; }
	const json& config() const { return m_params; }

private:
	json m_params;
};

template <typename T>
Loss<T>* create_loss(const json& params) { return new Loss<T>(params); }

} // namespace tcnn
