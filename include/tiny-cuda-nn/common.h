/* tiny-cuda-nn C++ API facade over the tcnn_b200 C ABI (header-only; link libtcnn_b200.so and cudart).
 *
 * Mirrors the names a host of the reference uses on the HashGrid + FullyFusedMLP path:
 *   include/tiny-cuda-nn/common.h      BATCH_SIZE_GRANULARITY (:235), MatrixLayout / RM / CM (:157-167), default_loss_scale (:229-233)
 *   include/tiny-cuda-nn/common_host.h CUDA_CHECK_THROW (:93-110), std::runtime_error with file:line text
 * Errors of the C ABI (status code + tcnn_b200_last_error) become std::runtime_error here, as in the reference. */
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <stdexcept>
#include <string>

#include "../tcnn_b200.h"

namespace tcnn {

static constexpr uint32_t BATCH_SIZE_GRANULARITY = 256;

enum class MatrixLayout { RowMajor = 0, SoA = 0, ColumnMajor = 1, AoS = 1 };
static constexpr MatrixLayout RM = MatrixLayout::RowMajor;
static constexpr MatrixLayout CM = MatrixLayout::ColumnMajor;

using network_precision_t = __half;

template <typename T> constexpr float default_loss_scale() { return 1.0f; }
template <> constexpr float default_loss_scale<__half>() { return 128.0f; }

template <typename T> constexpr T next_multiple(T v, T divisor) { return (v + divisor - 1) / divisor * divisor; }
template <typename T> constexpr T div_round_up(T v, T divisor) { return (v + divisor - 1) / divisor; }

#define TCNN_B200_STR_(x) #x
#define TCNN_B200_STR(x) TCNN_B200_STR_(x)
#define TCNN_B200_FILE_LINE __FILE__ ":" TCNN_B200_STR(__LINE__)

#define CUDA_CHECK_THROW(x)                                                                                             \
	do {                                                                                                                \
		cudaError_t _result = (x);                                                                                      \
		if (_result != cudaSuccess) throw std::runtime_error(std::string(TCNN_B200_FILE_LINE " " #x " failed: ") + cudaGetErrorString(_result)); \
	} while (0)

/* status of a C-ABI call -> exception */
#define TCNN_B200_CHECK_THROW(x)                                                                                        \
	do {                                                                                                                \
		if ((x) != 0) throw std::runtime_error(std::string(TCNN_B200_FILE_LINE " ") + tcnn_b200_last_error());         \
	} while (0)

inline void check_batch_size(uint32_t batch_size) {
	if (batch_size % BATCH_SIZE_GRANULARITY != 0)
		throw std::runtime_error("Batch size " + std::to_string(batch_size) + " must be a multiple of " + std::to_string(BATCH_SIZE_GRANULARITY) + ".");
}

} // namespace tcnn
