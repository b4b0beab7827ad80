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
#include "json.h"

namespace tcnn {

/* the reference's hosts test `cuda_compute_capability() < MIN_GPU_ARCH` first thing (samples/mlp_learning_an_image.cu:102-108) */
static constexpr uint32_t MIN_GPU_ARCH = 100;
inline int cuda_device() { int d = 0; if (cudaGetDevice(&d) != cudaSuccess) throw std::runtime_error("cudaGetDevice failed"); return d; }
inline void set_cuda_device(int device) { if (cudaSetDevice(device) != cudaSuccess) throw std::runtime_error("cudaSetDevice failed"); }
inline uint32_t cuda_compute_capability(int device) {
	cudaDeviceProp props;
	if (cudaGetDeviceProperties(&props, device) != cudaSuccess) throw std::runtime_error("cudaGetDeviceProperties failed");
	return (uint32_t)(props.major * 10 + props.minor);
}
inline uint32_t cuda_compute_capability() { return cuda_compute_capability(cuda_device()); }

/* object.h:115-119 */
enum class GradientMode { Ignore, Overwrite, Accumulate };
/* object.h:46-55: what a forward pass leaves for the backward pass */
struct Context { virtual ~Context() {} };

static constexpr uint32_t BATCH_SIZE_GRANULARITY = 256;

enum class MatrixLayout { RowMajor = 0, SoA = 0, ColumnMajor = 1, AoS = 1 };
static constexpr MatrixLayout RM = MatrixLayout::RowMajor;
static constexpr MatrixLayout CM = MatrixLayout::ColumnMajor;

using network_precision_t = __half;

template <typename T> constexpr float default_loss_scale() { return 1.0f; }
template <> constexpr float default_loss_scale<__half>() { return 128.0f; }

template <typename T> constexpr T next_multiple(T v, T divisor) { return (v + divisor - 1) / divisor * divisor; }
template <typename T> constexpr T div_round_up(T v, T divisor) { return (v + divisor - 1) / divisor; }

/* common_host.h:312-341: launch helpers hosts of the reference use for their own kernels (CUDA translation units only) */
static constexpr uint32_t N_THREADS_LINEAR = 128;
template <typename T> constexpr uint32_t n_blocks_linear(T n_elements) { return (uint32_t)div_round_up(n_elements, (T)N_THREADS_LINEAR); }
#if defined(__CUDACC__)
template <typename K, typename T, typename... Types>
inline void linear_kernel(K kernel, uint32_t shmem_size, cudaStream_t stream, T n_elements, Types... args) {
	if (n_elements <= 0) return;
	kernel<<<n_blocks_linear(n_elements), N_THREADS_LINEAR, shmem_size, stream>>>(n_elements, args...);
}
#endif

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
