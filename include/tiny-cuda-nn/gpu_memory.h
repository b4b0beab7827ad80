/* GPUMemory<T>: RAII device buffer with the accessors hosts of the reference use (include/tiny-cuda-nn/gpu_memory.h:99-424:
 * data(), size(), resize(), enlarge(), memset(), copy_from_host(), copy_to_host(), free_memory()).  Plain cudaMalloc: the
 * tcnn_b200 path has no per-step temporaries, so the reference's stream arena (gpu_memory.h:426-697) has no counterpart. */
#pragma once
#include <memory>
#include <vector>

#include "common.h"

namespace tcnn {

template <typename T>
class GPUMemory {
public:
	GPUMemory() = default;
	explicit GPUMemory(size_t n) { resize(n); }
	GPUMemory(const GPUMemory&) = delete;
	GPUMemory& operator=(const GPUMemory&) = delete;
	GPUMemory(GPUMemory&& o) noexcept : m_data(o.m_data), m_size(o.m_size) { o.m_data = nullptr; o.m_size = 0; }
	GPUMemory& operator=(GPUMemory&& o) noexcept {
		if (this != &o) { free_memory(); m_data = o.m_data; m_size = o.m_size; o.m_data = nullptr; o.m_size = 0; }
		return *this;
	}
	~GPUMemory() { if (m_data) cudaFree(m_data); }

	void resize(size_t n) {
		if (n == m_size) return;
		free_memory();
		if (n) CUDA_CHECK_THROW(cudaMalloc(reinterpret_cast<void**>(&m_data), n * sizeof(T)));
		m_size = n;
	}
	void enlarge(size_t n) { if (n > m_size) resize(n); }
	void free_memory() { if (m_data) { CUDA_CHECK_THROW(cudaFree(m_data)); m_data = nullptr; } m_size = 0; }
	void memset(int value) { if (m_size) CUDA_CHECK_THROW(cudaMemset(m_data, value, m_size * sizeof(T))); }
	void memset_async(cudaStream_t stream, int value) { if (m_size) CUDA_CHECK_THROW(cudaMemsetAsync(m_data, value, m_size * sizeof(T), stream)); }

	void copy_from_host(const T* host, size_t n) {
		if (n > m_size) throw std::runtime_error("GPUMemory::copy_from_host: more elements than allocated");
		CUDA_CHECK_THROW(cudaMemcpy(m_data, host, n * sizeof(T), cudaMemcpyHostToDevice));
	}
	void copy_from_host(const std::vector<T>& host) { enlarge(host.size()); copy_from_host(host.data(), host.size()); }
	void copy_from_host(const T* host) { copy_from_host(host, m_size); } /* gpu_memory.h:262-264: as many elements as are allocated */
	void copy_to_host(T* host) const { copy_to_host(host, m_size); }
	void copy_to_host(T* host, size_t n) const {
		if (n > m_size) throw std::runtime_error("GPUMemory::copy_to_host: more elements than allocated");
		CUDA_CHECK_THROW(cudaMemcpy(host, m_data, n * sizeof(T), cudaMemcpyDeviceToHost));
	}
	void copy_to_host(std::vector<T>& host) const { host.resize(m_size); copy_to_host(host.data(), m_size); }

	T* data() const { return m_data; }
	size_t size() const { return m_size; }
	size_t get_num_elements() const { return m_size; }
	size_t bytes() const { return m_size * sizeof(T); }

private:
	T* m_data = nullptr;
	size_t m_size = 0;
};

/* free_gpu_memory_arena / free_all_gpu_memory_arenas (gpu_memory.h:743-756).  There is no arena here: the only temporaries are the
 * scratch buffers the network / trainer objects keep for the C ABI's workspace argument; they register themselves and are
 * released by these calls (and grow again on the next use). */
inline std::vector<std::weak_ptr<GPUMemory<char>>>& scratch_registry() { static std::vector<std::weak_ptr<GPUMemory<char>>> r; return r; }
inline std::shared_ptr<GPUMemory<char>> make_scratch() { auto p = std::make_shared<GPUMemory<char>>(); scratch_registry().push_back(p); return p; }
inline void free_all_gpu_memory_arenas() {
	auto& r = scratch_registry();
	for (size_t i = 0; i < r.size();) { if (auto p = r[i].lock()) { p->free_memory(); ++i; } else { r[i] = r.back(); r.pop_back(); } }
}
inline void free_gpu_memory_arena(cudaStream_t) { free_all_gpu_memory_arenas(); }

} // namespace tcnn
