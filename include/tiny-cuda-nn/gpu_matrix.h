/* GPUMatrix / GPUMatrixDynamic with the reference's layout semantics (include/tiny-cuda-nn/gpu_matrix.h:105-130,222-224,
 * 417-433; common.h:157-167): element (r, c) of a column-major (CM, "AoS") matrix is at c * stride + r, so the [D x B]
 * CM input of Trainer::training_step is B contiguous D-vectors; row-major (RM, "SoA") is r * stride + c.  Owning
 * (rows, cols) or non-owning (data, rows, cols) as in the reference. */
#pragma once
#include <memory>

#include "gpu_memory.h"

namespace tcnn {

template <typename T>
class GPUMatrixDynamic {
public:
	GPUMatrixDynamic(uint32_t m, uint32_t n, MatrixLayout layout = CM)
	: m_rows(m), m_cols(n), m_layout(layout), m_owned(std::make_shared<GPUMemory<T>>((size_t)m * n)) { m_data = m_owned->data(); }
	GPUMatrixDynamic(T* data, uint32_t m, uint32_t n, MatrixLayout layout = CM) : m_rows(m), m_cols(n), m_data(data), m_layout(layout) {}
	GPUMatrixDynamic() = default;
	virtual ~GPUMatrixDynamic() = default;

	T* data() const { return m_data; }
	uint32_t rows() const { return m_rows; }
	uint32_t cols() const { return m_cols; }
	uint32_t m() const { return m_rows; }
	uint32_t n() const { return m_cols; }
	uint32_t n_elements() const { return m_rows * m_cols; }
	size_t n_bytes() const { return (size_t)n_elements() * sizeof(T); }
	uint32_t stride() const { return m_layout == CM ? m_rows : m_cols; }
	MatrixLayout layout() const { return m_layout; }

	void memset(int value) { CUDA_CHECK_THROW(cudaMemset(m_data, value, n_bytes())); }
	void memset_async(cudaStream_t stream, int value) { CUDA_CHECK_THROW(cudaMemsetAsync(m_data, value, n_bytes(), stream)); }
	std::vector<T> to_cpu_vector() const {
		std::vector<T> v(n_elements());
		CUDA_CHECK_THROW(cudaMemcpy(v.data(), m_data, n_bytes(), cudaMemcpyDeviceToHost));
		return v;
	}
	void copy_from_host(const T* host) { CUDA_CHECK_THROW(cudaMemcpy(m_data, host, n_bytes(), cudaMemcpyHostToDevice)); }

protected:
	uint32_t m_rows = 0, m_cols = 0;
	T* m_data = nullptr;
	MatrixLayout m_layout = CM;
	std::shared_ptr<GPUMemory<T>> m_owned;
};

template <typename T, MatrixLayout _layout = MatrixLayout::ColumnMajor>
class GPUMatrix : public GPUMatrixDynamic<T> {
public:
	static const MatrixLayout static_layout = _layout;
	GPUMatrix(uint32_t m, uint32_t n) : GPUMatrixDynamic<T>(m, n, _layout) {}
	GPUMatrix(T* data, uint32_t m, uint32_t n) : GPUMatrixDynamic<T>(data, m, n, _layout) {}
	GPUMatrix() : GPUMatrixDynamic<T>() { this->m_layout = _layout; }
};

} // namespace tcnn
