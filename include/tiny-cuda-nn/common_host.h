/* include/tiny-cuda-nn/common_host.h: CUDA_CHECK_THROW, linear_kernel, n_blocks_linear live in common.h of this facade. */
#pragma once
#include "common.h"
