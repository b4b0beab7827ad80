/* include/tiny-cuda-nn/common_device.h of the reference is the first include of its samples; here it only has to bring in the
 * host-visible names those samples use from it (default_rng_t :331, the launch helpers through common_host.h). */
#pragma once
#include "common.h"
#include "random.h"
