// Training kernels of the diffusion head (q_sample, MSE + VLB loss, backward elementwise parts).
#pragma once
#include "common.cuh"
namespace uva {
}  // namespace uva
