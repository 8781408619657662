// kernels.h — host-side launchers of the CUDA kernels (internal to the library).
#pragma once
#include <cstddef>
#include <cuda_runtime.h>

#include "common.cuh"

namespace voda {

// FfDLOptimizer (ffdl.cu)
int ffdl_grid(int sms, int rounds);
size_t ffdl_scratch_bytes(const AllocArgs& a, int grid);
size_t ffdl_smem_bytes(const AllocArgs& a);
cudaError_t launch_ffdl(const AllocArgs& a, int grid, void* scratch, size_t scratch_bytes,
                        cudaStream_t stream);

}  // namespace voda
