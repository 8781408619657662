// kernels.h — host-side launchers of the CUDA kernels (internal to the library).
#pragma once
#include <cstddef>
#include <cuda_runtime.h>

#include "common.cuh"

namespace voda {

// FfDLOptimizer (ffdl.cu)
size_t ffdl_scratch_per_round(const AllocArgs& a);
size_t ffdl_smem_bytes(const AllocArgs& a);
// rounds [r0, r0+count): DP kernel + backtrack kernel (2 launches)
cudaError_t launch_ffdl(const AllocArgs& a, int r0, int count, int sms, void* scratch, cudaStream_t stream);

// FIFO / SRJF / ElasticFIFO / ElasticSRJF (fill.cu)
size_t fill_smem_bytes(const AllocArgs& a);
cudaError_t launch_fill(const AllocArgs& a, int sms, cudaStream_t stream);

// Tiresias / ElasticTiresias (tiresias.cu)
size_t tiresias_smem_bytes(const AllocArgs& a, bool* table_in_smem);
cudaError_t launch_tiresias(const AllocArgs& a, int sms, cudaStream_t stream);

// AFS-L (afsl.cu)
size_t afsl_smem_bytes(const AllocArgs& a);
cudaError_t launch_afsl(const AllocArgs& a, int sms, cudaStream_t stream);
// rounds one full wave of the launch processes (0: no preference); host batches are cut into multiples
int afsl_wave_rounds(const AllocArgs& a, int sms);

// batched ComputeMunkresMax (munkres.cu)
size_t munkres_scratch_bytes(int n, int batch, int sms);
cudaError_t launch_munkres(int n, int batch, const int64_t* A, int32_t* out, int32_t* status, void* scratch,
                           int sms, cudaStream_t stream, int* launched);

// bestFit + score + munkres placement rounds (place.cu)
size_t place_smem_bytes(const PlaceArgs& a);
size_t place_scratch_bytes(const PlaceArgs& a, int sms, int rounds = -1);
cudaError_t launch_place(const PlaceArgs& a, int sms, void* scratch, cudaStream_t stream, int* launched, int first = 0,
                         int last = -1);

// speedup-curve builder (curves.cu)
cudaError_t launch_curves(int jobs, const int32_t* rec_off, const int32_t* workers, const double* t, int W,
                          double* out_time, double* out_speedup, double* out_eff, int32_t* out_status,
                          cudaStream_t stream);

// scheduler bookkeeping (tick.cu)
cudaError_t launch_tick(const TickArgs& a, cudaStream_t stream);
cudaError_t launch_compare(int jobs, const int32_t* old_gpu, const int32_t* new_gpu, const int32_t* status,
                           int32_t* action, cudaStream_t stream);

}  // namespace voda
