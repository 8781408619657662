// munkres.cu — batched ComputeMunkresMax (pkg/placement/placement_manager.go:505-512; module
// github.com/heyfey/munkres) : one warp-sized CTA per n x n int64 score matrix.
//
// The matrix is read once from HBM (coalesced 256-byte row segments), turned into
// C[i][j] = rowmax_i - A[i][j] and parked in the NARROWEST cell type that holds every C value.
// C is only read again by step 6 (a handful of times per matrix), so it lives in a per-CTA
// slab of global scratch that stays L2-resident (16 KB per matrix as uint8), NOT in shared
// memory: the solve is a chain of dependent warp steps, throughput comes from how many
// matrices an SM works on at once, and with 5.9 KB of shared memory per matrix that is the
// hardware limit of 32 one-warp CTAs per SM.  The cell type is found by trying: the uint8
// kernel runs over the whole batch and marks the matrices whose range does not fit; the
// uint16 / int32 / int64 kernels then pick up only what was marked (each is an almost-empty
// launch when nothing was).  A matrix with a row that mixes negative and non-negative scores goes to
// the widest kernel whatever its range: the reference's MaxInt64 - score wraps there, and only that
// kernel computes in the same wrapping arithmetic.  The solve itself is munkres.cuh.
#include <type_traits>

#include "kernels.h"
#include "munkres.cuh"

namespace voda {
namespace {

template <typename CellT>
struct CellMax;
template <> struct CellMax<uint8_t> { static constexpr long long v = 255; };
template <> struct CellMax<uint16_t> { static constexpr long long v = 65535; };
template <> struct CellMax<int32_t> { static constexpr long long v = 0x7fffffffLL; };
template <> struct CellMax<int64_t> { static constexpr long long v = 0x7fffffffffffffffLL; };

__device__ __forceinline__ long long warp_max_ll(long long v) {
  for (int d = 16; d > 0; d >>= 1) {
    long long o = __shfl_xor_sync(0xffffffffu, v, d);
    v = o > v ? o : v;
  }
  return v;
}

__device__ __forceinline__ long long warp_min_ll(long long v) {
  for (int d = 16; d > 0; d >>= 1) {
    long long o = __shfl_xor_sync(0xffffffffu, v, d);
    v = o < v ? o : v;
  }
  return v;
}

template <typename CellT>
#ifndef VODA_MK_MINBLOCKS
#define VODA_MK_MINBLOCKS 32  // 64 registers: all 32 one-warp CTAs of an SM resident (74 registers held 27: 8.1 -> 6.1 ms per 20k matrices)
#endif
__global__ void __launch_bounds__(32, VODA_MK_MINBLOCKS) munkres_kernel(int n, int batch, const int64_t* __restrict__ A,
                                                     int32_t* __restrict__ out, int32_t* __restrict__ status,
                                                     unsigned char* __restrict__ need, int level,
                                                     unsigned char* __restrict__ slabs, size_t slabBytes) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  using Pot = typename MunkresTypes<CellT>::Pot;
  CellT* C = reinterpret_cast<CellT*>(slabs + (size_t)blockIdx.x * slabBytes);
  size_t off = 0;
  MunkresSmem sm;
  sm.ra = smem_raw + off; off += 128 * sizeof(Pot);
  sm.cs = smem_raw + off; off += 128 * sizeof(Pot);
  sm.starCol = reinterpret_cast<short*>(smem_raw + off); off += 128 * sizeof(short);
  sm.starRow = reinterpret_cast<short*>(smem_raw + off); off += 128 * sizeof(short);
  sm.primeCol = smem_raw + off; off += 128 * sizeof(short);
  sm.zt = reinterpret_cast<uint32_t*>(smem_raw + off); off += 128 * 4 * sizeof(uint32_t);
  sm.zr = reinterpret_cast<uint32_t*>(smem_raw + off);
  const int lane = threadIdx.x;
  const unsigned full = 0xffffffffu;
  const long long NEG = (long long)0x8000000000000000ULL;

  for (int b = blockIdx.x; b < batch; b += gridDim.x) {
    if (need[b] != level) continue;
    const int64_t* M = A + (size_t)b * n * n;
    bool overflow = false;
    for (int x = lane; x < 128 * 4; x += 32) {
      sm.zt[x] = 0;
      sm.zr[x] = 0;
    }
    __syncwarp();
    // cells wider than 32 bits only exist in the int64 kernel; the others saturate a difference at 2^32 - 1
    using Diff = typename std::conditional<sizeof(CellT) == 8, unsigned long long, uint32_t>::type;
    uint32_t acc[4] = {0, 0, 0, 0};  // zeros of the current block of 32 rows in columns lane + 32 t
    for (int l = 0; l < 32 && 4 * l < n; l++) {
      // 4 rows x 4 column blocks in flight
      long long v[4][4];
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int i = 4 * l + q;
#pragma unroll
        for (int t = 0; t < 4; t++) {
          const int col = lane + 32 * t;
          // streamed (evict-first): 13 GB of input per 100k matrices would otherwise push the C slabs out of L2
          v[q][t] = (i < n && col < n) ? __ldcs(reinterpret_cast<const long long*>(M) + (size_t)i * n + col) : NEG;
        }
      }
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int i = 4 * l + q;
        if (i >= n) break;
        Diff c[4];
        // rows whose valid entries share one high word (the usual case: small scores) are handled on
        // the low words alone (unsigned order == numeric order there): one redux, 32-bit differences
        const int hi0 = __shfl_sync(full, (int)(v[q][0] >> 32), 0);  // column 0 is always valid
        bool same = true;
#pragma unroll
        for (int t = 0; t < 4; t++) same = same && (lane + 32 * t >= n || (int)(v[q][t] >> 32) == hi0);
        if (__all_sync(full, same)) {
          uint32_t lo[4], mx = 0;
#pragma unroll
          for (int t = 0; t < 4; t++) {
            lo[t] = lane + 32 * t < n ? (uint32_t)(unsigned long long)v[q][t] : 0u;
            mx = max(mx, lo[t]);
          }
          mx = __reduce_max_sync(full, mx);
#pragma unroll
          for (int t = 0; t < 4; t++) c[t] = (Diff)(mx - lo[t]);
        } else if (sizeof(CellT) == 8) {
          // The reference forms MaxInt64 - A[i][j] and subtracts the row minimum in wrapping int64
          // arithmetic (ComputeMunkresMax, then step 1).  For a row of one sign that is rowmax - A; a
          // row holding negative AND non-negative scores wraps, its cells come out negative or small,
          // and the assignment the steps then find is still the reference's answer: the widest kernel
          // computes exactly that, for every row that reaches it.
          long long ap[4], mn = 0x7fffffffffffffffLL;
#pragma unroll
          for (int t = 0; t < 4; t++) {
            ap[t] = lane + 32 * t < n ? (long long)(0x7fffffffffffffffULL - (unsigned long long)v[q][t]) : 0x7fffffffffffffffLL;
            mn = ap[t] < mn ? ap[t] : mn;
          }
          mn = warp_min_ll(mn);
#pragma unroll
          for (int t = 0; t < 4; t++) c[t] = (Diff)((unsigned long long)ap[t] - (unsigned long long)mn);
        } else {
          bool neg = false, pos = false;
#pragma unroll
          for (int t = 0; t < 4; t++) {
            if (lane + 32 * t < n) {
              if (v[q][t] < 0) neg = true;
              else pos = true;
            }
          }
          if (__any_sync(full, neg) && __any_sync(full, pos)) overflow = true;  // mixed signs: the widest kernel's case
          long long mx = v[q][0];
#pragma unroll
          for (int t = 1; t < 4; t++) mx = v[q][t] > mx ? v[q][t] : mx;
          mx = warp_max_ll(mx);
#pragma unroll
          for (int t = 0; t < 4; t++) {
            const unsigned long long d = (unsigned long long)mx - (unsigned long long)v[q][t];  // one sign: no wrap
            c[t] = (Diff)(d > 0xFFFFFFFFull ? 0xFFFFFFFFull : d);
          }
        }
        const uint32_t rowBit = 1u << (i & 31);
#pragma unroll
        for (int t = 0; t < 4; t++) {
          const int col = lane + 32 * t;
          const bool valid = col < n;
          if (sizeof(CellT) < 8 && valid && c[t] > (Diff)CellMax<CellT>::v) overflow = true;
          if (valid) C[(size_t)i * n + col] = (CellT)c[t];
          const bool z = valid && c[t] == 0;
          const unsigned bz = __ballot_sync(full, z);
          if (lane == 0) sm.zr[i * 4 + t] = bz;
          acc[t] |= z ? rowBit : 0u;  // this lane owns column col
        }
      }
      if ((l & 7) == 7 || 4 * (l + 1) >= n) {  // the block of 32 rows is complete
#pragma unroll
        for (int t = 0; t < 4; t++) {
          if (lane + 32 * t < n) sm.zt[(lane + 32 * t) * 4 + (l >> 3)] = acc[t];
          acc[t] = 0;
        }
      }
    }
    if (__any_sync(full, overflow)) {
      if (level < 3) {
        if (lane == 0) need[b] = (unsigned char)(level + 1);  // a wider kernel takes it
      } else {  // (not reached: the widest kernel's arithmetic is the reference's own, wrap included)
        for (int i = lane; i < n; i += 32) out[(size_t)b * n + i] = -1;
        if (lane == 0) status[b] = VODA_ST_BOUND;
      }
      continue;
    }
    __syncwarp();
    munkres_solve<CellT>(C, n, sm);
    for (int i = lane; i < n; i += 32) out[(size_t)b * n + i] = sm.starCol[i];
    if (lane == 0) status[b] = VODA_ST_OK;
    __syncwarp();
  }
}

template <typename CellT>
size_t munkres_smem() {
  using Pot = typename MunkresTypes<CellT>::Pot;
  return 2 * 128 * sizeof(Pot) + 3 * 128 * sizeof(short) + 2 * 128 * 4 * sizeof(uint32_t) + 16;
}

constexpr int kCtasPerSm = 32;  // one-warp CTAs: the hardware limit of resident CTAs per SM

// every level shares one slab region: narrower cells -> more resident matrices
size_t slab_region_bytes(int n, int sms) { return (size_t)sms * kCtasPerSm * (((size_t)n * n + 255) & ~(size_t)255); }

template <typename CellT>
cudaError_t launch_level(int n, int batch, const int64_t* A, int32_t* out, int32_t* status, unsigned char* need,
                         int level, int sms, unsigned char* slabs, cudaStream_t stream) {
  const size_t smem = munkres_smem<CellT>();
  auto kern = munkres_kernel<CellT>;
  const size_t slabBytes = (((size_t)n * n + 255) & ~(size_t)255) * sizeof(CellT);
  int grid = (int)(slab_region_bytes(n, sms) / slabBytes);  // sms * 32 / sizeof(CellT)
  if (grid < 1) grid = 1;
  if (batch < grid) grid = batch;
  kern<<<grid, 32, smem, stream>>>(n, batch, A, out, status, need, level, slabs, slabBytes);
  return cudaGetLastError();
}

}  // namespace

size_t munkres_scratch_bytes(int n, int batch, int sms) {
  // + the slack the last slab's vector loads may touch (munkres.cuh: kMunkresSlackRows rows of 16-bit cells)
  return (((size_t)batch + 255) & ~(size_t)255) + slab_region_bytes(n, sms) + (size_t)kMunkresSlackRows * 128 * 2 + 256;
}

// Enqueues memset + up to four kernels; returns the number of kernels launched in *launched.
cudaError_t launch_munkres(int n, int batch, const int64_t* A, int32_t* out, int32_t* status, void* scratch,
                           int sms, cudaStream_t stream, int* launched) {
  unsigned char* need = static_cast<unsigned char*>(scratch);
  unsigned char* slabs = need + (((size_t)batch + 255) & ~(size_t)255);
  cudaError_t e = cudaMemsetAsync(need, 0, (size_t)batch, stream);
  if (e != cudaSuccess) return e;
  *launched = 0;
  if ((e = launch_level<uint8_t>(n, batch, A, out, status, need, 0, sms, slabs, stream)) != cudaSuccess) return e;
  if ((e = launch_level<uint16_t>(n, batch, A, out, status, need, 1, sms, slabs, stream)) != cudaSuccess) return e;
  if ((e = launch_level<int32_t>(n, batch, A, out, status, need, 2, sms, slabs, stream)) != cudaSuccess) return e;
  if ((e = launch_level<int64_t>(n, batch, A, out, status, need, 3, sms, slabs, stream)) != cudaSuccess) return e;
  *launched = 4;
  return cudaSuccess;
}

}  // namespace voda
