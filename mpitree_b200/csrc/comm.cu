// K5 — u32 sum all-reduce over peer-mapped (NVLink / NVSwitch) buffers.
//
// The reference has no counterpart: its only collective is a pickled allgather of finished
// subtrees (mpitree/tree/decision_tree.py:456-466).  In this framework the cross-rank
// operation is the sum of count histograms; the fused split-search kernel does it inline
// (build.cu), and these stand-alone kernels are the same protocols exposed as a collective
// for the 4 KB - 256 MB sweep against ncclAllReduce:
//
//   one-shot : every rank pulls all peers' inputs and sums them (1 barrier in, 1 out);
//              latency-optimal, moves (P-1) x bytes per GPU.
//   two-shot : reduce-scatter by slice (each rank pulls its slice from all peers, sums) then
//              pushes the reduced slice into every peer's result buffer; moves
//              2 x (P-1)/P x bytes per GPU.
//
// Integer sums are order independent, so results are bit-identical on every rank.
// Each rank's symmetric area is [in: cap u32][res: cap u32]; flags are u32[world] per rank.
// Kernels spin on peers: every rank must launch on its own GPU (never as several waiting
// kernels on one device).
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;
using namespace mtb;

namespace {

struct ArArgs {
  uint32_t* bufs[MTB_MAX_WORLD];   // rank q's symmetric area as mapped here
  uint32_t* flags[MTB_MAX_WORLD];  // rank q's flag words as mapped here
  int rank, world;
  int64_t count, cap;
  uint32_t epoch;
  uint32_t* out;
};

// kGrid = false: single-CTA launch (small messages), block barriers instead of grid syncs
template <bool kGrid>
__device__ void ar_barrier(const ArArgs& a, uint32_t epoch) {
  if (kGrid) cg::this_grid().sync(); else __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x < (unsigned)a.world) {
    __threadfence_system();
    st_release_sys_u32(a.flags[threadIdx.x] + a.rank, epoch);
    const uint32_t* mine = a.flags[a.rank] + threadIdx.x;
    while ((int32_t)(ld_acquire_sys_u32(mine) - epoch) < 0) {
    }
    __threadfence_system();
  }
  if (kGrid) cg::this_grid().sync(); else __syncthreads();
}

__device__ __forceinline__ uint4 add4(uint4 a, uint4 b) {
  return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}

template <bool kGrid>
__global__ void __launch_bounds__(512) k_allreduce_oneshot(const __grid_constant__ ArArgs a) {
  ar_barrier<kGrid>(a, a.epoch + 1);  // every rank's input is in place
  const int64_t n4 = a.count / 4;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    uint4 acc = reinterpret_cast<const uint4*>(a.bufs[a.rank])[i];
    for (int p = 1; p < a.world; ++p) {
      const int q = (a.rank + p) % a.world;  // stagger peers so links are used evenly
      acc = add4(acc, ld_nc_u4(reinterpret_cast<const uint4*>(a.bufs[q]) + i));
    }
    reinterpret_cast<uint4*>(a.out)[i] = acc;
  }
  for (int64_t i = n4 * 4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.count; i += stride) {
    uint32_t acc = 0;
    for (int q = 0; q < a.world; ++q) acc += a.bufs[q][i];
    a.out[i] = acc;
  }
  ar_barrier<kGrid>(a, a.epoch + 2);  // nobody may overwrite its input before all have read it
}

template <bool kGrid>
__global__ void __launch_bounds__(512) k_allreduce_twoshot(const __grid_constant__ ArArgs a) {
  ar_barrier<kGrid>(a, a.epoch + 1);
  // slice owned by this rank, in 16-byte units
  const int64_t n4 = (a.count + 3) / 4;
  const int64_t per = (n4 + a.world - 1) / a.world;
  const int64_t lo = per * a.rank, hi = min(n4, lo + per);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
    uint4 acc = reinterpret_cast<const uint4*>(a.bufs[a.rank])[i];
    for (int p = 1; p < a.world; ++p) {
      const int q = (a.rank + p) % a.world;
      acc = add4(acc, ld_nc_u4(reinterpret_cast<const uint4*>(a.bufs[q]) + i));
    }
    // reduced slice -> every rank's result area (own one included)
    for (int p = 0; p < a.world; ++p) {
      const int q = (a.rank + p) % a.world;
      st_na_u4(reinterpret_cast<uint4*>(a.bufs[q] + a.cap) + i, acc);
    }
  }
  ar_barrier<kGrid>(a, a.epoch + 2);  // all slices have landed in my result area
  if (a.out != a.bufs[a.rank] + a.cap) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.count; i += stride)
      a.out[i] = a.bufs[a.rank][a.cap + i];
  }
}

}  // namespace

extern "C" int mtb_allreduce_u32(int algo, int rank, int world, void* const* bufs, void* const* flags, int64_t count,
                                 int64_t cap, uint32_t epoch, uint32_t* out, void* stream) {
  if (world < 1 || world > MTB_MAX_WORLD) {
    mtb_set_error("allreduce", "world size must be in [1, 16]", __FILE__, __LINE__);
    return -2;
  }
  if (count > cap || (cap % 4) != 0) {
    mtb_set_error("allreduce", "count exceeds the symmetric buffer capacity (or capacity not a multiple of 4)", __FILE__,
                  __LINE__);
    return -2;
  }
  ArArgs a;
  for (int q = 0; q < world; ++q) {
    a.bufs[q] = (uint32_t*)bufs[q];
    a.flags[q] = (uint32_t*)flags[q];
  }
  a.rank = rank;
  a.world = world;
  a.count = count;
  a.cap = cap;
  a.epoch = epoch;
  a.out = out;
  int dev = 0, sms = 0;
  MTB_CHECK(cudaGetDevice(&dev));
  MTB_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  // small messages: few CTAs (the two barriers dominate); large: fill the machine
  int64_t want = (count / 4 + 4 * 512 - 1) / (4 * 512);  // >= 4 x 16 B per thread before adding CTAs
  int blocks = (int)(want < 1 ? 1 : (want > 2 * sms ? 2 * sms : want));
  if (blocks == 1) {  // latency path: plain launch, no grid-wide syncs
    if (algo == 0) k_allreduce_oneshot<false><<<1, 512, 0, (cudaStream_t)stream>>>(a);
    else k_allreduce_twoshot<false><<<1, 512, 0, (cudaStream_t)stream>>>(a);
    MTB_CHECK(cudaGetLastError());
    return 0;
  }
  void* args[] = {&a};
  void* fn = algo == 0 ? (void*)k_allreduce_oneshot<true> : (void*)k_allreduce_twoshot<true>;
  MTB_CHECK(cudaLaunchCooperativeKernel(fn, dim3(blocks), dim3(512), args, 0, (cudaStream_t)stream));
  return 0;
}
