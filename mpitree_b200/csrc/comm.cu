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
// Each rank's symmetric area is [in: cap u32][res: cap u32] (cap a multiple of 4, inputs are read
// in 16-byte units, so the pad words of the last unit must be readable); flags are
// u32[512][16] per rank (one row per CTA index).
// Kernels spin on peers: every rank must launch on its own GPU (never as several waiting
// kernels on one device).
#include "common.cuh"

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

constexpr int kArMaxCtas = 512;  // flag words per rank: [kArMaxCtas][MTB_MAX_WORLD]

// Barrier between the CTAs with the same index on every rank (all ranks launch the same grid).
// No grid-wide sync is needed anywhere: a peer's input is ready as soon as any of its CTAs runs
// (the kernel is stream-ordered after the producer), and a kernel only completes when all of its
// CTAs have seen their counterparts finish, which together covers every peer CTA.
__device__ void cta_peer_barrier(const ArArgs& a, uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x < (unsigned)a.world) {
    __threadfence_system();
    st_release_sys_u32(a.flags[threadIdx.x] + (size_t)blockIdx.x * MTB_MAX_WORLD + a.rank, epoch);
    const uint32_t* mine = a.flags[a.rank] + (size_t)blockIdx.x * MTB_MAX_WORLD + threadIdx.x;
    while ((int32_t)(ld_acquire_sys_u32(mine) - epoch) < 0) {
    }
    __threadfence_system();
  }
  __syncthreads();
}

__device__ __forceinline__ uint4 add4(uint4 a, uint4 b) {
  return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}

// sum of element i over all ranks, peers' loads issued four at a time
__device__ __forceinline__ uint4 gather_sum(const ArArgs& a, int64_t i) {
  uint4 acc = reinterpret_cast<const uint4*>(a.bufs[a.rank])[i];
#pragma unroll
  for (int p0 = 1; p0 < MTB_MAX_WORLD; p0 += 4) {
    if (p0 < a.world) {
      uint4 y[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int p = p0 + j;
        const int q = (a.rank + p) % a.world;  // stagger peers so links are used evenly
        y[j] = p < a.world ? ld_nc_u4(reinterpret_cast<const uint4*>(a.bufs[q]) + i) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) acc = add4(acc, y[j]);
    }
  }
  return acc;
}

__global__ void __launch_bounds__(512) k_allreduce_oneshot(const __grid_constant__ ArArgs a) {
  cta_peer_barrier(a, a.epoch + 1);  // every rank's input is in place
  const int64_t n4 = (a.count + 3) / 4;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const uint4 v = gather_sum(a, i);
    if (i * 4 + 4 <= a.count) {
      reinterpret_cast<uint4*>(a.out)[i] = v;
    } else {  // last, partial unit: the caller's buffer may end at `count`
      const uint32_t e[4] = {v.x, v.y, v.z, v.w};
      for (int j = 0; i * 4 + j < a.count; ++j) a.out[i * 4 + j] = e[j];
    }
  }
  cta_peer_barrier(a, a.epoch + 2);  // nobody may overwrite its input before all have read it
}

// sums of elements i0 and i1 (i1 < 0: only i0) over all ranks: both elements' loads of four peers are in
// flight together (eight remote loads per thread), which is what fills NVLink between 256 KB and 4 MB,
// where a slice is only a few thousand 16-byte units per rank
__device__ __forceinline__ void gather_sum2(const ArArgs& a, int64_t i0, int64_t i1, uint4& s0, uint4& s1) {
  const bool two = i1 >= 0;
  s0 = reinterpret_cast<const uint4*>(a.bufs[a.rank])[i0];
  s1 = two ? reinterpret_cast<const uint4*>(a.bufs[a.rank])[i1] : make_uint4(0, 0, 0, 0);
#pragma unroll
  for (int p0 = 1; p0 < MTB_MAX_WORLD; p0 += 4) {
    if (p0 < a.world) {
      uint4 y0[4], y1[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int p = p0 + j;
        const int q = (a.rank + p) % a.world;  // stagger peers so links are used evenly
        const uint4* src = reinterpret_cast<const uint4*>(a.bufs[q]);
        y0[j] = p < a.world ? ld_nc_u4(src + i0) : make_uint4(0, 0, 0, 0);
        y1[j] = (p < a.world && two) ? ld_nc_u4(src + i1) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        s0 = add4(s0, y0[j]);
        s1 = add4(s1, y1[j]);
      }
    }
  }
}

__global__ void __launch_bounds__(512) k_allreduce_twoshot(const __grid_constant__ ArArgs a) {
  cta_peer_barrier(a, a.epoch + 1);
  // slice owned by this rank, in 16-byte units
  const int64_t n4 = (a.count + 3) / 4;
  const int64_t per = (n4 + a.world - 1) / a.world;
  const int64_t lo = per * a.rank, hi = min(n4, lo + per);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += 2 * stride) {
    const int64_t i1 = i + stride < hi ? i + stride : -1;
    uint4 s0, s1;
    gather_sum2(a, i, i1, s0, s1);
    // reduced slice -> every rank's result area (own one included)
    for (int p = 0; p < a.world; ++p) {
      const int q = (a.rank + p) % a.world;
      st_na_u4(reinterpret_cast<uint4*>(a.bufs[q] + a.cap) + i, s0);
      if (i1 >= 0) st_na_u4(reinterpret_cast<uint4*>(a.bufs[q] + a.cap) + i1, s1);
    }
  }
  cta_peer_barrier(a, a.epoch + 2);  // the slices of all peers' CTAs of this index have landed
}

// ---- NVLS variant ---------------------------------------------------------------------------------
// The symmetric area is also mapped through an NVSwitch multicast object (parallel/nvls.py), so one
// address names the same word on every rank.  A rank takes its slice, asks the switch for the sum
// over all ranks (multimem.ld_reduce, SASS LDGMC.ADD) and writes it to every rank's result area
// with one multicast store.  Per GPU this moves bytes/P in and bytes/P out instead of
// (P-1)/P x bytes each way.  Integer ld_reduce is scalar only, so counts travel as u64 pairs:
// exact while every global sum is below 2^32, which the u32 histograms require anyway.
__device__ __forceinline__ uint64_t mm_ld_reduce_add_u64(const uint64_t* mc) {
  uint64_t v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];" : "=l"(v) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mm_st_u64(uint64_t* mc, uint64_t v) {
  asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" ::"l"(mc), "l"(v) : "memory");
}

__global__ void __launch_bounds__(512) k_allreduce_nvls(const __grid_constant__ ArArgs a, uint32_t* mc) {
  cta_peer_barrier(a, a.epoch + 1);  // every rank's input is in place
  const int64_t n2 = (a.count + 1) / 2;  // 8-byte units (the pad word of an odd count is readable: cap % 4 == 0)
  const int64_t per = (n2 + a.world - 1) / a.world;
  const int64_t lo = per * a.rank, hi = min(n2, lo + per);
  const uint64_t* in = reinterpret_cast<const uint64_t*>(mc);
  uint64_t* res = reinterpret_cast<uint64_t*>(mc + a.cap);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < hi; i += 4 * stride) {  // four independent switch round trips in flight per lane
    uint64_t v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = mm_ld_reduce_add_u64(in + i + j * stride);
#pragma unroll
    for (int j = 0; j < 4; ++j) mm_st_u64(res + i + j * stride, v[j]);
  }
  for (; i < hi; i += stride) mm_st_u64(res + i, mm_ld_reduce_add_u64(in + i));
  cta_peer_barrier(a, a.epoch + 2);  // the slices of all peers' CTAs of this index have landed
}

}  // namespace

extern "C" int mtb_allreduce_u32_nvls(int rank, int world, void* const* bufs, void* const* flags, void* mc_buf,
                                      int64_t count, int64_t cap, uint32_t epoch, uint32_t* out, void* stream) {
  if (world < 1 || world > MTB_MAX_WORLD || mc_buf == nullptr) {
    mtb_set_error("allreduce_nvls", "world size must be in [1, 16] and the multicast pointer non-null", __FILE__,
                  __LINE__);
    return -2;
  }
  if (count > cap || (cap % 4) != 0) {
    mtb_set_error("allreduce_nvls", "count exceeds the symmetric buffer capacity (or capacity not a multiple of 4)",
                  __FILE__, __LINE__);
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
  // this rank only touches count / world words: size the grid for its slice (same on every rank)
  int64_t want = ((count + 1) / 2 / world + 511) / 512;
  int cap_ctas = 2 * sms < kArMaxCtas ? 2 * sms : kArMaxCtas;
  int blocks = (int)(want < 1 ? 1 : (want > cap_ctas ? cap_ctas : want));
  k_allreduce_nvls<<<blocks, 512, 0, (cudaStream_t)stream>>>(a, (uint32_t*)mc_buf);
  MTB_CHECK(cudaGetLastError());
  if (out != nullptr && out != a.bufs[rank] + cap)
    MTB_CHECK(cudaMemcpyAsync(out, a.bufs[rank] + cap, (size_t)count * 4, cudaMemcpyDeviceToDevice,
                              (cudaStream_t)stream));
  return 0;
}

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
  // small messages: few CTAs (the two flag round trips dominate); large: fill the machine.
  // Every rank computes the same grid (same count, same GPU model).
  int64_t want = (count / 4 + 511) / 512;  // one 16-byte unit per thread until the machine is full (remote latency bound)
  // two-shot: a rank only touches its 1/world slice, two units per thread; CTAs beyond that would only add
  // flag round trips (every CTA index runs its own peer barrier)
  if (algo != 0) want = ((count / 4 + world - 1) / world + 1023) / 1024;
  int cap_ctas = 2 * sms < kArMaxCtas ? 2 * sms : kArMaxCtas;
  int blocks = (int)(want < 1 ? 1 : (want > cap_ctas ? cap_ctas : want));
  if (algo == 0) {
    if (out == nullptr) {
      mtb_set_error("allreduce", "one-shot needs an output pointer", __FILE__, __LINE__);
      return -2;
    }
    k_allreduce_oneshot<<<blocks, 512, 0, (cudaStream_t)stream>>>(a);
  } else {
    k_allreduce_twoshot<<<blocks, 512, 0, (cudaStream_t)stream>>>(a);
    if (out != nullptr && out != a.bufs[rank] + cap)
      MTB_CHECK(cudaMemcpyAsync(out, a.bufs[rank] + cap, (size_t)count * 4, cudaMemcpyDeviceToDevice,
                                (cudaStream_t)stream));
  }
  MTB_CHECK(cudaGetLastError());
  return 0;
}
