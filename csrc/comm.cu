// Symmetric-memory communicator and the gradient all-reduce kernels that replace DDP's Reducer +
// NCCL all-reduce + "divide by world size" + foreach-SGD (reference train_ddp.py:34,199,200;
// SURVEY.md X13,X16,K14) with ONE kernel per gradient bucket.
//
// Every rank (one process per GPU) cudaMalloc's an identically laid out buffer and exports it
// with CUDA IPC; after the handle exchange (done by the Python side over torch.distributed)
// every rank holds a device pointer to every peer's buffer.  On an NVSwitch box those pointers
// are plain NVLink loads/stores at full bandwidth to any peer.
//
// Kernels (all: in-kernel flag barrier, no host synchronisation, CUDA-graph safe):
//   one-shot  : wait until every peer's bucket is ready, read all W copies, sum in rank order
//               0..W-1 (bit-identical on every rank), apply p -= lr/W * sum to the LOCAL params.
//               One flag round.  Best when latency-bound.
//   two-shot  : rank r reduces slice r only, computes the updated parameters of that slice and
//               stores them into EVERY rank's parameter arena, then a second flag round.
//               Moves 2(W-1)/W*S bytes per GPU instead of (W-1)*S.
//   plain all-reduce (fp32 / bf16, one-shot or two-shot) for the size sweep, and a broadcast used
//   for the initial parameter sync and for resume.
//
// Flag protocol: flag word [slot][block][src_rank] lives in the header of each rank's buffer and
// is written only by src_rank (st.release.sys) with a per-(slot,block) monotonically increasing
// epoch kept in device memory (so replaying a captured graph keeps counting).  Waiters spin on
// their own copy with ld.acquire.sys and a wall-clock bound: a dead peer becomes a trap (a CUDA
// error on the host), never a hung GPU.  Gradient buckets are double-buffered by step parity, so
// a single flag round per bucket also protects buffer reuse.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include <new>

#include "ptx.cuh"
#include "layout.cuh"
#include "api_internal.cuh"
#include "launch.cuh"

namespace ddpb {

constexpr int kMaxWorld = 8;
constexpr int kSlots = 8;
constexpr int kMaxBlocks = 160;
constexpr size_t kFlagWords = (size_t)kSlots * kMaxBlocks * kMaxWorld;
constexpr size_t kEpochWords = (size_t)kSlots * kMaxBlocks;
// Flag-carrying ("LL") staging lines for the latency-bound tiny buckets: per slot, two epochs
// (double buffer) x source rank x kLLMaxN 8-byte {epoch, fp32 payload} words.
constexpr int kLLMaxN = 32 * 608;  // 19,456 values: covers the conv2 wgrad accumulator image (18,944)
constexpr size_t kLLWordsPerSlot = 2ull * kMaxWorld * kLLMaxN;
// after the status word: per slot {epoch, finished-block counter} of the flag-carrying kernels
constexpr size_t kLLCtlWord = kFlagWords + kEpochWords + 16;
constexpr size_t kLLOffset = ((kFlagWords + kEpochWords + 64) * 4 + 255) / 256 * 256;
static_assert(kLLCtlWord + 2 * kSlots <= kFlagWords + kEpochWords + 64, "LL control words overflow");
// header: flags | epochs | status word | LL lines, rounded up
constexpr size_t kHeaderBytes = (kLLOffset + (size_t)kSlots * kLLWordsPerSlot * 8 + 4095) / 4096 * 4096;
constexpr int kCommThreads = 512;

struct Comm {
  int rank = 0, world = 1, device = 0;
  size_t bytes = 0;  // total buffer bytes incl. header
  uint8_t* local = nullptr;
  uint8_t* peer[kMaxWorld] = {};
  bool opened[kMaxWorld] = {};
  bool connected = false;
  unsigned long long timeout_ns = 20ull * 1000 * 1000 * 1000;
  // VMM + multicast (NVLS) variant: the buffer is a cuMemCreate allocation mapped by every peer
  // through an exported file descriptor and bound to one multicast object; `mc` is the multicast
  // address of the same bytes (multimem.ld_reduce / multimem.st go through the NVSwitch).
  bool vmm = false;
  size_t map_bytes = 0, align = 0;
  CUmemGenericAllocationHandle mem = 0, peer_mem[kMaxWorld] = {}, mc_handle = 0;
  bool mc_created = false, mc_bound = false;
  uint8_t* mc = nullptr;
};

struct PeerPtrs {
  uint8_t* p[kMaxWorld];
};

__device__ __forceinline__ uint32_t* flag_word(uint8_t* base, int slot, int block, int src) {
  return reinterpret_cast<uint32_t*>(base) + ((size_t)slot * kMaxBlocks + block) * kMaxWorld + src;
}
__device__ __forceinline__ uint32_t* epoch_word(uint8_t* base, int slot, int block) {
  return reinterpret_cast<uint32_t*>(base) + kFlagWords + (size_t)slot * kMaxBlocks + block;
}
__device__ __forceinline__ uint32_t* status_word(uint8_t* base) {
  return reinterpret_cast<uint32_t*>(base) + kFlagWords + kEpochWords;
}
__device__ __forceinline__ uint32_t* ll_ctl(uint8_t* base, int slot) {
  return reinterpret_cast<uint32_t*>(base) + kLLCtlWord + 2 * slot;
}
__device__ __forceinline__ uint64_t* ll_line(uint8_t* base, int slot, uint32_t epoch, int src) {
  return reinterpret_cast<uint64_t*>(base + kLLOffset) +
         ((size_t)slot * 2 + (epoch & 1)) * kMaxWorld * kLLMaxN + (size_t)src * kLLMaxN;
}

// One flag round for this block: tell every peer "block `blockIdx.x` of rank `rank` reached
// epoch e", then wait until every peer said the same.  Orders all prior global writes of the
// whole block before the signal and all later reads after the wait.
template <int WORLD>
__device__ __forceinline__ void flag_round(const PeerPtrs& pp, int rank, int slot, uint32_t e,
                                           unsigned long long timeout_ns) {
  __syncthreads();
  if (threadIdx.x < WORLD) {
    const int r = threadIdx.x;
    // st.release.sys is cumulative over everything ordered before it, which the bar.sync above
    // makes include the whole block's earlier writes: no separate system fence is needed.
    st_release_sys(flag_word(pp.p[r], slot, blockIdx.x, rank), e);
    const uint32_t* mine = flag_word(pp.p[rank], slot, blockIdx.x, r);
    if (static_cast<int32_t>(ld_acquire_sys(mine) - e) < 0) {
      const uint64_t t0 = global_timer_ns();
      uint32_t spins = 0;
      while (static_cast<int32_t>(ld_acquire_sys(mine) - e) < 0) {
        if ((++spins & 0xff) == 0 && global_timer_ns() - t0 > timeout_ns) {
          *status_word(pp.p[rank]) = 0xdead0000u | static_cast<uint32_t>(r);
          __threadfence_system();
          __trap();
        }
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_f4(float* p, const float4& v) {
  asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

// Grid-stride iterations whose peer loads are issued back to back (registers: U*WORLD float4).
template <int WORLD>
constexpr int kPeerUnroll = WORLD <= 2 ? 4 : (WORLD <= 4 ? 2 : 1);

// ---------------------------------------------------------------- fused all-reduce + SGD
// n4 = number of float4 elements in the bucket (buckets are 16-byte aligned slices).
template <int WORLD>
__global__ void __launch_bounds__(kCommThreads)
allreduce_sgd_oneshot_kernel(const __grid_constant__ PeerPtrs pp, int rank, int slot, size_t grad_off, size_t param_off,
                             long long n4, const float* __restrict__ lr,
                             unsigned long long timeout_ns) {
  pdl_sync();
  uint32_t* ep = epoch_word(pp.p[rank], slot, blockIdx.x);
  const uint32_t e = *ep + 1;
  flag_round<WORLD>(pp, rank, slot, e, timeout_ns);
  const float scale = -(*lr) * (1.0f / WORLD);
  float* params = reinterpret_cast<float*>(pp.p[rank] + param_off);
  // U grid-stride iterations per round: all U*WORLD peer loads are issued before any is used, so
  // a thread pays one NVLink round trip per U elements instead of one per element
  constexpr int U = kPeerUnroll<WORLD>;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < n4; i0 += stride * U) {
    float4 v[U][WORLD];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < n4) {
#pragma unroll
        for (int r = 0; r < WORLD; ++r)
          v[u][r] = ld_f4(reinterpret_cast<const float*>(pp.p[r] + grad_off) + i * 4);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < n4) {
        float4 s = v[u][0];
#pragma unroll
        for (int r = 1; r < WORLD; ++r) {
          s.x += v[u][r].x; s.y += v[u][r].y; s.z += v[u][r].z; s.w += v[u][r].w;
        }
        float4 p = *reinterpret_cast<float4*>(params + i * 4);
        p.x = fmaf(scale, s.x, p.x);
        p.y = fmaf(scale, s.y, p.y);
        p.z = fmaf(scale, s.z, p.z);
        p.w = fmaf(scale, s.w, p.w);
        *reinterpret_cast<float4*>(params + i * 4) = p;
      }
    }
  }
  if (threadIdx.x == 0) *ep = e;
}

template <int WORLD>
__global__ void __launch_bounds__(kCommThreads)
allreduce_sgd_twoshot_kernel(const __grid_constant__ PeerPtrs pp, int rank, int slot, size_t grad_off, size_t param_off,
                             long long n4, const float* __restrict__ lr,
                             unsigned long long timeout_ns) {
  pdl_sync();
  uint32_t* ep = epoch_word(pp.p[rank], slot, blockIdx.x);
  const uint32_t e = *ep + 1;
  flag_round<WORLD>(pp, rank, slot, e, timeout_ns);
  const float scale = -(*lr) * (1.0f / WORLD);
  const long long per = (n4 + WORLD - 1) / WORLD;
  const long long lo = per * rank;
  const long long hi = lo + per < n4 ? lo + per : n4;
  const float* params = reinterpret_cast<const float*>(pp.p[rank] + param_off);
  constexpr int U = kPeerUnroll<WORLD>;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < hi; i0 += stride * U) {
    float4 v[U][WORLD];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < hi) {
#pragma unroll
        for (int r = 0; r < WORLD; ++r)
          v[u][r] = ld_f4(reinterpret_cast<const float*>(pp.p[r] + grad_off) + i * 4);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < hi) {
        float4 s = v[u][0];
#pragma unroll
        for (int r = 1; r < WORLD; ++r) {
          s.x += v[u][r].x; s.y += v[u][r].y; s.z += v[u][r].z; s.w += v[u][r].w;
        }
        float4 p = *reinterpret_cast<const float4*>(params + i * 4);
        p.x = fmaf(scale, s.x, p.x);
        p.y = fmaf(scale, s.y, p.y);
        p.z = fmaf(scale, s.z, p.z);
        p.w = fmaf(scale, s.w, p.w);
#pragma unroll
        for (int r = 0; r < WORLD; ++r)
          st_f4(reinterpret_cast<float*>(pp.p[r] + param_off) + i * 4, p);
      }
    }
  }
  flag_round<WORLD>(pp, rank, slot, e + 1, timeout_ns);
  if (threadIdx.x == 0) *ep = e + 1;
}

// ---------------------------------------------------------------- split-K reduce + all-reduce + SGD
// The last gradient bucket of the step (conv1, 320 floats) sits on the critical path, and its
// gradients come out of the tensor-core kernel as one partial per CTA.  This kernel does, in one
// launch: (1) the fixed-order sum of the local partials, (2) publishes the local gradient in the
// peer-mapped arena, (3) the flag round, (4) the rank-ordered sum over all peers' gradients and
// (5) the SGD update — compute + collective + optimizer, no intermediate kernels.
// 256 threads = 32 consecutive outputs x 8 partial groups; one block per 32 outputs.
template <int WORLD>
__global__ void __launch_bounds__(256)
reduce_allreduce_sgd_kernel(const __grid_constant__ PeerPtrs pp, int rank, int slot, const float* __restrict__ partials,
                            int nparts, int n, size_t grad_off, size_t param_off,
                            const float* __restrict__ lr, unsigned long long timeout_ns) {
  __shared__ float sm[8][32];
  pdl_sync();
  const int lane = threadIdx.x & 31, pg = threadIdx.x >> 5;
  const int idx = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (idx < n) {
    int p = pg;
    for (; p + 24 < nparts; p += 32) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = partials[(long long)(p + 8 * u) * n + idx];
#pragma unroll
      for (int u = 0; u < 4; ++u) s += v[u];
    }
    for (; p < nparts; p += 8) s += partials[(long long)p * n + idx];
  }
  sm[pg][lane] = s;
  __syncthreads();
  float* my_grads = reinterpret_cast<float*>(pp.p[rank] + grad_off);
  if (pg == 0 && idx < n) {
#pragma unroll
    for (int u = 1; u < 8; ++u) s += sm[u][lane];
    my_grads[idx] = s;
  }
  uint32_t* ep = epoch_word(pp.p[rank], slot, blockIdx.x);
  const uint32_t e = *ep + 1;
  flag_round<WORLD>(pp, rank, slot, e, timeout_ns);
  if (threadIdx.x < 8) {
    const int i4 = blockIdx.x * 8 + threadIdx.x;  // float4 index
    if (i4 * 4 < n) {
      const float scale = -(*lr) * (1.0f / WORLD);
      float4 acc = ld_f4(reinterpret_cast<const float*>(pp.p[0] + grad_off) + i4 * 4);
#pragma unroll
      for (int r = 1; r < WORLD; ++r) {
        const float4 v = ld_f4(reinterpret_cast<const float*>(pp.p[r] + grad_off) + i4 * 4);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      float4* pr = reinterpret_cast<float4*>(pp.p[rank] + param_off) + i4;
      float4 p = *pr;
      p.x = fmaf(scale, acc.x, p.x);
      p.y = fmaf(scale, acc.y, p.y);
      p.z = fmaf(scale, acc.z, p.z);
      p.w = fmaf(scale, acc.w, p.w);
      *pr = p;
    }
  }
  if (threadIdx.x == 0) *ep = e;
}

// Same operation, push-based with flag-carrying words (the idea of NCCL's LL protocol): every
// rank writes {epoch, value} 8-byte words straight into each peer's staging line for this epoch
// and then polls its own lines.  No fence, no separate flag, no remote load: the critical path
// is one one-way NVLink store latency instead of a flag hop plus a remote-load round trip.
// Staging lines are double-buffered by epoch parity; a rank can only be one call ahead of a peer
// (call e+1 cannot finish without that peer's words of e+1, which the peer sends after it has
// finished reading e), so a line is never overwritten while it is still being read.
// The epoch is one word per slot (all blocks read it on entry; the last block to finish bumps
// it), so the grid is not limited by the per-block flag words of the pull kernels.
//
// Forward progress.  Block i of this rank polls words that block i of every peer sends, so the
// kernel needs, for every i, block i to become resident on ALL ranks at some point — not all blocks
// at once.  That holds because (a) a block sends before it polls and never waits for another block
// of its own grid, so resident blocks always finish once their peers' same-index blocks have run,
// and (b) the hardware dispatches the blocks of a grid in ascending order as SMs free up, on every
// rank alike, so the lowest unfinished index is always resident (or next) everywhere.  Up to 592
// blocks of 256 threads and ~1 KB of shared memory each: they fit beside the 220 KB convolution CTAs
// (which hold an SM's shared memory, not its thread slots).  Exercised under SM pressure by the
// step itself (this kernel runs beside the conv2 data gradient and the conv1 weight gradient in every
// multi-GPU test and bench run, per-rank batch up to 4096) and by tests/mp_workers.py: stress.
// Every wait is bounded: a peer that never sends becomes a trap, not a hang.
//
// MAP selects where output idx of the partial vectors lives in the gradient / parameter slices:
//   0  identity
//   1  the conv3x3 <CI=32, CO=64> weight-gradient accumulator image (csrc/conv_tc.cu):
//      per output channel [row0: 96 taps | 8 bias columns][row1: 96][row2: 96] ->
//      net.2.weight[co][ci][r][s] followed by net.2.bias[co]; unused bias columns are dropped.
template <int MAP>
__device__ __forceinline__ int ll_dest(int idx) {
  if (MAP == 0) return idx;
  constexpr int CI = 32, CO = 64, n0 = 3 * CI + 8, n1 = 3 * CI, ncols = n0 + 2 * n1;
  const int co = idx / ncols;
  int col = idx - co * ncols, r = 0;
  if (col < n0) {
    if (col >= n1) return col == n1 ? CO * CI * 9 + co : -1;
  } else if (col < n0 + n1) {
    r = 1;
    col -= n0;
  } else {
    r = 2;
    col -= n0 + n1;
  }
  const int sx = col / CI, ci = col - sx * CI;
  return (co * CI + ci) * 9 + r * 3 + sx;
}

template <int WORLD, int MAP>
__global__ void __launch_bounds__(256)
reduce_allreduce_sgd_ll_kernel(const __grid_constant__ PeerPtrs pp, int rank, int slot,
                               const float* __restrict__ partials, int nparts, int n, size_t grad_off,
                               size_t param_off, const float* __restrict__ lr,
                               __nv_bfloat16* __restrict__ pack, const float* __restrict__ bias_part,
                               int bp_nsplit, int bp_nbx, int bp_hw, unsigned long long timeout_ns) {
  __shared__ float sm[8][32];
  pdl_sync();
  const int lane = threadIdx.x & 31, pg = threadIdx.x >> 5;
  const int idx = blockIdx.x * 32 + lane;
  uint32_t* ctl = ll_ctl(pp.p[rank], slot);
  const uint32_t e = ctl[0] + 1;
  float s = 0.f;
  if (idx < n) {
    int p = pg;
    for (; p + 24 < nparts; p += 32) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = partials[(long long)(p + 8 * u) * n + idx];
#pragma unroll
      for (int u = 0; u < 4; ++u) s += v[u];
    }
    for (; p < nparts; p += 8) s += partials[(long long)p * n + idx];
  }
  sm[pg][lane] = s;
  __syncthreads();
  // every warp forms the same fixed-order total; warp r sends it to rank r
  float tot = sm[0][lane];
#pragma unroll
  for (int u = 1; u < 8; ++u) tot += sm[u][lane];
  const int dest = idx < n ? ll_dest<MAP>(idx) : -1;
  // conv2 bias gradient: the exact-fp32 sums of the linear backward (layout.cuh) replace the
  // tensor-core ones-column, whose operands carry 16 bits (every warp computes the same value)
  if (MAP == 1 && bias_part != nullptr && dest >= 64 * 32 * 9)
    tot = sum_bias_parts(bias_part, bp_nsplit, bp_nbx, bp_hw, dest - 64 * 32 * 9);
  if (pg < WORLD && dest >= 0)
    st_relaxed_sys_u64(ll_line(pp.p[pg], slot, e, rank) + idx,
                       (static_cast<uint64_t>(e) << 32) | __float_as_uint(tot));
  if (pg == 0 && dest >= 0) reinterpret_cast<float*>(pp.p[rank] + grad_off)[dest] = tot;
  __syncthreads();  // sm is reused below
  if (pg < WORLD && dest >= 0) {
    const uint64_t* src = ll_line(pp.p[rank], slot, e, pg) + idx;
    uint64_t v = ld_relaxed_sys_u64(src);
    if (static_cast<uint32_t>(v >> 32) != e) {
      const uint64_t t0 = global_timer_ns();
      uint32_t spins = 0;
      while (static_cast<uint32_t>((v = ld_relaxed_sys_u64(src)) >> 32) != e) {
        if ((++spins & 0xff) == 0 && global_timer_ns() - t0 > timeout_ns) {
          *status_word(pp.p[rank]) = 0xdead0000u | static_cast<uint32_t>(pg);
          __threadfence_system();
          __trap();
        }
      }
    }
    sm[pg][lane] = __uint_as_float(static_cast<uint32_t>(v));
  }
  __syncthreads();
  if (pg == 0 && dest >= 0) {
    float acc = sm[0][lane];
#pragma unroll
    for (int r = 1; r < WORLD; ++r) acc += sm[r][lane];
    float* pr = reinterpret_cast<float*>(pp.p[rank] + param_off) + dest;
    const float nv = fmaf(-(*lr) * (1.0f / WORLD), acc, *pr);
    *pr = nv;
    if (MAP == 1 && pack != nullptr && dest < 64 * 32 * 9) {
      // keep the split (hi + lo bf16) tensor-core tile images of net.2.weight (csrc/conv_tc.cu:
      // [fwd image | dgrad image], hi/lo rows adjacent) in step with the master weights: the next forward /
      // data gradient bulk-copy them without any conversion
      constexpr int n = 9 * 32 * 64;
      const int co = dest / (32 * 9), rem = dest - co * 32 * 9, ci = rem / 9, tap = rem - ci * 9;
      float hi, lo;
      split_bf16(nv, hi, lo);
      pack[conv3x3_fwd_split_index<32, 64>(co, ci, tap, 0)] = __float2bfloat16_rn(hi);
      pack[conv3x3_fwd_split_index<32, 64>(co, ci, tap, 1)] = __float2bfloat16_rn(lo);
      pack[2 * n + conv3x3_dgrad_split_index<32, 64>(co, ci, tap, 0)] = __float2bfloat16_rn(hi);
      pack[2 * n + conv3x3_dgrad_split_index<32, 64>(co, ci, tap, 1)] = __float2bfloat16_rn(lo);
    }
  }
  // the last block to finish publishes the new epoch for the next launch on this slot
  if (threadIdx.x == 0) {
    const uint32_t done = atomicAdd(ctl + 1, 1u);
    if (done == gridDim.x - 1) {
      ctl[1] = 0;
      ctl[0] = e;
    }
  }
}

// ---------------------------------------------------------------- plain all-reduce (sum)
template <typename T>
struct Vec16;
template <>
struct Vec16<float> {
  static constexpr int N = 4;
  __device__ static void add(float4& a, const float4& b) {
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
  }
};
template <>
struct Vec16<__nv_bfloat16> {
  static constexpr int N = 8;
  // accumulate in fp32, round once per hop in rank order
  __device__ static void add(float4& a, const float4& b) {
    const uint32_t* ua = reinterpret_cast<const uint32_t*>(&a);
    const uint32_t* ub = reinterpret_cast<const uint32_t*>(&b);
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      o[k] = pack_bf16x2(bf16_lo(ua[k]) + bf16_lo(ub[k]), bf16_hi(ua[k]) + bf16_hi(ub[k]));
    a = *reinterpret_cast<float4*>(o);
  }
};

// mode 0: one-shot (every rank reduces everything into its own `out`)
// mode 1: two-shot (rank r reduces slice r and stores it into every rank's `out`)
template <typename T, int WORLD, int MODE>
__global__ void __launch_bounds__(kCommThreads)
allreduce_kernel(const __grid_constant__ PeerPtrs pp, int rank, int slot, size_t in_off, size_t out_off, long long nvec,
                 unsigned long long timeout_ns) {
  uint32_t* ep = epoch_word(pp.p[rank], slot, blockIdx.x);
  const uint32_t e = *ep + 1;
  flag_round<WORLD>(pp, rank, slot, e, timeout_ns);
  long long lo = 0, hi = nvec;
  if (MODE == 1) {
    const long long per = (nvec + WORLD - 1) / WORLD;
    lo = per * rank;
    hi = lo + per < nvec ? lo + per : nvec;
  }
  for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi;
       i += (long long)gridDim.x * blockDim.x) {
    float4 v[WORLD];
#pragma unroll
    for (int r = 0; r < WORLD; ++r)
      v[r] = ld_f4(reinterpret_cast<const float*>(pp.p[r] + in_off) + i * 4);
    float4 s = v[0];
#pragma unroll
    for (int r = 1; r < WORLD; ++r) Vec16<T>::add(s, v[r]);
    if (MODE == 0) {
      st_f4(reinterpret_cast<float*>(pp.p[rank] + out_off) + i * 4, s);
    } else {
#pragma unroll
      for (int r = 0; r < WORLD; ++r) st_f4(reinterpret_cast<float*>(pp.p[r] + out_off) + i * 4, s);
    }
  }
  // second round: (one-shot) peers finished reading my input, so the caller may overwrite it;
  // (two-shot) every rank's slice has landed in my output.
  flag_round<WORLD>(pp, rank, slot, e + 1, timeout_ns);
  if (threadIdx.x == 0) *ep = e + 1;
}

// Every rank copies root's region into its own buffer (pull), bracketed by flag rounds.
template <int WORLD>
__global__ void __launch_bounds__(kCommThreads)
broadcast_kernel(const __grid_constant__ PeerPtrs pp, int rank, int slot, int root, size_t off, long long nvec,
                 unsigned long long timeout_ns) {
  uint32_t* ep = epoch_word(pp.p[rank], slot, blockIdx.x);
  const uint32_t e = *ep + 1;
  flag_round<WORLD>(pp, rank, slot, e, timeout_ns);
  if (rank != root) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
         i += (long long)gridDim.x * blockDim.x)
      st_f4(reinterpret_cast<float*>(pp.p[rank] + off) + i * 4,
            ld_f4(reinterpret_cast<const float*>(pp.p[root] + off) + i * 4));
  }
  flag_round<WORLD>(pp, rank, slot, e + 1, timeout_ns);
  if (threadIdx.x == 0) *ep = e + 1;
}

// ---------------------------------------------------------------- NVLS (multimem) kernels
// With the buffer bound to a multicast object, `mc + off` addresses the same offset of EVERY
// rank's buffer at once: multimem.ld_reduce returns the sum over all ranks computed inside the
// NVSwitch (one request instead of W-1 peer loads), multimem.st stores to all ranks at once.
// Each element is reduced by exactly one rank and broadcast, so all replicas stay bit-identical
// (the switch's summation order is its own, but there is only one sum per element).
__device__ __forceinline__ float4 multimem_ld_reduce_f32x4(const uint8_t* mc_addr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc_addr)
               : "memory");
  return v;
}
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const uint8_t* mc_addr) {
  uint4 v;
  asm volatile(
      "multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
      : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
      : "l"(mc_addr)
      : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f32x4(uint8_t* mc_addr, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr),
               "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void multimem_st_bf16x8(uint8_t* mc_addr, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr),
               "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// two-shot through the switch: rank r owns slice r; in-switch sum -> SGD -> multicast store of
// the updated parameters into every rank's arena.
template <int WORLD>
__global__ void __launch_bounds__(kCommThreads)
allreduce_sgd_nvls_kernel(const __grid_constant__ PeerPtrs pp, uint8_t* mc, int rank, int slot,
                          size_t grad_off, size_t param_off, long long n4,
                          const float* __restrict__ lr, unsigned long long timeout_ns) {
  pdl_sync();
  uint32_t* ep = epoch_word(pp.p[rank], slot, blockIdx.x);
  const uint32_t e = *ep + 1;
  flag_round<WORLD>(pp, rank, slot, e, timeout_ns);
  const float scale = -(*lr) * (1.0f / WORLD);
  const long long per = (n4 + WORLD - 1) / WORLD;
  const long long lo = per * rank;
  const long long hi = lo + per < n4 ? lo + per : n4;
  const float* params = reinterpret_cast<const float*>(pp.p[rank] + param_off);
  for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi;
       i += (long long)gridDim.x * blockDim.x) {
    const float4 s = multimem_ld_reduce_f32x4(mc + grad_off + i * 16);
    float4 p = *reinterpret_cast<const float4*>(params + i * 4);
    p.x = fmaf(scale, s.x, p.x);
    p.y = fmaf(scale, s.y, p.y);
    p.z = fmaf(scale, s.z, p.z);
    p.w = fmaf(scale, s.w, p.w);
    multimem_st_f32x4(mc + param_off + i * 16, p);
  }
  flag_round<WORLD>(pp, rank, slot, e + 1, timeout_ns);
  if (threadIdx.x == 0) *ep = e + 1;
}

// plain sum all-reduce through the switch (fp32 or bf16 with fp32 accumulation in the switch)
template <typename T, int WORLD>
__global__ void __launch_bounds__(kCommThreads)
allreduce_nvls_kernel(const __grid_constant__ PeerPtrs pp, uint8_t* mc, int rank, int slot,
                      size_t in_off, size_t out_off, long long nvec, unsigned long long timeout_ns) {
  uint32_t* ep = epoch_word(pp.p[rank], slot, blockIdx.x);
  const uint32_t e = *ep + 1;
  flag_round<WORLD>(pp, rank, slot, e, timeout_ns);
  const long long per = (nvec + WORLD - 1) / WORLD;
  const long long lo = per * rank;
  const long long hi = lo + per < nvec ? lo + per : nvec;
  for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi;
       i += (long long)gridDim.x * blockDim.x) {
    if (sizeof(T) == 4) {
      multimem_st_f32x4(mc + out_off + i * 16, multimem_ld_reduce_f32x4(mc + in_off + i * 16));
    } else {
      multimem_st_bf16x8(mc + out_off + i * 16, multimem_ld_reduce_bf16x8(mc + in_off + i * 16));
    }
  }
  flag_round<WORLD>(pp, rank, slot, e + 1, timeout_ns);
  if (threadIdx.x == 0) *ep = e + 1;
}

// Every dependent round of peer loads costs a full NVLink round trip (~4.5 us), so a thread should
// need as few rounds as possible — but these kernels also run beside the convolution backward on
// other streams, so a small bucket must not flood the SMs with spinning blocks either.  Buckets of
// up to 8 MB get at most kSmallMaxBlocks blocks (the fused kernels issue kPeerUnroll iterations of
// loads per round); DDPB_COMM_MAXBLOCKS overrides the cap.
constexpr int kSmallMaxBlocks = 48;
static int small_block_cap() {
  static int v = 0;
  if (v == 0) {
    const char* e = getenv("DDPB_COMM_MAXBLOCKS");
    v = e ? atoi(e) : kSmallMaxBlocks;
    if (v < 1) v = 1;
    if (v > kMaxBlocks) v = kMaxBlocks;
  }
  return v;
}
static int blocks_for(long long nvec, int world, bool twoshot) {
  long long work = twoshot ? (nvec + world - 1) / world : nvec;
  long long b = (work + kCommThreads - 1) / kCommThreads;
  if (b < 1) b = 1;
  const long long cap = nvec * 16 <= (8ll << 20) ? small_block_cap() : kMaxBlocks;
  if (b > cap) b = cap;
  return static_cast<int>(b);
}

static PeerPtrs peer_ptrs(const Comm* c) {
  PeerPtrs pp;
  for (int r = 0; r < kMaxWorld; ++r) pp.p[r] = r < c->world ? c->peer[r] : c->peer[c->rank];
  return pp;
}

// ---------------------------------------------------------------- driver API (VMM + multicast)
// Resolved at run time through the runtime's cudaGetDriverEntryPoint so the library has no link
// dependency on libcuda (it must still load on a machine without a driver, e.g. for CPU tests).
struct DrvApi {
  decltype(&cuDeviceGet) DeviceGet = nullptr;
  decltype(&cuDeviceGetAttribute) DeviceGetAttribute = nullptr;
  decltype(&cuMemCreate) MemCreate = nullptr;
  decltype(&cuMemRelease) MemRelease = nullptr;
  decltype(&cuMemAddressReserve) AddressReserve = nullptr;
  decltype(&cuMemAddressFree) AddressFree = nullptr;
  decltype(&cuMemMap) Map = nullptr;
  decltype(&cuMemUnmap) Unmap = nullptr;
  decltype(&cuMemSetAccess) SetAccess = nullptr;
  decltype(&cuMemGetAllocationGranularity) GetGranularity = nullptr;
  decltype(&cuMemExportToShareableHandle) Export = nullptr;
  decltype(&cuMemImportFromShareableHandle) Import = nullptr;
  decltype(&cuMulticastCreate) McCreate = nullptr;
  decltype(&cuMulticastAddDevice) McAddDevice = nullptr;
  decltype(&cuMulticastBindMem) McBindMem = nullptr;
  decltype(&cuMulticastUnbind) McUnbind = nullptr;
  decltype(&cuMulticastGetGranularity) McGetGranularity = nullptr;
  bool ok = false;
};

template <class F>
static bool drv_load(const char* name, F* out) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || fn == nullptr) {
    cudaGetLastError();
    return false;
  }
  *out = reinterpret_cast<F>(fn);
  return true;
}

static const DrvApi* drv() {
  static DrvApi api;
  static bool tried = false;
  if (!tried) {
    tried = true;
    DrvApi a;
    bool ok = drv_load("cuDeviceGet", &a.DeviceGet);
    ok = ok && drv_load("cuDeviceGetAttribute", &a.DeviceGetAttribute);
    ok = ok && drv_load("cuMemCreate", &a.MemCreate);
    ok = ok && drv_load("cuMemRelease", &a.MemRelease);
    ok = ok && drv_load("cuMemAddressReserve", &a.AddressReserve);
    ok = ok && drv_load("cuMemAddressFree", &a.AddressFree);
    ok = ok && drv_load("cuMemMap", &a.Map);
    ok = ok && drv_load("cuMemUnmap", &a.Unmap);
    ok = ok && drv_load("cuMemSetAccess", &a.SetAccess);
    ok = ok && drv_load("cuMemGetAllocationGranularity", &a.GetGranularity);
    ok = ok && drv_load("cuMemExportToShareableHandle", &a.Export);
    ok = ok && drv_load("cuMemImportFromShareableHandle", &a.Import);
    ok = ok && drv_load("cuMulticastCreate", &a.McCreate);
    ok = ok && drv_load("cuMulticastAddDevice", &a.McAddDevice);
    ok = ok && drv_load("cuMulticastBindMem", &a.McBindMem);
    ok = ok && drv_load("cuMulticastUnbind", &a.McUnbind);
    ok = ok && drv_load("cuMulticastGetGranularity", &a.McGetGranularity);
    a.ok = ok;
    api = a;
  }
  return api.ok ? &api : nullptr;
}

#define DDPB_CU_OK(expr)                                 \
  do {                                                   \
    CUresult _r = (expr);                                \
    if (_r != CUDA_SUCCESS) return static_cast<int>(_r); \
  } while (0)

static CUmemAllocationProp vmm_alloc_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

// Reserve a VA range, map `handle` into it and give THIS device read/write access.
static int vmm_map(const DrvApi* d, CUmemGenericAllocationHandle handle, size_t bytes, size_t align,
                   int device, uint8_t** out) {
  CUdeviceptr va = 0;
  DDPB_CU_OK(d->AddressReserve(&va, bytes, align, 0, 0));
  CUresult r = d->Map(va, bytes, 0, handle, 0);
  if (r == CUDA_SUCCESS) {
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = device;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = d->SetAccess(va, bytes, &acc, 1);
    if (r != CUDA_SUCCESS) d->Unmap(va, bytes);
  }
  if (r != CUDA_SUCCESS) {
    d->AddressFree(va, bytes);
    return static_cast<int>(r);
  }
  *out = reinterpret_cast<uint8_t*>(va);
  return 0;
}

static void vmm_unmap(const DrvApi* d, uint8_t* p, size_t bytes) {
  if (!p) return;
  d->Unmap(reinterpret_cast<CUdeviceptr>(p), bytes);
  d->AddressFree(reinterpret_cast<CUdeviceptr>(p), bytes);
}

#define DDPB_DISPATCH_WORLD(W, ...)                   \
  switch (W) {                                        \
    case 1: { constexpr int WORLD = 1; __VA_ARGS__; } break; \
    case 2: { constexpr int WORLD = 2; __VA_ARGS__; } break; \
    case 3: { constexpr int WORLD = 3; __VA_ARGS__; } break; \
    case 4: { constexpr int WORLD = 4; __VA_ARGS__; } break; \
    case 5: { constexpr int WORLD = 5; __VA_ARGS__; } break; \
    case 6: { constexpr int WORLD = 6; __VA_ARGS__; } break; \
    case 7: { constexpr int WORLD = 7; __VA_ARGS__; } break; \
    case 8: { constexpr int WORLD = 8; __VA_ARGS__; } break; \
    default: return DDPB_ERR_SHAPE;                   \
  }

}  // namespace ddpb

using namespace ddpb;

extern "C" {

size_t ddpb_comm_header_bytes() { return kHeaderBytes; }
int ddpb_comm_handle_bytes() { return static_cast<int>(sizeof(cudaIpcMemHandle_t)); }
int ddpb_comm_max_world() { return kMaxWorld; }
int ddpb_comm_num_slots() { return kSlots; }

// Allocates this rank's symmetric buffer (header + payload_bytes, zero-filled) on the current
// device.  world: 1 .. 8 (any size; the kernels are instantiated per world size).
int ddpb_comm_create(int rank, int world, size_t payload_bytes, void** out) {
  if (world < 1 || world > kMaxWorld || rank < 0 || rank >= world) return DDPB_ERR_SHAPE;
  Comm* c = new (std::nothrow) Comm();
  if (!c) return static_cast<int>(cudaErrorMemoryAllocation);
  c->rank = rank;
  c->world = world;
  DDPB_CUDA_OK(cudaGetDevice(&c->device));
  c->bytes = kHeaderBytes + (payload_bytes + 255) / 256 * 256;
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&c->local), c->bytes);
  if (e != cudaSuccess) {
    delete c;
    return static_cast<int>(e);
  }
  DDPB_CUDA_OK(cudaMemset(c->local, 0, c->bytes));
  DDPB_CUDA_OK(cudaDeviceSynchronize());
  c->peer[rank] = c->local;
  c->connected = (world == 1);
  *out = c;
  return 0;
}

// ---------------------------------------------------------------- VMM / multicast communicator
// Same buffer layout and kernels as the IPC communicator, but the memory is a cuMemCreate
// allocation shared through POSIX file descriptors (the caller moves the descriptors between
// the processes, e.g. over an AF_UNIX socket) so that it can ALSO be bound to a multicast
// object.  Call order on every rank:
//   create_vmm -> export_fd / import_peer_fd (all peers) ->
//   [rank 0: mc_create, others: mc_import] -> mc_add_device -> (barrier) -> mc_bind -> (barrier)
//   -> set_connected.
// Every step returns 0 or a CUDA/driver error code; nothing here aborts, so the caller can agree
// with its peers to fall back to the IPC communicator.

// 1 if the current device can take part in a multicast object, 0 if not, < 0 if the driver API
// is not available.
int ddpb_multicast_supported() {
  const DrvApi* d = drv();
  if (!d) return -1;
  int dev = 0, flag = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  CUdevice cudev;
  if (d->DeviceGet(&cudev, dev) != CUDA_SUCCESS) return -1;
  if (d->DeviceGetAttribute(&flag, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev) != CUDA_SUCCESS)
    return -1;
  return flag ? 1 : 0;
}

int ddpb_comm_create_vmm(int rank, int world, size_t payload_bytes, void** out) {
  if (world < 1 || world > kMaxWorld || rank < 0 || rank >= world) return DDPB_ERR_SHAPE;
  const DrvApi* d = drv();
  if (!d) return DDPB_ERR_UNSUPPORTED;
  Comm* c = new (std::nothrow) Comm();
  if (!c) return static_cast<int>(cudaErrorMemoryAllocation);
  c->rank = rank;
  c->world = world;
  c->vmm = true;
  int rc = static_cast<int>(cudaGetDevice(&c->device));
  if (rc == 0) rc = static_cast<int>(cudaFree(nullptr));  // make sure the primary context exists
  size_t gran = 0, mc_gran = 0;
  const CUmemAllocationProp prop = vmm_alloc_prop(c->device);
  if (rc == 0) rc = static_cast<int>(d->GetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (rc == 0 && ddpb_multicast_supported() == 1) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = static_cast<unsigned>(world);
    mp.size = gran;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t rec = 0, mn = 0;
    if (d->McGetGranularity(&mn, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) == CUDA_SUCCESS) mc_gran = mn;
    // the recommended granularity can be very large; take it only while it stays modest
    if (d->McGetGranularity(&rec, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS &&
        rec <= (32u << 20) && rec > mc_gran)
      mc_gran = rec;
  }
  if (rc == 0) {
    size_t align = gran > mc_gran ? gran : mc_gran;
    if (align == 0) align = 2u << 20;
    c->bytes = kHeaderBytes + (payload_bytes + 255) / 256 * 256;
    c->map_bytes = (c->bytes + align - 1) / align * align;
    c->align = align;
    rc = static_cast<int>(d->MemCreate(&c->mem, c->map_bytes, &prop, 0));
  }
  if (rc == 0) {
    rc = vmm_map(d, c->mem, c->map_bytes, c->align, c->device, &c->local);
    if (rc != 0) d->MemRelease(c->mem);
  }
  if (rc == 0) rc = static_cast<int>(cudaMemset(c->local, 0, c->map_bytes));
  if (rc == 0) rc = static_cast<int>(cudaDeviceSynchronize());
  if (rc != 0) {
    cudaGetLastError();
    delete c;
    return rc;
  }
  c->peer[rank] = c->local;
  c->connected = false;
  *out = c;
  return 0;
}

// A new file descriptor naming this rank's allocation; the caller sends it to the peers and
// closes it.
int ddpb_comm_export_fd(void* comm, int* fd_out) {
  Comm* c = static_cast<Comm*>(comm);
  const DrvApi* d = drv();
  if (!d || !c->vmm) return DDPB_ERR_STATE;
  int fd = -1;
  DDPB_CU_OK(d->Export(&fd, c->mem, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd_out = fd;
  return 0;
}

// Maps rank r's allocation from a descriptor received from r (the descriptor stays the caller's).
int ddpb_comm_import_peer_fd(void* comm, int r, int fd) {
  Comm* c = static_cast<Comm*>(comm);
  const DrvApi* d = drv();
  if (!d || !c->vmm || r < 0 || r >= c->world || r == c->rank || c->opened[r]) return DDPB_ERR_STATE;
  CUmemGenericAllocationHandle h = 0;
  DDPB_CU_OK(d->Import(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                       CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  uint8_t* p = nullptr;
  const int rc = vmm_map(d, h, c->map_bytes, c->align, c->device, &p);
  if (rc != 0) {
    d->MemRelease(h);
    return rc;
  }
  c->peer_mem[r] = h;
  c->peer[r] = p;
  c->opened[r] = true;
  return 0;
}

// Rank 0 creates the multicast object and exports a descriptor for the others.
int ddpb_comm_mc_create(void* comm, int* fd_out) {
  Comm* c = static_cast<Comm*>(comm);
  const DrvApi* d = drv();
  if (!d || !c->vmm || c->mc_created) return DDPB_ERR_STATE;
  if (ddpb_multicast_supported() != 1) return DDPB_ERR_UNSUPPORTED;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = static_cast<unsigned>(c->world);
  mp.size = c->map_bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  DDPB_CU_OK(d->McCreate(&c->mc_handle, &mp));
  c->mc_created = true;
  int fd = -1;
  DDPB_CU_OK(d->Export(&fd, c->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd_out = fd;
  return 0;
}

int ddpb_comm_mc_import(void* comm, int fd) {
  Comm* c = static_cast<Comm*>(comm);
  const DrvApi* d = drv();
  if (!d || !c->vmm || c->mc_created) return DDPB_ERR_STATE;
  if (ddpb_multicast_supported() != 1) return DDPB_ERR_UNSUPPORTED;
  DDPB_CU_OK(d->Import(&c->mc_handle, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                       CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  c->mc_created = true;
  return 0;
}

int ddpb_comm_mc_add_device(void* comm) {
  Comm* c = static_cast<Comm*>(comm);
  const DrvApi* d = drv();
  if (!d || !c->mc_created) return DDPB_ERR_STATE;
  CUdevice cudev;
  DDPB_CU_OK(d->DeviceGet(&cudev, c->device));
  DDPB_CU_OK(d->McAddDevice(c->mc_handle, cudev));
  return 0;
}

// Binds this rank's memory to the multicast object and maps the multicast address range.  The
// driver blocks here until EVERY rank has called mc_add_device: only call it once all ranks
// reported success for that step.
int ddpb_comm_mc_bind(void* comm) {
  Comm* c = static_cast<Comm*>(comm);
  const DrvApi* d = drv();
  if (!d || !c->mc_created || c->mc_bound) return DDPB_ERR_STATE;
  DDPB_CU_OK(d->McBindMem(c->mc_handle, 0, c->mem, 0, c->map_bytes, 0));
  c->mc_bound = true;
  uint8_t* p = nullptr;
  const int rc = vmm_map(d, c->mc_handle, c->map_bytes, c->align, c->device, &p);
  if (rc != 0) return rc;
  c->mc = p;
  return 0;
}

int ddpb_comm_set_connected(void* comm) {
  Comm* c = static_cast<Comm*>(comm);
  for (int r = 0; r < c->world; ++r)
    if (!c->peer[r]) return DDPB_ERR_STATE;
  c->connected = true;
  return 0;
}

int ddpb_comm_has_multicast(void* comm) { return static_cast<Comm*>(comm)->mc != nullptr ? 1 : 0; }
// multicast alias of ddpb_comm_local_ptr (nullptr without multicast)
void* ddpb_comm_multicast_ptr(void* comm) {
  Comm* c = static_cast<Comm*>(comm);
  return c->mc ? c->mc + kHeaderBytes : nullptr;
}

int ddpb_comm_get_handle(void* comm, void* handle_out) {
  Comm* c = static_cast<Comm*>(comm);
  if (c->vmm) return DDPB_ERR_STATE;  // VMM allocations are shared by file descriptor, not IPC
  cudaIpcMemHandle_t h;
  DDPB_CUDA_OK(cudaIpcGetMemHandle(&h, c->local));
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}

// all_handles: world * ddpb_comm_handle_bytes() bytes (host), rank-major.
int ddpb_comm_connect(void* comm, const void* all_handles) {
  Comm* c = static_cast<Comm*>(comm);
  if (c->vmm) return DDPB_ERR_STATE;
  const uint8_t* hs = static_cast<const uint8_t*>(all_handles);
  for (int r = 0; r < c->world; ++r) {
    if (r == c->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, hs + (size_t)r * sizeof(h), sizeof(h));
    void* p = nullptr;
    DDPB_CUDA_OK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->peer[r] = static_cast<uint8_t*>(p);
    c->opened[r] = true;
  }
  c->connected = true;
  return 0;
}

void* ddpb_comm_local_ptr(void* comm) { return static_cast<Comm*>(comm)->local + kHeaderBytes; }
void* ddpb_comm_peer_ptr(void* comm, int r) {
  Comm* c = static_cast<Comm*>(comm);
  return (r >= 0 && r < c->world && c->peer[r]) ? c->peer[r] + kHeaderBytes : nullptr;
}
size_t ddpb_comm_payload_bytes(void* comm) { return static_cast<Comm*>(comm)->bytes - kHeaderBytes; }
void ddpb_comm_set_timeout_ms(void* comm, long long ms) {
  static_cast<Comm*>(comm)->timeout_ns = static_cast<unsigned long long>(ms) * 1000000ull;
}
// Last status word written by a kernel that timed out waiting for a peer (0 = healthy).
int ddpb_comm_status(void* comm, unsigned int* status_out) {
  Comm* c = static_cast<Comm*>(comm);
  return static_cast<int>(cudaMemcpy(status_out,
                                     reinterpret_cast<uint32_t*>(c->local) + kFlagWords + kEpochWords,
                                     4, cudaMemcpyDeviceToHost));
}

int ddpb_comm_destroy(void* comm) {
  Comm* c = static_cast<Comm*>(comm);
  if (!c) return 0;
  cudaDeviceSynchronize();
  if (c->vmm) {
    const DrvApi* d = drv();
    if (d) {
      if (c->mc) vmm_unmap(d, c->mc, c->map_bytes);
      if (c->mc_bound) {
        CUdevice cudev;
        if (d->DeviceGet(&cudev, c->device) == CUDA_SUCCESS)
          d->McUnbind(c->mc_handle, cudev, 0, c->map_bytes);
      }
      if (c->mc_created) d->MemRelease(c->mc_handle);
      for (int r = 0; r < c->world; ++r)
        if (c->opened[r]) {
          vmm_unmap(d, c->peer[r], c->map_bytes);
          d->MemRelease(c->peer_mem[r]);
        }
      if (c->local) {
        vmm_unmap(d, c->local, c->map_bytes);
        d->MemRelease(c->mem);
      }
    }
    delete c;
    return 0;
  }
  for (int r = 0; r < c->world; ++r)
    if (c->opened[r]) cudaIpcCloseMemHandle(c->peer[r]);
  if (c->local) cudaFree(c->local);
  delete c;
  return 0;
}

// Offsets below are byte offsets into the payload (i.e. relative to ddpb_comm_local_ptr),
// 16-byte aligned, identical on every rank.  algo: 1 = one-shot, 2 = two-shot, 4 = two-shot through
// the switch (multimem; needs a multicast-bound communicator), 0 = by size.
int ddpb_allreduce_sgd(void* comm, size_t grad_off, size_t param_off, long long n_floats,
                       const float* lr_dev, int algo, int slot, void* stream) {
  Comm* c = static_cast<Comm*>(comm);
  if (!c->connected) return DDPB_ERR_STATE;
  if ((n_floats & 3) || (grad_off & 15) || (param_off & 15) || slot < 0 || slot >= kSlots)
    return DDPB_ERR_SHAPE;
  const long long n4 = n_floats / 4;
  if (algo == 0) algo = (c->world <= 2 || n_floats * 4 <= (256 << 10)) ? 1 : (c->mc ? 4 : 2);
  if (algo == 4 && !c->mc) return DDPB_ERR_STATE;
  const bool two = algo == 2 || algo == 4;
  const int blocks = blocks_for(n4, c->world, two);
  const PeerPtrs pp = peer_ptrs(c);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t go = grad_off + kHeaderBytes, po = param_off + kHeaderBytes;
  DDPB_DISPATCH_WORLD(c->world, {
    if (algo == 4)
      return static_cast<int>(launch_k(allreduce_sgd_nvls_kernel<WORLD>, dim3(blocks),
                                       dim3(kCommThreads), 0, st, pp, c->mc, c->rank, slot, go, po, n4,
                                       lr_dev, c->timeout_ns));
    if (two)
      return static_cast<int>(launch_k(allreduce_sgd_twoshot_kernel<WORLD>, dim3(blocks),
                                       dim3(kCommThreads), 0, st, pp, c->rank, slot, go, po, n4, lr_dev,
                                       c->timeout_ns));
    return static_cast<int>(launch_k(allreduce_sgd_oneshot_kernel<WORLD>, dim3(blocks),
                                     dim3(kCommThreads), 0, st, pp, c->rank, slot, go, po, n4, lr_dev,
                                     c->timeout_ns));
  });
  return static_cast<int>(cudaGetLastError());
}

// grads[slice] = fixed-order sum of `nparts` partial vectors (each n floats, n % 4 == 0), then
// params[slice] -= lr/world * sum_over_ranks(grads[slice]); one kernel.
// algo: 0 or 3 = push with flag-carrying words (lowest latency; n <= 19,456), 1 = flag round + pull
// (n <= 5,120).  layout: 0 = partial element i is slice element i; 1 = the partials are conv3x3
// <32,64> weight-gradient accumulator images (n = 18,944) and the slice is net.2.weight followed by
// net.2.bias (push only).
// ddpb_reduce_allreduce_sgd_pack: same, and with layout 1 also rewrites `pack` (the split bf16 tile
// images of net.2.weight, ddpb_conv3x3_pack_split_bytes(32, 64) bytes) from the updated weights.
static int reduce_allreduce_sgd_impl(void* comm, const float* partials, int nparts, long long n,
                                     size_t grad_off, size_t param_off, const float* lr_dev, int algo,
                                     int layout, int slot, __nv_bfloat16* pack, const float* bias_part,
                                     int bp_nsplit, int bp_nbx, int bp_hw, void* stream);

int ddpb_reduce_allreduce_sgd(void* comm, const float* partials, int nparts, long long n,
                              size_t grad_off, size_t param_off, const float* lr_dev, int algo,
                              int layout, int slot, void* stream) {
  return reduce_allreduce_sgd_impl(comm, partials, nparts, n, grad_off, param_off, lr_dev, algo, layout,
                                   slot, nullptr, nullptr, 0, 0, 0, stream);
}
// bias_part (optional, layout 1): exact-fp32 bias-gradient parts written by ddpb_linear_bwd_f32
// (extents bp_nsplit x bp_nbx, bp_hw = H*W); their fixed-order sum replaces the bias column of the
// accumulator images.
int ddpb_reduce_allreduce_sgd_pack(void* comm, const float* partials, int nparts, long long n,
                                   size_t grad_off, size_t param_off, const float* lr_dev, int algo,
                                   int layout, int slot, void* pack, const float* bias_part,
                                   int bp_nsplit, int bp_nbx, int bp_hw, void* stream) {
  if ((pack != nullptr || bias_part != nullptr) && layout != 1) return DDPB_ERR_SHAPE;
  return reduce_allreduce_sgd_impl(comm, partials, nparts, n, grad_off, param_off, lr_dev, algo, layout,
                                   slot, static_cast<__nv_bfloat16*>(pack), bias_part, bp_nsplit, bp_nbx,
                                   bp_hw, stream);
}

static int reduce_allreduce_sgd_impl(void* comm, const float* partials, int nparts, long long n,
                                     size_t grad_off, size_t param_off, const float* lr_dev, int algo,
                                     int layout, int slot, __nv_bfloat16* pack, const float* bias_part,
                                     int bp_nsplit, int bp_nbx, int bp_hw, void* stream) {
  Comm* c = static_cast<Comm*>(comm);
  if (!c->connected) return DDPB_ERR_STATE;
  const long long blocks = (n + 31) / 32;
  if ((n & 3) || (grad_off & 15) || (param_off & 15) || slot < 0 || slot >= kSlots || nparts < 1 ||
      !(algo == 0 || algo == 1 || algo == 3) || !(layout == 0 || layout == 1))
    return DDPB_ERR_SHAPE;
  if (layout == 1 && (algo == 1 || n != 64 * (9 * 32 + 8))) return DDPB_ERR_SHAPE;
  if (algo == 1 ? blocks > kMaxBlocks : n > kLLMaxN) return DDPB_ERR_LIMIT;
  const PeerPtrs pp = peer_ptrs(c);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t go = grad_off + kHeaderBytes, po = param_off + kHeaderBytes;
  const dim3 grid(static_cast<int>(blocks));
  DDPB_DISPATCH_WORLD(c->world, {
    if (algo == 1)
      return static_cast<int>(launch_k(reduce_allreduce_sgd_kernel<WORLD>, grid, dim3(256), 0, st, pp,
                                       c->rank, slot, partials, nparts, static_cast<int>(n), go, po,
                                       lr_dev, c->timeout_ns));
    if (layout == 1)
      return static_cast<int>(launch_k(reduce_allreduce_sgd_ll_kernel<WORLD, 1>, grid, dim3(256), 0, st,
                                       pp, c->rank, slot, partials, nparts, static_cast<int>(n), go, po,
                                       lr_dev, pack, bias_part, bp_nsplit, bp_nbx, bp_hw, c->timeout_ns));
    return static_cast<int>(launch_k(reduce_allreduce_sgd_ll_kernel<WORLD, 0>, grid, dim3(256), 0, st, pp,
                                     c->rank, slot, partials, nparts, static_cast<int>(n), go, po,
                                     lr_dev, static_cast<__nv_bfloat16*>(nullptr),
                                     static_cast<const float*>(nullptr), 0, 0, 0, c->timeout_ns));
  });
  return static_cast<int>(cudaGetLastError());
}

// dtype: 0 = fp32, 1 = bf16.  out_off must not overlap in_off.
int ddpb_allreduce(void* comm, size_t in_off, size_t out_off, long long nbytes, int dtype, int algo,
                   int slot, void* stream) {
  Comm* c = static_cast<Comm*>(comm);
  if (!c->connected) return DDPB_ERR_STATE;
  if ((nbytes & 15) || (in_off & 15) || (out_off & 15) || slot < 0 || slot >= kSlots)
    return DDPB_ERR_SHAPE;
  const long long nvec = nbytes / 16;
  if (algo == 0) algo = (c->world <= 2 || nbytes <= (256 << 10)) ? 1 : (c->mc ? 4 : 2);
  if (algo == 4 && !c->mc) return DDPB_ERR_STATE;
  const bool two = algo == 2 || algo == 4;
  const int blocks = blocks_for(nvec, c->world, two);
  const PeerPtrs pp = peer_ptrs(c);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t io = in_off + kHeaderBytes, oo = out_off + kHeaderBytes;
  DDPB_DISPATCH_WORLD(c->world, {
    if (algo == 4) {
      if (dtype == 0) allreduce_nvls_kernel<float, WORLD><<<blocks, kCommThreads, 0, st>>>(pp, c->mc, c->rank, slot, io, oo, nvec, c->timeout_ns);
      else if (dtype == 1) allreduce_nvls_kernel<__nv_bfloat16, WORLD><<<blocks, kCommThreads, 0, st>>>(pp, c->mc, c->rank, slot, io, oo, nvec, c->timeout_ns);
      else return DDPB_ERR_SHAPE;
    } else if (dtype == 0) {
      if (two) allreduce_kernel<float, WORLD, 1><<<blocks, kCommThreads, 0, st>>>(pp, c->rank, slot, io, oo, nvec, c->timeout_ns);
      else allreduce_kernel<float, WORLD, 0><<<blocks, kCommThreads, 0, st>>>(pp, c->rank, slot, io, oo, nvec, c->timeout_ns);
    } else if (dtype == 1) {
      if (two) allreduce_kernel<__nv_bfloat16, WORLD, 1><<<blocks, kCommThreads, 0, st>>>(pp, c->rank, slot, io, oo, nvec, c->timeout_ns);
      else allreduce_kernel<__nv_bfloat16, WORLD, 0><<<blocks, kCommThreads, 0, st>>>(pp, c->rank, slot, io, oo, nvec, c->timeout_ns);
    } else {
      return DDPB_ERR_SHAPE;
    }
  });
  return static_cast<int>(cudaGetLastError());
}

int ddpb_broadcast(void* comm, size_t off, long long nbytes, int root, int slot, void* stream) {
  Comm* c = static_cast<Comm*>(comm);
  if (!c->connected) return DDPB_ERR_STATE;
  if ((nbytes & 15) || (off & 15) || slot < 0 || slot >= kSlots || root < 0 || root >= c->world)
    return DDPB_ERR_SHAPE;
  const long long nvec = nbytes / 16;
  const int blocks = blocks_for(nvec, c->world, false);
  const PeerPtrs pp = peer_ptrs(c);
  DDPB_DISPATCH_WORLD(c->world, {
    broadcast_kernel<WORLD><<<blocks, kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        pp, c->rank, slot, root, off + kHeaderBytes, nvec, c->timeout_ns);
  });
  return static_cast<int>(cudaGetLastError());
}

}  // extern "C"
