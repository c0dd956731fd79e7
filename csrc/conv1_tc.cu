// First convolution (Cin = 1, 3x3, pad 1) on the tensor cores.
//
// K = 9 is far too thin for a normal tile, but a tcgen05.mma costs ~45 cycles whatever its
// shape (tools/mma_bench.py), while the CUDA-core version needs 72 FMAs per pixel per 8 channels
// and was instruction-bound (profiles/).  So:
//   forward : per 128-pixel tile the producer warps build the im2col rows [128 x 16] (9 taps +
//             zero padding) in shared memory as TF32 — the precision cuDNN uses for the
//             reference's convolutions by default — two K=8 MMAs against W1 [32 x 16], epilogue
//             = bias + ReLU + bf16 planes (same epilogue as conv2).
//   wgrad   : dW1[c][tap] = sum_p dact1[p][c] * x[p + tap]:  A = the dact1 planes as an MN-major
//             operand (bulk-copied), B = im2col rows built by the producers as TWO bf16 terms
//             (x = hi + lo, so the bf16 MMA keeps ~16 mantissa bits of x) plus a column of ones
//             that yields the bias gradient; M = 64 (32 real channels), N = 32, K = 128 pixels
//             per tile; per-CTA partial in TMEM, fixed-order reduce.
// Replaces cuDNN conv fwd / wgrad + ATen relu for reference model.py:9-10 (SURVEY.md K1,K2,K12,K13).
#include <stdlib.h>

#include "ptx.cuh"
#include "layout.cuh"
#include "api_internal.cuh"
#include "launch.cuh"

namespace ddpb {

constexpr int kC1Threads = 288;     // wgrad: warps 0-3 producers A, 4 MMA/TMEM, 5-8 producers B + epilogue
constexpr int kC1FwdThreads = 416;  // fwd: warps 0-3 producers A, 4 MMA/TMEM, 5-8 epilogue, 9-12 producers B
constexpr int kC1Stages = 3;

// Input-image staging: each producer group owns a private ring of kXStagesG stages (so the threads
// that wait on a stage barrier are the ones that consumed its previous phase — a ring shared by both
// groups let a group that ran ahead mistake an old phase of the other group's tile for its own).
constexpr int kXStagesG = 2;              // prefetch distance, in tiles of one group
constexpr int kXStages = 2 * kXStagesG;   // stages (and barriers) in total
constexpr uint32_t kXStageBytes = 8192;  // two images of up to 32x32 fp32

// A 128-row tile touches at most two consecutive images.  One bulk async copy stages both
// (clamped to the batch) into shared memory; returns the first staged image index.
__device__ __forceinline__ int x_stage_first_image(const Geom& g, int tile) {
  const uint32_t row0 = static_cast<uint32_t>(tile * kTileM);
  const int img0 = g.mul_phw ? static_cast<int>(__umulhi(row0, g.mul_phw) >> g.sh_phw)
                             : static_cast<int>(row0 / g.PHW);
  const int last_start = g.batch >= 2 ? g.batch - 2 : 0;
  return img0 < last_start ? img0 : last_start;
}
__device__ __forceinline__ void x_stage_issue(const float* __restrict__ x, const Geom& g, int tile,
                                              uint32_t dst_smem, uint32_t bar) {
  const int first = x_stage_first_image(g, tile);
  const uint32_t img_bytes = g.H * g.W * 4;
  const uint32_t bytes = (g.batch >= 2 ? 2 : 1) * img_bytes;
  mbar_arrive_expect_tx(bar, bytes);
  bulk_g2s(dst_smem, x + (long long)first * g.H * g.W, bytes, bar);
}
__device__ __forceinline__ void taps_from_stage(const float* xs, const Geom& g, int li, int h, int w,
                                                float (&t)[9]) {
  const float* xp = xs + li * g.H * g.W;
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int hh = h + r - 1, ww = w + s - 1;
      t[r * 3 + s] = (hh >= 0 && hh < g.H && ww >= 0 && ww < g.W) ? xp[hh * g.W + ww] : 0.f;
    }
}
__device__ __forceinline__ void producer_group_sync(int grp) {  // 128 threads of producer group grp
  if (grp == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
  else asm volatile("bar.sync 2, 128;" ::: "memory");
}

// ------------------------------------------------------------------------------------- forward
// X3 = false: one TF32 GEMM, K = 16 (9 taps + padding), two K=8 MMAs per tile.
// X3 = true : "3xTF32" — x = hi + lo and w = hi + lo with all four terms tf32 (22-bit significands):
//             A row = [x_hi taps 0-8 | x_lo taps 0-8 | 6 zeros] (K = 24), B1 = [w_hi | w_hi | 0],
//             B2 = [w_lo | 0]; three MMAs against B1 (x_hi.w_hi + x_lo.w_hi) and two against B2
//             (x_hi.w_lo; columns 9-15 of B2 are zero).  fp32-level accuracy: cuDNN runs this Cin = 1
//             layer in plain fp32 for some batch sizes (profiles/precision.md), so TF32 alone would be
//             below the reference there.
template <bool X3>
struct C1FwdSmem {
  static constexpr uint32_t NKA = X3 ? 6 : 4;             // 4-wide k-chunks of an A row
  static constexpr uint32_t b1_bytes = NKA * 32 * 16;     // W1 as tf32 [kchunks][32 n][4]
  static constexpr uint32_t b_bytes = b1_bytes + (X3 ? 4 * 32 * 16 : 0);
  static constexpr uint32_t a_plane = kTileM * 16;        // one k-chunk of all 128 rows
  static constexpr uint32_t a_stage = NKA * a_plane;      // 8 / 12 KB
  static constexpr uint32_t off_a = b_bytes;
  static constexpr uint32_t off_x = off_a + kC1Stages * a_stage;
  static constexpr uint32_t off_bias = off_x + kXStages * kXStageBytes;
  static constexpr uint32_t off_bar = off_bias + 32 * 4;
  static constexpr uint32_t off_tmem = off_bar + (2 * kC1Stages + 4 + kXStages) * 8;
  static constexpr uint32_t total = off_tmem + 16;
};

// P selects the output planes: bf16, split (hi + lo bf16 planes, the reference-precision mode) or
// fp32 rounded to tf32 (for tf32 GEMM consumers: the rounding cuDNN applies to its input).
template <class P>
__global__ void __launch_bounds__(kC1FwdThreads, 2)
conv1_fwd_tc_kernel(const float* __restrict__ x, const float* __restrict__ w,
                    const float* __restrict__ bias, uint4* __restrict__ out, Geom g) {
  constexpr bool X3 = P::kSets == 2;  // the reference-precision mode
  using S = C1FwdSmem<X3>;
  constexpr int CO = 32;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t s_a = s_base + S::off_a;
  float* s_bias = reinterpret_cast<float*>(smem + S::off_bias);
  const uint32_t bar_full = s_base + S::off_bar;
  const uint32_t bar_empty = bar_full + kC1Stages * 8;
  const uint32_t bar_tfull = bar_empty + kC1Stages * 8;
  const uint32_t bar_tempty = bar_tfull + 16;
  const uint32_t bar_x = bar_tempty + 16;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + S::off_tmem);

  if (tid == 0) {
    for (int i = 0; i < kC1Stages; ++i) {
      mbar_init(bar_full + i * 8, 128);
      mbar_init(bar_empty + i * 8, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_tfull + i * 8, 1);
      mbar_init(bar_tempty + i * 8, 128);
    }
    for (int i = 0; i < kXStages; ++i) mbar_init(bar_x + i * 8, 1);
    fence_mbar_init();
  }
  if (warp == 4) {
    tmem_alloc(smem_u32(s_tmem), 64);
    tmem_relinquish();
  }
  // The stream predecessor is the previous step's parameter update, which writes w and bias:
  // nothing global may be read before it has completed.
  pdl_sync();
  // W1 [32][9] -> tf32 K-major canonical tile(s): word (k/4)*128 + n*4 + k%4 ; zero padding elsewhere
  if constexpr (X3) {
    float* b1 = reinterpret_cast<float*>(smem);
    float* b2 = reinterpret_cast<float*>(smem + S::b1_bytes);
    for (int i = tid; i < 32 * 24; i += kC1FwdThreads) {
      const int n = i / 24, k = i - n * 24;
      const int tap = k < 9 ? k : k - 9;
      float hi = 0.f, lo = 0.f;
      if (k < 18) {
        const float wv = w[n * 9 + tap];
        hi = round_tf32(wv);
        lo = round_tf32(wv - hi);
      }
      b1[(k >> 2) * 128 + n * 4 + (k & 3)] = hi;
      if (k < 16) b2[(k >> 2) * 128 + n * 4 + (k & 3)] = k < 9 ? lo : 0.f;
    }
  } else {
    for (int i = tid; i < 32 * 16; i += kC1FwdThreads) {
      const int n = i >> 4, k = i & 15;
      reinterpret_cast<float*>(smem)[(k >> 2) * 128 + n * 4 + (k & 3)] =
          k < 9 ? round_tf32(w[n * 9 + k]) : 0.f;
    }
  }
  if (tid < CO) s_bias[tid] = bias[tid];
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp < 4 || warp >= 9) {
    // ---- producers: one im2col row per thread; the tile's input images arrive by bulk copy.
    // Two groups of 128 threads take alternate tiles, so two tiles are being built at any time.
    const int grp = warp < 4 ? 0 : 1;
    const int ptid = grp ? tid - 288 : tid;
    if (ptid == 0) {  // this group's first tiles: it = grp, grp + 2, ...
      for (int j = 0; j < kXStagesG; ++j) {
        const long long t = (long long)blockIdx.x + (long long)(grp + 2 * j) * gridDim.x;
        if (t < g.ntiles)
          x_stage_issue(x, g, static_cast<int>(t),
                        s_base + S::off_x + (grp * kXStagesG + j) * kXStageBytes,
                        bar_x + (grp * kXStagesG + j) * 8);
      }
    }
    for (int it = grp;; it += 2) {
      const long long tl = (long long)blockIdx.x + (long long)it * gridDim.x;
      if (tl >= g.ntiles) break;
      const int t = static_cast<int>(tl);
      const int stage = it % kC1Stages;
      const uint32_t phase = (it / kC1Stages) & 1;
      const int j = it >> 1;  // this group's tile counter
      const int xst = grp * kXStagesG + j % kXStagesG;
      int img, h, wc;
      const bool valid = decode_row(g, t * kTileM + ptid, img, h, wc);
      float tp[9];
#pragma unroll
      for (int i = 0; i < 9; ++i) tp[i] = 0.f;
      mbar_wait(bar_x + xst * 8, (j / kXStagesG) & 1);
      if (valid)
        taps_from_stage(reinterpret_cast<const float*>(smem + S::off_x + xst * kXStageBytes), g,
                        img - x_stage_first_image(g, t), h, wc, tp);
      producer_group_sync(grp);  // the whole group has read this x stage: refill it
      if (ptid == 0) {
        const long long tn = tl + 2ll * kXStagesG * gridDim.x;  // this group's tile j + kXStagesG
        if (tn < g.ntiles)
          x_stage_issue(x, g, static_cast<int>(tn), s_base + S::off_x + xst * kXStageBytes,
                        bar_x + xst * 8);
      }
      mbar_wait(bar_empty + stage * 8, phase ^ 1);
      uint8_t* a = smem + S::off_a + stage * S::a_stage + ptid * 16;
      float lo[9];
#pragma unroll
      for (int i = 0; i < 9; ++i) {
        const float hi = round_tf32(tp[i]);  // nearest, not the MMA's truncation
        lo[i] = round_tf32(tp[i] - hi);
        tp[i] = hi;
      }
      *reinterpret_cast<float4*>(a) = make_float4(tp[0], tp[1], tp[2], tp[3]);
      *reinterpret_cast<float4*>(a + S::a_plane) = make_float4(tp[4], tp[5], tp[6], tp[7]);
      if constexpr (X3) {
        *reinterpret_cast<float4*>(a + 2 * S::a_plane) = make_float4(tp[8], lo[0], lo[1], lo[2]);
        *reinterpret_cast<float4*>(a + 3 * S::a_plane) = make_float4(lo[3], lo[4], lo[5], lo[6]);
        *reinterpret_cast<float4*>(a + 4 * S::a_plane) = make_float4(lo[7], lo[8], 0.f, 0.f);
        *reinterpret_cast<float4*>(a + 5 * S::a_plane) = make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
        *reinterpret_cast<float4*>(a + 2 * S::a_plane) = make_float4(tp[8], 0.f, 0.f, 0.f);
        *reinterpret_cast<float4*>(a + 3 * S::a_plane) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      fence_proxy_async_smem();
      mbar_arrive(bar_full + stage * 8);
    }
  } else if (warp == 4) {
    constexpr uint32_t idesc = umma_idesc_tf32(kTileM, CO);
    int it = 0;
    for (int t = blockIdx.x; t < g.ntiles; t += gridDim.x, ++it) {
      const int stage = it % kC1Stages;
      const uint32_t phase = (it / kC1Stages) & 1;
      const int acc = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(bar_tempty + acc * 8, aphase ^ 1);
      mbar_wait(bar_full + stage * 8, phase);
      tc_fence_after();
      const uint32_t a_stage = s_a + stage * S::a_stage;
#pragma unroll
      for (int kk = 0; kk < (int)S::NKA / 2; ++kk) {
        const uint64_t adesc = umma_smem_desc(a_stage + kk * 2 * S::a_plane, S::a_plane, 128);
        const uint64_t bdesc = umma_smem_desc(s_base + kk * 2 * 512, 512, 128);
        if (elect_one()) umma_tf32(tmem_base + acc * CO, adesc, bdesc, idesc, kk != 0);
      }
      if constexpr (X3) {  // x_hi . w_lo
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          const uint64_t adesc = umma_smem_desc(a_stage + kk * 2 * S::a_plane, S::a_plane, 128);
          const uint64_t bdesc = umma_smem_desc(s_base + S::b1_bytes + kk * 2 * 512, 512, 128);
          if (elect_one()) umma_tf32(tmem_base + acc * CO, adesc, bdesc, idesc, 1);
        }
      }
      if (elect_one()) {
        umma_commit(bar_empty + stage * 8);
        umma_commit(bar_tfull + acc * 8);
      }
      __syncwarp();
    }
  } else {
    const int q = warp & 3;
    int it = 0;
    for (int t = blockIdx.x; t < g.ntiles; t += gridDim.x, ++it) {
      const int acc = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(bar_tfull + acc * 8, aphase);
      tc_fence_after();
      const int rel = t * kTileM + q * 32 + lane;
      int img, h, wc;
      const bool valid = decode_row(g, rel, img, h, wc);
      uint32_t v[32];
      tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * CO, v);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(bar_tempty + acc * 8);
      float xo[32];
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        const float t = __uint_as_float(v[c]) + s_bias[c];
        xo[c] = valid ? fmaxf(t, 0.f) : 0.f;
      }
      const long long row = (long long)kGuard + rel;
      if constexpr (P::kTf32) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          out[(long long)j * g.plane_rows + row] =
              make_uint4(__float_as_uint(round_tf32(xo[4 * j])), __float_as_uint(round_tf32(xo[4 * j + 1])),
                         __float_as_uint(round_tf32(xo[4 * j + 2])), __float_as_uint(round_tf32(xo[4 * j + 3])));
      } else if constexpr (P::kSets == 2) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float hi[8], lo[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) split_bf16(xo[8 * j + e], hi[e], lo[e]);
          out[(long long)j * g.plane_rows + row] =
              make_uint4(pack_bf16x2(hi[0], hi[1]), pack_bf16x2(hi[2], hi[3]), pack_bf16x2(hi[4], hi[5]),
                         pack_bf16x2(hi[6], hi[7]));
          out[(long long)(4 + j) * g.plane_rows + row] =
              make_uint4(pack_bf16x2(lo[0], lo[1]), pack_bf16x2(lo[2], lo[3]), pack_bf16x2(lo[4], lo[5]),
                         pack_bf16x2(lo[6], lo[7]));
        }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          out[(long long)j * g.plane_rows + row] =
              make_uint4(pack_bf16x2(xo[8 * j], xo[8 * j + 1]), pack_bf16x2(xo[8 * j + 2], xo[8 * j + 3]),
                         pack_bf16x2(xo[8 * j + 4], xo[8 * j + 5]), pack_bf16x2(xo[8 * j + 6], xo[8 * j + 7]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, 64);
}

// ------------------------------------------------------------------------------------- wgrad
// Accumulator [64 rows (32 real channels)][32 columns]:
//   columns 0..8 : sum dact * bf16_hi(x tap)     9 : sum dact (bias gradient)
//   columns 16..24: sum dact * bf16_lo(x tap)    others: zero
//
// Split planes: dact = hi + lo.  The four lo planes sit where the bf16 layout has its four zero planes
// (rows 32..63 of the M = 64 operand): the MMA loop runs twice per tile, once with the A descriptor on
// the hi planes and once on the lo planes, accumulating into rows 0..31 both times.  Rows 32..63 then
// see whatever follows in shared memory — accumulator rows are independent and those are never read.
template <class P>
struct C1WgradSmem {
  static_assert(!P::kTf32, "no MN-major tf32 operand in the no-swizzle layout");
  static constexpr uint32_t a_plane = kTileM * 16;   // 2048
  static constexpr uint32_t a_real = 4;              // planes that hold the 32 real channels
  static constexpr uint32_t a_bytes = 8 * a_plane;   // bf16: planes 4..7 stay zero; split: the lo planes
  static constexpr uint32_t b_planes = 4;            // N = 32 columns
  static constexpr uint32_t b_bytes = b_planes * a_plane;
  static constexpr uint32_t stage = a_bytes + b_bytes;
  static constexpr uint32_t off_x = kC1Stages * stage;
  static constexpr uint32_t off_bar = off_x + kXStages * kXStageBytes;
  static constexpr uint32_t off_tmem = off_bar + (3 * kC1Stages + 1 + kXStages) * 8;
  static constexpr uint32_t total = off_tmem + 16;
};

template <class P>
__global__ void __launch_bounds__(kC1Threads, 1)
conv1_wgrad_tc_kernel(const uint4* __restrict__ dact, const float* __restrict__ x,
                      float* __restrict__ partial, Geom g) {
  using S = C1WgradSmem<P>;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t bar_fa = s_base + S::off_bar;            // dact planes landed (TMA tx)
  const uint32_t bar_fb = bar_fa + kC1Stages * 8;         // im2col rows written (128 arrivals)
  const uint32_t bar_empty = bar_fb + kC1Stages * 8;
  const uint32_t bar_done = bar_empty + kC1Stages * 8;
  const uint32_t bar_x = bar_done + 8;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + S::off_tmem);

  for (int i = tid; i < (int)(kC1Stages * S::stage / 16); i += kC1Threads)
    reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (tid == 0) {
    for (int i = 0; i < kC1Stages; ++i) {
      mbar_init(bar_fa + i * 8, 1);
      mbar_init(bar_fb + i * 8, 128);
      mbar_init(bar_empty + i * 8, 1);
    }
    mbar_init(bar_done, 1);
    for (int i = 0; i < kXStages; ++i) mbar_init(bar_x + i * 8, 1);
    fence_mbar_init();
  }
  if (warp == 4) {
    tmem_alloc(smem_u32(s_tmem), 32);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  pdl_sync();

  if (warp != 4) {
    // ---- producers: TWO groups of 128 threads (warps 0-3 and 5-8) take alternate tiles, so two
    // tiles' im2col rows are being built at any time; group 1 also runs the epilogue at the end.
    const int grp = warp < 4 ? 0 : 1;
    const int ptid = grp ? tid - 160 : tid;
    if (ptid == 0) {  // this group's first tiles: it = grp, grp + 2, ...
      for (int j = 0; j < kXStagesG; ++j) {
        const long long t = (long long)blockIdx.x + (long long)(grp + 2 * j) * gridDim.x;
        if (t < g.ntiles)
          x_stage_issue(x, g, static_cast<int>(t),
                        s_base + S::off_x + (grp * kXStagesG + j) * kXStageBytes,
                        bar_x + (grp * kXStagesG + j) * 8);
      }
    }
    for (int it = grp;; it += 2) {
      const long long tl = (long long)blockIdx.x + (long long)it * gridDim.x;
      if (tl >= g.ntiles) break;
      const int t = static_cast<int>(tl);
      const int stage = it % kC1Stages;
      const uint32_t phase = (it / kC1Stages) & 1;
      const int j = it >> 1;  // this group's tile counter
      const int xst = grp * kXStagesG + j % kXStagesG;
      int img, h, wc;
      const bool valid = decode_row(g, t * kTileM + ptid, img, h, wc);
      float tp[9];
#pragma unroll
      for (int i = 0; i < 9; ++i) tp[i] = 0.f;
      mbar_wait(bar_x + xst * 8, (j / kXStagesG) & 1);
      if (valid)
        taps_from_stage(reinterpret_cast<const float*>(smem + S::off_x + xst * kXStageBytes), g,
                        img - x_stage_first_image(g, t), h, wc, tp);
      producer_group_sync(grp);  // the whole group has read this x stage: refill it
      if (ptid == 0) {
        const long long tn = tl + 2ll * kXStagesG * gridDim.x;  // this group's tile j + kXStagesG
        if (tn < g.ntiles)
          x_stage_issue(x, g, static_cast<int>(tn), s_base + S::off_x + xst * kXStageBytes,
                        bar_x + xst * 8);
      }
      // x = hi + lo with both terms bf16; column 9 of the hi block = ones -> bias gradient
      uint32_t hi[5], lo[5];
      float hf[9], lf[9];
#pragma unroll
      for (int i = 0; i < 9; ++i) split_bf16(tp[i], hf[i], lf[i]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        hi[i] = pack_bf16x2(hf[2 * i], hf[2 * i + 1]);
        lo[i] = pack_bf16x2(lf[2 * i], lf[2 * i + 1]);
      }
      hi[4] = pack_bf16x2(hf[8], 1.0f);
      lo[4] = pack_bf16x2(lf[8], 0.f);
      mbar_wait(bar_empty + stage * 8, phase ^ 1);
      if (ptid == 0) {
        constexpr int NP = P::kSets * 4;  // hi planes (and, split, the lo planes right after them)
        mbar_arrive_expect_tx(bar_fa + stage * 8, NP * S::a_plane);
        const long long rowa = (long long)kGuard + (long long)t * kTileM;
#pragma unroll
        for (int p = 0; p < NP; ++p)
          bulk_g2s(s_base + stage * S::stage + p * S::a_plane,
                   dact + ((long long)p * g.plane_rows + rowa), S::a_plane, bar_fa + stage * 8);
      }
      uint8_t* b = smem + stage * S::stage + S::a_bytes + ptid * 16;
      *reinterpret_cast<uint4*>(b) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
      *reinterpret_cast<uint4*>(b + S::a_plane) = make_uint4(hi[4], 0, 0, 0);
      *reinterpret_cast<uint4*>(b + 2 * S::a_plane) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
      *reinterpret_cast<uint4*>(b + 3 * S::a_plane) = make_uint4(lo[4], 0, 0, 0);
      fence_proxy_async_smem();
      mbar_arrive(bar_fb + stage * 8);
    }
    if (grp == 1) {
      // ---- epilogue (after this group's last tile): accumulator -> this CTA's partial
      const int q = warp & 3;
      mbar_wait(bar_done, 0);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16), v);
      tmem_ld_wait();
      // M = 64: row m lives in lane (m % 16) + 32 * (m / 16); channels 0..31 are rows 0..31
      if (q < 2 && lane < 16) {
        const int c = q * 16 + lane;
        float* dst = partial + (long long)blockIdx.x * (32 * 9 + 32);
#pragma unroll
        for (int tap = 0; tap < 9; ++tap)
          dst[c * 9 + tap] = __uint_as_float(v[tap]) + __uint_as_float(v[16 + tap]);
        dst[32 * 9 + c] = __uint_as_float(v[9]);
      }
    }
  } else {
    constexpr uint32_t idesc = umma_idesc_bf16(64, 32, 1, 1);
    int it = 0;
    for (int t = blockIdx.x; t < g.ntiles; t += gridDim.x, ++it) {
      const int stage = it % kC1Stages;
      const uint32_t phase = (it / kC1Stages) & 1;
      mbar_wait(bar_fa + stage * 8, phase);
      mbar_wait(bar_fb + stage * 8, phase);
      tc_fence_after();
      const uint32_t sa = s_base + stage * S::stage;
      const uint32_t sb = sa + S::a_bytes;
#pragma unroll
      for (int set = 0; set < P::kSets; ++set) {  // dact hi, then (split) dact lo
#pragma unroll
        for (int ks = 0; ks < kTileM / 16; ++ks) {
          const uint64_t adesc = umma_smem_desc(sa + set * 4 * S::a_plane + ks * 256, 128, S::a_plane);
          const uint64_t bdesc = umma_smem_desc(sb + ks * 256, 128, S::a_plane);
          if (elect_one()) umma_bf16(tmem_base, adesc, bdesc, idesc, (it | ks | set) != 0);
        }
      }
      if (elect_one()) umma_commit(bar_empty + stage * 8);
      __syncwarp();
    }
    if (elect_one()) umma_commit(bar_done);
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, 32);
}

}  // namespace ddpb

using namespace ddpb;

// Both kernels are latency-bound with one 9-warp CTA per SM (ncu: issue slots 19-33 % busy, 2.25
// active warps per scheduler); their shared memory (51 / 98 KB) and TMEM (64 / 32 columns) leave
// room for two co-resident CTAs, which doubles the loads in flight.  DDPB_C1_CTAS=1 restores one.
constexpr int kC1MaxCtasPerSm = 2;
static int c1_ctas_per_sm() {
  static int v = 0;
  if (v == 0) {
    const char* e = getenv("DDPB_C1_CTAS");
    v = (e && e[0] == '1') ? 1 : kC1MaxCtasPerSm;
  }
  return v;
}
// Two CTAs per SM only when every CTA still gets a few tiles (measured: B=4096 step 1.243 ->
// 1.122 ms, but at B=64 = 450 tiles the extra start-up cost makes the step slightly slower).
static int c1_grid(int ntiles, bool allow_two) {
  const int sms = ddpb_num_sms();
  const int per_sm = (allow_two && ntiles >= 4 * kC1MaxCtasPerSm * sms) ? c1_ctas_per_sm() : 1;
  const int slots = sms * per_sm;
  return ntiles < slots ? ntiles : slots;
}

template <class P>
static int conv1_fwd_tc_launch(const float* x, const float* w, const float* b, void* out, int cin,
                               int cout, int batch, int batch_cap, int H, int W, void* stream) {
  if (!(cin == 1 && cout == 32)) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  if (H * W * 4 * 2 > (int)kXStageBytes || (H * W * 4) % 16 != 0) return DDPB_ERR_SHAPE;
  static bool configured[64] = {};
  int dev = 0;
  DDPB_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    DDPB_CUDA_OK(cudaFuncSetAttribute(conv1_fwd_tc_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C1FwdSmem<P::kSets == 2>::total));
    configured[dev & 63] = true;
  }
  const int grid = c1_grid(g.ntiles, true);
  return static_cast<int>(launch_k(conv1_fwd_tc_kernel<P>, dim3(grid), dim3(kC1FwdThreads),
                                   C1FwdSmem<P::kSets == 2>::total, static_cast<cudaStream_t>(stream), x, w, b,
                                   static_cast<uint4*>(out), g));
}

template <class P>
static int conv1_wgrad_tc_launch(const void* dact, const float* x, float* scratch, int* nparts, int cin,
                                 int cout, int batch, int batch_cap, int H, int W, void* stream) {
  if (!(cin == 1 && cout == 32)) return DDPB_ERR_SHAPE;
  if (H * W * 4 * 2 > (int)kXStageBytes || (H * W * 4) % 16 != 0) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  using S = C1WgradSmem<P>;
  static bool configured[64] = {};
  int dev = 0;
  DDPB_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    DDPB_CUDA_OK(cudaFuncSetAttribute(conv1_wgrad_tc_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      S::total));
    configured[dev & 63] = true;
  }
  const int grid = c1_grid(g.ntiles, 2 * S::total <= 220 * 1024);
  DDPB_CUDA_OK(launch_k(conv1_wgrad_tc_kernel<P>, dim3(grid), dim3(kC1Threads), S::total,
                        static_cast<cudaStream_t>(stream), static_cast<const uint4*>(dact), x, scratch, g));
  *nparts = grid;
  return 0;
}

extern "C" {

int ddpb_reduce_partials(const float* part, int nparts, long long n, float* out, void* stream);

int ddpb_conv1_fwd_tc(const float* x, const float* w, const float* b, void* out, int cin, int cout,
                      int batch, int batch_cap, int H, int W, void* stream) {
  return conv1_fwd_tc_launch<Bf16Planes>(x, w, b, out, cin, cout, batch, batch_cap, H, W, stream);
}
// Same with split output planes (hi + lo bf16: the reference-precision mode).
int ddpb_conv1_fwd_tc_split(const float* x, const float* w, const float* b, void* out, int cin, int cout,
                            int batch, int batch_cap, int H, int W, void* stream) {
  return conv1_fwd_tc_launch<SplitPlanes>(x, w, b, out, cin, cout, batch, batch_cap, H, W, stream);
}
// Same, fp32 output planes (values rounded to tf32).
int ddpb_conv1_fwd_tc_f32(const float* x, const float* w, const float* b, void* out, int cin, int cout,
                          int batch, int batch_cap, int H, int W, void* stream) {
  return conv1_fwd_tc_launch<F32Planes>(x, w, b, out, cin, cout, batch, batch_cap, H, W, stream);
}

size_t ddpb_conv1_wgrad_tc_scratch_bytes(int cin, int cout) {
  return static_cast<size_t>(ddpb_num_sms()) * kC1MaxCtasPerSm * (cout * cin * 9 + cout) * sizeof(float);
}

// Partials only: scratch[nparts][cout*cin*9 + cout]; *nparts_out (host) receives the partial count.
// Feed them to ddpb_reduce_allreduce_sgd (fused split-K reduce + all-reduce + SGD).
int ddpb_conv1_wgrad_tc_partials(const void* dact, const float* x, float* scratch, int* nparts_out,
                                 int cin, int cout, int batch, int batch_cap, int H, int W,
                                 void* stream) {
  return conv1_wgrad_tc_launch<Bf16Planes>(dact, x, scratch, nparts_out, cin, cout, batch, batch_cap, H,
                                           W, stream);
}
int ddpb_conv1_wgrad_tc_split_partials(const void* dact, const float* x, float* scratch, int* nparts_out,
                                     int cin, int cout, int batch, int batch_cap, int H, int W,
                                     void* stream) {
  return conv1_wgrad_tc_launch<SplitPlanes>(dact, x, scratch, nparts_out, cin, cout, batch, batch_cap, H,
                                          W, stream);
}

// dw_db: cout*cin*9 weight gradients immediately followed by cout bias gradients.
int ddpb_conv1_wgrad_tc(const void* dact, const float* x, float* dw_db, float* scratch, int cin,
                        int cout, int batch, int batch_cap, int H, int W, void* stream) {
  int grid = 0;
  const int rc = conv1_wgrad_tc_launch<Bf16Planes>(dact, x, scratch, &grid, cin, cout, batch, batch_cap,
                                                   H, W, stream);
  if (rc != 0) return rc;
  return ddpb_reduce_partials(scratch, grid, (long long)cout * cin * 9 + cout, dw_db, stream);
}
int ddpb_conv1_wgrad_tc_split(const void* dact, const float* x, float* dw_db, float* scratch, int cin,
                            int cout, int batch, int batch_cap, int H, int W, void* stream) {
  int grid = 0;
  const int rc = conv1_wgrad_tc_launch<SplitPlanes>(dact, x, scratch, &grid, cin, cout, batch, batch_cap,
                                                  H, W, stream);
  if (rc != 0) return rc;
  return ddpb_reduce_partials(scratch, grid, (long long)cout * cin * 9 + cout, dw_db, stream);
}

}  // extern "C"
