// 3x3 / pad-1 convolution forward, data-gradient and weight-gradient as implicit GEMMs on the
// 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM), operands fetched by 1-D bulk
// async copies (TMA) from the padded planar-chunk layout described in layout.cuh.
//
// Replaces what the reference delegates to cuDNN + ATen for its second convolution:
//   forward   conv2 + bias + ReLU          reference model.py:11-12   (SURVEY.md K3,K4)
//   backward  ReLU' + dgrad + wgrad + db   reference train_ddp.py:199 (SURVEY.md K10,K11)
//
// Warp roles (192 threads, one CTA per SM, persistent over tiles):
//   warp 0    bulk-copy producer (one lane)
//   warp 1    TMEM allocator + MMA issuer (one lane issues tcgen05.mma)
//   warps 2-9 epilogue: tcgen05.ld the accumulator tile, fuse bias/ReLU/mask, store the planes
//             (two warps per TMEM lane quadrant, each takes half of the columns: the epilogue,
//             not the MMA, was the per-tile bottleneck with four warps — profiles/)
//
// Every kernel is a template over the plane formats of layout.cuh.  The reference-precision mode
// runs them on split planes (x = hi + lo, three MMAs per K step, "bf16x3"): at least the operand
// precision of the TF32 convolutions cuDNN runs for the reference, fp32 accumulation.
#include <stdlib.h>

#include "ptx.cuh"
#include "layout.cuh"
#include "api_internal.cuh"
#include "launch.cuh"

namespace ddpb {

constexpr int kConvThreads = 320;   // fwd / dgrad: 1 producer + 1 MMA + 8 epilogue warps
constexpr int kWgradThreads = 192;  // wgrad: 1 producer + 1 MMA + 4 epilogue warps
constexpr uint32_t kRowBytes = 16;
constexpr uint32_t kStagePlaneBytes = kStageRows * kRowBytes;  // 3072

template <class P, int CI, int CO, int STAGES>
struct ConvSmem {
  static constexpr uint32_t KP = CI / P::kCh;      // planes per set
  static constexpr uint32_t NPL = P::kSets * KP;   // planes staged per tile (hi set, then lo set)
  static constexpr bool kPrepacked = P::kTf32 || P::kSets == 2;
  // B tile of one tap: [KP k-chunks][CO rows][16 B].  Plain bf16: built per CTA from the fp32 master
  // weights; the k-chunk stride (UMMA LBO) is skewed by 16 B so that the conversion writes shared
  // memory without bank conflicts.  Split / tf32: the tile images (one per set) are pre-packed in
  // global memory (pack kernels below / the fused bucket update) and arrive with one bulk copy.
  // Split: the hi and lo rows of a K chunk are adjacent ([W_hi | W_lo] along N), so ONE MMA with
  // N = 2*CO multiplies an operand by both terms; w_rows = B rows per K chunk.
  static constexpr uint32_t w_rows = (P::kSets == 2 ? 2 : 1) * CO;
  static constexpr uint32_t w_lbo = w_rows * 16 + (kPrepacked ? 0 : 16);
  static constexpr uint32_t w_tap_bytes = KP * w_lbo;
  static constexpr uint32_t w_img_bytes = 9 * w_tap_bytes;   // whole image (split: both sets)
  static constexpr uint32_t w_bytes = (w_img_bytes + 127) / 128 * 128;
  static constexpr uint32_t acc_cols = w_rows;               // accumulator columns per TMEM buffer
  static constexpr uint32_t a_stage_bytes = NPL * kStagePlaneBytes;
  static constexpr uint32_t off_a = w_bytes;
  // plain bf16: fp32 master weights are bulk-copied (TMA, one copy) into a staging area and converted
  // from there.  It aliases the A-stage ring when that is large enough (the ring is idle during setup).
  static constexpr uint32_t wstage_bytes = kPrepacked ? 0 : 9 * CI * CO * 4;
  static constexpr uint32_t ring_bytes = STAGES * a_stage_bytes;
  static constexpr bool stage_aliases_ring = ring_bytes >= wstage_bytes;
  static constexpr uint32_t off_wstage = stage_aliases_ring ? off_a : off_a + ring_bytes;
  static constexpr uint32_t off_bias = off_a + ring_bytes + (stage_aliases_ring ? 0 : wstage_bytes);
  static constexpr uint32_t off_bar = off_bias + CO * 4;
  static constexpr uint32_t off_tmem = off_bar + (2 * STAGES + 5) * 8;
  static constexpr uint32_t total = off_tmem + 16;
  static_assert(total <= 232448, "conv CTA exceeds the 227 KB shared-memory limit");
};

// One MMA of the format's kind (bf16 / split: kind::f16, K = 16; fp32 planes: kind::tf32, K = 8).
template <class P>
__device__ __forceinline__ void umma_p(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                       uint32_t accumulate) {
  if constexpr (P::kTf32) umma_tf32(d_tmem, adesc, bdesc, idesc, accumulate);
  else umma_bf16(d_tmem, adesc, bdesc, idesc, accumulate);
}
template <class P>
__host__ __device__ constexpr uint32_t umma_idesc_p(int M, int N, int a_mn, int b_mn) {
  return P::kTf32 ? umma_idesc_tf32(M, N, a_mn, b_mn) : umma_idesc_bf16(M, N, a_mn, b_mn);
}

// Store CH*NV consecutive channels (starting at channel c0, a multiple of 8) of one pixel row in
// the output format PO.  `nch` = channels of the whole tensor (locates the lo plane set).
template <class PO, int N>
__device__ __forceinline__ void store_channels(uint4* __restrict__ out, long long plane_rows, long long row,
                                               int c0, int nch, const float (&x)[N], int round_out) {
  if constexpr (PO::kTf32) {
#pragma unroll
    for (int j = 0; j < N / 4; ++j) {
      float v[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) v[e] = round_out ? round_tf32(x[4 * j + e]) : x[4 * j + e];
      out[(long long)(c0 / 4 + j) * plane_rows + row] =
          make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3]));
    }
  } else if constexpr (PO::kSets == 2) {
#pragma unroll
    for (int j = 0; j < N / 8; ++j) {
      float hi[8], lo[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) split_bf16(x[8 * j + e], hi[e], lo[e]);
      out[(long long)(c0 / 8 + j) * plane_rows + row] =
          make_uint4(pack_bf16x2(hi[0], hi[1]), pack_bf16x2(hi[2], hi[3]), pack_bf16x2(hi[4], hi[5]),
                     pack_bf16x2(hi[6], hi[7]));
      out[(long long)(nch / 8 + c0 / 8 + j) * plane_rows + row] =
          make_uint4(pack_bf16x2(lo[0], lo[1]), pack_bf16x2(lo[2], lo[3]), pack_bf16x2(lo[4], lo[5]),
                     pack_bf16x2(lo[6], lo[7]));
    }
  } else {
#pragma unroll
    for (int j = 0; j < N / 8; ++j)
      out[(long long)(c0 / 8 + j) * plane_rows + row] =
          make_uint4(pack_bf16x2(x[8 * j], x[8 * j + 1]), pack_bf16x2(x[8 * j + 2], x[8 * j + 3]),
                     pack_bf16x2(x[8 * j + 4], x[8 * j + 5]), pack_bf16x2(x[8 * j + 6], x[8 * j + 7]));
  }
}

// Optional in-kernel phase timestamps (SM clock) for bring-up/tuning: prof[block*16 + slot].
__device__ __forceinline__ void prof_stamp(unsigned long long* prof, int slot) {
  if (prof) prof[blockIdx.x * 16 + slot] = clock64();
}

constexpr uint32_t pow2_cols(uint32_t c) {
  return c <= 32 ? 32 : c <= 64 ? 64 : c <= 128 ? 128 : c <= 256 ? 256 : 512;
}

// MODE 0: out = relu(conv(in, w) + bias)                 (GEMM K = CI input channels, N = CO)
// MODE 1: out = relu'(mask) * conv_transpose(in, w)      (in = grad wrt conv output with CI
//         channels, w = the layer's weight [CI][CO][3][3], out = grad wrt conv input, CO chans)
// Planes are addressed as 16-byte rows (uint4) whatever the format.  P = format of `in` (and of the
// tensor-core operands), PO = format of `out`, PM = format of `mask` (MODE 1).  For split / fp32
// input planes `w` is the pre-packed tile image of this MODE (ddpb_conv3x3_pack_split / _tf32);
// `round_out` != 0 rounds an fp32 result to tf32 (for when the next consumer is a tf32 GEMM).
template <class P, class PO, class PM, int CI, int CO, int MODE, int STAGES>
__global__ void __launch_bounds__(kConvThreads, 1)
conv3x3_tc_kernel(const uint4* __restrict__ in, const float* __restrict__ w,
                  const float* __restrict__ bias, const uint4* __restrict__ mask,
                  uint4* __restrict__ out, Geom g, int round_out, unsigned long long* prof) {
  using S = ConvSmem<P, CI, CO, STAGES>;
  constexpr uint32_t KP = CI / P::kCh;
  constexpr uint32_t TMEM_COLS = pow2_cols(2 * S::acc_cols);
  extern __shared__ __align__(128) uint8_t smem[];

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t s_w = s_base;
  const uint32_t s_a = s_base + S::off_a;
  float* s_bias = reinterpret_cast<float*>(smem + S::off_bias);
  const uint32_t bar_full = s_base + S::off_bar;
  const uint32_t bar_empty = bar_full + STAGES * 8;
  const uint32_t bar_tfull = bar_empty + STAGES * 8;
  const uint32_t bar_tempty = bar_tfull + 16;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + S::off_tmem);

  const uint32_t bar_w = bar_tempty + 16;
  if (tid == 0) prof_stamp(prof, 0);
  if (tid == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(bar_full + i * 8, 1);
      mbar_init(bar_empty + i * 8, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_tfull + i * 8, 1);
      mbar_init(bar_tempty + i * 8, 256);
    }
    mbar_init(bar_w, 1);
    fence_mbar_init();
    if constexpr (S::kPrepacked) {
      // pre-packed tile images (all sets): one bulk async copy straight into the B-tile area
      mbar_arrive_expect_tx(bar_w, S::w_img_bytes);
      bulk_g2s(s_w, w, S::w_img_bytes, bar_w);
    } else {
      // fp32 master weights: one bulk async copy into the staging area
      mbar_arrive_expect_tx(bar_w, S::wstage_bytes);
      bulk_g2s(s_base + S::off_wstage, w, S::wstage_bytes, bar_w);
    }
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(s_tmem), TMEM_COLS);
    tmem_relinquish();
  }
  if (MODE == 0)
    for (int i = tid; i < CO; i += kConvThreads) s_bias[i] = bias[i];
  __syncthreads();  // barrier inits visible to everyone
  mbar_wait(bar_w, 0);
  if (tid == 0) prof_stamp(prof, 1);
  // ---- staged fp32 weights -> bf16 UMMA canonical (K-major, no swizzle) B tiles.
  // B operand of tap t: rows n (GEMM N), 8-element K chunks; byte offset
  //   t*w_tap + (k/8)*w_lbo + n*16 + (k%8)*2          => LBO = w_lbo, SBO = 128
  // Thread->element maps are chosen so that both the staged reads and the tile writes of a warp
  // fall into 32 different banks.
  if constexpr (!S::kPrepacked) {
    const float* ws = reinterpret_cast<const float*>(smem + S::off_wstage);
    if (MODE == 0) {
      // source w[n][k][tap] (tap fastest, then k): a warp covers 16 k-pairs x 2 taps of one n
      constexpr int kPairs = CI / 2;
      constexpr int total = CO * 5 * kPairs * 2;  // taps handled in pairs (0,1)..(8,9): 9 is a dummy
      for (int idx = tid; idx < total; idx += kConvThreads) {
        const int tl = idx & 1;
        const int kp = (idx >> 1) % kPairs;
        const int rest = (idx >> 1) / kPairs;
        const int tp = rest % 5;
        const int n = rest / 5;
        const int tap = tp * 2 + tl;
        if (tap >= 9) continue;
        const int k = kp * 2;
        const float lo = ws[(n * CI + k) * 9 + tap];
        const float hi = ws[(n * CI + k + 1) * 9 + tap];
        *reinterpret_cast<uint32_t*>(smem + tap * S::w_tap_bytes + (k >> 3) * S::w_lbo + n * 16 +
                                     (k & 7) * 2) = pack_bf16x2(lo, hi);
      }
    } else {
      // source w[k = layer out][n = layer in][8 - tap] (taps mirrored): a warp covers 32 n of one
      // (tap, k-chunk) and each thread writes a whole 16-byte row (8 k values)
      constexpr int total = 9 * KP * CO;
      for (int idx = tid; idx < total; idx += kConvThreads) {
        const int n = idx % CO;
        const int rest = idx / CO;
        const int kc = rest % KP;
        const int tap = rest / KP;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = ws[((kc * 8 + i) * CO + n) * 9 + (8 - tap)];
        *reinterpret_cast<uint4*>(smem + tap * S::w_tap_bytes + kc * S::w_lbo + n * 16) =
            make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                       pack_bf16x2(v[6], v[7]));
      }
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  if (tid == 0) prof_stamp(prof, 2);
  // Everything above touched only this CTA's shared memory / TMEM and the layer's weights (last
  // written by the previous step's update, complete before our stream predecessor even started).
  // From here on we read what the predecessor wrote.
  pdl_sync();

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int t = blockIdx.x; t < g.ntiles; t += gridDim.x, ++it) {
        const int stage = it % STAGES;
        const uint32_t phase = (it / STAGES) & 1;
        mbar_wait(bar_empty + stage * 8, phase ^ 1);
        mbar_arrive_expect_tx(bar_full + stage * 8, S::a_stage_bytes);
        const long long row0 = (long long)kGuard + (long long)t * kTileM - kHalo;
#pragma unroll
        for (uint32_t p = 0; p < S::NPL; ++p)  // hi planes, then (split) lo planes
          bulk_g2s(s_a + stage * S::a_stage_bytes + p * kStagePlaneBytes,
                   in + ((long long)p * g.plane_rows + row0), kStagePlaneBytes,
                   bar_full + stage * 8);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // MMA issuer: the whole warp stays converged, one elected lane issues (see elect_one()).
    constexpr uint32_t idesc = umma_idesc_p<P>(kTileM, CO, 0, 0);
    constexpr uint32_t idesc2 = umma_idesc_p<P>(kTileM, S::acc_cols, 0, 0);  // split: N = 2*CO
    (void)idesc2;
    uint32_t tap_off[9];
#pragma unroll
    for (int tap = 0; tap < 9; ++tap)
      tap_off[tap] = (kHalo + (tap / 3 - 1) * g.PW + (tap % 3 - 1)) * kRowBytes;
    int it = 0;
    for (int t = blockIdx.x; t < g.ntiles; t += gridDim.x, ++it) {
      const int stage = it % STAGES;
      const uint32_t phase = (it / STAGES) & 1;
      const int acc = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(bar_tempty + acc * 8, aphase ^ 1);
      mbar_wait(bar_full + stage * 8, phase);
      tc_fence_after();
      if (it == 0 && lane == 0) prof_stamp(prof, 4);
      const uint32_t a_stage = s_a + stage * S::a_stage_bytes;
      const uint32_t d_tmem = tmem_base + acc * S::acc_cols;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
#pragma unroll
        for (int kk = 0; kk < CI / P::kMmaK; ++kk) {
          const uint64_t adesc = umma_smem_desc(a_stage + tap_off[tap] + kk * 2 * kStagePlaneBytes,
                                                kStagePlaneBytes, 128);
          const uint64_t bdesc =
              umma_smem_desc(s_w + tap * S::w_tap_bytes + kk * 2 * S::w_lbo, S::w_lbo, 128);
          if constexpr (P::kSets == 2) {
            // split operands, two MMAs per K step instead of three: A_hi x [W_hi | W_lo] (N = 2*CO:
            // columns [0,CO) += hi.hi, [CO,2CO) += hi.lo) and A_lo x W_hi (N = CO: columns [0,CO) +=
            // lo.hi); the epilogue adds the two column halves.  (A tcgen05.mma costs ~max(45, N/2)
            // cycles whatever M is, so the wider N is nearly free.)
            const uint64_t adesc_lo =
                umma_smem_desc(a_stage + KP * kStagePlaneBytes + tap_off[tap] + kk * 2 * kStagePlaneBytes,
                               kStagePlaneBytes, 128);
            if (elect_one()) {
              umma_p<P>(d_tmem, adesc, bdesc, idesc2, (tap | kk) != 0);
              umma_p<P>(d_tmem, adesc_lo, bdesc, idesc, 1);
            }
          } else {
            if (elect_one()) umma_p<P>(d_tmem, adesc, bdesc, idesc, (tap | kk) != 0);
          }
        }
      }
      if (elect_one()) {
        umma_commit(bar_empty + stage * 8);
        umma_commit(bar_tfull + acc * 8);
      }
      __syncwarp();
    }
    if (lane == 0) prof_stamp(prof, 5);
  } else {
    const int q = warp & 3;            // TMEM lane quadrant this warp may read
    const int half = (warp - 2) >> 2;  // which half of the CO columns this warp handles
    constexpr int COLS = CO / 2;       // 32 (fwd) or 16 (dgrad)
    const int c0 = half * COLS;
    int it = 0;
    for (int t = blockIdx.x; t < g.ntiles; t += gridDim.x, ++it) {
      const int acc = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const int rel = t * kTileM + q * 32 + lane;
      int img, h, wc;
      const bool valid = decode_row(g, rel, img, h, wc);
      const long long row = (long long)kGuard + rel;
      constexpr int MCH = PM::kCh;        // mask: channels per 16-byte row (hi planes when split)
      constexpr int NMV = COLS / MCH;
      uint4 mk[NMV];
      if (MODE == 1) {  // issue the ReLU-mask loads before blocking on the accumulator
#pragma unroll
        for (int j = 0; j < NMV; ++j) mk[j] = mask[(long long)(c0 / MCH + j) * g.plane_rows + row];
      }
      mbar_wait(bar_tfull + acc * 8, aphase);
      tc_fence_after();
      if (it == 0 && tid == 64) prof_stamp(prof, 6);
      uint32_t v[COLS];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * S::acc_cols + c0;
      if constexpr (COLS == 32) tmem_ld32(taddr, v);
      else tmem_ld16(taddr, v);
      if constexpr (P::kSets == 2) {  // second column half: the hi.lo products
        uint32_t v2[COLS];
        if constexpr (COLS == 32) tmem_ld32(taddr + CO, v2);
        else tmem_ld16(taddr + CO, v2);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < COLS; ++c) v[c] = __float_as_uint(__uint_as_float(v[c]) + __uint_as_float(v2[c]));
      }
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(bar_tempty + acc * 8);  // accumulator is in registers: release it early
      float x[COLS];
      if (MODE == 0) {
#pragma unroll
        for (int c = 0; c < COLS; ++c) {
          const float t = __uint_as_float(v[c]) + s_bias[c0 + c];
          x[c] = valid ? fmaxf(t, 0.f) : 0.f;
        }
      } else {
        const uint32_t* mp = reinterpret_cast<const uint32_t*>(mk);
#pragma unroll
        for (int c = 0; c < COLS; ++c) {
          float m;  // sign of the forward activation (for split planes its hi term decides)
          if constexpr (PM::kTf32) m = __uint_as_float(mp[c]);
          else m = (c & 1) ? bf16_hi(mp[c >> 1]) : bf16_lo(mp[c >> 1]);
          x[c] = (valid && m > 0.f) ? __uint_as_float(v[c]) : 0.f;
        }
      }
      store_channels<PO, COLS>(out, g.plane_rows, row, c0, CO, x, round_out);
    }
    if (tid == 64) prof_stamp(prof, 7);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
  if (tid == 0) prof_stamp(prof, 8);
}

// ------------------------------------------------------------------------------------------
// Weight gradient:  dW[co][ci][tap] = sum_p dout[p][co] * in[p + d(tap)][ci],  db[co] = sum_p dout
// GEMM: M = CO (dout channels), K = pixels, both operands MN-major views of the planar buffers.
//
// Measured on B200 (tools/mma_bench.py): a tcgen05.mma costs max(~45, N/2) cycles whatever M is,
// so nine N=32 MMAs per K-step waste 3x.  Instead the in-planes of a tile are staged THREE times,
// shifted by -1/0/+1 rows (the three horizontal taps), as 3*CI/8 consecutive planes; one MMA
// with N = 3*CI then covers a whole filter row, and the vertical tap is the descriptor's start
// row.  A constant plane of ones appended after them makes the bias gradient 8 more columns of the
// first MMA (N = 3*CI + 8).  Accumulator columns: [row0: 3*CI taps | 8 bias][row1: 3*CI][row2: 3*CI].
// Each CTA accumulates its share of pixel tiles in TMEM and writes one partial; wgrad_reduce_kernel
// sums the partials in a fixed order (deterministic).
// Split planes: the tile's hi and lo sets are both staged and every K step issues three MMAs per
// filter row, (dout_hi, in_hi) + (dout_lo, in_hi) + (dout_hi, in_lo); the ones plane belongs to the
// hi set only, so the bias gradient is sum(dout_hi) + sum(dout_lo).
// (There is no fp32-plane instantiation: tcgen05 defines an MN-major tf32 operand only for the
// 128B_BASE32B swizzled layout, not for the no-swizzle layout of the planes — tools/umma_probe.cu.)
template <class P, int CI, int CO, int STAGES>
struct WgradSmem {
  static_assert(!P::kTf32, "no MN-major tf32 operand in the no-swizzle layout");
  static constexpr uint32_t AP = CO / P::kCh, BP = CI / P::kCh;
  static constexpr uint32_t a_plane = kTileM * kRowBytes;  // 2048
  static constexpr uint32_t a_set = AP * a_plane;
  static constexpr uint32_t a_bytes = P::kSets * a_set;
  static constexpr uint32_t b_set = (3 * BP + 1) * kStagePlaneBytes;  // 3 shifted copies + ones plane
  static constexpr uint32_t b_lo_set = 3 * BP * kStagePlaneBytes;     // the lo set has no ones plane
  static constexpr uint32_t b_bytes = b_set + (P::kSets == 2 ? b_lo_set : 0);
  static constexpr uint32_t stage_bytes = a_bytes + b_bytes;
  static constexpr uint32_t tx_bytes = a_bytes + P::kSets * 3 * BP * kStagePlaneBytes;
  static constexpr uint32_t off_bar = STAGES * stage_bytes;
  static constexpr uint32_t off_tmem = off_bar + (2 * STAGES + 1) * 8;
  static constexpr uint32_t total = off_tmem + 16;
  static constexpr uint32_t n0 = 3 * CI + 8;  // first filter row + bias block
  static constexpr uint32_t n1 = 3 * CI;
  static constexpr uint32_t ncols = n0 + 2 * n1;  // 9*CI + 8
  static_assert(total <= 232448, "wgrad CTA exceeds the 227 KB shared-memory limit");
};

template <class P, int CI, int CO, int STAGES>
__global__ void __launch_bounds__(kWgradThreads, 1)
conv3x3_wgrad_tc_kernel(const uint4* __restrict__ dout, const uint4* __restrict__ in,
                        float* __restrict__ partial, Geom g, int want_bias, unsigned long long* prof) {
  using S = WgradSmem<P, CI, CO, STAGES>;
  static_assert(CO == 64 || CO == 128, "UMMA M must be 64 or 128");
  static_assert(S::n0 <= 128, "one filter row must fit a 128-column TMEM allocation");
  // Three separate 128-column allocations (one per filter row) instead of one 512-column block:
  // 384 columns leave room for a concurrently resident dgrad CTA (64 columns) on the same SM.
  constexpr uint32_t TMEM_COLS = 128;
  extern __shared__ __align__(128) uint8_t smem[];

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t bar_full = s_base + S::off_bar;
  const uint32_t bar_empty = bar_full + STAGES * 8;
  const uint32_t bar_done = bar_empty + STAGES * 8;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + S::off_tmem);

  if (tid == 0) prof_stamp(prof, 0);
  // the constant ones plane of every stage (never touched by the bulk copies)
  for (int st = 0; st < STAGES; ++st) {
    uint32_t* ones = reinterpret_cast<uint32_t*>(smem + st * S::stage_bytes + S::a_bytes +
                                                 3 * S::BP * kStagePlaneBytes);
    for (int i = tid; i < (int)kStagePlaneBytes / 4; i += kWgradThreads) ones[i] = 0x3f803f80u;
  }
  if (tid == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(bar_full + i * 8, 1);
      mbar_init(bar_empty + i * 8, 1);
    }
    mbar_init(bar_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    for (int r = 0; r < 3; ++r) tmem_alloc(smem_u32(s_tmem + r), TMEM_COLS);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_row[3] = {s_tmem[0], s_tmem[1], s_tmem[2]};
  if (tid == 0) prof_stamp(prof, 2);
  pdl_sync();

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int t = blockIdx.x; t < g.ntiles; t += gridDim.x, ++it) {
        const int stage = it % STAGES;
        const uint32_t phase = (it / STAGES) & 1;
        mbar_wait(bar_empty + stage * 8, phase ^ 1);
        mbar_arrive_expect_tx(bar_full + stage * 8, S::tx_bytes);
        const uint32_t sa = s_base + stage * S::stage_bytes;
        const uint32_t sb = sa + S::a_bytes;
        const long long rowa = (long long)kGuard + (long long)t * kTileM;
#pragma unroll
        for (uint32_t p = 0; p < P::kSets * S::AP; ++p)  // hi planes, then (split) lo planes
          bulk_g2s(sa + p * S::a_plane, dout + ((long long)p * g.plane_rows + rowa), S::a_plane,
                   bar_full + stage * 8);
#pragma unroll
        for (int set = 0; set < P::kSets; ++set)
#pragma unroll
          for (int sh = 0; sh < 3; ++sh)  // horizontal tap s: rows shifted by s-1
#pragma unroll
            for (uint32_t p = 0; p < S::BP; ++p)
              bulk_g2s(sb + set * S::b_set + (sh * S::BP + p) * kStagePlaneBytes,
                       in + ((long long)(set * S::BP + p) * g.plane_rows + rowa - kHalo + (sh - 1)),
                       kStagePlaneBytes, bar_full + stage * 8);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    constexpr uint32_t idesc0 = umma_idesc_bf16(CO, S::n0, 1, 1);
    constexpr uint32_t idesc1 = umma_idesc_bf16(CO, S::n1, 1, 1);
    constexpr uint32_t idesc_s = umma_idesc_bf16(2 * CO, S::n1, 1, 1);  // split: stacked hi/lo rows
    constexpr uint32_t idesc_b = umma_idesc_bf16(2 * CO, 16, 1, 1);
    (void)idesc_s; (void)idesc_b; (void)idesc0; (void)idesc1;
    uint32_t row_off[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) row_off[r] = (kHalo + (r - 1) * g.PW) * kRowBytes;
    int it = 0;
    for (int t = blockIdx.x; t < g.ntiles; t += gridDim.x, ++it) {
      const int stage = it % STAGES;
      const uint32_t phase = (it / STAGES) & 1;
      mbar_wait(bar_full + stage * 8, phase);
      tc_fence_after();
      if (it == 0 && lane == 0) prof_stamp(prof, 4);
      const uint32_t sa = s_base + stage * S::stage_bytes;
      const uint32_t sb = sa + S::a_bytes;
#pragma unroll
      for (int ks = 0; ks < kTileM / 16; ++ks) {
        const uint32_t accum = (it | ks) != 0;
        const uint64_t adesc = umma_smem_desc(sa + ks * 16 * kRowBytes, 128, S::a_plane);
#pragma unroll
        for (int r = 0; r < 3; ++r) {  // vertical tap r: start row shifted by (r-1)*PW
          const uint64_t bdesc =
              umma_smem_desc(sb + row_off[r] + ks * 16 * kRowBytes, 128, kStagePlaneBytes);
          if constexpr (P::kSets == 2) {
            // The hi and lo plane sets of dout are adjacent in the stage, so ONE descriptor with
            // M = 2*CO = 128 spans [dout_hi ; dout_lo]: two MMAs per filter row (x in_hi, x in_lo) give
            // all four products — rows 0..63 accumulate hi.hi + hi.lo, rows 64..127 lo.hi + lo.lo — and
            // the epilogue adds the two row halves.  (M = 128 costs the same issue time as M = 64.)
            const uint64_t bdesc_lo =
                umma_smem_desc(sb + S::b_set + row_off[r] + ks * 16 * kRowBytes, 128, kStagePlaneBytes);
            if (elect_one()) {
              umma_bf16(tmem_row[r], adesc, bdesc, idesc_s, accum);
              umma_bf16(tmem_row[r], adesc, bdesc_lo, idesc_s, 1);
              if (r == 0 && want_bias) {
                // bias block: [dout_hi ; dout_lo] x ones, 16 columns at column n1 (the second 8 read
                // the plane after the ones plane and are never used)
                const uint64_t bdesc_1 = umma_smem_desc(sb + 3 * S::BP * kStagePlaneBytes + row_off[0] +
                                                        ks * 16 * kRowBytes, 128, kStagePlaneBytes);
                umma_bf16(tmem_row[0] + S::n1, adesc, bdesc_1, idesc_b, accum);
              }
            }
          } else {
            if (elect_one()) umma_bf16(tmem_row[r], adesc, bdesc, r == 0 ? idesc0 : idesc1, accum);
          }
        }
      }
      if (elect_one()) umma_commit(bar_empty + stage * 8);
      __syncwarp();
    }
    if (elect_one()) umma_commit(bar_done);
    __syncwarp();
    if (lane == 0) prof_stamp(prof, 5);
  } else {
    const int q = warp & 3;
    mbar_wait(bar_done, 0);
    tc_fence_after();
    if (tid == 64) prof_stamp(prof, 6);
    if constexpr (P::kSets == 2) {
      // M = 128: accumulator row m sits in TMEM lane m; rows m and m + 64 belong to output channel m.
      // Warps of lanes 64..127 park their rows in shared memory (the stage ring is idle now), the
      // warps of lanes 0..63 add them to their own and write the partial image.
      static_assert(CO == 64, "stacked hi/lo rows need 2*CO == 128");
      float* park = reinterpret_cast<float*>(smem);  // [64 rows][ncols]
      const int m = (q & 1) * 32 + lane;             // output channel of this thread's row
      const bool upper = q >= 2;
      float* dst = partial + ((long long)blockIdx.x * CO + m) * S::ncols;
      float* prow = park + m * S::ncols;
      for (int pass = 0; pass < 2; ++pass) {
        if ((pass == 0) == upper) {
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            const uint32_t tb = tmem_row[r] + (static_cast<uint32_t>(q * 32) << 16);
            const int cbase = r == 0 ? 0 : S::n0 + (r - 1) * S::n1;
            const int ncol = (r == 0 && want_bias) ? S::n1 + 8 : S::n1;
#pragma unroll 1
            for (int c0 = 0; c0 < ncol; c0 += 8) {
              uint32_t v[8];
              tmem_ld8(tb + c0, v);
              tmem_ld_wait();
              float4 a = make_float4(__uint_as_float(v[0]), __uint_as_float(v[1]), __uint_as_float(v[2]),
                                     __uint_as_float(v[3]));
              float4 b = make_float4(__uint_as_float(v[4]), __uint_as_float(v[5]), __uint_as_float(v[6]),
                                     __uint_as_float(v[7]));
              if (upper) {
                *reinterpret_cast<float4*>(prow + cbase + c0) = a;
                *reinterpret_cast<float4*>(prow + cbase + c0 + 4) = b;
              } else {
                const float4 pa = *reinterpret_cast<const float4*>(prow + cbase + c0);
                const float4 pb = *reinterpret_cast<const float4*>(prow + cbase + c0 + 4);
                a.x += pa.x; a.y += pa.y; a.z += pa.z; a.w += pa.w;
                b.x += pb.x; b.y += pb.y; b.z += pb.z; b.w += pb.w;
                *reinterpret_cast<float4*>(dst + cbase + c0) = a;
                *reinterpret_cast<float4*>(dst + cbase + c0 + 4) = b;
              }
            }
          }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps
      }
    } else {
    // M=64: accumulator row m sits in TMEM lane (m%16) + 32*(m/16); M=128: lane m.
    const int rows_per_q = CO / 4;
    const int m = q * rows_per_q + lane;
    const bool has_row = lane < rows_per_q;
    float* dst = partial + ((long long)blockIdx.x * CO + m) * S::ncols;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const uint32_t tb = tmem_row[r] + (static_cast<uint32_t>(q * 32) << 16);
      float* drow = dst + (r == 0 ? 0 : S::n0 + (r - 1) * S::n1);
#pragma unroll 1
      for (int c0 = 0; c0 < (int)S::n1; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tb + c0, v);
        tmem_ld_wait();
        if (has_row) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<uint4*>(drow + c0 + j * 4) =
                make_uint4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
        }
      }
      if (r == 0) {  // bias block: 8 columns after the first filter row
        uint32_t v[8];
        tmem_ld8(tb + S::n1, v);
        tmem_ld_wait();
        if (has_row) {
          *reinterpret_cast<uint4*>(drow + S::n1) = make_uint4(v[0], v[1], v[2], v[3]);
          *reinterpret_cast<uint4*>(drow + S::n1 + 4) = make_uint4(v[4], v[5], v[6], v[7]);
        }
      }
    }
    }
    if (tid == 64) prof_stamp(prof, 7);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1)
    for (int r = 0; r < 3; ++r) tmem_dealloc(tmem_row[r], TMEM_COLS);
  if (tid == 0) prof_stamp(prof, 8);
}

// Partial column layout -> dW[co][ci][tap], db[co]; see the accumulator layout above.
template <int CI, int CO>
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const float* __restrict__ partial, int nparts, float* __restrict__ dw,
                    float* __restrict__ db, const float* __restrict__ bias_part, int nsplit, int nbx,
                    int HW) {
  // 256 threads = 32 consecutive outputs x 8 partial groups; group pg sums parts pg, pg+8, ...
  constexpr int n0 = 3 * CI + 8, n1 = 3 * CI;
  constexpr int ncols = n0 + 2 * n1;
  constexpr int total = CO * ncols;
  __shared__ float sm[8][32];
  pdl_sync();
  const int lane = threadIdx.x & 31, pg = threadIdx.x >> 5;
  const int idx = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (idx < total) {
    int p = pg;
    for (; p + 24 < nparts; p += 32) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = partial[(long long)(p + 8 * u) * total + idx];
#pragma unroll
      for (int u = 0; u < 4; ++u) s += v[u];
    }
    for (; p < nparts; p += 8) s += partial[(long long)p * total + idx];
  }
  sm[pg][lane] = s;
  __syncthreads();
  if (pg != 0 || idx >= total) return;
#pragma unroll
  for (int u = 1; u < 8; ++u) s += sm[u][lane];
  const int co = idx / ncols;
  int col = idx - co * ncols;
  int r;
  if (col < n0) {
    r = 0;
    if (col >= n1) {  // bias block: only its first column is used
      // bias_part: the exact-fp32 sums of the linear backward replace the tensor-core column
      if (col == n1) db[co] = bias_part ? sum_bias_parts(bias_part, nsplit, nbx, HW, co) : s;
      return;
    }
  } else if (col < n0 + n1) {
    r = 1;
    col -= n0;
  } else {
    r = 2;
    col -= n0 + n1;
  }
  const int sx = col / CI;
  const int ci = col - sx * CI;
  dw[(co * CI + ci) * 9 + r * 3 + sx] = s;
}

// ------------------------------------------------------------------------------------------
// fp32 master weights w[CO_L][CI_L][3][3] -> the pre-packed B-tile images the split / fp32-plane
// kernels bulk-copy into shared memory (index maps: layout.cuh):
//   split : [fwd image | dgrad image], each 2*9*CI_L*CO_L bf16, hi and lo rows of a K chunk adjacent
//           (layout.cuh: conv3x3_{fwd,dgrad}_split_index; w = hi + lo)
//   tf32  : [fwd | dgrad], each 9*CI_L*CO_L floats rounded to nearest tf32
// The fused bucket update (comm.cu, MAP 1) rewrites the split images in place after every step; these
// kernels are for initialisation, checkpoint load and the non-fused update modes.
template <int CI_L, int CO_L>
__global__ void __launch_bounds__(256)
pack_conv3x3_split_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ packed) {
  pdl_sync();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  constexpr int n = CO_L * CI_L * 9;
  if (idx >= n) return;
  const int t = idx % 9;
  const int ci = (idx / 9) % CI_L;
  const int co = idx / (9 * CI_L);
  float hi, lo;
  split_bf16(w[idx], hi, lo);
  packed[conv3x3_fwd_split_index<CI_L, CO_L>(co, ci, t, 0)] = __float2bfloat16_rn(hi);
  packed[conv3x3_fwd_split_index<CI_L, CO_L>(co, ci, t, 1)] = __float2bfloat16_rn(lo);
  packed[2 * n + conv3x3_dgrad_split_index<CI_L, CO_L>(co, ci, t, 0)] = __float2bfloat16_rn(hi);
  packed[2 * n + conv3x3_dgrad_split_index<CI_L, CO_L>(co, ci, t, 1)] = __float2bfloat16_rn(lo);
}

template <int CI_L, int CO_L>
__global__ void __launch_bounds__(256)
pack_conv3x3_tf32_kernel(const float* __restrict__ w, float* __restrict__ fwd, float* __restrict__ dgr) {
  pdl_sync();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= CO_L * CI_L * 9) return;
  const int t = idx % 9;
  const int ci = (idx / 9) % CI_L;
  const int co = idx / (9 * CI_L);
  const float v = round_tf32(w[idx]);
  fwd[conv3x3_fwd_tile_index<CI_L, CO_L, 4>(co, ci, t)] = v;
  dgr[conv3x3_dgrad_tile_index<CI_L, CO_L, 4>(co, ci, t)] = v;
}

static unsigned long long* g_prof = nullptr;  // set by ddpb_set_prof_buffer (tuning only)

// Backward grids.  The split-plane data-gradient and weight-gradient CTAs each take ~220 KB of shared
// memory, so two of them cannot share an SM the way the bf16 ones do.  Launched on different streams
// with full grids they serialise; with each capped to about half of the SMs they run side by side on
// disjoint SMs.  DDPB_BWD_GRID (CTAs per kernel, 0 = all SMs) overrides the per-mode default.
static int bwd_grid_cap(int sms, bool split, int ntiles) {
  static int env = -1;
  if (env < 0) {
    const char* e = getenv("DDPB_BWD_GRID");
    env = e ? atoi(e) : 0;
    if (env < 0) env = 0;
  }
  if (env > 0) return env < sms ? env : sms;
  // Measured at batch 64 (450 tiles, B200): 148 CTAs each 110.1 us/step, 111: 103.9, 96: 104.6, 74: 108.0.
  // Three quarters of the SMs per kernel lets the two grids overlap partly and trims per-CTA set-up.
  if (split && ntiles <= 8 * sms) return sms * 3 / 4;
  return sms;
}

template <class KernT>
static int configure_smem(KernT kern, uint32_t bytes, bool (&configured)[64]) {
  int dev = 0;
  DDPB_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    DDPB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    configured[dev & 63] = true;
  }
  return 0;
}

template <class P, class PO, class PM, int CI, int CO, int MODE>
static int launch_conv(const void* in, const void* w, const float* bias, const void* mask,
                       void* out, const Geom& g, int round_out, int num_sms, cudaStream_t st) {
  // dgrad: 3 stages (bf16: its ring then exactly holds the fp32 weight staging area) keeps the bf16
  // CTA at ~111 KB so it can share an SM with a weight-gradient CTA running on another stream.
  constexpr int STAGES = MODE == 0 ? 4 : 3;
  using S = ConvSmem<P, CI, CO, STAGES>;
  auto kern = conv3x3_tc_kernel<P, PO, PM, CI, CO, MODE, STAGES>;
  static bool configured[64] = {};
  const int rc = configure_smem(kern, S::total, configured);
  if (rc != 0) return rc;
  const int cap = MODE == 1 ? bwd_grid_cap(num_sms, P::kSets == 2, g.ntiles) : num_sms;
  const int grid = g.ntiles < cap ? g.ntiles : cap;
  return static_cast<int>(launch_k(kern, dim3(grid), dim3(kConvThreads), S::total, st,
                                   static_cast<const uint4*>(in), static_cast<const float*>(w), bias,
                                   static_cast<const uint4*>(mask), static_cast<uint4*>(out), g,
                                   round_out, g_prof));
}

template <class P>
static int conv3x3_wgrad_launch(const void* dout, const void* in, float* scratch, int* nparts, int cin,
                                int cout, int batch, int batch_cap, int H, int W, void* stream,
                                int want_bias = 1) {
  const Geom g = make_geom(batch, batch_cap, H, W);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (W + 3 > kHalo) return DDPB_ERR_SHAPE;
  if (!(cin == 32 && cout == 64)) return DDPB_ERR_SHAPE;
  constexpr int CI = 32, CO = 64, STAGES = 2;
  using S = WgradSmem<P, CI, CO, STAGES>;
  auto kern = conv3x3_wgrad_tc_kernel<P, CI, CO, STAGES>;
  static bool configured[64] = {};
  const int rc = configure_smem(kern, S::total, configured);
  if (rc != 0) return rc;
  const int sms = bwd_grid_cap(ddpb_num_sms(), P::kSets == 2, g.ntiles);
  const int grid = g.ntiles < sms ? g.ntiles : sms;
  DDPB_CUDA_OK(launch_k(kern, dim3(grid), dim3(kWgradThreads), S::total, st,
                        static_cast<const uint4*>(dout), static_cast<const uint4*>(in), scratch, g,
                        want_bias, g_prof));
  *nparts = grid;
  return 0;
}

template <class P>
static int conv3x3_wgrad_full(const void* dout, const void* in, float* dw, float* db, float* scratch,
                              int cin, int cout, int batch, int batch_cap, int H, int W, void* stream,
                              const float* bias_part = nullptr, int nsplit = 0, int nbx = 0) {
  int grid = 0;
  // with bias parts from the linear backward the tensor-core bias column is not needed
  const int rc = conv3x3_wgrad_launch<P>(dout, in, scratch, &grid, cin, cout, batch, batch_cap, H, W, stream,
                                         bias_part == nullptr);
  if (rc != 0) return rc;
  constexpr int CI = 32, CO = 64;
  const int n = CO * (9 * CI + 8);
  return static_cast<int>(launch_k(wgrad_reduce_kernel<CI, CO>, dim3((n + 31) / 32), dim3(256), 0,
                                   static_cast<cudaStream_t>(stream), static_cast<const float*>(scratch),
                                   grid, dw, db, bias_part, nsplit, nbx, H * W));
}

}  // namespace ddpb

using namespace ddpb;

extern "C" {

// Tuning hook: when non-NULL, the conv kernels write SM-clock phase stamps to
// buf[block*16 + slot] (needs >= gridDim*16 entries).  NULL (default) disables it.
void ddpb_set_prof_buffer(void* buf) { g_prof = static_cast<unsigned long long*>(buf); }

// ---- bf16 planes (fast mode): fp32 master weights, converted per CTA
int ddpb_conv3x3_fwd(const void* in, const float* w, const float* bias, void* out, int cin,
                     int cout, int batch, int batch_cap, int H, int W, void* stream) {
  const Geom g = make_geom(batch, batch_cap, H, W);
  const int sms = ddpb_num_sms();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (W + 3 > kHalo) return DDPB_ERR_SHAPE;
  if (cin == 32 && cout == 64)
    return launch_conv<Bf16Planes, Bf16Planes, Bf16Planes, 32, 64, 0>(in, w, bias, nullptr, out, g, 0, sms, st);
  return DDPB_ERR_SHAPE;
}

// dout has `cout` channels, result din (and mask = the forward activation that fed the conv)
// have `cin` channels; w is the layer weight [cout][cin][3][3].
int ddpb_conv3x3_dgrad(const void* dout, const float* w, const void* mask, void* din, int cin,
                       int cout, int batch, int batch_cap, int H, int W, void* stream) {
  const Geom g = make_geom(batch, batch_cap, H, W);
  const int sms = ddpb_num_sms();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (W + 3 > kHalo) return DDPB_ERR_SHAPE;
  if (cin == 32 && cout == 64)
    return launch_conv<Bf16Planes, Bf16Planes, Bf16Planes, 64, 32, 1>(dout, w, nullptr, mask, din, g, 0, sms, st);
  return DDPB_ERR_SHAPE;
}

size_t ddpb_conv3x3_wgrad_scratch_bytes(int cin, int cout) {
  return static_cast<size_t>(ddpb_num_sms()) * cout * (9 * cin + 8) * sizeof(float);
}

int ddpb_conv3x3_wgrad(const void* dout, const void* in, float* dw, float* db, float* scratch,
                       int cin, int cout, int batch, int batch_cap, int H, int W, void* stream) {
  return conv3x3_wgrad_full<Bf16Planes>(dout, in, dw, db, scratch, cin, cout, batch, batch_cap, H, W,
                                        stream);
}

// Partials only: scratch[nparts][cout * (9*cin + 8)] accumulator images (one per CTA) and their count
// in *nparts_out (host).  Feed them to ddpb_reduce_allreduce_sgd with layout = 1 (fused split-K
// reduce + all-reduce + SGD).
int ddpb_conv3x3_wgrad_partials(const void* dout, const void* in, float* scratch, int* nparts_out,
                                int cin, int cout, int batch, int batch_cap, int H, int W,
                                void* stream) {
  return conv3x3_wgrad_launch<Bf16Planes>(dout, in, scratch, nparts_out, cin, cout, batch, batch_cap, H,
                                          W, stream);
}

// ---- split planes (x = hi + lo, bf16x3): the reference-precision mode.
// packed: [fwd hi | fwd lo | dgrad hi | dgrad lo] bf16 tile images, ddpb_conv3x3_pack_split_bytes bytes.
size_t ddpb_conv3x3_pack_split_bytes(int cin, int cout) {
  return static_cast<size_t>(4) * 9 * cin * cout * sizeof(__nv_bfloat16);
}
int ddpb_conv3x3_pack_split(const float* w, void* packed, int cin, int cout, void* stream) {
  if (!(cin == 32 && cout == 64)) return DDPB_ERR_SHAPE;
  const int n = 9 * cin * cout;
  return static_cast<int>(launch_k(pack_conv3x3_split_kernel<32, 64>, dim3((n + 255) / 256), dim3(256), 0,
                                   static_cast<cudaStream_t>(stream), w,
                                   static_cast<__nv_bfloat16*>(packed)));
}
// in: split planes (cin channels); out: fp32 planes when out_f32 != 0 (the consumer is the fp32
// linear layer), else split planes.
int ddpb_conv3x3_fwd_split(const void* in, const void* packed, const float* bias, void* out, int cin,
                           int cout, int batch, int batch_cap, int H, int W, int out_f32, void* stream) {
  const Geom g = make_geom(batch, batch_cap, H, W);
  if (W + 3 > kHalo) return DDPB_ERR_SHAPE;
  if (!(cin == 32 && cout == 64)) return DDPB_ERR_SHAPE;
  const int sms = ddpb_num_sms();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (out_f32)
    return launch_conv<SplitPlanes, F32Planes, SplitPlanes, 32, 64, 0>(in, packed, bias, nullptr, out, g, 0, sms, st);
  return launch_conv<SplitPlanes, SplitPlanes, SplitPlanes, 32, 64, 0>(in, packed, bias, nullptr, out, g, 0, sms, st);
}
// dout and mask (the forward activation that fed the conv) are split planes; din = unrounded fp32
// planes when out_f32 != 0 (its consumer is the exact-fp32 conv1 weight gradient), else split planes.
int ddpb_conv3x3_dgrad_split(const void* dout, const void* packed, const void* mask, void* din, int cin,
                             int cout, int batch, int batch_cap, int H, int W, int out_f32,
                             void* stream) {
  const Geom g = make_geom(batch, batch_cap, H, W);
  if (W + 3 > kHalo) return DDPB_ERR_SHAPE;
  if (!(cin == 32 && cout == 64)) return DDPB_ERR_SHAPE;
  const __nv_bfloat16* dg = static_cast<const __nv_bfloat16*>(packed) + 2 * 9 * cin * cout;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (out_f32)
    return launch_conv<SplitPlanes, F32Planes, SplitPlanes, 64, 32, 1>(dout, dg, nullptr, mask, din, g, 0,
                                                                      ddpb_num_sms(), st);
  return launch_conv<SplitPlanes, SplitPlanes, SplitPlanes, 64, 32, 1>(dout, dg, nullptr, mask, din, g, 0,
                                                                      ddpb_num_sms(), st);
}
// bias_part (optional): exact-fp32 bias-gradient parts from ddpb_linear_bwd_f32 with their extents;
// when given, db is their fixed-order sum instead of the tensor-core ones-column (16-bit operands).
int ddpb_conv3x3_wgrad_split(const void* dout, const void* in, float* dw, float* db, float* scratch,
                             const float* bias_part, int nsplit, int nbx, int cin, int cout, int batch,
                             int batch_cap, int H, int W, void* stream) {
  return conv3x3_wgrad_full<SplitPlanes>(dout, in, dw, db, scratch, cin, cout, batch, batch_cap, H, W,
                                         stream, bias_part, nsplit, nbx);
}
int ddpb_conv3x3_wgrad_split_partials(const void* dout, const void* in, float* scratch, int* nparts_out,
                                      int cin, int cout, int batch, int batch_cap, int H, int W,
                                      void* stream) {
  return conv3x3_wgrad_launch<SplitPlanes>(dout, in, scratch, nparts_out, cin, cout, batch, batch_cap, H,
                                           W, stream);
}

// ---- fp32 planes, tf32 tensor-core operands: forward and data gradient only (K-major operands).
size_t ddpb_conv3x3_pack_tf32_bytes(int cin, int cout) {
  return static_cast<size_t>(2) * 9 * cin * cout * sizeof(float);
}
// packed: [fwd image | dgrad image], each 9*cin*cout floats (see pack_conv3x3_tf32_kernel).
int ddpb_conv3x3_pack_tf32(const float* w, void* packed, int cin, int cout, void* stream) {
  if (!(cin == 32 && cout == 64)) return DDPB_ERR_SHAPE;
  float* fwd = static_cast<float*>(packed);
  float* dgr = fwd + 9 * cin * cout;
  const int n = 9 * cin * cout;
  return static_cast<int>(launch_k(pack_conv3x3_tf32_kernel<32, 64>, dim3((n + 255) / 256), dim3(256), 0,
                                   static_cast<cudaStream_t>(stream), w, fwd, dgr));
}
int ddpb_conv3x3_fwd_tf32(const void* in, const void* packed, const float* bias, void* out, int cin,
                          int cout, int batch, int batch_cap, int H, int W, int round_out,
                          void* stream) {
  const Geom g = make_geom(batch, batch_cap, H, W);
  if (W + 3 > kHalo) return DDPB_ERR_SHAPE;
  if (cin == 32 && cout == 64)
    return launch_conv<F32Planes, F32Planes, F32Planes, 32, 64, 0>(
        in, packed, bias, nullptr, out, g, round_out, ddpb_num_sms(), static_cast<cudaStream_t>(stream));
  return DDPB_ERR_SHAPE;
}
int ddpb_conv3x3_dgrad_tf32(const void* dout, const void* packed, const void* mask, void* din, int cin,
                            int cout, int batch, int batch_cap, int H, int W, int round_out,
                            void* stream) {
  const Geom g = make_geom(batch, batch_cap, H, W);
  if (W + 3 > kHalo) return DDPB_ERR_SHAPE;
  if (cin == 32 && cout == 64)
    return launch_conv<F32Planes, F32Planes, F32Planes, 64, 32, 1>(
        dout, static_cast<const float*>(packed) + 9 * cin * cout, nullptr, mask, din, g, round_out,
        ddpb_num_sms(), static_cast<cudaStream_t>(stream));
  return DDPB_ERR_SHAPE;
}

}  // extern "C"
