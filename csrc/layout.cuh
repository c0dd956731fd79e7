// Activation layout shared by every kernel of the step ("padded planar-chunk" layout).
//
// An activation tensor with C channels over a batch of HxW images is stored as C/8 planes.
// A plane is a dense array of rows, 16 bytes per row = the 8 bf16 channels of one pixel.
// Rows enumerate *padded* pixels: image b, padded coordinates (hp, wp) in [0,H+2)x[0,W+2)
// live at row  GUARD + b*(H+2)*(W+2) + hp*(W+2) + wp.   Border rows hold zeros.
//
// Why: with this layout a 3x3 tap (r,s) of a pad-1 convolution is a constant row shift
// d = (r-1)*(W+2) + (s-1).  A 128-pixel GEMM tile plus a HALO of rows on both sides is ONE
// contiguous range per plane, so it is fetched by plain 1-D bulk async copies (TMA, UBLKCP),
// and every tap's A operand is the same shared-memory buffer addressed with a shifted UMMA
// descriptor start address (no-swizzle canonical layout: a core matrix is 8 consecutive rows
// x 16 B, exactly 8 consecutive rows of a plane).  The very same buffer is a valid *MN-major*
// operand (channels along M/N, pixels along K) for the weight-gradient GEMM.
//
// Replaces the NCHW fp32 tensors the reference hands to cuDNN (reference model.py:8-14).
#pragma once
#include <stdint.h>

namespace ddpb {

// Three storage formats share the layout; a plane row is 16 bytes in all of them:
//   Bf16Planes  : 8 bf16 channels per row; kind::f16 MMAs (K = 16 = two planes).        Fast mode.
//   SplitPlanes : every value as TWO bf16 terms, x = hi + lo (hi = nearest bf16, lo = the
//                 remainder rounded to bf16: a 16-bit significand, 4 bytes per value like fp32):
//                 C/8 "hi" planes followed by C/8 "lo" planes.  A GEMM over split operands issues
//                 three kind::f16 MMAs per K step (hi.hi + lo.hi + hi.lo, fp32 accumulation; the
//                 dropped lo.lo term is < 2^-16 relative) — "bf16x3".  This is how the
//                 reference-precision mode feeds the tensor cores: it carries more operand bits than
//                 the TF32 (11-bit) convolutions cuDNN runs for the reference (model.py:8-16), and,
//                 unlike tf32, both operand majors exist for it in the no-swizzle layout this file
//                 relies on (a tf32 MN-major operand — the weight gradient's view of a plane — is
//                 only defined for the 128B_BASE32B swizzled layout: tools/umma_probe.cu,
//                 profiles/umma_probe_tf32_mn.txt).
//   F32Planes   : 4 fp32 channels per row; kind::tf32 MMAs (K = 8 = two planes), K-major operands
//                 only.  Used for tensors the fp32 linear layer reads (act2) and by the tf32
//                 forward / data-gradient building blocks.
struct Bf16Planes {
  static constexpr int kCh = 8;    // channels per 16-byte row
  static constexpr int kMmaK = 16;
  static constexpr int kSets = 1;  // plane sets per tensor (2 = hi + lo)
  static constexpr bool kTf32 = false;
};
struct SplitPlanes {
  static constexpr int kCh = 8;
  static constexpr int kMmaK = 16;
  static constexpr int kSets = 2;
  static constexpr bool kTf32 = false;
};
struct F32Planes {
  static constexpr int kCh = 4;
  static constexpr int kMmaK = 8;
  static constexpr int kSets = 1;
  static constexpr bool kTf32 = true;
};

constexpr int kTileM = 128;  // pixels (rows) per GEMM tile
constexpr int kHalo = 32;    // rows fetched before/after a tile; >= (W+2)+1, i.e. W <= 29
constexpr int kGuard = 64;   // zero rows before the first / after the last tile
constexpr int kStageRows = kTileM + 2 * kHalo;  // 192 rows per plane per pipeline stage

struct Geom {
  int batch;      // images in this step (<= capacity the buffers were sized for)
  int H, W;       // image size
  int PW, PHW;    // padded width (W+2) and padded pixels per image (H+2)*(W+2)
  int ntiles;     // ceil(batch*PHW / kTileM)
  long long plane_rows;  // rows per plane (capacity), see plane_rows_for()
  // exact division by PHW / PW as multiply-high (decode_row runs once per row in every producer
  // and epilogue; the generic integer division was ~70 instructions of it): for n < 2^31,
  // n / PHW == umulhi(n, mul_phw) >> sh_phw ; for n < PHW, n / PW == umulhi(n, mul_pw).
  // mul_phw == 0 (PHW a power of two) selects the plain division.
  uint32_t mul_phw, sh_phw, mul_pw;
};

__host__ __device__ inline int tiles_for(int batch, int H, int W) {
  return (batch * (H + 2) * (W + 2) + kTileM - 1) / kTileM;
}
__host__ __device__ inline long long plane_rows_for(int batch_cap, int H, int W) {
  return (long long)kGuard + (long long)tiles_for(batch_cap, H, W) * kTileM + kGuard;
}
__host__ inline Geom make_geom(int batch, int batch_cap, int H, int W) {
  Geom g;
  g.batch = batch;
  g.H = H;
  g.W = W;
  g.PW = W + 2;
  g.PHW = (H + 2) * (W + 2);
  g.ntiles = tiles_for(batch, H, W);
  g.plane_rows = plane_rows_for(batch_cap, H, W);
  uint32_t sh = 0;
  while ((2u << sh) <= (uint32_t)g.PHW) ++sh;  // floor(log2(PHW))
  const unsigned long long m = ((1ull << (32 + sh)) + g.PHW - 1) / g.PHW;
  g.sh_phw = sh;
  g.mul_phw = m < (1ull << 32) ? (uint32_t)m : 0u;
  g.mul_pw = (uint32_t)(((1ull << 32) + g.PW - 1) / g.PW);
  return g;
}

// Decode a row index relative to kGuard.  Returns true for a real (interior) pixel of an image
// that is part of this step's batch.
__device__ __forceinline__ bool decode_row(const Geom& g, int rel, int& img, int& h, int& w) {
#ifdef __CUDA_ARCH__
  img = g.mul_phw ? (int)(__umulhi((uint32_t)rel, g.mul_phw) >> g.sh_phw) : rel / g.PHW;
  int pos = rel - img * g.PHW;
  int hp = (int)__umulhi((uint32_t)pos, g.mul_pw);
#else
  img = rel / g.PHW;
  int pos = rel - img * g.PHW;
  int hp = pos / g.PW;
#endif
  int wp = pos - hp * g.PW;
  h = hp - 1;
  w = wp - 1;
  return img < g.batch && hp >= 1 && hp <= g.H && wp >= 1 && wp <= g.W;
}

// Exact-fp32 bias gradient of the convolution that precedes the linear layer.  The fp32 linear backward
// (simt_ops.cu) sums, per block of kBiasBlock feature vectors (4 channels of one pixel each) and per
// image split, the masked dX values it produces while they are still fp32 registers:
//   bias_part[(split * nbx + bx) * 8 + slot * 4 + e],  slot = (4-channel group of the vector) -
//   (group of the block's first vector) in {0, 1},  e = channel within the group.
// This sums the parts of channel ch in a fixed order.
constexpr int kBiasBlock = 256;
__device__ __forceinline__ float sum_bias_parts(const float* __restrict__ bp, int nsplit, int nbx, int HW,
                                                int ch) {
  const int c4 = ch >> 2, e = ch & 3;
  const int b_lo = (c4 * HW) / kBiasBlock;
  int b_hi = ((c4 + 1) * HW - 1) / kBiasBlock;
  if (b_hi >= nbx) b_hi = nbx - 1;
  float s = 0.f;
  for (int sp = 0; sp < nsplit; ++sp)
    for (int bx = b_lo; bx <= b_hi; ++bx) {
      const int slot = c4 - (bx * kBiasBlock) / HW;
      if (slot == 0 || slot == 1) s += bp[((long long)sp * nbx + bx) * 8 + slot * 4 + e];
    }
  return s;
}

// Element index of layer weight w[co][ci][tap] inside the pre-packed B-tile images of a 3x3
// convolution with CI_L input / CO_L output channels; E = elements per 16-byte row (4 tf32 / 8 bf16):
//   forward image   [tap][ci/E][co][ci%E]
//   data-grad image [8-tap][co/E][ci][co%E]   (GEMM K = layer output channels, taps mirrored)
template <int CI_L, int CO_L, int E>
__host__ __device__ constexpr int conv3x3_fwd_tile_index(int co, int ci, int tap) {
  return ((tap * (CI_L / E) + ci / E) * CO_L + co) * E + ci % E;
}
template <int CI_L, int CO_L, int E>
__host__ __device__ constexpr int conv3x3_dgrad_tile_index(int co, int ci, int tap) {
  return (((8 - tap) * (CO_L / E) + co / E) * CI_L + ci) * E + co % E;
}
// Split (hi + lo bf16) tile images keep BOTH terms of a K chunk next to each other along N:
//   forward   [tap][ci/8][set][co][ci%8],   data-grad [8-tap][co/8][set][ci][co%8]   (set 0 = hi, 1 = lo)
// so one MMA with N = 2 x (rows of one set) multiplies an operand by [W_hi | W_lo] at once.
template <int CI_L, int CO_L>
__host__ __device__ constexpr int conv3x3_fwd_split_index(int co, int ci, int tap, int set) {
  return (((tap * (CI_L / 8) + ci / 8) * 2 + set) * CO_L + co) * 8 + ci % 8;
}
template <int CI_L, int CO_L>
__host__ __device__ constexpr int conv3x3_dgrad_split_index(int co, int ci, int tap, int set) {
  return ((((8 - tap) * (CO_L / 8) + co / 8) * 2 + set) * CI_L + ci) * 8 + co % 8;
}

}  // namespace ddpb
