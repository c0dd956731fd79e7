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

}  // namespace ddpb
