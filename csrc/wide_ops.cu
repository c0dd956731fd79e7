// fp32 CUDA-core kernels for the widened CNN of BASELINE.json config 4 (models/wide_cnn.py), all in
// channels-last (NHWC) activation layout so that no layout transposes surround them:
//   conv3x3_small_cin_fwd   first convolution (3 -> 64 channels, K = 27: too thin for the tensor cores)
//                           + bias + ReLU, NCHW input image -> NHWC activation
//   conv3x3_small_cin_wgrad its weight and bias gradient (ReLU mask applied on load), as per-block
//                           partials summed in fixed order by ddpb_reduce_partials
//   relu_maxpool2_nhwc      bias + ReLU + 2x2 max pool forward / backward on NHWC tensors
// The reference's model has no such layers (model.py:8-16: two 3x3 convolutions at 28x28, no pooling);
// they exist for the scaled configuration, where torch would run cuDNN/ATen kernels instead.
#include <stdint.h>

#include "ptx.cuh"
#include "api_internal.cuh"
#include "launch.cuh"

namespace ddpb {

constexpr int kWideW = 32;       // image width these kernels are tiled for
constexpr int kWideRows = 4;     // output rows per block tile (128 pixels)
constexpr int kWideXs = 36;      // padded row length of the staged input tile (cols -1 .. 32, + pad)

// Stage rows [h0-1, h0+kWideRows] x cols [-1, 32] of image b (NCHW, CIN channels), zero outside.
template <int CIN>
__device__ __forceinline__ void stage_input_tile(const float* __restrict__ x, float (*xs)[kWideRows + 2][kWideXs],
                                                 int b, int h0, int H, int tid, int nthreads) {
  // all loads of a thread are issued before the first store (the loop below is fully unrolled): one global
  // round trip per tile instead of one per element
  constexpr int N = CIN * (kWideRows + 2) * 34;
  constexpr int PER = (N + 127) / 128;  // nthreads is 128 at every call site
  float v[PER];
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    const int i = tid + k * nthreads;
    const int col = i % 34, row = (i / 34) % (kWideRows + 2), ci = i / (34 * (kWideRows + 2));
    const int h = h0 - 1 + row, w = col - 1;
    v[k] = 0.f;
    if (i < N && h >= 0 && h < H && w >= 0 && w < kWideW) v[k] = x[(((long long)b * CIN + ci) * H + h) * kWideW + w];
  }
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    const int i = tid + k * nthreads;
    const int col = i % 34, row = (i / 34) % (kWideRows + 2), ci = i / (34 * (kWideRows + 2));
    if (i < N) xs[ci][row][col] = v[k];
  }
}

// Thread = 4 horizontally adjacent pixels x 16 output channels {16 i + 4 cg + e}: the 54 input values
// of the 3x6 window stay in registers, every weight float4 read from shared memory feeds 16 FMAs, and
// the four lanes cg = 0..3 of a pixel store 64 contiguous bytes per instruction.
template <int CIN, int COUT>
__global__ void __launch_bounds__(128)
conv3x3_small_cin_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                             const float* __restrict__ bias, float* __restrict__ out, int batch, int H) {
  static_assert(COUT % 16 == 0, "channels are handled in groups of 16");
  __shared__ __align__(16) float ws[CIN * 9][COUT];
  __shared__ float xs[CIN][kWideRows + 2][kWideXs];
  __shared__ __align__(16) float bs[COUT];
  const int tid = threadIdx.x;
  const int rbs = H / kWideRows;
  pdl_sync();  // the weights may have been written by the stream predecessor (an optimizer update)
  for (int i = tid; i < CIN * 9 * COUT; i += 128) {  // ws[k][co] = w[co][k]; consecutive threads -> consecutive co
    const int co = i % COUT, k = i / COUT;
    ws[k][co] = w[co * (CIN * 9) + k];
  }
  for (int i = tid; i < COUT; i += 128) bs[i] = bias ? bias[i] : 0.f;
  const int cg = tid & 3, quad = tid >> 2;
  const int r = quad >> 3, c0 = (quad & 7) * 4;
  // a block walks its share of the 128-pixel tiles: the weights are staged once
  for (int t = blockIdx.x; t < batch * rbs; t += gridDim.x) {
  const int b = t / rbs, h0 = (t % rbs) * kWideRows;
  __syncthreads();  // the previous tile's window values are in registers
  stage_input_tile<CIN>(x, xs, b, h0, H, tid, 128);
  __syncthreads();
  float xin[CIN][3][6];
#pragma unroll
  for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
    for (int dy = 0; dy < 3; ++dy)
#pragma unroll
      for (int dx = 0; dx < 6; ++dx) xin[ci][dy][dx] = xs[ci][r + dy][c0 + dx];
  float* orow = out + (((long long)b * H + h0 + r) * kWideW + c0) * COUT;
#pragma unroll 1
  for (int i = 0; i < COUT / 16; ++i) {
    const int ch = 16 * i + 4 * cg;
    float acc[4][4];
    const float4 bv = *reinterpret_cast<const float4*>(&bs[ch]);
#pragma unroll
    for (int p = 0; p < 4; ++p) { acc[p][0] = bv.x; acc[p][1] = bv.y; acc[p][2] = bv.z; acc[p][3] = bv.w; }
#pragma unroll
    for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
      for (int dy = 0; dy < 3; ++dy)
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
          const float4 wv = *reinterpret_cast<const float4*>(&ws[ci * 9 + dy * 3 + dx][ch]);
#pragma unroll
          for (int p = 0; p < 4; ++p) {
            const float xv = xin[ci][dy][dx + p];
            acc[p][0] = fmaf(xv, wv.x, acc[p][0]);
            acc[p][1] = fmaf(xv, wv.y, acc[p][1]);
            acc[p][2] = fmaf(xv, wv.z, acc[p][2]);
            acc[p][3] = fmaf(xv, wv.w, acc[p][3]);
          }
        }
#pragma unroll
    for (int p = 0; p < 4; ++p)
      *reinterpret_cast<float4*>(orow + p * COUT + ch) =
          make_float4(fmaxf(acc[p][0], 0.f), fmaxf(acc[p][1], 0.f), fmaxf(acc[p][2], 0.f), fmaxf(acc[p][3], 0.f));
  }
  }
}

// dW[co][ci][dy][dx] = sum_pixels g[p][co] * x[ci][h+dy-1][w+dx-1],  db[co] = sum_pixels g[p][co], with
// g = dact where act > 0 (ReLU backward fused into the tile load).  Thread = 4 output channels x 4 of the
// 28 "slots" (27 taps + the bias as a tap whose input is 1); a block walks its share of 128-pixel tiles
// accumulating in registers and writes one partial image [COUT][28].
constexpr int kWgradSlots = 28;
template <int CIN, int COUT>
__global__ void __launch_bounds__(128)
conv3x3_small_cin_wgrad_kernel(const float* __restrict__ g, const float* __restrict__ act,
                               const float* __restrict__ x, float* __restrict__ partial, int batch, int H) {
  static_assert(CIN * 9 + 1 <= kWgradSlots && COUT == 64, "thread map: 16 channel groups x 7 slot groups");
  __shared__ __align__(16) float gs[kWideRows * kWideW][COUT];
  __shared__ float xs[CIN + 1][kWideRows + 2][kWideXs];  // channel CIN is all ones: the bias slot's input
  const int tid = threadIdx.x;
  const int cog = tid & 15, tg = tid >> 4;
  int s_off[4];  // offset of slot j's input for pixel (0, 0) inside xs
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int slot = tg * 4 + j;
    s_off[j] = slot < CIN * 9 ? ((slot / 9) * (kWideRows + 2) + (slot % 9) / 3) * kWideXs + slot % 3
                              : CIN * (kWideRows + 2) * kWideXs;
  }
  for (int i = tid; i < (kWideRows + 2) * kWideXs; i += 128) (&xs[CIN][0][0])[i] = 1.f;
  float acc[4][4];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[j][e] = 0.f;
  pdl_sync();
  const int rbs = H / kWideRows;
  for (int t = blockIdx.x; t < batch * rbs; t += gridDim.x) {
    const int b = t / rbs, h0 = (t % rbs) * kWideRows;
    __syncthreads();  // the previous tile is consumed
    stage_input_tile<CIN>(x, xs, b, h0, H, tid, 128);  // fills channels [0, CIN) only
    const long long base = (((long long)b * H + h0) * kWideW) * COUT;
    constexpr int kVec = kWideRows * kWideW * COUT / 4 / 128;  // float4 per thread (16)
#pragma unroll
    for (int k0 = 0; k0 < kVec; k0 += 8) {  // eight gradient (and mask) loads in flight per thread
      float4 gv[8], av[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = tid + (k0 + u) * 128;
        gv[u] = *reinterpret_cast<const float4*>(g + base + 4 * i);
        if (act) av[u] = *reinterpret_cast<const float4*>(act + base + 4 * i);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = tid + (k0 + u) * 128;
        if (act) {
          gv[u].x = av[u].x > 0.f ? gv[u].x : 0.f; gv[u].y = av[u].y > 0.f ? gv[u].y : 0.f;
          gv[u].z = av[u].z > 0.f ? gv[u].z : 0.f; gv[u].w = av[u].w > 0.f ? gv[u].w : 0.f;
        }
        *reinterpret_cast<float4*>(&gs[0][0] + 4 * i) = gv[u];
      }
    }
    __syncthreads();
    if (tg < 7) {
#pragma unroll 4
      for (int p = 0; p < kWideRows * kWideW; ++p) {
        const float* xp = &xs[0][0][0] + (p >> 5) * kWideXs + (p & 31);
        const float4 gv = *reinterpret_cast<const float4*>(&gs[p][cog * 4]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float xv = xp[s_off[j]];
          acc[j][0] = fmaf(xv, gv.x, acc[j][0]);
          acc[j][1] = fmaf(xv, gv.y, acc[j][1]);
          acc[j][2] = fmaf(xv, gv.z, acc[j][2]);
          acc[j][3] = fmaf(xv, gv.w, acc[j][3]);
        }
      }
    }
  }
  if (tg < 7) {
    float* dst = partial + (long long)blockIdx.x * COUT * kWgradSlots;
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) dst[(cog * 4 + e) * kWgradSlots + tg * 4 + j] = acc[j][e];
  }
}

// out[i] = sum_p part[p*stride + i], i < n, for a short vector and many parts: 1024 threads = 32 consecutive outputs x
// 32 part groups; group g adds parts g, g+32, ... (loads independent, all in flight), then the 32 group
// sums are added in group order.  Fixed order => deterministic.
__global__ void __launch_bounds__(1024)
reduce_parts32_kernel(const float* __restrict__ part, int nparts, long long stride, int n,
                      float* __restrict__ out) {
  __shared__ float sm[32][33];
  pdl_sync();
  const int lane = threadIdx.x & 31, pg = threadIdx.x >> 5;
  const int idx = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (idx < n) {
    int p = pg;
    for (; p + 96 < nparts; p += 128) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = part[(long long)(p + 32 * u) * stride + idx];
#pragma unroll
      for (int u = 0; u < 4; ++u) s += v[u];
    }
    for (; p < nparts; p += 32) s += part[(long long)p * stride + idx];
  }
  sm[pg][lane] = s;
  __syncthreads();
  if (pg == 0 && idx < n) {
    float t = 0.f;
#pragma unroll
    for (int u = 0; u < 32; ++u) t += sm[u][lane];
    out[idx] = t;
  }
}

// NHWC: out[b][oh][ow][c] = relu(max over the 2x2 window of in[b][2oh+dy][2ow+dx][c] + bias[c]); code = the
// winner (first maximum in row-major order) or 4 where the result is <= 0.  Thread = 4 channels.
__global__ void __launch_bounds__(256)
relu_maxpool2_nhwc_fwd_kernel(const float* __restrict__ in, const float* __restrict__ bias, float* __restrict__ out,
                              uint8_t* __restrict__ code, long long n4, int C, int H, int W) {
  pdl_sync();
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= n4) return;
  const int c4 = C / 4, OW = W / 2, OH = H / 2;
  const int cq = (int)(idx % c4);
  const int ow = (int)((idx / c4) % OW);
  const int oh = (int)((idx / ((long long)c4 * OW)) % OH);
  const long long b = idx / ((long long)c4 * OW * OH);
  const float* p = in + (((b * H + 2 * oh) * W + 2 * ow) * C) + 4 * cq;
  const float4 v[4] = {*reinterpret_cast<const float4*>(p), *reinterpret_cast<const float4*>(p + C),
                       *reinterpret_cast<const float4*>(p + (long long)W * C),
                       *reinterpret_cast<const float4*>(p + (long long)W * C + C)};
  const float4 bv = bias ? *reinterpret_cast<const float4*>(bias + 4 * cq) : make_float4(0.f, 0.f, 0.f, 0.f);
  const float bb[4] = {bv.x, bv.y, bv.z, bv.w};
  float o[4];
  uint8_t cd[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float t[4] = {reinterpret_cast<const float*>(&v[0])[e], reinterpret_cast<const float*>(&v[1])[e],
                        reinterpret_cast<const float*>(&v[2])[e], reinterpret_cast<const float*>(&v[3])[e]};
    float best = t[0];
    int arg = 0;
#pragma unroll
    for (int k = 1; k < 4; ++k)
      if (t[k] > best) { best = t[k]; arg = k; }
    best += bb[e];
    o[e] = fmaxf(best, 0.f);
    cd[e] = best > 0.f ? (uint8_t)arg : (uint8_t)4;
  }
  *reinterpret_cast<float4*>(out + 4 * idx) = make_float4(o[0], o[1], o[2], o[3]);
  *reinterpret_cast<uchar4*>(code + 4 * idx) = make_uchar4(cd[0], cd[1], cd[2], cd[3]);
}

__global__ void __launch_bounds__(256)
relu_maxpool2_nhwc_bwd_kernel(const float* __restrict__ dout, const uint8_t* __restrict__ code,
                              float* __restrict__ din, long long n4, int C, int H, int W) {
  pdl_sync();
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= n4) return;
  const int c4 = C / 4, OW = W / 2, OH = H / 2;
  const int cq = (int)(idx % c4);
  const int ow = (int)((idx / c4) % OW);
  const int oh = (int)((idx / ((long long)c4 * OW)) % OH);
  const long long b = idx / ((long long)c4 * OW * OH);
  const float4 d = *reinterpret_cast<const float4*>(dout + 4 * idx);
  const uchar4 cd = *reinterpret_cast<const uchar4*>(code + 4 * idx);
  float* p = din + (((b * H + 2 * oh) * W + 2 * ow) * C) + 4 * cq;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float4 r = make_float4(cd.x == k ? d.x : 0.f, cd.y == k ? d.y : 0.f, cd.z == k ? d.z : 0.f,
                                 cd.w == k ? d.w : 0.f);
    *reinterpret_cast<float4*>(p + (long long)(k >> 1) * W * C + (k & 1) * C) = r;
  }
}

// ------------------------------------------------------------------------------------------
// Linear layers with a handful of outputs (a classifier head: NOUT <= 16), exact fp32.  The tensor-core
// GEMM of gemm_tc.cu pads such a layer to a 128-wide tile and needs ten launches per step (three packs,
// three GEMMs, split-K reduces, a column sum); these are four.
constexpr int kSmallOutMax = 16;

// y[b][o] = sum_k x[b][k] W[o][k] + bias[o]; one block per batch row, its 256 threads stride over k (a few
// independent iterations each: the loads of all of them are in flight together), then a fixed-order
// reduction over lanes and warps.
__global__ void __launch_bounds__(256)
small_linear_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                        float* __restrict__ y, int batch, int K, int nout) {
  __shared__ float red[8][kSmallOutMax];
  pdl_sync();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.x;
  float acc[kSmallOutMax];
#pragma unroll
  for (int o = 0; o < kSmallOutMax; ++o) acc[o] = 0.f;
  const float* xr = x + (long long)b * K;
#pragma unroll 4
  for (int k = threadIdx.x; k < K; k += 256) {
    const float xv = xr[k];
#pragma unroll
    for (int o = 0; o < kSmallOutMax; ++o)
      if (o < nout) acc[o] = fmaf(xv, w[(long long)o * K + k], acc[o]);
  }
#pragma unroll
  for (int o = 0; o < kSmallOutMax; ++o) {
    float v = acc[o];
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
    if (lane == 0) red[warp][o] = v;
  }
  __syncthreads();
  if (threadIdx.x < nout) {
    float v = 0.f;
#pragma unroll
    for (int u = 0; u < 8; ++u) v += red[u][threadIdx.x];
    y[(long long)b * nout + threadIdx.x] = v + (bias ? bias[threadIdx.x] : 0.f);
  }
}

// dx[b][k] = sum_o dy[b][o] W[o][k]; thread per (b, k).
__global__ void __launch_bounds__(256)
small_linear_dx_kernel(const float* __restrict__ dy, const float* __restrict__ w, float* __restrict__ dx, int batch,
                       int K, int nout) {
  pdl_sync();
  const int k = blockIdx.x * 256 + threadIdx.x;
  const int b = blockIdx.y;
  if (k >= K) return;
  float s = 0.f;
  for (int o = 0; o < nout; ++o) s = fmaf(dy[(long long)b * nout + o], w[(long long)o * K + k], s);
  dx[(long long)b * K + k] = s;
}

// partial[part][o*K + k] = sum over the part's batch rows of dy[b][o] x[b][k]; partial[part][nout*K + o] =
// sum of dy[b][o] (the bias gradient).  Thread per k; blockIdx.y = part (32 batch rows).
__global__ void __launch_bounds__(256)
small_linear_dw_kernel(const float* __restrict__ dy, const float* __restrict__ x, float* __restrict__ partial,
                       int batch, int K, int nout) {
  __shared__ float dys[32][kSmallOutMax];
  pdl_sync();
  const int k = blockIdx.x * 256 + threadIdx.x;
  const int b0 = blockIdx.y * 32;
  const int nb = min(32, batch - b0);
  for (int i = threadIdx.x; i < 32 * kSmallOutMax; i += 256) {
    const int bb = i / kSmallOutMax, o = i % kSmallOutMax;
    dys[bb][o] = (bb < nb && o < nout) ? dy[(long long)(b0 + bb) * nout + o] : 0.f;
  }
  __syncthreads();
  float* dst = partial + (long long)blockIdx.y * ((long long)nout * K + nout);
  if (k < K) {
    float acc[kSmallOutMax];
#pragma unroll
    for (int o = 0; o < kSmallOutMax; ++o) acc[o] = 0.f;
    for (int bb = 0; bb < nb; ++bb) {
      const float xv = x[(long long)(b0 + bb) * K + k];
#pragma unroll
      for (int o = 0; o < kSmallOutMax; ++o) acc[o] = fmaf(xv, dys[bb][o], acc[o]);
    }
#pragma unroll
    for (int o = 0; o < kSmallOutMax; ++o)
      if (o < nout) dst[(long long)o * K + k] = acc[o];
  }
  if (blockIdx.x == 0 && threadIdx.x < nout) {
    float s = 0.f;
    for (int bb = 0; bb < nb; ++bb) s += dys[bb][threadIdx.x];
    dst[(long long)nout * K + threadIdx.x] = s;
  }
}

}  // namespace ddpb

using namespace ddpb;

extern "C" {

// out[b][h][w][co] (NHWC) = relu(conv3x3_pad1(x[b][ci][h][w] (NCHW), w[co][ci][3][3]) + bias[co]); fp32 FMA.
// Supported: cin = 3, cout = 64, W = 32, H a multiple of 4.
int ddpb_conv3x3_small_cin_fwd(const float* x, const float* w, const float* bias, float* out, int cin, int cout,
                               int batch, int H, int W, void* stream) {
  if (cin != 3 || cout != 64 || W != kWideW || H < kWideRows || H % kWideRows || batch <= 0) return DDPB_ERR_SHAPE;
  if (reinterpret_cast<uintptr_t>(out) & 15) return DDPB_ERR_SHAPE;
  const int tiles = batch * (H / kWideRows), cap = 8 * ddpb_num_sms();
  DDPB_CUDA_OK(launch_k(conv3x3_small_cin_fwd_kernel<3, 64>, dim3(tiles < cap ? tiles : cap), dim3(128), 0,
                        (cudaStream_t)stream, x, w, bias, out, batch, H));
  return 0;
}

// Number of partial images (cout * 28 floats each) ddpb_conv3x3_small_cin_wgrad needs scratch for.
int ddpb_conv3x3_small_cin_wgrad_parts(int batch, int H) {
  const int tiles = batch * (H / kWideRows);
  const int cap = 4 * ddpb_num_sms();  // ~36 KB of shared memory per block: up to six blocks per SM
  return tiles < cap ? tiles : cap;
}

// dw_db[cout][28] (per output channel the 27 weight-gradient taps in [ci][dy][dx] order, then the bias
// gradient) from dact (NHWC gradient wrt the convolution's ReLU output), act (NHWC forward output, the ReLU
// mask; NULL = no mask) and the NCHW input x.  partial = scratch of parts * cout * 28 floats.
int ddpb_conv3x3_small_cin_wgrad(const float* dact, const float* act, const float* x, float* partial,
                                 float* dw_db, int cin, int cout, int batch, int H, int W, void* stream) {
  if (cin != 3 || cout != 64 || W != kWideW || H < kWideRows || H % kWideRows || batch <= 0) return DDPB_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(dact) & 15) || (reinterpret_cast<uintptr_t>(act) & 15)) return DDPB_ERR_SHAPE;
  const int parts = ddpb_conv3x3_small_cin_wgrad_parts(batch, H);
  DDPB_CUDA_OK(launch_k(conv3x3_small_cin_wgrad_kernel<3, 64>, dim3(parts), dim3(128), 0, (cudaStream_t)stream,
                        dact, act, x, partial, batch, H));
  const int n = cout * kWgradSlots;
  DDPB_CUDA_OK(launch_k(reduce_parts32_kernel, dim3((n + 31) / 32), dim3(1024), 0, (cudaStream_t)stream,
                        (const float*)partial, parts, (long long)n, n, dw_db));
  return 0;
}

// NHWC fused bias + ReLU + 2x2 max pool (C a multiple of 4, H and W even); see ddpb_relu_maxpool2_fwd.
int ddpb_relu_maxpool2_nhwc_fwd(const float* in, const float* bias, float* out, unsigned char* code, int batch,
                                int C, int H, int W, void* stream) {
  if (batch <= 0 || C <= 0 || (C & 3) || H < 2 || W < 2 || (H & 1) || (W & 1)) return DDPB_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(in) & 15) || (reinterpret_cast<uintptr_t>(out) & 15) ||
      (reinterpret_cast<uintptr_t>(code) & 3) || (bias && (reinterpret_cast<uintptr_t>(bias) & 15)))
    return DDPB_ERR_SHAPE;
  const long long n4 = (long long)batch * (H / 2) * (W / 2) * (C / 4);
  DDPB_CUDA_OK(launch_k(relu_maxpool2_nhwc_fwd_kernel, dim3((unsigned)((n4 + 255) / 256)), dim3(256), 0,
                        (cudaStream_t)stream, in, bias, out, code, n4, C, H, W));
  return 0;
}

int ddpb_relu_maxpool2_nhwc_bwd(const float* dout, const unsigned char* code, float* din, int batch, int C, int H,
                                int W, void* stream) {
  if (batch <= 0 || C <= 0 || (C & 3) || H < 2 || W < 2 || (H & 1) || (W & 1)) return DDPB_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(dout) & 15) || (reinterpret_cast<uintptr_t>(din) & 15) ||
      (reinterpret_cast<uintptr_t>(code) & 3))
    return DDPB_ERR_SHAPE;
  const long long n4 = (long long)batch * (H / 2) * (W / 2) * (C / 4);
  DDPB_CUDA_OK(launch_k(relu_maxpool2_nhwc_bwd_kernel, dim3((unsigned)((n4 + 255) / 256)), dim3(256), 0,
                        (cudaStream_t)stream, dout, code, din, n4, C, H, W));
  return 0;
}

// ---- linear layers with at most 16 outputs, exact fp32 (x [batch][k], w [nout][k] row-major)
int ddpb_small_linear_max_out() { return kSmallOutMax; }

int ddpb_small_linear_fwd(const float* x, const float* w, const float* bias, float* y, int batch, int k, int nout,
                          void* stream) {
  if (batch <= 0 || k <= 0 || nout <= 0 || nout > kSmallOutMax) return DDPB_ERR_SHAPE;
  DDPB_CUDA_OK(launch_k(small_linear_fwd_kernel, dim3(batch), dim3(256), 0, (cudaStream_t)stream, x, w, bias, y,
                        batch, k, nout));
  return 0;
}

// Floats of scratch ddpb_small_linear_bwd needs.
size_t ddpb_small_linear_bwd_scratch_floats(int batch, int k, int nout) {
  return (size_t)((batch + 31) / 32) * ((size_t)nout * k + nout);
}

// dx [batch][k] (NULL = skip), dw [nout][k], db [nout] (NULL = skip) from dy [batch][nout], x and w.
int ddpb_small_linear_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw, float* db,
                          float* scratch, int batch, int k, int nout, void* stream) {
  if (batch <= 0 || k <= 0 || nout <= 0 || nout > kSmallOutMax) return DDPB_ERR_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  if (dx)
    DDPB_CUDA_OK(launch_k(small_linear_dx_kernel, dim3((k + 255) / 256, batch), dim3(256), 0, st, dy, w, dx, batch, k,
                          nout));
  const int parts = (batch + 31) / 32;
  if (parts > 65535) return DDPB_ERR_LIMIT;
  DDPB_CUDA_OK(launch_k(small_linear_dw_kernel, dim3((k + 255) / 256, parts), dim3(256), 0, st, dy, x, scratch, batch,
                        k, nout));
  const long long stride = (long long)nout * k + nout;
  DDPB_CUDA_OK(launch_k(reduce_parts32_kernel, dim3((nout * k + 31) / 32), dim3(1024), 0, st,
                        (const float*)scratch, parts, stride, nout * k, dw));
  if (db)
    DDPB_CUDA_OK(launch_k(reduce_parts32_kernel, dim3(1), dim3(1024), 0, st,
                          (const float*)(scratch + (long long)nout * k), parts, stride, nout, db));
  return 0;
}

}  // extern "C"
