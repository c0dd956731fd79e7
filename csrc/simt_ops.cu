// CUDA-core kernels of the training step: the ops whose shapes are too thin for a tensor-core
// tile (Cin=1 convolution, N=10 linear layer) or are pure streaming/reduction work.
//
//   conv1 fwd + bias + ReLU            reference model.py:9-10    (SURVEY.md K1,K2)
//   conv1 weight/bias gradient         reference train_ddp.py:199 (SURVEY.md K12,K13)
//   linear fwd                         reference model.py:16,19   
//   log-softmax + NLL(mean) + dlogits  reference train_ddp.py:40,198-199 (SURVEY.md K7,K8)
//   linear bwd (dX with ReLU', dW, db) reference train_ddp.py:199 (SURVEY.md K9,K10)
//   SGD update                         reference train_ddp.py:41,200
//   dataset gather + u8->f32 scale     reference data.py:11-14 (ToTensor) + DataLoader collate
//
// All activations use the padded planar-chunk bf16 layout of layout.cuh; parameters and
// gradients are fp32 in the reference's canonical (state_dict) layouts.
#include "ptx.cuh"
#include "layout.cuh"
#include "api_internal.cuh"
#include "launch.cuh"
#include <stdlib.h>

namespace ddpb {

// The setting belongs to the calling host thread (kernels are launched by the thread that asked for
// them): two engines driven from two threads of one process do not see each other's scoped overrides.
thread_local int g_pdl = -1;  // -1: take the default from the environment (DDPB_PDL, default on)
int pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("DDPB_PDL");
    g_pdl = (e && e[0] == '0') ? 0 : 1;
  }
  return g_pdl;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Reduce 80 per-lane values across the 32 lanes of a warp with recursive halving: each stage
// swaps half of the values with the partner lane, so 80 shuffles do the work of 400.  On return
// v[0..4] of lane l hold the warp totals of elements  40*b4 + 20*b3 + 10*b2 + 5*b1 + (0..4)
// (b_k = bit k of l); lanes l and l^1 hold the same five totals.
__device__ __forceinline__ void warp_reduce80(float (&v)[80], int lane) {
#define DDPB_HALVE(H, BIT)                                          \
  _Pragma("unroll") for (int i = 0; i < H; ++i) {                   \
    const bool up = (lane & BIT) != 0;                              \
    const float send = up ? v[i] : v[i + H];                        \
    const float keep = up ? v[i + H] : v[i];                        \
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, BIT);          \
  }
  DDPB_HALVE(40, 16)
  DDPB_HALVE(20, 8)
  DDPB_HALVE(10, 4)
  DDPB_HALVE(5, 2)
#undef DDPB_HALVE
#pragma unroll
  for (int i = 0; i < 5; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], 1);
}
// Same scheme for 40 values: v[0..4] of lane l end up as the totals of elements
// 20*b4 + 10*b3 + 5*b2 + (0..4); lanes l, l^1, l^2, l^3 hold the same five totals.
__device__ __forceinline__ void warp_reduce40(float (&v)[40], int lane) {
#define DDPB_HALVE(H, BIT)                                          \
  _Pragma("unroll") for (int i = 0; i < H; ++i) {                   \
    const bool up = (lane & BIT) != 0;                              \
    const float send = up ? v[i] : v[i + H];                        \
    const float keep = up ? v[i + H] : v[i];                        \
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, BIT);          \
  }
  DDPB_HALVE(20, 16)
  DDPB_HALVE(10, 8)
  DDPB_HALVE(5, 4)
#undef DDPB_HALVE
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    v[i] += __shfl_xor_sync(0xffffffffu, v[i], 2);
    v[i] += __shfl_xor_sync(0xffffffffu, v[i], 1);
  }
}
__device__ __forceinline__ int warp_reduce40_base(int lane) {
  return ((lane >> 4) & 1) * 20 + ((lane >> 3) & 1) * 10 + ((lane >> 2) & 1) * 5;
}
__device__ __forceinline__ int warp_reduce80_base(int lane) {
  return ((lane >> 4) & 1) * 40 + ((lane >> 3) & 1) * 20 + ((lane >> 2) & 1) * 10 +
         ((lane >> 1) & 1) * 5;
}

__device__ __forceinline__ void unpack8(const uint4& a, float (&f)[8]) {
  f[0] = bf16_lo(a.x); f[1] = bf16_hi(a.x);
  f[2] = bf16_lo(a.y); f[3] = bf16_hi(a.y);
  f[4] = bf16_lo(a.z); f[5] = bf16_hi(a.z);
  f[6] = bf16_lo(a.w); f[7] = bf16_hi(a.w);
}

// ------------------------------------------------------------------ dataset gather
// x[i] = float(images[idx[i]]) / 255 ; y[i] = labels[idx[i]]     (ToTensor semantics)
__global__ void gather_u8_kernel(const uint8_t* __restrict__ images,
                                 const long long* __restrict__ labels,
                                 const long long* __restrict__ idx, float* __restrict__ x,
                                 long long* __restrict__ y, int batch, int pix) {
  const int i = blockIdx.x;
  if (i >= batch) return;
  const long long src = idx[i];
  const uint8_t* p = images + src * pix;
  for (int k = threadIdx.x; k < pix; k += blockDim.x)
    x[(long long)i * pix + k] = static_cast<float>(p[k]) / 255.0f;  // true division: bit-equal to ToTensor
  if (threadIdx.x == 0) y[i] = labels[src];
}

// ------------------------------------------------------------------ conv1 forward
// Block (bx, chunk) handles a contiguous range of images for 8 output channels.  Per image the
// (zero-bordered) input tile is staged in shared memory so the 3x3 window needs no bounds tests;
// the chunk's 8x(CIN0*9) weights live in registers.  Border rows are written as zeros.
constexpr int kConv1Group = 8;  // images staged in shared memory per round
constexpr int kConv1Guard = 64;  // zeroed floats before/after the staged tiles (>= padded width + 1)

// Stage `ng` zero-bordered images [CIN0][PHW] each into shared memory.
template <int CIN0>
__device__ __forceinline__ void stage_padded_images(float* xs, const float* __restrict__ x, int img0,
                                                    int ng, const Geom& g) {
  const int per = CIN0 * g.PHW;
  for (int i = threadIdx.x; i < ng * per; i += blockDim.x) {
    int li, ci, pos, hp;
    if (CIN0 == 1 && g.mul_phw) {  // exact multiply-high division (layout.cuh), ~10x fewer instructions
      li = static_cast<int>(__umulhi(static_cast<uint32_t>(i), g.mul_phw) >> g.sh_phw);
      ci = 0;
      pos = i - li * g.PHW;
      hp = static_cast<int>(__umulhi(static_cast<uint32_t>(pos), g.mul_pw));
    } else {
      li = i / per;
      const int rem = i - li * per;
      ci = rem / g.PHW;
      pos = rem - ci * g.PHW;
      hp = pos / g.PW;
    }
    const int wp = pos - hp * g.PW;
    const bool in = hp >= 1 && hp <= g.H && wp >= 1 && wp <= g.W;
    xs[i] = in ? x[((long long)(img0 + li) * CIN0 + ci) * g.H * g.W + (hp - 1) * g.W + (wp - 1)] : 0.f;
  }
}

template <int CIN0, int C1>
__global__ void __launch_bounds__(256)
conv1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                 const float* __restrict__ b, __nv_bfloat16* __restrict__ out, Geom g) {
  extern __shared__ float xs[];  // [kConv1Group][CIN0][PHW]
  constexpr int NW = CIN0 * 9;
  const int chunk = blockIdx.y;
  float wr[8][NW], br[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    br[c] = b[chunk * 8 + c];
#pragma unroll
    for (int j = 0; j < NW; ++j) wr[c][j] = w[(chunk * 8 + c) * NW + j];
  }
  const int ipb = (g.batch + gridDim.x - 1) / gridDim.x;
  const int lo = blockIdx.x * ipb;
  const int hi = min(g.batch, lo + ipb);
  __nv_bfloat16* plane = out + ((long long)chunk * g.plane_rows + kGuard) * 8;
  for (int img0 = lo; img0 < hi; img0 += kConv1Group) {
    const int ng = min(kConv1Group, hi - img0);
    __syncthreads();
    stage_padded_images<CIN0>(xs, x, img0, ng, g);
    __syncthreads();
    for (int idx = threadIdx.x; idx < ng * g.PHW; idx += blockDim.x) {
      const int li = idx / g.PHW;
      const int pos = idx - li * g.PHW;
      const int hp = pos / g.PW, wp = pos - hp * g.PW;
      uint4 o = make_uint4(0, 0, 0, 0);
      if (hp >= 1 && hp <= g.H && wp >= 1 && wp <= g.W) {
        const float* xi = xs + li * CIN0 * g.PHW;
        float in[NW];
#pragma unroll
        for (int ci = 0; ci < CIN0; ++ci)
#pragma unroll
          for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int s2 = 0; s2 < 3; ++s2)
              in[ci * 9 + r * 3 + s2] = xi[ci * g.PHW + (hp + r - 1) * g.PW + (wp + s2 - 1)];
        float acc[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          float a = br[c];
#pragma unroll
          for (int j = 0; j < NW; ++j) a = fmaf(wr[c][j], in[j], a);
          acc[c] = fmaxf(a, 0.f);
        }
        o.x = pack_bf16x2(acc[0], acc[1]);
        o.y = pack_bf16x2(acc[2], acc[3]);
        o.z = pack_bf16x2(acc[4], acc[5]);
        o.w = pack_bf16x2(acc[6], acc[7]);
      }
      *reinterpret_cast<uint4*>(plane + ((long long)img0 * g.PHW + idx) * 8) = o;
    }
  }
  // rows between the end of the batch and the end of its last tile
  if (blockIdx.x == gridDim.x - 1) {
    const int first = g.batch * g.PHW, last = g.ntiles * kTileM;
    for (int r = first + threadIdx.x; r < last; r += blockDim.x)
      *reinterpret_cast<uint4*>(plane + (long long)r * 8) = make_uint4(0, 0, 0, 0);
  }
}

// ------------------------------------------------------------------ conv1 weight gradient
// dW[c][ci][tap] = sum_p dact[p][c] * x[p + tap shift][ci],  db[c] = sum_p dact[p][c].
// gridDim.y = C1/8 (channel chunk); block (bx, chunk) accumulates a contiguous range of images and
// writes one partial per blockIdx.x: partial[bx][C1*CIN0*9 + C1] (weights then biases, canonical
// order); reduce_partials_kernel sums them in a fixed order.
// F32IN: dact is fp32 planes (4 channels per row; the chunk's 8 channels are planes 2*chunk and
// 2*chunk+1) and the whole computation is exact fp32 FMA — the reference-precision mode uses this
// for the Cin = 1 layer, which cuDNN also runs in plain fp32 (profiles/precision.md); otherwise bf16
// planes.
// A thread accumulates CH channels x 9 taps (+ CH bias sums): 8 channels from bf16 planes (80
// accumulators), 4 — one fp32 plane — in the exact-fp32 variant (40 accumulators: three CTAs per SM
// instead of one, a cheaper final reduction).
template <int CIN0, int C1, bool F32IN>
__global__ void __launch_bounds__(256, F32IN ? 3 : 1)
conv1_wgrad_kernel(const void* __restrict__ dact_v, const float* __restrict__ x,
                   float* __restrict__ partial, Geom g) {
  constexpr int NW = CIN0 * 9;
  constexpr int CH = F32IN ? 4 : 8;  // channels per thread (= per blockIdx.y)
  constexpr int NV = CH * NW + CH;
  constexpr int PF = 4;          // gradient vectors kept in flight per thread
  extern __shared__ float xs_raw[];  // [guard][kConv1Group][CIN0][PHW][guard], guards zeroed once
  float* xs = xs_raw + kConv1Guard;
  __shared__ float red[8][NV];
  // everything a border row's window can touch must be finite (it is multiplied by zero): clear the
  // guards and the tiles once; afterwards tiles only ever hold staged images
  for (int i = threadIdx.x; i < kConv1Group * CIN0 * g.PHW + 2 * kConv1Guard; i += blockDim.x) xs_raw[i] = 0.f;
  // launched through launch_k in the step (programmatic dependent launch): nothing the predecessor
  // wrote (dact) may be read before this; a no-op for plain launches
  pdl_sync();
  const int chunk = blockIdx.y;
  static_assert(NV == 80 || NV == 40, "warp reductions exist for 80 and 40 accumulators (CIN0 == 1)");
  float acc[NV];  // [c*NW + j] weight taps, then [CH*NW + c] bias
#pragma unroll
  for (int i = 0; i < NV; ++i) acc[i] = 0.f;
  const int ipb = (g.batch + gridDim.x - 1) / gridDim.x;
  const int lo = blockIdx.x * ipb;
  const int hi = min(g.batch, lo + ipb);
  // 16-byte rows: plane `chunk` holds this thread's channels in either format
  const uint4* plane = static_cast<const uint4*>(dact_v) + (long long)chunk * g.plane_rows + kGuard;
  for (int img0 = lo; img0 < hi; img0 += kConv1Group) {
    const int ng = min(kConv1Group, hi - img0);
    const int nidx = ng * g.PHW;
    __syncthreads();
    stage_padded_images<CIN0>(xs, x, img0, ng, g);
    __syncthreads();
    // border rows of dact are zero, so every row of the group can be accumulated unconditionally;
    // the window read stays inside the image's padded tile for interior rows only -> test those.
    for (int base = threadIdx.x; base < nidx; base += PF * blockDim.x) {
      uint4 dv[PF];
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        const int idx = base + u * blockDim.x;
        dv[u] = idx < nidx ? plane[(long long)img0 * g.PHW + idx] : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        const int idx = base + u * blockDim.x;
        if (idx >= nidx) break;
        // No row decode and no interior test: the staged tiles have the planes' row structure (row idx
        // of the group = xs[idx]), border rows of dact are zero, and the 3x3 window of a border row
        // only reaches into a neighbouring tile or the zeroed guard around the staging area, i.e.
        // finite values that are multiplied by zero.
        float d[8];
        if (F32IN) {
          d[0] = __uint_as_float(dv[u].x); d[1] = __uint_as_float(dv[u].y);
          d[2] = __uint_as_float(dv[u].z); d[3] = __uint_as_float(dv[u].w);
        } else {
          unpack8(dv[u], d);
        }
        static_assert(CIN0 == 1, "window addressing below assumes one input channel");
        const float* xi = xs + idx;
        float in[NW];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int s2 = 0; s2 < 3; ++s2) in[r * 3 + s2] = xi[(r - 1) * g.PW + (s2 - 1)];
#pragma unroll
        for (int c = 0; c < CH; ++c) {
          acc[CH * NW + c] += d[c];
#pragma unroll
          for (int j = 0; j < NW; ++j) acc[c * NW + j] = fmaf(d[c], in[j], acc[c * NW + j]);
        }
      }
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if constexpr (NV == 80) {
    warp_reduce80(acc, lane);
    if ((lane & 1) == 0) {
      const int base = warp_reduce80_base(lane);
#pragma unroll
      for (int i = 0; i < 5; ++i) red[warp][base + i] = acc[i];
    }
  } else {
    warp_reduce40(acc, lane);
    if ((lane & 3) == 0) {
      const int base = warp_reduce40_base(lane);
#pragma unroll
      for (int i = 0; i < 5; ++i) red[warp][base + i] = acc[i];
    }
  }
  __syncthreads();
  float* dst = partial + (long long)blockIdx.x * (C1 * NW + C1);
  for (int i = threadIdx.x; i < NV; i += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int wdx = 0; wdx < 8; ++wdx) s += red[wdx][i];
    if (i < CH * NW) dst[chunk * CH * NW + i] = s;
    else dst[C1 * NW + chunk * CH + (i - CH * NW)] = s;
  }
}

// ------------------------------------------------------------------ generic partial reduce
// out[i] = sum_p part[p*n + i], fixed order => deterministic.
__global__ void reduce_partials_kernel(const float* __restrict__ part, int nparts, long long n,
                                       float* __restrict__ out) {
  pdl_sync();
  const long long i4 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 >= n) return;
  if (i4 + 3 < n && (n & 3) == 0) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    int p = 0;
    for (; p + 4 <= nparts; p += 4) {  // 4 independent loads in flight, summed in order
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        v[u] = *reinterpret_cast<const float4*>(part + (long long)(p + u) * n + i4);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w;
      }
    }
    for (; p < nparts; ++p) {
      const float4 v = *reinterpret_cast<const float4*>(part + (long long)p * n + i4);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    *reinterpret_cast<float4*>(out + i4) = s;
  } else {
    for (long long i = i4; i < n && i < i4 + 4; ++i) {
      float s = 0.f;
      for (int p = 0; p < nparts; ++p) s += part[(long long)p * n + i];
      out[i] = s;
    }
  }
}

// Same sum (fixed order) when there are many partials of a short vector: 256 threads = 32
// consecutive outputs x 8 partial groups, group pg adds parts pg, pg+8, ... and the 8 group sums
// are combined in group order.  Turns a chain of nparts dependent loads into nparts/8.
__global__ void __launch_bounds__(256)
reduce_partials_wide_kernel(const float* __restrict__ part, int nparts, long long n,
                            float* __restrict__ out) {
  pdl_sync();
  __shared__ float sm[8][32];
  const int lane = threadIdx.x & 31, pg = threadIdx.x >> 5;
  const long long idx = (long long)blockIdx.x * 32 + lane;
  float s = 0.f;
  if (idx < n) {
    int p = pg;
    for (; p + 24 < nparts; p += 32) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = part[(long long)(p + 8 * u) * n + idx];
#pragma unroll
      for (int u = 0; u < 4; ++u) s += v[u];
    }
    for (; p < nparts; p += 8) s += part[(long long)p * n + idx];
  }
  sm[pg][lane] = s;
  __syncthreads();
  if (pg == 0 && idx < n) {
#pragma unroll
    for (int u = 1; u < 8; ++u) s += sm[u][lane];
    out[idx] = s;
  }
}

// ------------------------------------------------------------------ linear forward
// logits[img][j] = sum_k act[img][k] * W[j][k]; k = c*H*W + h*W + w (NCHW flatten order of the
// reference, model.py:13,16).  grid = (S feature slices, image groups of IMG); writes
// partial[s][img][j]; the loss kernel adds the slices and the bias.
template <int NOUT, int IMG>
__global__ void __launch_bounds__(256, 1)
linear_fwd_kernel(const __nv_bfloat16* __restrict__ act, const float* __restrict__ W,
                  float* __restrict__ partial, Geom g, int C, int vec_per_slice) {
  pdl_sync();
  __shared__ float red[8][IMG * NOUT];
  const int HW = g.H * g.W;
  const int nvec = (C / 8) * HW;
  const long long K = (long long)C * HW;
  const int img0 = blockIdx.y * IMG;
  const int v_begin = blockIdx.x * vec_per_slice;
  const int v_end = min(nvec, v_begin + vec_per_slice);
  static_assert(IMG * NOUT == 80, "warp_reduce80 expects 80 accumulators");
  float acc[IMG * NOUT];
#pragma unroll
  for (int i = 0; i < IMG * NOUT; ++i) acc[i] = 0.f;

  for (int v = v_begin + threadIdx.x; v < v_end; v += blockDim.x) {
    const int chunk = v / HW;
    const int pix = v - chunk * HW;
    const int h = pix / g.W, wc = pix - h * g.W;
    const long long rowi = (long long)(h + 1) * g.PW + (wc + 1);
    float wv[NOUT][8];
#pragma unroll
    for (int j = 0; j < NOUT; ++j)
#pragma unroll
      for (int c = 0; c < 8; ++c) wv[j][c] = W[j * K + (long long)(chunk * 8 + c) * HW + pix];
    uint4 av[IMG];
#pragma unroll
    for (int i = 0; i < IMG; ++i) {
      const int img = img0 + i;
      av[i] = make_uint4(0, 0, 0, 0);
      if (img < g.batch)
        av[i] = *reinterpret_cast<const uint4*>(
            act + ((long long)chunk * g.plane_rows + kGuard + (long long)img * g.PHW + rowi) * 8);
    }
#pragma unroll
    for (int i = 0; i < IMG; ++i) {
      float a[8];
      unpack8(av[i], a);
#pragma unroll
      for (int j = 0; j < NOUT; ++j)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[i * NOUT + j] = fmaf(wv[j][c], a[c], acc[i * NOUT + j]);
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  warp_reduce80(acc, lane);
  if ((lane & 1) == 0) {
    const int base = warp_reduce80_base(lane);
#pragma unroll
    for (int i = 0; i < 5; ++i) red[warp][base + i] = acc[i];
  }
  __syncthreads();
  for (int t = threadIdx.x; t < IMG * NOUT; t += blockDim.x) {
    const int i = t / NOUT, j = t - i * NOUT;
    const int img = img0 + i;
    if (img >= g.batch) continue;
    float s = 0.f;
#pragma unroll
    for (int wdx = 0; wdx < 8; ++wdx) s += red[wdx][t];
    partial[((long long)blockIdx.x * g.batch + img) * NOUT + j] = s;
  }
}

// ------------------------------------------------------------------ loss
// z = bias + sum_s partial[s]; loss = mean_i(logsumexp(z_i) - z_i[y_i]);
// dlogits = (softmax(z) - onehot(y)) / batch; db[j] = sum_i dlogits[i][j].
// Single block: the work is batch*NOUT*S floats.
// Multi-block: block b handles kXentImgs images.  Phase 1 spreads the sum over the S feature-slice
// partials across kXentSG thread groups (the tensor-core linear forward produces up to one slice per
// SM).  Block sums of the loss and of dbias go to a small library-resident array; the block that
// finishes last (ticket counter) adds them up in a fixed order, so the result is deterministic.
// The ticket resets itself; calls on one device must not overlap (the step engine issues one at
// a time).
constexpr int kXentImgs = 8;
constexpr int kXentSG = 8;
constexpr int kXentMaxBlocks = 8192;  // batches up to 65536
static __device__ float g_xent_block_sums[kXentMaxBlocks * 12];
static __device__ unsigned int g_xent_ticket;

template <int NOUT>
__global__ void __launch_bounds__(kXentImgs * NOUT * kXentSG)
softmax_xent_kernel(const float* __restrict__ partial, int S, const float* __restrict__ bias,
                    const long long* __restrict__ labels, float* __restrict__ logits,
                    float* __restrict__ dlogits, float* __restrict__ dbias,
                    float* __restrict__ loss, int batch, uint4* __restrict__ dl_packed) {
  static_assert(NOUT + 1 <= 12, "block-sum row too small");
  constexpr int NP = kXentImgs * NOUT;  // (image, class) pairs per block
  __shared__ float sums[kXentSG][NP];
  __shared__ float red[NOUT + 1][kXentImgs];
  __shared__ bool is_last;
  pdl_sync();
  const int img0 = blockIdx.x * kXentImgs;
  const int nimg = max(0, min(kXentImgs, batch - img0));  // 0 for the zero-row blocks past the batch
  const int tidx = static_cast<int>(threadIdx.x);         // signed: compared with nimg-derived counts
  const long long npair_all = (long long)batch * NOUT;
  // phase 1: thread (pair, sg) adds partial slices sg, sg+SG, ...
  {
    const int pair = threadIdx.x % NP, sg = threadIdx.x / NP;
    float z = 0.f;
    if (pair < nimg * NOUT) {
      const long long gt = (long long)img0 * NOUT + pair;
      int sidx = sg;
      for (; sidx + 3 * kXentSG < S; sidx += 4 * kXentSG) {
        float v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = partial[(long long)(sidx + u * kXentSG) * npair_all + gt];
#pragma unroll
        for (int u = 0; u < 4; ++u) z += v[u];
      }
      for (; sidx < S; sidx += kXentSG) z += partial[(long long)sidx * npair_all + gt];
    }
    sums[sg][pair] = z;
  }
  __syncthreads();
  if (tidx < nimg * NOUT) {
    float z = bias[tidx % NOUT];
#pragma unroll
    for (int u = 0; u < kXentSG; ++u) z += sums[u][tidx];
    sums[0][tidx] = z;
    logits[(long long)img0 * NOUT + tidx] = z;
  }
  __syncthreads();
  // phase 2: one thread per image: log-softmax, loss and gradient
  if (threadIdx.x < kXentImgs) {
    const int li = threadIdx.x;
    float lsum = 0.f, d[NOUT];
#pragma unroll
    for (int j = 0; j < NOUT; ++j) d[j] = 0.f;
    if (li < nimg) {
      const long long i = img0 + li;
      const float inv = 1.0f / static_cast<float>(batch);
      float z[NOUT];
#pragma unroll
      for (int j = 0; j < NOUT; ++j) z[j] = sums[0][li * NOUT + j];
      float m = z[0];
#pragma unroll
      for (int j = 1; j < NOUT; ++j) m = fmaxf(m, z[j]);
      float se = 0.f;
#pragma unroll
      for (int j = 0; j < NOUT; ++j) se += expf(z[j] - m);
      const float lse = m + logf(se);
      const int y = static_cast<int>(labels[i]);
#pragma unroll
      for (int j = 0; j < NOUT; ++j) {
        const float p = expf(z[j] - lse);
        d[j] = (p - (j == y ? 1.f : 0.f)) * inv;
        dlogits[i * NOUT + j] = d[j];
        if (j == y) lsum = lse - z[j];
      }
    }
    // bf16 hi/lo tiles of dlogits for the tensor-core linear backward:
    // tile (image/128): [4 groups: hi j0-7, hi j8-15, lo j0-7, lo j8-15][128 images][8]
    if (dl_packed != nullptr) {
      const long long i = img0 + li;
      const long long padded = ((long long)batch + 127) / 128 * 128;
      if (i < padded) {  // rows past the batch (up to the end of its last tile) are zero
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int qd = 0; qd < 8; ++qd) {
          const float a = (li < nimg && 2 * qd < NOUT) ? d[2 * qd < NOUT ? 2 * qd : 0] : 0.f;
          const float b = (li < nimg && 2 * qd + 1 < NOUT) ? d[2 * qd + 1 < NOUT ? 2 * qd + 1 : 0] : 0.f;
          const float ah = __bfloat162float(__float2bfloat16(a)), bh = __bfloat162float(__float2bfloat16(b));
          hi[qd] = pack_bf16x2(ah, bh);
          lo[qd] = pack_bf16x2(a - ah, b - bh);
        }
        uint4* t = dl_packed + (i >> 7) * 512 + (i & 127);
        t[0] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        t[128] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
        t[256] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        t[384] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
      }
    }
#pragma unroll
    for (int j = 0; j < NOUT; ++j) red[j][li] = d[j];
    red[NOUT][li] = lsum;
  }
  __syncthreads();
  if (threadIdx.x <= NOUT) {
    float s = 0.f;
#pragma unroll
    for (int li = 0; li < kXentImgs; ++li) s += red[threadIdx.x][li];
    g_xent_block_sums[blockIdx.x * 12 + threadIdx.x] = s;
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(&g_xent_ticket, 1u);
    is_last = (t == gridDim.x - 1);
    if (is_last) g_xent_ticket = 0;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // warp w (<= NOUT) adds column w of the block sums: lanes take blocks lane, lane+32, ... then a
  // fixed-shape shuffle tree => deterministic
  const int wq = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (wq <= NOUT) {
    float s = 0.f;
    for (unsigned int bk = lane; bk < gridDim.x; bk += 32)
      s += *reinterpret_cast<volatile float*>(&g_xent_block_sums[bk * 12 + wq]);
    s = warp_sum(s);
    if (lane == 0) {
      if (wq < NOUT) dbias[wq] = s;
      else loss[0] = s / static_cast<float>(batch);
    }
  }
}

// Small batches (batch * NOUT <= 1024 threads, no packed output): ONE block, no ticket and no
// second phase across blocks — at per-rank batch 64 the multi-block version above spends ~9.6 us on
// 640 numbers, most of it in its cross-block hand-over.  Thread (image, class) adds its S slices (in
// slice order) and the bias; one thread per image does the softmax; NOUT threads add the bias-gradient
// columns and one the loss, each in image order (deterministic).
template <int NOUT>
__global__ void __launch_bounds__(1024)
softmax_xent_small_kernel(const float* __restrict__ partial, int S, const float* __restrict__ bias,
                          const long long* __restrict__ labels, float* __restrict__ logits,
                          float* __restrict__ dlogits, float* __restrict__ dbias, float* __restrict__ loss,
                          int batch) {
  __shared__ float zs[1024];
  __shared__ float ls[128];
  pdl_sync();
  const int npair = batch * NOUT;
  const int p = threadIdx.x;
  if (p < npair) {
    float z = 0.f;
    int sl = 0;
    for (; sl + 4 <= S; sl += 4) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = partial[(long long)(sl + u) * npair + p];
#pragma unroll
      for (int u = 0; u < 4; ++u) z += v[u];
    }
    for (; sl < S; ++sl) z += partial[(long long)sl * npair + p];
    z += bias[p % NOUT];
    zs[p] = z;
    logits[p] = z;
  }
  __syncthreads();
  if (p < batch) {
    const float inv = 1.0f / static_cast<float>(batch);
    float z[NOUT];
#pragma unroll
    for (int j = 0; j < NOUT; ++j) z[j] = zs[p * NOUT + j];
    float m = z[0];
#pragma unroll
    for (int j = 1; j < NOUT; ++j) m = fmaxf(m, z[j]);
    float se = 0.f;
#pragma unroll
    for (int j = 0; j < NOUT; ++j) se += expf(z[j] - m);
    const float lse = m + logf(se);
    const int y = static_cast<int>(labels[p]);
    float lsum = 0.f;
#pragma unroll
    for (int j = 0; j < NOUT; ++j) {
      const float pr = expf(z[j] - lse);
      const float d = (pr - (j == y ? 1.f : 0.f)) * inv;
      dlogits[p * NOUT + j] = d;
      zs[p * NOUT + j] = d;
      if (j == y) lsum = lse - z[j];
    }
    ls[p] = lsum;
  }
  __syncthreads();
  if (p < NOUT) {
    float sacc = 0.f;
    for (int i = 0; i < batch; ++i) sacc += zs[i * NOUT + p];
    dbias[p] = sacc;
  } else if (p == 32) {
    float sacc = 0.f;
    for (int i = 0; i < batch; ++i) sacc += ls[i];
    loss[0] = sacc / static_cast<float>(batch);
  }
}

// ------------------------------------------------------------------ linear backward
// dact[img][k] = relu'(act[img][k]) * sum_j dlogits[img][j] W[j][k]     (bf16 planes)
// dW[j][k]    = sum_img dlogits[img][j] act[img][k]                      (fp32, via partials)
// grid = (feature-vector blocks, image splits); partial[split][NOUT][K].
template <int NOUT>
__global__ void __launch_bounds__(256, 2)
linear_bwd_kernel(const __nv_bfloat16* __restrict__ act, const float* __restrict__ dlogits,
                  const float* __restrict__ W, __nv_bfloat16* __restrict__ dact,
                  float* __restrict__ dw_partial, Geom g, int C, int imgs_per_split) {
  pdl_sync();
  extern __shared__ float sdl[];  // [imgs_per_split][NOUT]
  constexpr int PF = 8;           // image vectors kept in flight per thread
  const int HW = g.H * g.W;
  const int nhv = (C / 4) * HW;   // half-vectors: 4 channels (8 bytes) of one pixel
  const long long K = (long long)C * HW;
  const int img_begin = blockIdx.y * imgs_per_split;
  const int img_end = min(g.batch, img_begin + imgs_per_split);
  for (int t = threadIdx.x; t < (img_end - img_begin) * NOUT; t += blockDim.x)
    sdl[t] = dlogits[(long long)img_begin * NOUT + t];
  __syncthreads();
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nhv) return;
  const int hc = v / HW;  // half-chunk index: channels [hc*4, hc*4+4)
  const int pix = v - hc * HW;
  const int h = pix / g.W, wc = pix - h * g.W;
  const long long rowi = (long long)(h + 1) * g.PW + (wc + 1);
  float wv[NOUT][4], dwv[NOUT][4];
#pragma unroll
  for (int j = 0; j < NOUT; ++j)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      wv[j][c] = W[j * K + (long long)(hc * 4 + c) * HW + pix];
      dwv[j][c] = 0.f;
    }
  const long long base =
      ((long long)(hc >> 1) * g.plane_rows + kGuard + rowi) * 8 + (hc & 1) * 4;
  const long long img_stride = (long long)g.PHW * 8;
  for (int i0 = img_begin; i0 < img_end; i0 += PF) {
    uint2 av[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u)
      av[u] = (i0 + u < img_end)
                  ? *reinterpret_cast<const uint2*>(act + base + (long long)(i0 + u) * img_stride)
                  : make_uint2(0, 0);
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      if (i0 + u >= img_end) break;
      const float a[4] = {bf16_lo(av[u].x), bf16_hi(av[u].x), bf16_lo(av[u].y), bf16_hi(av[u].y)};
      float gr[4] = {0.f, 0.f, 0.f, 0.f};
      const float* dl = sdl + (i0 + u - img_begin) * NOUT;
#pragma unroll
      for (int j = 0; j < NOUT; ++j) {
        const float d = dl[j];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          gr[c] = fmaf(d, wv[j][c], gr[c]);
          dwv[j][c] = fmaf(d, a[c], dwv[j][c]);
        }
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) gr[c] = a[c] > 0.f ? gr[c] : 0.f;
      uint2 o;
      o.x = pack_bf16x2(gr[0], gr[1]);
      o.y = pack_bf16x2(gr[2], gr[3]);
      *reinterpret_cast<uint2*>(dact + base + (long long)(i0 + u) * img_stride) = o;
    }
  }
  float* dst = dw_partial + (long long)blockIdx.y * NOUT * K;
#pragma unroll
  for (int j = 0; j < NOUT; ++j)
#pragma unroll
    for (int c = 0; c < 4; ++c) dst[j * K + (long long)(hc * 4 + c) * HW + pix] = dwv[j][c];
}

// Zero the rows of every plane between the end of the batch and the end of its last tile, so
// the weight-gradient GEMM (which reduces over whole tiles) never sees stale rows after a
// short tail batch (reference data.py:21-25 uses drop_last=False).
__global__ void zero_tail_rows_kernel(uint4* __restrict__ planes, int nplanes, Geom g) {
  pdl_sync();
  const int first = g.batch * g.PHW;
  const int nrows = g.ntiles * kTileM - first;
  const long long total = (long long)nplanes * nrows;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int p = static_cast<int>(idx / nrows);
    const int r = static_cast<int>(idx - (long long)p * nrows);
    planes[(long long)p * g.plane_rows + kGuard + first + r] = make_uint4(0, 0, 0, 0);
  }
}

// ------------------------------------------------------------------ linear layer, fp32 planes
// The reference's linear layer is plain fp32 (cuBLAS addmm/mm with TF32 off, reference
// model.py:16,19): in the precision-parity mode it therefore runs as fp32 FMAs on the CUDA cores,
// reading the fp32 planes (4 channels per 16-byte row) the tf32 convolutions write.  N = 10 makes it
// a GEMV-shaped, bandwidth-bound op; the products are exact fp32 and every sum has a fixed order.
//
// forward: same decomposition as linear_fwd_kernel (a thread owns one 4-channel vector of a pixel
// for 8 images), partial[s][img][j].
// Optional fused loss tail (small batches, where a separate 9 us loss kernel is a tenth of the step):
// the block that finishes last (ticket counter, like softmax_xent_kernel) adds the slices and the bias
// in slice order and computes loss, dlogits and the bias gradient for the whole batch — the same
// arithmetic as softmax_xent_kernel with its own fixed summation order (deterministic; the two routes
// agree to fp32 rounding).  Calls on one device must not overlap (one ticket word).
constexpr int kFusedXentMaxBatch = 128;
static __device__ unsigned int g_linfwd_ticket;
struct XentArgs {
  const float* bias;
  const long long* labels;
  float* logits;
  float* dlogits;
  float* dbias;
  float* loss;
};

// IMG = images per thread: 8 (80 accumulators, one CTA per SM) for small batches, 4 (two CTAs per SM:
// twice the loads in flight) for large ones, where the kernel is bound by DRAM latency, not by FMAs.
template <int NOUT, int IMG>
__global__ void __launch_bounds__(256, IMG == 8 ? 1 : 2)
linear_fwd_f32_kernel(const float4* __restrict__ act, const float* __restrict__ W,
                      float* __restrict__ partial, Geom g, int C, int vec_per_slice, XentArgs xe) {
  pdl_sync();
  __shared__ float red[8][IMG * NOUT];
  const int HW = g.H * g.W;
  const int nvec = (C / 4) * HW;
  const long long K = (long long)C * HW;
  const int img0 = blockIdx.y * IMG;
  const int v_begin = blockIdx.x * vec_per_slice;
  const int v_end = min(nvec, v_begin + vec_per_slice);
  static_assert(IMG * NOUT == 80 || IMG * NOUT == 40, "warp reductions exist for 80 and 40 accumulators");
  float acc[IMG * NOUT];
#pragma unroll
  for (int i = 0; i < IMG * NOUT; ++i) acc[i] = 0.f;

  for (int v = v_begin + threadIdx.x; v < v_end; v += blockDim.x) {
    const int chunk = v / HW;
    const int pix = v - chunk * HW;
    const int h = pix / g.W, wc = pix - h * g.W;
    const long long rowi = (long long)chunk * g.plane_rows + kGuard + (long long)(h + 1) * g.PW + (wc + 1);
    float4 av[IMG];
#pragma unroll
    for (int i = 0; i < IMG; ++i) {
      const int img = img0 + i;
      av[i] = img < g.batch ? act[rowi + (long long)img * g.PHW] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float wv[NOUT][4];
#pragma unroll
    for (int j = 0; j < NOUT; ++j)
#pragma unroll
      for (int c = 0; c < 4; ++c) wv[j][c] = W[j * K + (long long)(chunk * 4 + c) * HW + pix];
#pragma unroll
    for (int i = 0; i < IMG; ++i) {
      const float a[4] = {av[i].x, av[i].y, av[i].z, av[i].w};
#pragma unroll
      for (int j = 0; j < NOUT; ++j)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[i * NOUT + j] = fmaf(wv[j][c], a[c], acc[i * NOUT + j]);
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if constexpr (IMG * NOUT == 80) {
    warp_reduce80(acc, lane);
    if ((lane & 1) == 0) {
      const int base = warp_reduce80_base(lane);
#pragma unroll
      for (int i = 0; i < 5; ++i) red[warp][base + i] = acc[i];
    }
  } else {
    warp_reduce40(acc, lane);
    if ((lane & 3) == 0) {
      const int base = warp_reduce40_base(lane);
#pragma unroll
      for (int i = 0; i < 5; ++i) red[warp][base + i] = acc[i];
    }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < IMG * NOUT; t += blockDim.x) {
    const int i = t / NOUT, j = t - i * NOUT;
    const int img = img0 + i;
    if (img >= g.batch) continue;
    float s = 0.f;
#pragma unroll
    for (int wdx = 0; wdx < 8; ++wdx) s += red[wdx][t];
    partial[((long long)blockIdx.x * g.batch + img) * NOUT + j] = s;
  }
  if (xe.logits == nullptr) return;
  // ---- fused loss tail
  __shared__ bool is_last;
  __shared__ float zs[kFusedXentMaxBatch * NOUT];
  __shared__ float ls[kFusedXentMaxBatch];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int nblk = gridDim.x * gridDim.y;
    const unsigned int t = atomicAdd(&g_linfwd_ticket, 1u);
    is_last = (t == nblk - 1);
    if (is_last) g_linfwd_ticket = 0;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const int S = gridDim.x, batch = g.batch;
  const int npair = batch * NOUT;
  for (int p = threadIdx.x; p < npair; p += blockDim.x) {
    float z = 0.f;
    for (int sl = 0; sl < S; ++sl) z += __ldcg(partial + (long long)sl * npair + p);
    z += xe.bias[p % NOUT];
    zs[p] = z;
    xe.logits[p] = z;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < batch; i += blockDim.x) {
    const float inv = 1.0f / static_cast<float>(batch);
    float z[NOUT];
#pragma unroll
    for (int j = 0; j < NOUT; ++j) z[j] = zs[i * NOUT + j];
    float m = z[0];
#pragma unroll
    for (int j = 1; j < NOUT; ++j) m = fmaxf(m, z[j]);
    float se = 0.f;
#pragma unroll
    for (int j = 0; j < NOUT; ++j) se += expf(z[j] - m);
    const float lse = m + logf(se);
    const int y = static_cast<int>(xe.labels[i]);
    float lsum = 0.f;
#pragma unroll
    for (int j = 0; j < NOUT; ++j) {
      const float pr = expf(z[j] - lse);
      const float d = (pr - (j == y ? 1.f : 0.f)) * inv;
      xe.dlogits[i * NOUT + j] = d;
      zs[i * NOUT + j] = d;
      if (j == y) lsum = lse - z[j];
    }
    ls[i] = lsum;
  }
  __syncthreads();
  if (threadIdx.x < NOUT) {  // bias gradient: column sums in image order
    float sacc = 0.f;
    for (int i = 0; i < batch; ++i) sacc += zs[i * NOUT + threadIdx.x];
    xe.dbias[threadIdx.x] = sacc;
  } else if (threadIdx.x == 32) {
    float sacc = 0.f;
    for (int i = 0; i < batch; ++i) sacc += ls[i];
    xe.loss[0] = sacc / static_cast<float>(batch);
  }
}

// backward: dact[img][k] = relu'(act[img][k]) * sum_j dlogits[img][j] W[j][k]; dW[j][k] =
// sum_img dlogits[img][j] act[img][k] through per-image-split partials.  A thread owns one 4-channel
// vector of one pixel.  dact is what the conv2 gradient GEMMs consume: SPLIT_OUT writes it as split
// planes (hi + lo bf16, half a 16-byte row per thread and set), otherwise as fp32 planes rounded to
// tf32.
// DO_DX / DO_DW select the halves: the step engine runs the dX half (with the bias sums) on the main
// stream — only dact is needed by the convolution backward that follows — and the dW half on the fl
// bucket's side stream, where it overlaps the convolution backward.
template <int NOUT, bool SPLIT_OUT, bool DO_DX, bool DO_DW>
__global__ void __launch_bounds__(256, (DO_DX && DO_DW) ? 2 : 3)
linear_bwd_f32_kernel(const float4* __restrict__ act, const float* __restrict__ dlogits,
                      const float* __restrict__ W, float4* __restrict__ dact,
                      float* __restrict__ dw_partial, float* __restrict__ bias_part, Geom g, int C,
                      int imgs_per_split) {
  pdl_sync();
  extern __shared__ float sdl[];  // [imgs_per_split][NOUT]
  __shared__ float bred[8][8];
  static_assert(kBiasBlock == 256, "one bias block per thread block");
  constexpr int PF = DO_DX ? 4 : 8;  // image vectors kept in flight per thread (register budget)
  const int HW = g.H * g.W;
  const int nvec = (C / 4) * HW;
  const long long K = (long long)C * HW;
  const int img_begin = blockIdx.y * imgs_per_split;
  const int img_end = min(g.batch, img_begin + imgs_per_split);
  for (int t = threadIdx.x; t < (img_end - img_begin) * NOUT; t += blockDim.x)
    sdl[t] = dlogits[(long long)img_begin * NOUT + t];
  __syncthreads();
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = v < nvec;
  const int vv = active ? v : nvec - 1;
  const int chunk = vv / HW;
  const int pix = vv - chunk * HW;
  const int h = pix / g.W, wc = pix - h * g.W;
  float wv[NOUT][4], dwv[NOUT][4];
#pragma unroll
  for (int j = 0; j < NOUT; ++j)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      wv[j][c] = DO_DX ? W[j * K + (long long)(chunk * 4 + c) * HW + pix] : 0.f;
      dwv[j][c] = 0.f;
    }
  float bsum[4] = {0.f, 0.f, 0.f, 0.f};  // sum over this split's images of the masked dX (exact fp32)
  const long long base = (long long)chunk * g.plane_rows + kGuard + (long long)(h + 1) * g.PW + (wc + 1);
  if (active) {
    for (int i0 = img_begin; i0 < img_end; i0 += PF) {
      float4 av[PF];
#pragma unroll
      for (int u = 0; u < PF; ++u)
        av[u] = (i0 + u < img_end) ? act[base + (long long)(i0 + u) * g.PHW] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        if (i0 + u >= img_end) break;
        const float a[4] = {av[u].x, av[u].y, av[u].z, av[u].w};
        float gr[4] = {0.f, 0.f, 0.f, 0.f};
        const float* dl = sdl + (i0 + u - img_begin) * NOUT;
#pragma unroll
        for (int j = 0; j < NOUT; ++j) {
          const float d = dl[j];
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            if (DO_DX) gr[c] = fmaf(d, wv[j][c], gr[c]);
            if (DO_DW) dwv[j][c] = fmaf(d, a[c], dwv[j][c]);
          }
        }
        if constexpr (!DO_DX) continue;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          gr[c] = a[c] > 0.f ? gr[c] : 0.f;
          bsum[c] += gr[c];
        }
        if constexpr (SPLIT_OUT) {
          float hi[4], lo[4];
#pragma unroll
          for (int c = 0; c < 4; ++c) split_bf16(gr[c], hi[c], lo[c]);
          // bf16 plane chunk/2, row = the same pixel row, 8-byte half (chunk & 1); lo set C/8 planes on
          uint2* d2 = reinterpret_cast<uint2*>(dact);
          const long long r16 = (long long)(chunk >> 1) * g.plane_rows + kGuard +
                                (long long)(i0 + u) * g.PHW + (long long)(h + 1) * g.PW + (wc + 1);
          d2[r16 * 2 + (chunk & 1)] = make_uint2(pack_bf16x2(hi[0], hi[1]), pack_bf16x2(hi[2], hi[3]));
          d2[(r16 + (long long)(C / 8) * g.plane_rows) * 2 + (chunk & 1)] =
              make_uint2(pack_bf16x2(lo[0], lo[1]), pack_bf16x2(lo[2], lo[3]));
        } else {
          dact[base + (long long)(i0 + u) * g.PHW] =
              make_float4(round_tf32(gr[0]), round_tf32(gr[1]), round_tf32(gr[2]), round_tf32(gr[3]));
        }
      }
    }
    if constexpr (DO_DW) {
      float* dst = dw_partial + (long long)blockIdx.y * NOUT * K;
#pragma unroll
      for (int j = 0; j < NOUT; ++j)
#pragma unroll
        for (int c = 0; c < 4; ++c) dst[j * K + (long long)(chunk * 4 + c) * HW + pix] = dwv[j][c];
    }
  }
  if (!DO_DX || bias_part == nullptr) return;
  // fixed-shape reduction of the block's bias sums: a block spans at most two 4-channel groups
  const int slot = chunk - (blockIdx.x * kBiasBlock) / HW;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int sl = 0; sl < 2; ++sl)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float t = warp_sum((active && slot == sl) ? bsum[c] : 0.f);
      if (lane == 0) bred[warp][sl * 4 + c] = t;
    }
  __syncthreads();
  if (threadIdx.x < 8) {
    float t = 0.f;
#pragma unroll
    for (int wdx = 0; wdx < 8; ++wdx) t += bred[wdx][threadIdx.x];
    bias_part[((long long)blockIdx.y * gridDim.x + blockIdx.x) * 8 + threadIdx.x] = t;
  }
}

// ------------------------------------------------------------------ SGD (single-GPU path)
// p -= lr * g.  lr lives in device memory so a captured graph can be replayed with a new value.
__global__ void sgd_kernel(float* __restrict__ p, const float* __restrict__ grad,
                           const float* __restrict__ lr, long long n) {
  pdl_sync();
  const float l = *lr;
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(grad)) & 15) == 0;
  const long long n4 = vec ? n / 4 : 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float4 a = reinterpret_cast<float4*>(p)[i];
    const float4 g4 = reinterpret_cast<const float4*>(grad)[i];
    a.x = fmaf(-l, g4.x, a.x);
    a.y = fmaf(-l, g4.y, a.y);
    a.z = fmaf(-l, g4.z, a.z);
    a.w = fmaf(-l, g4.w, a.w);
    reinterpret_cast<float4*>(p)[i] = a;
  }
  for (long long i = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    p[i] = fmaf(-l, grad[i], p[i]);
}

}  // namespace ddpb

using namespace ddpb;

static int g_num_sms[64] = {};

extern "C" {

int ddpb_num_sms() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (g_num_sms[dev & 63] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    g_num_sms[dev & 63] = n;
  }
  return g_num_sms[dev & 63];
}

long long ddpb_plane_rows(int batch_cap, int H, int W) { return plane_rows_for(batch_cap, H, W); }
int ddpb_guard_rows() { return kGuard; }
int ddpb_tile_rows() { return kTileM; }

int ddpb_gather_u8(const void* images, const void* labels, const void* idx, float* x, void* y,
                   int batch, int pix, void* stream) {
  if (batch <= 0) return 0;
  gather_u8_kernel<<<batch, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(images), static_cast<const long long*>(labels),
      static_cast<const long long*>(idx), x, static_cast<long long*>(y), batch, pix);
  return static_cast<int>(cudaGetLastError());
}

int ddpb_conv1_fwd(const float* x, const float* w, const float* b, void* out, int cin, int cout,
                   int batch, int batch_cap, int H, int W, void* stream) {
  if (!(cin == 1 && cout == 32)) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  int bx = batch < ddpb_num_sms() ? batch : ddpb_num_sms();
  const size_t smem = (size_t)kConv1Group * cin * g.PHW * sizeof(float);
  conv1_fwd_kernel<1, 32><<<dim3(bx, cout / 8), 256, smem, static_cast<cudaStream_t>(stream)>>>(
      x, w, b, static_cast<__nv_bfloat16*>(out), g);
  return static_cast<int>(cudaGetLastError());
}

static int conv1_wgrad_blocks(int batch, int H, int W) {
  // one block per image up to 64 blocks: every block ends in an 80-value reduction and its
  // partial must be re-read, so more blocks than that only adds reduction work
  (void)H; (void)W;
  int bx = batch < 64 ? batch : 64;
  if (bx < 1) bx = 1;
  return bx;
}
size_t ddpb_conv1_wgrad_scratch_bytes(int cin, int cout, int batch_cap, int H, int W) {
  return static_cast<size_t>(conv1_wgrad_blocks(batch_cap, H, W)) * (cout * cin * 9 + cout) *
         sizeof(float);
}
// dw and db must be contiguous (dw[cout*cin*9] immediately followed by db[cout]); the gradient
// arena lays parameters out that way.
int ddpb_conv1_wgrad(const void* dact, const float* x, float* dw_db, float* scratch, int cin,
                     int cout, int batch, int batch_cap, int H, int W, void* stream) {
  if (!(cin == 1 && cout == 32)) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int bx = conv1_wgrad_blocks(batch, H, W);
  conv1_wgrad_kernel<1, 32, false><<<dim3(bx, cout / 8), 256, ((size_t)kConv1Group * cin * g.PHW + 2 * kConv1Guard) * sizeof(float), st>>>(
      dact, x, scratch, g);
  DDPB_CUDA_OK(cudaGetLastError());
  const long long n = (long long)cout * cin * 9 + cout;
  reduce_partials_kernel<<<static_cast<int>((n / 4 + 255) / 256 + 1), 256, 0, st>>>(scratch, bx, n,
                                                                                    dw_db);
  return static_cast<int>(cudaGetLastError());
}

int ddpb_reduce_partials(const float* part, int nparts, long long n, float* out, void* stream);

// ---- exact-fp32 conv1 weight gradient from fp32 planes (reference-precision mode).
// Partials: scratch[nparts][cout*cin*9 + cout] (same layout as the tensor-core version, so the fused
// reduce + all-reduce + SGD kernel consumes them unchanged); scratch size:
// ddpb_conv1_wgrad_f32_scratch_bytes.
static int conv1_wgrad_f32_blocks(int batch) {
  // x-blocks (each times 8 channel groups in y).  Three CTAs fit an SM (444 slots on 148 SMs): small
  // batches get one image per block, large ones 55 x 8 = 440 blocks with an even share of images each
  // (every block ends in a 40-value reduction whose partial the update kernel re-reads).
  if (batch <= 74) return batch < 1 ? 1 : batch;
  const int ipb = (batch + 54) / 55;
  return (batch + ipb - 1) / ipb;
}
size_t ddpb_conv1_wgrad_f32_scratch_bytes(int cin, int cout) {
  return static_cast<size_t>(74) * (cout * cin * 9 + cout) * sizeof(float);
}
int ddpb_conv1_wgrad_f32_partials(const void* dact, const float* x, float* scratch, int* nparts_out,
                                  int cin, int cout, int batch, int batch_cap, int H, int W,
                                  void* stream) {
  if (!(cin == 1 && cout == 32)) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  const int bx = conv1_wgrad_f32_blocks(batch);
  if (g.PW + 1 > kConv1Guard) return DDPB_ERR_SHAPE;
  DDPB_CUDA_OK(launch_k(conv1_wgrad_kernel<1, 32, true>, dim3(bx, cout / 4), dim3(256),
                        ((size_t)kConv1Group * cin * g.PHW + 2 * kConv1Guard) * sizeof(float),
                        static_cast<cudaStream_t>(stream), dact, x, scratch, g));
  *nparts_out = bx;
  return 0;
}
int ddpb_conv1_wgrad_f32(const void* dact, const float* x, float* dw_db, float* scratch, int cin,
                         int cout, int batch, int batch_cap, int H, int W, void* stream) {
  int bx = 0;
  const int rc = ddpb_conv1_wgrad_f32_partials(dact, x, scratch, &bx, cin, cout, batch, batch_cap, H, W,
                                               stream);
  if (rc != 0) return rc;
  return ddpb_reduce_partials(scratch, bx, (long long)cout * cin * 9 + cout, dw_db, stream);
}

// ---- linear layer launch geometry (shared by scratch-size queries and launches)
static void linear_fwd_geom(int batch, int nvec, int& S, int& per, int& groups) {
  constexpr int IMG = 8;
  groups = (batch + IMG - 1) / IMG;
  const int want = ddpb_num_sms() / groups;  // one CTA per SM (255 registers): stay within one wave
  const int maxS = (nvec + 255) / 256;
  S = want < 1 ? 1 : (want > maxS ? maxS : want);
  per = ((nvec + S - 1) / S + 255) / 256 * 256;
  S = (nvec + per - 1) / per;
}
size_t ddpb_linear_fwd_scratch_bytes(int nout, int C, int batch_cap, int H, int W) {
  // S is largest for the smallest batch; bound it by the feature-vector block count.
  const int nvec = (C / 8) * H * W;
  const int maxS = (nvec + 255) / 256;
  return static_cast<size_t>(maxS) * batch_cap * nout * sizeof(float);
}
int ddpb_linear_fwd(const void* act, const float* Wt, float* partial, int* S_out, int nout, int C,
                    int batch, int batch_cap, int H, int W, void* stream) {
  if (nout != 10 || (C % 8) != 0) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  const int nvec = (C / 8) * H * W;
  int S, per, groups;
  linear_fwd_geom(batch, nvec, S, per, groups);
  if (S_out) *S_out = S;
  return static_cast<int>(launch_k(linear_fwd_kernel<10, 8>, dim3(S, groups), dim3(256), 0,
                                   static_cast<cudaStream_t>(stream),
                                   static_cast<const __nv_bfloat16*>(act), Wt, partial, g, C, per));
}

static int softmax_xent_launch(const float* partial, int S, const float* bias, const void* labels,
                               float* logits, float* dlogits, float* dbias, float* loss, int nout,
                               int batch, void* dl_packed, void* stream);

int ddpb_softmax_xent(const float* partial, int S, const float* bias, const void* labels,
                      float* logits, float* dlogits, float* dbias, float* loss, int nout,
                      int batch, void* stream) {
  return softmax_xent_launch(partial, S, bias, labels, logits, dlogits, dbias, loss, nout, batch,
                             nullptr, stream);
}

// Same, and also writes dlogits as bf16 hi/lo tiles ((batch rounded up to 128) * 64 bytes) for
// ddpb_linear_bwd_tc.
int ddpb_softmax_xent_packed(const float* partial, int S, const float* bias, const void* labels,
                             float* logits, float* dlogits, float* dbias, float* loss, int nout,
                             int batch, void* dl_packed, void* stream) {
  return softmax_xent_launch(partial, S, bias, labels, logits, dlogits, dbias, loss, nout, batch,
                             dl_packed, stream);
}

static int softmax_xent_launch(const float* partial, int S, const float* bias, const void* labels,
                               float* logits, float* dlogits, float* dbias, float* loss, int nout,
                               int batch, void* dl_packed, void* stream) {
  if (nout != 10 || batch <= 0 || logits == nullptr) return DDPB_ERR_SHAPE;
  if (dl_packed == nullptr && batch * 10 <= 1024) {  // one block does it all
    const int threads = (batch * 10 + 31) / 32 * 32 < 64 ? 64 : (batch * 10 + 31) / 32 * 32;
    return static_cast<int>(launch_k(softmax_xent_small_kernel<10>, dim3(1), dim3(threads), 0,
                                     static_cast<cudaStream_t>(stream), partial, S, bias,
                                     static_cast<const long long*>(labels), logits, dlogits, dbias, loss,
                                     batch));
  }
  // with packed output the grid also covers the zero rows up to the end of the last 128-image tile
  const int covered = dl_packed ? (batch + 127) / 128 * 128 : batch;
  const int blocks = (covered + kXentImgs - 1) / kXentImgs;
  if (blocks > kXentMaxBlocks) return DDPB_ERR_LIMIT;
  return static_cast<int>(launch_k(softmax_xent_kernel<10>, dim3(blocks), dim3(kXentImgs * 10 * kXentSG), 0,
                                   static_cast<cudaStream_t>(stream), partial, S, bias,
                                   static_cast<const long long*>(labels), logits, dlogits, dbias, loss,
                                   batch, static_cast<uint4*>(dl_packed)));
}

int ddpb_reduce_partials(const float* part, int nparts, long long n, float* out, void* stream);

static int linear_bwd_imgs_per_split(int batch) {
  // enough splits to fill the GPU at small batch, few enough that partial traffic stays small
  if (batch <= 64) return 16;  // (8 was tried: 392 blocks no longer fit one wave, +8 us on the step)
  if (batch <= 512) return 32;
  return 128;
}
size_t ddpb_linear_bwd_scratch_bytes(int nout, int C, int batch_cap, int H, int W) {
  size_t worst = 0;
  const int cands[3] = {64 < batch_cap ? 64 : batch_cap, 512 < batch_cap ? 512 : batch_cap, batch_cap};
  for (int i = 0; i < 3; ++i) {
    const int ips = linear_bwd_imgs_per_split(cands[i]);
    const size_t splits = (cands[i] + ips - 1) / ips;
    if (splits > worst) worst = splits;
  }
  return worst * nout * static_cast<size_t>(C) * H * W * sizeof(float);
}
int ddpb_linear_bwd(const void* act, const float* dlogits, const float* Wt, void* dact, float* dW,
                    float* scratch, int nout, int C, int batch, int batch_cap, int H, int W,
                    void* stream) {
  if (nout != 10 || (C % 8) != 0) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nvec = (C / 4) * H * W;  // half-vectors
  const int ips = linear_bwd_imgs_per_split(batch);
  const int splits = (batch + ips - 1) / ips;
  if (g.ntiles * kTileM > batch * g.PHW) {
    DDPB_CUDA_OK(launch_k(zero_tail_rows_kernel, dim3(8), dim3(256), 0, st,
                          static_cast<uint4*>(dact), C / 8, g));
  }
  DDPB_CUDA_OK(launch_k(linear_bwd_kernel<10>, dim3((nvec + 255) / 256, splits), dim3(256),
                        ips * 10 * sizeof(float), st, static_cast<const __nv_bfloat16*>(act), dlogits,
                        Wt, static_cast<__nv_bfloat16*>(dact), scratch, g, C, ips));
  return ddpb_reduce_partials(scratch, splits, (long long)nout * C * H * W, dW, stream);
}

}  // extern "C"
namespace ddpb {
// Large-batch linear forward (exact fp32): the kernel above gives every block of IMG images the whole
// reduction, so at batch 4096 each of 1024 blocks streams the full 2 MB weight from L2 (~2 GB of L2 traffic
// next to 0.9 GB of activations; measured 385 us).  Here a block owns ONE 256-vector slice of the reduction,
// keeps that slice of the weight in shared memory (40 KB) and walks the image groups: a warp takes a group of
// four images, its lanes stride over the slice (8 iterations of 4 activation loads + 10 LDS.128 + 160 FMA),
// reduces its 40 sums inside the warp and writes partial[slice][img][j] — the layout the loss kernel sums.
constexpr int kLinBigSlice = 256;   // 16-byte activation vectors (4 channels of one pixel) per slice
template <int NOUT>
__global__ void __launch_bounds__(256, 2)
linear_fwd_f32_big_kernel(const float4* __restrict__ act, const float* __restrict__ W,
                          float* __restrict__ partial, Geom g, int C) {
  static_assert(NOUT == 10, "warp reduction over 4 x 10 accumulators");
  __shared__ float4 ws[NOUT][kLinBigSlice];   // ws[j][v] = W[j][4 channels of vector v]
  pdl_sync();
  const int HW = g.H * g.W;
  const int nvec = (C / 4) * HW;
  const long long K = (long long)C * HW;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int v0 = blockIdx.x * kLinBigSlice;
  {
    const int v = v0 + threadIdx.x;
    const int chunk = v / HW, pix = v - chunk * HW;
#pragma unroll
    for (int j = 0; j < NOUT; ++j) {
      float4 w4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (v < nvec) {
        const float* wp = W + j * K + (long long)(chunk * 4) * HW + pix;
        w4 = make_float4(wp[0], wp[HW], wp[2 * (long long)HW], wp[3 * (long long)HW]);
      }
      ws[j][threadIdx.x] = w4;
    }
  }
  // plane row (image 0) of every vector of the slice; -1 beyond the end of the reduction
  __shared__ int s_row[kLinBigSlice];
  {
    const int v = v0 + threadIdx.x;
    const int chunk = v / HW, pix = v - chunk * HW;
    const int h = pix / g.W, wc = pix - h * g.W;
    s_row[threadIdx.x] = v < nvec ? (int)((long long)chunk * g.plane_rows + kGuard + (h + 1) * g.PW + (wc + 1)) : -1;
  }
  __syncthreads();
  const int groups = (g.batch + 3) / 4;
  for (int grp = blockIdx.y * 8 + warp; grp < groups; grp += gridDim.y * 8) {
    const int img0 = grp * 4;
    float acc[4 * NOUT];
#pragma unroll
    for (int i = 0; i < 4 * NOUT; ++i) acc[i] = 0.f;
#pragma unroll 2
    for (int it = 0; it < kLinBigSlice / 32; ++it) {
      const int row = s_row[it * 32 + lane];
      float4 av[4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
        av[i] = (row >= 0 && img0 + i < g.batch) ? act[row + (long long)(img0 + i) * g.PHW]
                                                 : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int j = 0; j < NOUT; ++j) {
        const float4 w4 = ws[j][it * 32 + lane];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float s = acc[i * NOUT + j];
          s = fmaf(w4.x, av[i].x, s);
          s = fmaf(w4.y, av[i].y, s);
          s = fmaf(w4.z, av[i].z, s);
          s = fmaf(w4.w, av[i].w, s);
          acc[i * NOUT + j] = s;
        }
      }
    }
    warp_reduce40(acc, lane);
    if ((lane & 3) == 0) {
      const int base = warp_reduce40_base(lane);
#pragma unroll
      for (int ii = 0; ii < 5; ++ii) {
        const int tt = base + ii, i = tt / NOUT, j = tt - i * NOUT;
        if (img0 + i < g.batch) partial[((long long)blockIdx.x * g.batch + img0 + i) * NOUT + j] = acc[ii];
      }
    }
  }
}

}  // namespace ddpb
extern "C" {
// ---- fp32-plane (precision-parity) linear layer.  Scratch sizes: ddpb_linear_fwd_f32_scratch_bytes /
// ddpb_linear_bwd_scratch_bytes (the partial layout of the backward is the same as the bf16 one).
static int linear_fwd_f32_img(int batch) { return batch >= 512 ? 4 : 8; }
// batch >= 512: the sliced kernel with the weight slice in shared memory (DDPB_LINFWD_BIG=0: the old one)
static bool linear_fwd_f32_big(int batch) {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("DDPB_LINFWD_BIG");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on == 1 && batch >= 512;
}
static void linear_fwd_f32_geom(int batch, int nvec, int& S, int& per, int& groups) {
  const int IMG = linear_fwd_f32_img(batch);
  groups = (batch + IMG - 1) / IMG;
  const int want = (IMG == 8 ? 1 : 2) * ddpb_num_sms() / groups;  // one wave of resident CTAs
  const int maxS = (nvec + 255) / 256;
  S = want < 1 ? 1 : (want > maxS ? maxS : want);
  per = ((nvec + S - 1) / S + 255) / 256 * 256;
  S = (nvec + per - 1) / per;
}
size_t ddpb_linear_fwd_f32_scratch_bytes(int nout, int C, int batch_cap, int H, int W) {
  const int nvec = (C / 4) * H * W;
  const int maxS = (nvec + 255) / 256;
  return static_cast<size_t>(maxS) * batch_cap * nout * sizeof(float);
}
int ddpb_linear_fwd_f32(const void* act, const float* Wt, float* partial, int* S_out, int nout, int C,
                        int batch, int batch_cap, int H, int W, void* stream) {
  if (nout != 10 || (C % 4) != 0) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  const int nvec = (C / 4) * H * W;
  int S, per, groups;
  if (linear_fwd_f32_big(batch)) {
    S = (nvec + kLinBigSlice - 1) / kLinBigSlice;  // == the maxS the scratch is sized for
    if (S_out) *S_out = S;
    const int ngroups = (batch + 3) / 4;
    int by = (2 * ddpb_num_sms() + S - 1) / S;     // about two resident blocks per SM over all slices
    if (by * 8 > ngroups) by = (ngroups + 7) / 8;
    if (by < 1) by = 1;
    return static_cast<int>(launch_k(linear_fwd_f32_big_kernel<10>, dim3(S, by), dim3(256), 0,
                                     static_cast<cudaStream_t>(stream), static_cast<const float4*>(act), Wt,
                                     partial, g, C));
  }
  linear_fwd_f32_geom(batch, nvec, S, per, groups);
  if (S_out) *S_out = S;
  if (linear_fwd_f32_img(batch) == 4)
    return static_cast<int>(launch_k(linear_fwd_f32_kernel<10, 4>, dim3(S, groups), dim3(256), 0,
                                     static_cast<cudaStream_t>(stream), static_cast<const float4*>(act),
                                     Wt, partial, g, C, per, XentArgs{}));
  return static_cast<int>(launch_k(linear_fwd_f32_kernel<10, 8>, dim3(S, groups), dim3(256), 0,
                                   static_cast<cudaStream_t>(stream), static_cast<const float4*>(act),
                                   Wt, partial, g, C, per, XentArgs{}));
}
// Linear forward with the loss fused into its tail (what ddpb_linear_fwd_f32 followed by
// ddpb_softmax_xent compute, one launch): batch <= ddpb_linear_fwd_f32_xent_max_batch().
int ddpb_linear_fwd_f32_xent_max_batch() { return kFusedXentMaxBatch; }
int ddpb_linear_fwd_f32_xent(const void* act, const float* Wt, float* partial, const float* bias,
                             const void* labels, float* logits, float* dlogits, float* dbias, float* loss,
                             int nout, int C, int batch, int batch_cap, int H, int W, void* stream) {
  if (nout != 10 || (C % 4) != 0 || logits == nullptr) return DDPB_ERR_SHAPE;
  if (batch > kFusedXentMaxBatch) return DDPB_ERR_LIMIT;
  const Geom g = make_geom(batch, batch_cap, H, W);
  const int nvec = (C / 4) * H * W;
  int S, per, groups;
  linear_fwd_f32_geom(batch, nvec, S, per, groups);
  XentArgs xe{bias, static_cast<const long long*>(labels), logits, dlogits, dbias, loss};
  return static_cast<int>(launch_k(linear_fwd_f32_kernel<10, 8>, dim3(S, groups), dim3(256), 0,
                                   static_cast<cudaStream_t>(stream), static_cast<const float4*>(act),
                                   Wt, partial, g, C, per, xe));
}
// act: fp32 planes.  dact: split planes (split_out != 0: 2*C/8 bf16 planes) or fp32 planes rounded
// to tf32 (C/4 planes).  bias_part (optional, ddpb_linear_bwd_f32_bias_part_bytes bytes): receives the
// exact-fp32 partial sums of dact per channel (layout.cuh: sum_bias_parts) — the bias gradient of the
// convolution that produced act; *nsplit_out / *nbx_out (host) receive its two extents.
size_t ddpb_linear_bwd_f32_bias_part_bytes(int C, int batch_cap, int H, int W) {
  const int nvec = (C / 4) * H * W;
  size_t worst = 0;
  const int cands[3] = {64 < batch_cap ? 64 : batch_cap, 512 < batch_cap ? 512 : batch_cap, batch_cap};
  for (int i = 0; i < 3; ++i) {
    const int ips = linear_bwd_imgs_per_split(cands[i]);
    const size_t splits = (cands[i] + ips - 1) / ips;
    if (splits > worst) worst = splits;
  }
  // x2: the dX-only launch (mode 1) uses twice the image splits of the combined one
  return 2 * worst * ((nvec + kBiasBlock - 1) / kBiasBlock) * 8 * sizeof(float);
}
}  // extern "C"
template <bool DX, bool DW>
static int linear_bwd_f32_launch(bool split_out, dim3 grid, size_t smem, cudaStream_t st, const float4* act,
                                 const float* dl, const float* Wt, float4* dact, float* scratch,
                                 float* bias_part, const Geom& g, int C, int ips) {
  if (split_out)
    return static_cast<int>(launch_k(linear_bwd_f32_kernel<10, true, DX, DW>, grid, dim3(256), smem, st, act,
                                     dl, Wt, dact, scratch, bias_part, g, C, ips));
  return static_cast<int>(launch_k(linear_bwd_f32_kernel<10, false, DX, DW>, grid, dim3(256), smem, st, act,
                                   dl, Wt, dact, scratch, bias_part, g, C, ips));
}
extern "C" {
// mode 0: everything (dact, bias parts, dW).  mode 1: dact + bias parts only.  mode 2: the dW half
// only (partials in scratch; summed into dW unless dW == NULL).  Modes 1 and 2 together equal mode 0.
int ddpb_linear_bwd_f32_mode(const void* act, const float* dlogits, const float* Wt, void* dact, float* dW,
                             float* scratch, int nout, int C, int batch, int batch_cap, int H, int W,
                             int split_out, float* bias_part, int* nsplit_out, int* nbx_out, int mode,
                             void* stream) {
  if (nout != 10 || (C % 8) != 0 || mode < 0 || mode > 2) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nvec = (C / 4) * H * W;
  // the dX half writes no partials, so it can afford twice the image splits of the dW half
  const int ips_dw = linear_bwd_imgs_per_split(batch);
  const int ips = (mode == 1 && ips_dw >= 16) ? ips_dw / 2 : ips_dw;
  const int splits = (batch + ips - 1) / ips;
  const int nbx = (nvec + kBiasBlock - 1) / kBiasBlock;
  if (nsplit_out) *nsplit_out = splits;
  if (nbx_out) *nbx_out = nbx;
  if (mode != 2 && g.ntiles * kTileM > batch * g.PHW) {
    DDPB_CUDA_OK(launch_k(zero_tail_rows_kernel, dim3(8), dim3(256), 0, st, static_cast<uint4*>(dact),
                          C / 4, g));  // C/4 planes in either format (split: hi + lo sets of C/8)
  }
  const dim3 grid(nbx, splits);
  const size_t smem = ips * 10 * sizeof(float);
  const float4* a4 = static_cast<const float4*>(act);
  float4* d4 = static_cast<float4*>(dact);
  int rc;
  if (mode == 0) rc = linear_bwd_f32_launch<true, true>(split_out, grid, smem, st, a4, dlogits, Wt, d4, scratch, bias_part, g, C, ips);
  else if (mode == 1) rc = linear_bwd_f32_launch<true, false>(split_out, grid, smem, st, a4, dlogits, Wt, d4, scratch, bias_part, g, C, ips);
  else rc = linear_bwd_f32_launch<false, true>(split_out, grid, smem, st, a4, dlogits, Wt, d4, scratch, nullptr, g, C, ips);
  if (rc != 0) return rc;
  // dW == NULL: the per-split partials stay in scratch[*nsplit_out][nout*C*H*W] and the caller sums them
  // with ddpb_reduce_partials where it suits
  if (mode == 1 || dW == nullptr) return 0;
  return ddpb_reduce_partials(scratch, splits, (long long)nout * C * H * W, dW, stream);
}
int ddpb_linear_bwd_f32(const void* act, const float* dlogits, const float* Wt, void* dact, float* dW,
                        float* scratch, int nout, int C, int batch, int batch_cap, int H, int W,
                        int split_out, float* bias_part, int* nsplit_out, int* nbx_out, void* stream) {
  return ddpb_linear_bwd_f32_mode(act, dlogits, Wt, dact, dW, scratch, nout, C, batch, batch_cap, H, W,
                                  split_out, bias_part, nsplit_out, nbx_out, 0, stream);
}

// Zero the rows of `nplanes` planes between the end of the batch and the end of its last tile.
int ddpb_zero_tail_rows(void* planes, int nplanes, int batch, int batch_cap, int H, int W,
                        void* stream) {
  const Geom g = make_geom(batch, batch_cap, H, W);
  if (g.ntiles * kTileM <= batch * g.PHW) return 0;
  return static_cast<int>(launch_k(zero_tail_rows_kernel, dim3(8), dim3(256), 0,
                                   static_cast<cudaStream_t>(stream),
                                   static_cast<uint4*>(planes), nplanes, g));
}

// out[i] = sum_p part[p*n + i] in a fixed order (shared by the split-K style kernels).
int ddpb_reduce_partials(const float* part, int nparts, long long n, float* out, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (nparts >= 16 && n <= (1 << 16))
    return static_cast<int>(launch_k(reduce_partials_wide_kernel, dim3(static_cast<int>((n + 31) / 32)),
                                     dim3(256), 0, st, part, nparts, n, out));
  return static_cast<int>(launch_k(reduce_partials_kernel,
                                   dim3(static_cast<int>((n / 4 + 255) / 256 + 1)), dim3(256), 0, st, part,
                                   nparts, n, out));
}

int ddpb_sgd(float* p, const float* grad, const float* lr_dev, long long n, void* stream) {
  long long blocks = (n / 4 + 255) / 256;
  if (blocks > 592) blocks = 592;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(launch_k(sgd_kernel, dim3(static_cast<int>(blocks)), dim3(256), 0,
                                   static_cast<cudaStream_t>(stream), p, grad, lr_dev, n));
}

// Programmatic dependent launch for the step kernels: 1 = on, 0 = off, -1 = DDPB_PDL / default.
// It hides launch latency and kernel prologues (a win for small, latency-bound steps) at the price
// of early-resident CTAs competing for SM resources (a small loss for large, bandwidth-bound ones).
void ddpb_set_pdl(int on) { ddpb::g_pdl = on; }
int ddpb_get_pdl() { return ddpb::pdl_enabled(); }

const char* ddpb_error_string(int code) {
  switch (code) {
    case DDPB_ERR_SHAPE: return "ddp_b200: unsupported shape or argument";
    case DDPB_ERR_STATE: return "ddp_b200: call made in the wrong state";
    case DDPB_ERR_LIMIT: return "ddp_b200: size exceeds a fixed limit of this entry point";
    case DDPB_ERR_UNSUPPORTED: return "ddp_b200: not supported by this driver or system";
    default: return cudaGetErrorString(static_cast<cudaError_t>(code));
  }
}

}  // extern "C"

// =================================================================================================
// ddpb_conv1_wgrad_noalloc — the conv1 weight/bias gradient for hosts that cannot hand us scratch.
//
// Who this is for.  ddpb_conv1_wgrad() needs a caller-allocated scratch buffer for its per-block
// partial sums.  Some embedders have no way to provide one: a serving or training host that
// gives ALL device memory to its own pool at start-up (there is nothing left for cudaMalloc to
// return), or a plug-in running inside such a host under a "no driver allocations after
// initialisation" rule, because cudaMalloc/cudaFree synchronise the whole device and would stall
// every other stream of the host.  For them the partial sums live in the fixed array below, which
// is part of this library's device image: every GPU that loads the library holds these 80 KB until
// the process exits, whether or not this entry point is ever called.
//
// Limit.  The array holds kNoAllocMaxBlocks (64) partial blocks; a batch needs rows/2048 of them
// (rows = its padded-plane rows rounded up to whole 128-row tiles), so batches of up to 147 images
// of 28x28 fit.  Larger batches are refused with DDPB_ERR_LIMIT (nothing is launched); use
// ddpb_conv1_wgrad() for those.
//
// Rules of use.
//  * The first call on each device looks the array's device address up (cudaGetSymbolAddress;
//    the array's name is only a host-side stand-in in host code, never a pointer a kernel may
//    be given).  That look-up is set-up work: make the first call on a device before starting
//    any stream capture on it.  Two host threads that make that first call on one device at the
//    same time both store the same address into the per-device slot: harmless, but the slot is not
//    otherwise synchronised, so do not rely on more than that.
//  * No call allocates or frees anything (no cudaMalloc/cudaFree, no library that would).
//  * All calls on a device share the one array, and it keeps whatever the previous call left.
//    Nothing in it has to start at zero: every partial slot the reduction reads is written by the
//    same call first.  For the same reason two calls whose kernels overlap on the GPU (different
//    streams, no dependency) race on it — serialise them.
// =================================================================================================
namespace {
constexpr int kNoAllocMaxBlocks = 64;
constexpr int kNoAllocCin = 1, kNoAllocCout = 32;
constexpr int kNoAllocFloatsPerBlock = kNoAllocCout * kNoAllocCin * 9 + kNoAllocCout;  // 320
constexpr int kNoAllocFloats = kNoAllocMaxBlocks * kNoAllocFloatsPerBlock;            // 20480 = 80 KB
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocFloats];
static float* g_noalloc_devptr[64] = {};  // per device: resolved address of g_noalloc_scratch
}  // namespace

extern "C" int ddpb_conv1_wgrad_noalloc(const void* dact, const float* x, float* dw_db, int cin,
                                        int cout, int batch, int batch_cap, int H, int W,
                                        void* stream) {
  if (!(cin == kNoAllocCin && cout == kNoAllocCout) || batch <= 0 || batch > batch_cap)
    return DDPB_ERR_SHAPE;
  // partial blocks this batch is split into: one per 2048 padded-plane rows, at least 4
  const int nrows = tiles_for(batch, H, W) * kTileM;
  int bx = nrows / (256 * 8);
  if (bx < 4) bx = 4;
  if (bx > kNoAllocMaxBlocks) return DDPB_ERR_LIMIT;

  int dev = 0;
  DDPB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return DDPB_ERR_SHAPE;
  float* scratch = g_noalloc_devptr[dev];
  if (scratch == nullptr) {  // once per device, with that device current
    float* p = nullptr;
    DDPB_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
    if (p == nullptr) return DDPB_ERR_STATE;
    g_noalloc_devptr[dev] = p;
    scratch = p;
  }

  const Geom g = make_geom(batch, batch_cap, H, W);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  conv1_wgrad_kernel<kNoAllocCin, kNoAllocCout, false>
      <<<dim3(bx, cout / 8), 256, ((size_t)kConv1Group * cin * g.PHW + 2 * kConv1Guard) * sizeof(float), st>>>(
          dact, x, scratch, g);
  DDPB_CUDA_OK(cudaGetLastError());
  const long long n = kNoAllocFloatsPerBlock;
  reduce_partials_kernel<<<static_cast<int>((n / 4 + 255) / 256 + 1), 256, 0, st>>>(scratch, bx, n,
                                                                                    dw_db);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ddpb_conv1_wgrad_noalloc_max_batch(int H, int W) {
  // largest batch whose padded rows fit kNoAllocMaxBlocks partial blocks
  int b = 1;
  while (tiles_for(b + 1, H, W) * kTileM / (256 * 8) <= kNoAllocMaxBlocks) ++b;
  return b;
}
