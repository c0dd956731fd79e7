// The linear layer (flattened activation -> NOUT logits) on the tensor cores.
//
// The CUDA-core versions were instruction-bound (profiles/ncu_step_b4096.md: 0.2-0.5 of the
// measured HBM rate, zero tensor-pipe use).  Here the batch dimension is the GEMM M: for a tile of
// 128 images the producer warps gather, with 16-byte cp.async copies, the 8-channel vector of the
// same (channel chunk, pixel) of every image into one shared-memory row group — which is exactly a
// no-swizzle UMMA core matrix (8 images x 16 B).  The same gathered tile is
//   * a K-major A operand   (M = images,  K = features)  for  logits  = act . W^T          (forward)
//   * an MN-major A operand (M = features, K = images)   for  dW      = act^T . dlogits    (backward)
// Weights are pre-packed once per step (pack kernel) into bf16 UMMA tiles, as hi + lo terms in the
// forward so the fp32 master weights keep ~16 mantissa bits (the reference's linear layer is fp32).
//
// Replaces cuBLAS addmm / mm for reference model.py:16,19 and train_ddp.py:199.
#include "ptx.cuh"
#include "layout.cuh"
#include "api_internal.cuh"
#include "launch.cuh"

namespace ddpb {

constexpr int kLinThreads = 288;        // warps 0-3 gather, 4 MMA/TMEM, 5-8 epilogue
constexpr int kLinKB = 32;              // feature chunks (8 channels of one pixel) per pipeline stage
constexpr uint32_t kLinChunkPitch = 16 * 128 + 16;  // bytes between chunks of the gathered tile:
                                                    // 16 image groups x 128 B, +16 B bank skew
constexpr int kLinStages = 2;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const uint32_t n = valid ? 16u : 0u;  // src-size 0 => zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_arrive(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

// ------------------------------------------------------------------------------------- pack
// wf [nkc_pad][32][8]  bf16: row n < NOUT = hi(W[n][k]), row 16+n = lo(W[n][k]), k = (chunk*8+e)*HW+pix
// wb [4][nkc_pad][8][8] bf16: [g][kc][e][j%8]: g = 0,1: hi(W[(g&1)*8 + j%8][k]); g = 2,3: lo(...)   (j >= NOUT: 0)
template <int NOUT>
__global__ void __launch_bounds__(128)
pack_linear_weights_kernel(const float* __restrict__ W, __nv_bfloat16* __restrict__ wf,
                           __nv_bfloat16* __restrict__ wb, int C, int HW, int nkc, int nkc_pad) {
  pdl_sync();
  const int kc = blockIdx.x * blockDim.x + threadIdx.x;
  if (kc >= nkc_pad) return;
  uint4* wf_row = reinterpret_cast<uint4*>(wf + (long long)kc * 32 * 8);
  uint4* wbg[4];
#pragma unroll
  for (int gq = 0; gq < 4; ++gq)
    wbg[gq] = reinterpret_cast<uint4*>(wb + ((long long)gq * nkc_pad + kc) * 64);
  if (kc >= nkc) {
    for (int n = 0; n < 32; ++n) wf_row[n] = make_uint4(0, 0, 0, 0);
    for (int e = 0; e < 8; ++e)
      for (int gq = 0; gq < 4; ++gq) wbg[gq][e] = make_uint4(0, 0, 0, 0);
    return;
  }
  const int chunk = kc / HW, pix = kc - chunk * HW;
  const long long K = (long long)C * HW;
  float w[NOUT][8];
#pragma unroll
  for (int j = 0; j < NOUT; ++j)
#pragma unroll
    for (int e = 0; e < 8; ++e) w[j][e] = W[j * K + (long long)(chunk * 8 + e) * HW + pix];
#pragma unroll
  for (int n = 0; n < 32; ++n) {
    uint4 o = make_uint4(0, 0, 0, 0);
    const int j = n & 15;
    if (j < NOUT) {
      uint32_t p[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float a = w[j][2 * q], b = w[j][2 * q + 1];
        const float ah = __bfloat162float(__float2bfloat16(a)), bh = __bfloat162float(__float2bfloat16(b));
        p[q] = n < 16 ? pack_bf16x2(ah, bh) : pack_bf16x2(a - ah, b - bh);
      }
      o = make_uint4(p[0], p[1], p[2], p[3]);
    }
    wf_row[n] = o;
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    uint32_t hi[8], lo[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float a = 2 * q < NOUT ? w[2 * q < NOUT ? 2 * q : 0][e] : 0.f;
      const float b = 2 * q + 1 < NOUT ? w[2 * q + 1 < NOUT ? 2 * q + 1 : 0][e] : 0.f;
      const float ah = __bfloat162float(__float2bfloat16(a)), bh = __bfloat162float(__float2bfloat16(b));
      hi[q] = pack_bf16x2(ah, bh);
      lo[q] = pack_bf16x2(a - ah, b - bh);
    }
    wbg[0][e] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    wbg[1][e] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
    wbg[2][e] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    wbg[3][e] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
  }
}

// ------------------------------------------------------------------------------------- forward
struct LinFwdSmem {
  static constexpr uint32_t a_bytes = kLinKB * kLinChunkPitch;  // 66,048
  static constexpr uint32_t b_bytes = kLinKB * 512;             // 16,384
  static constexpr uint32_t stage = a_bytes + b_bytes;
  static constexpr uint32_t off_bar = kLinStages * stage;
  static constexpr uint32_t off_tmem = off_bar + (2 * kLinStages + 1) * 8;
  static constexpr uint32_t total = off_tmem + 16;
};

// Gather the 8-channel vectors of chunk kc (this lane's) for the 128 images of the tile:
// image i goes to  dst + (i/8)*128 + (i%8)*16  (one UMMA core matrix per 8 images).
__device__ __forceinline__ void gather_chunk(uint32_t dst_chunk, const __nv_bfloat16* __restrict__ act,
                                             const Geom& g, int kc, int nkc, int HW, int img0, int warp) {
  const bool kc_ok = kc < nkc;
  const int kcc = kc_ok ? kc : 0;
  const int chunk = kcc / HW, pix = kcc - chunk * HW;
  const int h = pix / g.W, w = pix - h * g.W;
  const __nv_bfloat16* src0 =
      act + ((long long)chunk * g.plane_rows + kGuard + (long long)(h + 1) * g.PW + (w + 1)) * 8;
#pragma unroll 8
  for (int i = warp; i < kTileM; i += 4) {
    const int img = img0 + i;
    const bool ok = kc_ok && img < g.batch;
    cp_async16(dst_chunk + (i >> 3) * 128 + (i & 7) * 16, src0 + (long long)(ok ? img : 0) * g.PHW * 8, ok);
  }
}

template <int NOUT>
__global__ void __launch_bounds__(kLinThreads, 1)
linear_fwd_tc_kernel(const __nv_bfloat16* __restrict__ act, const __nv_bfloat16* __restrict__ wf,
                     float* __restrict__ partial, Geom g, int HW, int nkc, int kc_per_slice) {
  using S = LinFwdSmem;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t bar_full = s_base + S::off_bar;
  const uint32_t bar_empty = bar_full + kLinStages * 8;
  const uint32_t bar_done = bar_empty + kLinStages * 8;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + S::off_tmem);
  const int img0 = blockIdx.y * kTileM;
  const int kc_lo = blockIdx.x * kc_per_slice;
  const int nblk = kc_per_slice / kLinKB;

  if (tid == 0) {
    for (int i = 0; i < kLinStages; ++i) {
      mbar_init(bar_full + i * 8, 129);  // 128 gather threads + the weight-tile bulk copy
      mbar_init(bar_empty + i * 8, 1);
    }
    mbar_init(bar_done, 1);
    fence_mbar_init();
  }
  if (warp == 4) {
    tmem_alloc(smem_u32(s_tmem), 32);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  pdl_sync();

  if (warp < 4) {
    for (int it = 0; it < nblk; ++it) {
      const int stage = it % kLinStages;
      const uint32_t phase = (it / kLinStages) & 1;
      const int kc0 = kc_lo + it * kLinKB;
      mbar_wait(bar_empty + stage * 8, phase ^ 1);
      const uint32_t sa = s_base + stage * S::stage;
      if (tid == 0) {
        mbar_arrive_expect_tx(bar_full + stage * 8, S::b_bytes);
        bulk_g2s(sa + S::a_bytes, wf + (long long)kc0 * 256, S::b_bytes, bar_full + stage * 8);
      }
      gather_chunk(sa + lane * kLinChunkPitch, act, g, kc0 + lane, nkc, HW, img0, warp);
      cp_async_arrive(bar_full + stage * 8);
    }
  } else if (warp == 4) {
    constexpr uint32_t idesc = umma_idesc_bf16(kTileM, 32, 0, 0);
    for (int it = 0; it < nblk; ++it) {
      const int stage = it % kLinStages;
      const uint32_t phase = (it / kLinStages) & 1;
      mbar_wait(bar_full + stage * 8, phase);
      fence_proxy_async_smem();  // cp.async (generic proxy) wrote the tile the MMA will read
      tc_fence_after();
      const uint32_t sa = s_base + stage * S::stage;
      const uint32_t sb = sa + S::a_bytes;
#pragma unroll
      for (int k = 0; k < kLinKB / 2; ++k) {
        const uint64_t adesc = umma_smem_desc(sa + 2 * k * kLinChunkPitch, kLinChunkPitch, 128);
        const uint64_t bdesc = umma_smem_desc(sb + 2 * k * 512, 512, 128);
        if (elect_one()) umma_bf16(tmem_base, adesc, bdesc, idesc, (it | k) != 0);
      }
      if (elect_one()) umma_commit(bar_empty + stage * 8);
      __syncwarp();
    }
    if (elect_one()) umma_commit(bar_done);
    __syncwarp();
  } else {
    const int q = warp & 3;
    mbar_wait(bar_done, 0);
    tc_fence_after();
    uint32_t v[32];
    tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16), v);
    tmem_ld_wait();
    const int img = img0 + q * 32 + lane;
    if (img < g.batch) {
      float* dst = partial + ((long long)blockIdx.x * g.batch + img) * NOUT;
#pragma unroll
      for (int j = 0; j < NOUT; ++j) dst[j] = __uint_as_float(v[j]) + __uint_as_float(v[16 + j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, 32);
}

}  // namespace ddpb

using namespace ddpb;

static void lin_fwd_geom(int batch, int nkc, int& S, int& per, int& mtiles) {
  mtiles = (batch + kTileM - 1) / kTileM;
  int want = ddpb_num_sms() / mtiles;
  if (want < 1) want = 1;
  per = ((nkc + want - 1) / want + kLinKB - 1) / kLinKB * kLinKB;
  S = (nkc + per - 1) / per;
}

extern "C" {

// bytes of the packed weight buffers (forward tile + backward tile), 16-byte aligned sub-buffers
size_t ddpb_linear_pack_bytes(int nout, int C, int H, int W) {
  (void)nout;
  const long long nkc_pad = ((long long)(C / 8) * H * W + kLinKB - 1) / kLinKB * kLinKB;
  return static_cast<size_t>(nkc_pad) * (512 + 512);
}

int ddpb_linear_pack(const float* Wfl, void* packed, int nout, int C, int H, int W, void* stream) {
  if (nout != 10 || (C % 8) != 0) return DDPB_ERR_SHAPE;
  const int nkc = (C / 8) * H * W;
  const int nkc_pad = (nkc + kLinKB - 1) / kLinKB * kLinKB;
  __nv_bfloat16* wf = static_cast<__nv_bfloat16*>(packed);
  __nv_bfloat16* wb = wf + (long long)nkc_pad * 256;
  return static_cast<int>(launch_k(pack_linear_weights_kernel<10>, dim3((nkc_pad + 127) / 128), dim3(128),
                                   0, static_cast<cudaStream_t>(stream), Wfl, wf, wb, C, H * W, nkc,
                                   nkc_pad));
}

size_t ddpb_linear_fwd_tc_scratch_bytes(int nout, int C, int batch_cap, int H, int W) {
  (void)C; (void)H; (void)W;
  const size_t mt = (batch_cap + kTileM - 1) / kTileM;
  return static_cast<size_t>(ddpb_num_sms() + 1) * (mt > 1 ? batch_cap / mt + kTileM : batch_cap) * nout *
             sizeof(float) + 1024;
}

// partial[S][batch][nout]; *S_out (host) receives S.  `packed` comes from ddpb_linear_pack.
int ddpb_linear_fwd_tc(const void* act, const void* packed, float* partial, int* S_out, int nout, int C,
                       int batch, int batch_cap, int H, int W, void* stream) {
  if (nout != 10 || (C % 8) != 0) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  const int nkc = (C / 8) * H * W;
  int S, per, mt;
  lin_fwd_geom(batch, nkc, S, per, mt);
  static bool configured[64] = {};
  int dev = 0;
  DDPB_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    DDPB_CUDA_OK(cudaFuncSetAttribute(linear_fwd_tc_kernel<10>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      LinFwdSmem::total));
    configured[dev & 63] = true;
  }
  if (S_out) *S_out = S;
  return static_cast<int>(launch_k(linear_fwd_tc_kernel<10>, dim3(S, mt), dim3(kLinThreads),
                                   LinFwdSmem::total, static_cast<cudaStream_t>(stream),
                                   static_cast<const __nv_bfloat16*>(act),
                                   static_cast<const __nv_bfloat16*>(packed), partial, g, H * W, nkc, per));
}

}  // extern "C"

// ===================================================================================== backward
// One persistent kernel, no partial buffers and no separate reduce:
//   dX[img][f] = relu'(act[img][f]) * sum_j dlogits[img][j] W[j][f]          -> bf16 planes
//   dW[j][f]   = sum_img dlogits[img][j] act[img][f]                          -> fp32, canonical
// A CTA owns feature blocks of 16 chunks (128 features) and, for each, walks over all image tiles:
//   * the gathered activation tile (see the forward) is the MN-major A operand of the dW GEMM
//     (M = 128 features, K = 128 images, N = 32 = dlogits hi|lo), accumulated in TMEM over ALL
//     image tiles, so each dW element is produced by exactly one CTA (deterministic);
//   * the bf16 dlogits tile [4 groups][128 images][8] is at once the K-major A operand of the dX
//     GEMM (K = 32 = hi|lo) and the MN-major B operand of the dW GEMM;
//   * the epilogue warps mask dX with the activation tile still in shared memory and store planes.
namespace ddpb {

constexpr int kLbFB = 16;                         // feature chunks per feature block (= 128 features)
constexpr int kLbStages = 3;
constexpr uint32_t kLbActBytes = kLbFB * kLinChunkPitch;  // 33,024
constexpr uint32_t kLbDlBytes = 4 * kTileM * 16;          // 8,192
constexpr uint32_t kLbStage = kLbActBytes + kLbDlBytes;
constexpr uint32_t kLbWbBytes = 4 * kLbFB * 128;          // 8,192 per feature block (hi, hi, lo, lo)

// Eight epilogue warps: two per TMEM lane quadrant, each owning half of the 128 accumulator columns.
// (With four, each scheduler had ONE epilogue warp running a ~550-instruction dependent chain per
// tile — tcgen05.ld, mask, pack, stage, store — and the whole kernel waited on it.)
constexpr int kLbThreads = 416;  // warps 0-3 gather, 4 MMA/TMEM, 5-12 epilogue
constexpr uint32_t kLbOutPitch = (kLbFB / 2) * 16 + 16;  // staged half row of one image: 128 B + 16 B bank skew

struct LinBwdSmem {
  static constexpr uint32_t off_wb = kLbStages * kLbStage;          // 2 buffers
  static constexpr uint32_t off_out = off_wb + 2 * kLbWbBytes;      // per epilogue warp: 32 half rows
  static constexpr uint32_t off_bar = off_out + 8 * 32 * kLbOutPitch;
  // full[ST], empty[ST], wbfull[2], wbfree[2], d1full[2], d1free[2], dwfull, dwfree
  static constexpr uint32_t nbar = 2 * kLbStages + 12;
  static constexpr uint32_t off_tmem = off_bar + nbar * 8;
  static constexpr uint32_t total = off_tmem + 16;
};

template <int NOUT>
__global__ void __launch_bounds__(kLbThreads, 1)
linear_bwd_tc_kernel(const __nv_bfloat16* __restrict__ act, const __nv_bfloat16* __restrict__ dlp,
                     const __nv_bfloat16* __restrict__ wb, __nv_bfloat16* __restrict__ dact,
                     float* __restrict__ dW, Geom g, int HW, int nkc, int nkc_pad, int C) {
  using S = LinBwdSmem;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t bar_full = s_base + S::off_bar;
  const uint32_t bar_empty = bar_full + kLbStages * 8;
  const uint32_t bar_wbfull = bar_empty + kLbStages * 8;
  const uint32_t bar_wbfree = bar_wbfull + 16;
  const uint32_t bar_d1full = bar_wbfree + 16;
  const uint32_t bar_d1free = bar_d1full + 16;
  const uint32_t bar_dwfull = bar_d1free + 16;
  const uint32_t bar_dwfree = bar_dwfull + 16;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + S::off_tmem);
  const int nfb = (nkc + kLbFB - 1) / kLbFB;
  const int mtiles = (g.batch + kTileM - 1) / kTileM;

  if (tid == 0) {
    for (int i = 0; i < kLbStages; ++i) {
      mbar_init(bar_full + i * 8, 129);   // 128 cp.async completions + the dlogits-tile bulk copy
      mbar_init(bar_empty + i * 8, 257);  // MMA commit + 256 epilogue threads (mask read)
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_wbfull + i * 8, 1);
      mbar_init(bar_wbfree + i * 8, 1);
      mbar_init(bar_d1full + i * 8, 1);
      mbar_init(bar_d1free + i * 8, 256);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_dwfull + i * 8, 1);
      mbar_init(bar_dwfree + i * 8, 128);
    }
    fence_mbar_init();
  }
  if (warp == 4) {
    tmem_alloc(smem_u32(s_tmem), 256);      // dX accumulator, two buffers of 128 columns
    tmem_alloc(smem_u32(s_tmem + 1), 64);   // dW accumulator, two buffers of 32 columns
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_dx = s_tmem[0], tmem_dw = s_tmem[1];
  pdl_sync();

  if (warp < 4) {
    // ---------------- producers
    int it = 0, fbi = 0;
    for (int fb = blockIdx.x; fb < nfb; fb += gridDim.x, ++fbi) {
      if (tid == 0) {  // the block's weight tile (B operand of the dX GEMM), double-buffered
        const int wbuf = fbi & 1;
        mbar_wait(bar_wbfree + wbuf * 8, ((fbi >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(bar_wbfull + wbuf * 8, kLbWbBytes);
        for (int jg = 0; jg < 4; ++jg)
          bulk_g2s(s_base + S::off_wb + wbuf * kLbWbBytes + jg * (kLbFB * 128),
                   wb + ((long long)jg * nkc_pad + (long long)fb * kLbFB) * 64, kLbFB * 128,
                   bar_wbfull + wbuf * 8);
      }
      for (int mt = 0; mt < mtiles; ++mt, ++it) {
        const int stage = it % kLbStages;
        const uint32_t phase = (it / kLbStages) & 1;
        mbar_wait(bar_empty + stage * 8, phase ^ 1);
        const uint32_t sa = s_base + stage * kLbStage;
        // activation tile: lanes 0..15 own one chunk each, two half-warps split the images
        {
          const int kl = lane & 15;
          const int kc = fb * kLbFB + kl;
          const bool kc_ok = kc < nkc;
          const int kcc = kc_ok ? kc : 0;
          const int chunk = kcc / HW, pix = kcc - chunk * HW;
          const int h = pix / g.W, w = pix - h * g.W;
          const __nv_bfloat16* src0 =
              act + ((long long)chunk * g.plane_rows + kGuard + (long long)(h + 1) * g.PW + (w + 1)) * 8;
#pragma unroll 8
          for (int i = warp * 2 + (lane >> 4); i < kTileM; i += 8) {
            const int img = mt * kTileM + i;
            const bool ok = kc_ok && img < g.batch;
            cp_async16(sa + kl * kLinChunkPitch + (i >> 3) * 128 + (i & 7) * 16,
                       src0 + (long long)(ok ? img : 0) * g.PHW * 8, ok);
          }
        }
        cp_async_arrive(bar_full + stage * 8);
        if (tid == 0) {  // dlogits tile, pre-packed by the loss kernel: [4 groups][128 images][8] bf16
          mbar_arrive_expect_tx(bar_full + stage * 8, kLbDlBytes);
          bulk_g2s(sa + kLbActBytes, dlp + (long long)mt * (kLbDlBytes / 2), kLbDlBytes,
                   bar_full + stage * 8);
        }
      }
    }
  } else if (warp == 4) {
    // ---------------- MMA issuer (converged warp, one elected lane issues)
    constexpr uint32_t idesc_dx = umma_idesc_bf16(kTileM, 128, 0, 0);  // dl (K-major) x wb (K-major)
    constexpr uint32_t idesc_dw = umma_idesc_bf16(128, 32, 1, 1);      // act (MN-major) x dl (MN-major)
    int it = 0, fbi = 0;
    for (int fb = blockIdx.x; fb < nfb; fb += gridDim.x, ++fbi) {
      const int wbuf = fbi & 1;
      mbar_wait(bar_wbfull + wbuf * 8, (fbi >> 1) & 1);
      mbar_wait(bar_dwfree + wbuf * 8, ((fbi >> 1) & 1) ^ 1);  // this dW buffer has been read out
      tc_fence_after();
      const uint32_t swb = s_base + S::off_wb + wbuf * kLbWbBytes;
      const uint32_t d_dw = tmem_dw + wbuf * 32;
      for (int mt = 0; mt < mtiles; ++mt, ++it) {
        const int stage = it % kLbStages;
        const uint32_t phase = (it / kLbStages) & 1;
        const int acc = it & 1;
        mbar_wait(bar_d1free + acc * 8, ((it >> 1) & 1) ^ 1);
        mbar_wait(bar_full + stage * 8, phase);
        fence_proxy_async_smem();
        tc_fence_after();
        const uint32_t sa = s_base + stage * kLbStage;
        const uint32_t sdl = sa + kLbActBytes;
        // dX = dl_hi.W_hi + dl_lo.W_hi + dl_hi.W_lo   (operand groups are 2 x [.][16 B], LBO 2048)
        {
          const uint64_t a_hi = umma_smem_desc(sdl, 2048, 128), a_lo = umma_smem_desc(sdl + 4096, 2048, 128);
          const uint64_t b_hi = umma_smem_desc(swb, kLbFB * 128, 128);
          const uint64_t b_lo = umma_smem_desc(swb + 2 * kLbFB * 128, kLbFB * 128, 128);
          if (elect_one()) {
            umma_bf16(tmem_dx + acc * 128, a_hi, b_hi, idesc_dx, 0);
            umma_bf16(tmem_dx + acc * 128, a_lo, b_hi, idesc_dx, 1);
            umma_bf16(tmem_dx + acc * 128, a_hi, b_lo, idesc_dx, 1);
            umma_commit(bar_d1full + acc * 8);
          }
        }
        // dW += act^T . dl   (K = images, 16 per MMA)
#pragma unroll
        for (int ks = 0; ks < kTileM / 16; ++ks) {
          const uint64_t adesc = umma_smem_desc(sa + ks * 256, 128, kLinChunkPitch);
          const uint64_t bdesc = umma_smem_desc(sdl + ks * 256, 128, 2048);
          if (elect_one()) umma_bf16(d_dw, adesc, bdesc, idesc_dw, (mt | ks) != 0);
        }
        if (elect_one()) umma_commit(bar_empty + stage * 8);
        __syncwarp();
      }
      if (elect_one()) {
        umma_commit(bar_wbfree + wbuf * 8);
        umma_commit(bar_dwfull + wbuf * 8);
      }
      __syncwarp();
    }
  } else {
    // ---------------- epilogue
    const int q = warp & 3;          // TMEM lane quadrant this warp may read
    const int half = (warp - 5) >> 2;  // which 64 of the 128 accumulator columns (8 of the 16 chunks)
    const int row = q * 32 + lane;  // image within the tile (dX) / feature within the block (dW)
    int it = 0, fbi = 0;
    for (int fb = blockIdx.x; fb < nfb; fb += gridDim.x, ++fbi) {
      // plane offset (in elements) of chunk kl of this block, computed once by lane kl
      long long my_off = -1;
      if (lane < kLbFB) {
        const int kc = fb * kLbFB + lane;
        if (kc < nkc) {
          const int chunk = kc / HW, pix = kc - chunk * HW;
          const int h = pix / g.W, w = pix - h * g.W;
          my_off = ((long long)chunk * g.plane_rows + kGuard + (long long)(h + 1) * g.PW + (w + 1)) * 8;
        }
      }
      // this lane's chunk in the coalesced write-out below
      const long long off = __shfl_sync(0xffffffffu, my_off, half * 8 + (lane & 7));
      for (int mt = 0; mt < mtiles; ++mt, ++it) {
        const int stage = it % kLbStages;
        const int acc = it & 1;
        mbar_wait(bar_d1full + acc * 8, (it >> 1) & 1);
        tc_fence_after();
        const uint8_t* sact = smem + stage * kLbStage + (row >> 3) * 128 + (row & 7) * 16;
        // masked dX half row of this image -> this warp's staging rows (128 B per image)
        uint8_t* sout = smem + S::off_out + (warp - 5) * 32 * kLbOutPitch;
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c0 = half * 64 + cc * 32;
          uint32_t v[32];
          tmem_ld32(tmem_dx + (static_cast<uint32_t>(q * 32) << 16) + acc * 128 + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int kl = c0 / 8 + j;
            const uint4 mk = *reinterpret_cast<const uint4*>(sact + kl * kLinChunkPitch);
            const uint32_t* mp = reinterpret_cast<const uint32_t*>(&mk);
            uint4 o;
            uint32_t* op = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float x0 = bf16_lo(mp[e]) > 0.f ? __uint_as_float(v[j * 8 + 2 * e]) : 0.f;
              const float x1 = bf16_hi(mp[e]) > 0.f ? __uint_as_float(v[j * 8 + 2 * e + 1]) : 0.f;
              op[e] = pack_bf16x2(x0, x1);
            }
            *reinterpret_cast<uint4*>(sout + lane * kLbOutPitch + (cc * 4 + j) * 16) = o;
          }
        }
        __syncwarp();
        // coalesced write-out: each instruction stores four images x 8 chunks = four 128 B runs
        {
          const int kl8 = lane & 7;
#pragma unroll 4
          for (int r4 = 0; r4 < 8; ++r4) {
            const int il = 4 * r4 + (lane >> 3);
            const int im = mt * kTileM + q * 32 + il;
            const uint4 o = *reinterpret_cast<const uint4*>(sout + il * kLbOutPitch + kl8 * 16);
            if (off >= 0 && im < g.batch)
              *reinterpret_cast<uint4*>(dact + (long long)im * g.PHW * 8 + off) = o;
          }
        }
        __syncwarp();
        tc_fence_before();
        mbar_arrive(bar_d1free + acc * 8);
        mbar_arrive(bar_empty + stage * 8);  // done reading the activation tile (mask)
      }
      if (half != 0) continue;  // the dW read-out below is the first four warps' job
      // dW of this feature block: row = feature kl*8 + e
      const int wbuf = fbi & 1;
      mbar_wait(bar_dwfull + wbuf * 8, (fbi >> 1) & 1);
      tc_fence_after();
      {
        uint32_t v[32];
        tmem_ld32(tmem_dw + wbuf * 32 + (static_cast<uint32_t>(q * 32) << 16), v);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(bar_dwfree + wbuf * 8);
        const int kc = fb * kLbFB + (row >> 3);
        if (kc < nkc) {
          const int chunk = kc / HW, pix = kc - chunk * HW;
          const long long kcanon = (long long)(chunk * 8 + (row & 7)) * HW + pix;
          const long long K = (long long)C * HW;
#pragma unroll
          for (int j = 0; j < NOUT; ++j)
            dW[j * K + kcanon] = __uint_as_float(v[j]) + __uint_as_float(v[16 + j]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tmem_dealloc(tmem_dx, 256);
    tmem_dealloc(tmem_dw, 64);
  }
}

}  // namespace ddpb

extern "C" int ddpb_zero_tail_rows(void* planes, int nplanes, int batch, int batch_cap, int H, int W,
                                   void* stream);

// dl_packed: the bf16 hi/lo dlogits tiles written by ddpb_softmax_xent (its dl_packed argument).
extern "C" int ddpb_linear_bwd_tc(const void* act, const void* dl_packed, const void* packed, void* dact,
                                  float* dW, int nout, int C, int batch, int batch_cap, int H, int W,
                                  void* stream) {
  using namespace ddpb;
  if (nout != 10 || (C % 8) != 0) return DDPB_ERR_SHAPE;
  const Geom g = make_geom(batch, batch_cap, H, W);
  const int nkc = (C / 8) * H * W;
  const int nkc_pad = (nkc + kLinKB - 1) / kLinKB * kLinKB;
  const __nv_bfloat16* wb = static_cast<const __nv_bfloat16*>(packed) + (long long)nkc_pad * 256;
  static bool configured[64] = {};
  int dev = 0;
  DDPB_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    DDPB_CUDA_OK(cudaFuncSetAttribute(linear_bwd_tc_kernel<10>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      LinBwdSmem::total));
    configured[dev & 63] = true;
  }
  if (g.ntiles * kTileM > batch * g.PHW) {
    const int rc = ddpb_zero_tail_rows(dact, C / 8, batch, batch_cap, H, W, stream);
    if (rc != 0) return rc;
  }
  const int nfb = (nkc + kLbFB - 1) / kLbFB;
  const int sms = ddpb_num_sms();
  const int grid = nfb < sms ? nfb : sms;
  return static_cast<int>(launch_k(linear_bwd_tc_kernel<10>, dim3(grid), dim3(kLbThreads),
                                   LinBwdSmem::total, static_cast<cudaStream_t>(stream),
                                   static_cast<const __nv_bfloat16*>(act),
                                   static_cast<const __nv_bfloat16*>(dl_packed), wb,
                                   static_cast<__nv_bfloat16*>(dact), dW, g, H * W, nkc, nkc_pad, C));
}
