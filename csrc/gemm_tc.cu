// General dense GEMM on the 5th-generation tensor cores for fully connected layers of ANY shape:
// split-precision ("bf16x3") operands, fp32 accumulation in TMEM, operand tiles fetched by tensor-map TMA
// from "row planes" — the same 16-byte-row plane idea as layout.cuh, without the image padding:
//
//   a matrix X[R][K] (R rows, K contiguous elements each) is stored as two sets (hi, lo; x = hi + lo,
//   both bf16) of K/8 planes; a plane is R_pad rows of 16 bytes = the 8 consecutive K elements
//   [8c, 8c+8) of each row.  Rows and planes are zero-padded to multiples of 256 elements.
//
// One stored tensor serves both operand majors of tcgen05.mma in the no-swizzle canonical layout:
//   K-major  (reduce over K):  rows are the M/N index, a plane is one 8-wide K chunk;
//   MN-major (reduce over R):  the K elements are the M/N index, rows are the reduction index.
// so the three GEMMs of a linear layer y = x W^T read the SAME three packed tensors (x, W, dy):
//   forward   y^T[o][b] = sum_k W[o][k] x[b][k]     A = W  (K-major),  B = x  (K-major)
//   data grad dx^T[k][b] = sum_o W[o][k] dy[b][o]   A = W  (MN-major), B = dy (K-major)
//   weight    dW[o][k]  = sum_b dy[b][o] x[b][k]    A = dy (MN-major), B = x  (MN-major)
// Every K step issues three kind::f16 MMAs (hi.hi + lo.hi + hi.lo; the dropped lo.lo term is below
// 2^-16 relative): more operand bits than TF32, fp32 accumulate.
//
// Replaces what the reference delegates to cuBLAS through nn.Linear (reference model.py:16,
// train_ddp.py:199) for layer shapes beyond the 10-class head the step engine hard-codes (simt_ops.cu):
// BASELINE.json config 4's fc 32768 -> 750 -> 10.
//
// Operand tiles arrive by TMA tensor copies (cp.async.bulk.tensor, one instruction per operand and
// stage: a 4-D box [set][plane][row block][32 rows x 16 B] lands exactly in the no-swizzle canonical
// layout, whichever major it is read in).  Measured on B200: the first version issued the same bytes as
// 16..96 one-dimensional bulk copies per stage and was bound by ~50 ns per copy per SM (weight gradient
// of fc 32768->750 at batch 256: 195 us); DDPB_GEMM_TMA=0 selects that path for comparison.
//
// A stage of the ring holds 32 reduction elements of both operands, hi and lo (48 KB at a 256-wide N tile):
// two stages and two CTAs per SM (one's epilogue under the other's main loop), or four stages when the whole
// grid is a single wave of CTAs.  Tuning notes with numbers: DESIGN.md §4 and the comments at gemm_tiling(),
// gemm_cluster_enabled() and encode_operand_map().
//
// Warp roles (192 threads, one output tile (x one K split) per CTA):
//   warp 0    TMA producer (one lane; the 1-D fallback path issues from all 32 lanes)
//   warp 1    TMEM allocator + MMA issuer (one elected lane issues tcgen05.mma)
//   warps 2-5 epilogue: tcgen05.ld the accumulator, store fp32 (transposed: a column at a time, 128 contiguous
//             bytes per warp; row-major: through shared memory so that stores fill whole 32-byte sectors)
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include "ptx.cuh"
#include "api_internal.cuh"
#include "launch.cuh"

namespace ddpb {

constexpr int kGemmThreads = 192;
constexpr int kGemmBM = 128;   // M tile
constexpr int kGemmKS = 32;    // reduction elements per pipeline stage (two MMA K steps)
constexpr int kGemmPad = 256;  // rows and K elements of a packed tensor are padded to this

struct GemmOperand {
  const uint4* base;     // [2 sets][planes][rows] of 16-byte rows
  long long rows;        // rows per plane (padded)
  long long set_stride;  // rows * planes: distance between the hi and the lo set, in rows
};
struct GemmArgs {
  GemmOperand a, b;
  float* out;
  long long ld;            // leading dimension of out
  long long split_stride;  // elements between the partial images of two K splits
  int M, N;                // real extents of the output (stores are masked to them)
  int k_stages;            // stages (kGemmKS reduction elements each) over the whole reduction
  int transposed;          // 0: out[m*ld + n]   1: out[n*ld + m]
};

template <int BN, int STAGES, int MT>
struct GemmSmem {
  static constexpr uint32_t a_set = MT * kGemmBM * kGemmKS * 2;  // 8192 B per M tile, either major
  static constexpr uint32_t b_set = BN * kGemmKS * 2;
  static constexpr uint32_t stage_bytes = 2 * a_set + 2 * b_set;
  static constexpr uint32_t off_bar = STAGES * stage_bytes;
  static constexpr uint32_t off_tmem = off_bar + (2 * STAGES + 1) * 8;
  static constexpr uint32_t total = off_tmem + 16;
  static_assert(total <= 232448, "GEMM CTA exceeds the 227 KB shared-memory limit");
};

// Shared-memory image of one operand tile (T = 128 or BN elements along M/N) of one stage, per set:
//   K-major : [KS/8 chunks][T rows][16 B]      one copy per chunk  (T*16 bytes)
//   MN-major: [T/8 planes][KS rows][16 B]      one copy per plane  (KS*16 bytes)
template <bool MN, int T>
struct OperandTile {
  static constexpr int copies = MN ? T / 8 : kGemmKS / 8;          // per set
  static constexpr uint32_t copy_bytes = MN ? kGemmKS * 16 : T * 16;
  // descriptor fields (ptx.cuh umma_smem_desc): K-major lbo = chunk stride, sbo = 128;
  // MN-major lbo = 128 (8-row K groups are contiguous), sbo = plane stride
  static constexpr uint32_t lbo = MN ? 128 : T * 16;
  static constexpr uint32_t sbo = MN ? kGemmKS * 16 : 128;
  static constexpr uint32_t kstep = MN ? 16 * 16 : 2 * T * 16;     // bytes per MMA K step (16 elements)
  // source row (16-byte units, within one set) of copy i of the stage starting at reduction index k0
  __device__ static __forceinline__ long long src_row(const GemmOperand& op, int tile, int k0, int i) {
    if (MN) return ((long long)tile * (T / 8) + i) * op.rows + k0;
    return ((long long)(k0 / 8) + i) * op.rows + (long long)tile * T;
  }
};

// 4-D tensor copy global -> shared (SASS: UTMALDG), completion reported to an mbarrier as tx bytes.
__device__ __forceinline__ void tma_load_4d(uint32_t dst_smem, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2];" ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// The same copy delivered to the same shared-memory offset (and signalled on the mbarrier at the same offset)
// of every CTA of the cluster named in cta_mask.  The source is always this GPU's own memory.
__device__ __forceinline__ void tma_load_4d_multicast(uint32_t dst_smem, const CUtensorMap* map, uint32_t bar, int c0,
                                                      int c1, int c2, int c3, uint16_t cta_mask) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster "
      "[%0], [%1, {%4, %5, %6, %7}], [%2], %3;" ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar),
      "h"(cta_mask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
// tcgen05.commit arriving on the mbarrier at the same offset in every CTA of cta_mask.
__device__ __forceinline__ void umma_commit_multicast(uint32_t bar, uint16_t cta_mask) {
  asm volatile(
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"(cta_mask)
      : "memory");
}
__device__ __forceinline__ uint32_t cluster_cta_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// TM: operand tiles by tensor-map TMA (ma / mb), else by 1-D bulk copies from g.a / g.b.
// CL = 2: the kernel runs as clusters of two CTAs that differ only in their M tile.  Each CTA fetches its own
// A tile and ONE of the two sets (hi / lo) of the common B tile, multicast into both CTAs' shared memory: the
// B tile crosses L2 -> SM once per pair.  A stage is free again when BOTH CTAs' MMAs have consumed it
// (tcgen05.commit multicast onto both "empty" barriers).
// MT: M tiles (of 128) per CTA.  MT = 2 multiplies two A tiles by the same B tile (two accumulators, all
// 512 TMEM columns at BN = 256, one CTA per SM): a third less operand traffic through L2 per FLOP; opt-in,
// see gemm_tiling().
template <bool A_MN, bool B_MN, int BN, int STAGES, bool TM, int MT, int CL>
__global__ void __launch_bounds__(kGemmThreads, MT == 1 ? 2 : 1)
gemm_split_tc_kernel(GemmArgs g, const __grid_constant__ CUtensorMap ma, const __grid_constant__ CUtensorMap mb) {
  static_assert(CL == 1 || (CL == 2 && TM), "the cluster variant needs the tensor-map path");
  using S = GemmSmem<BN, STAGES, MT>;
  using TA = OperandTile<A_MN, MT * kGemmBM>;
  using TB = OperandTile<B_MN, BN>;
  constexpr uint32_t TMEM_COLS = MT * BN < 32 ? 32 : MT * BN;
  constexpr uint32_t a_sub = A_MN ? (kGemmBM / 8) * kGemmKS * 16 : kGemmBM * 16;  // second M tile inside an A set
  static_assert(MT == 1 || (MT == 2 && BN * MT <= 512), "two accumulators need 2*BN TMEM columns");
  extern __shared__ __align__(128) uint8_t smem[];

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t bar_full = s_base + S::off_bar;
  const uint32_t bar_empty = bar_full + STAGES * 8;
  const uint32_t bar_done = bar_empty + STAGES * 8;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + S::off_tmem);

  const int m_tile = blockIdx.x, n_tile = blockIdx.y, split = blockIdx.z;
  const int per = (g.k_stages + gridDim.z - 1) / gridDim.z;
  const int st0 = split * per;
  const int st1 = min(g.k_stages, st0 + per);
  const int nst = max(0, st1 - st0);

  if (tid == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(bar_full + i * 8, 1);
      mbar_init(bar_empty + i * 8, CL);  // one arrival per CTA of the cluster that reads the stage's B tile
    }
    mbar_init(bar_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(s_tmem), TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();  // the peer's barriers exist before anything is multicast at them
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  const uint32_t crank = CL > 1 ? cluster_cta_rank() : 0;
  pdl_sync();

  if (warp == 0 && TM) {
    // packed tensor as a 4-D array [set][plane][rows/32][32 rows x 8 elements]; the box of one operand
    // tile: K-major {256, T/32, KS/8, 2} at {0, tile*T/32, k0/8, 0}; MN-major {256, 1, T/8, 2} at
    // {0, k0/32, tile*T/8, 0} (ddpb_gemm_split encodes the maps with these boxes)
    if (lane == 0) {
      for (int it = 0; it < nst; ++it) {
        const int stage = it % STAGES;
        const uint32_t phase = (it / STAGES) & 1;
        const uint32_t full = bar_full + stage * 8;
        mbar_wait(bar_empty + stage * 8, phase ^ 1);
        mbar_arrive_expect_tx(full, S::stage_bytes);
        const int k0 = (st0 + it) * kGemmKS;
        const uint32_t sa = s_base + stage * S::stage_bytes;
        const uint32_t sb = sa + 2 * S::a_set;
        if (A_MN) tma_load_4d(sa, &ma, full, 0, k0 / 32, m_tile * (MT * kGemmBM / 8), 0);
        else tma_load_4d(sa, &ma, full, 0, m_tile * (MT * kGemmBM / 32), k0 / 8, 0);
        if (CL > 1) {
          // mb's box holds ONE set here: this CTA fetches set `crank` for the whole cluster
          const uint32_t dst = sb + crank * S::b_set;
          if (B_MN) tma_load_4d_multicast(dst, &mb, full, 0, k0 / 32, n_tile * (BN / 8), crank, (1u << CL) - 1);
          else tma_load_4d_multicast(dst, &mb, full, 0, n_tile * (BN / 32), k0 / 8, crank, (1u << CL) - 1);
        } else {
          if (B_MN) tma_load_4d(sb, &mb, full, 0, k0 / 32, n_tile * (BN / 8), 0);
          else tma_load_4d(sb, &mb, full, 0, n_tile * (BN / 32), k0 / 8, 0);
        }
      }
    }
    __syncwarp();
  } else if (warp == 0) {
    constexpr int nA = 2 * TA::copies, nB = 2 * TB::copies;
    for (int it = 0; it < nst; ++it) {
      const int stage = it % STAGES;
      const uint32_t phase = (it / STAGES) & 1;
      const uint32_t full = bar_full + stage * 8;
      if (lane == 0) {
        mbar_wait(bar_empty + stage * 8, phase ^ 1);
        mbar_arrive_expect_tx(full, S::stage_bytes);
      }
      __syncwarp();
      const int k0 = (st0 + it) * kGemmKS;
      const uint32_t sa = s_base + stage * S::stage_bytes;
      const uint32_t sb = sa + 2 * S::a_set;
      for (int i = lane; i < nA + nB; i += 32) {
        if (i < nA) {
          const int set = i / TA::copies, c = i % TA::copies;
          bulk_g2s(sa + set * S::a_set + c * TA::copy_bytes,
                   g.a.base + set * g.a.set_stride + TA::src_row(g.a, m_tile, k0, c), TA::copy_bytes, full);
        } else {
          const int j = i - nA;
          const int set = j / TB::copies, c = j % TB::copies;
          bulk_g2s(sb + set * S::b_set + c * TB::copy_bytes,
                   g.b.base + set * g.b.set_stride + TB::src_row(g.b, n_tile, k0, c), TB::copy_bytes, full);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_bf16(kGemmBM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
    for (int it = 0; it < nst; ++it) {
      const int stage = it % STAGES;
      const uint32_t phase = (it / STAGES) & 1;
      mbar_wait(bar_full + stage * 8, phase);
      tc_fence_after();
      const uint32_t sa = s_base + stage * S::stage_bytes;
      const uint32_t sb = sa + 2 * S::a_set;
#pragma unroll
      for (int kk = 0; kk < kGemmKS / 16; ++kk) {
        const uint64_t a_hi = umma_smem_desc(sa + kk * TA::kstep, TA::lbo, TA::sbo);
        const uint64_t a_lo = umma_smem_desc(sa + S::a_set + kk * TA::kstep, TA::lbo, TA::sbo);
        const uint64_t b_hi = umma_smem_desc(sb + kk * TB::kstep, TB::lbo, TB::sbo);
        const uint64_t b_lo = umma_smem_desc(sb + S::b_set + kk * TB::kstep, TB::lbo, TB::sbo);
        if (elect_one()) {
          umma_bf16(tmem_base, a_hi, b_hi, idesc, (it | kk) != 0);
          umma_bf16(tmem_base, a_lo, b_hi, idesc, 1);
          umma_bf16(tmem_base, a_hi, b_lo, idesc, 1);
        }
        if constexpr (MT == 2) {  // second M tile, same B tile, second accumulator
          const uint64_t a2_hi = umma_smem_desc(sa + a_sub + kk * TA::kstep, TA::lbo, TA::sbo);
          const uint64_t a2_lo = umma_smem_desc(sa + S::a_set + a_sub + kk * TA::kstep, TA::lbo, TA::sbo);
          if (elect_one()) {
            umma_bf16(tmem_base + BN, a2_hi, b_hi, idesc, (it | kk) != 0);
            umma_bf16(tmem_base + BN, a2_lo, b_hi, idesc, 1);
            umma_bf16(tmem_base + BN, a2_hi, b_lo, idesc, 1);
          }
        }
      }
      if (elect_one()) {
        if (CL > 1) umma_commit_multicast(bar_empty + stage * 8, (1u << CL) - 1);
        else umma_commit(bar_empty + stage * 8);
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(bar_done);
    __syncwarp();
  } else {
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    const int n0 = n_tile * BN;
    float* out = g.out + (long long)split * g.split_stride;
    if (nst > 0) {
      mbar_wait(bar_done, 0);
      tc_fence_after();
    }
    const bool vec_ok = !g.transposed && (g.ld & 3) == 0 && (g.split_stride & 3) == 0 &&
                        (reinterpret_cast<uintptr_t>(g.out) & 15) == 0;
#pragma unroll 1
    for (int cc = 0; cc < MT * BN; cc += 32) {
      const int mt = cc / BN, c0 = cc - mt * BN;  // accumulator (M tile) and column inside it
      const int m = (m_tile * MT + mt) * kGemmBM + q * 32 + lane;
      uint32_t v[32];
      if (nst > 0) {
        tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + cc, v);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int c = 0; c < 32; ++c) v[c] = 0u;  // a K split without work still defines its partial image
      }
      if (g.transposed) {
        if (m < g.M) {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const int n = n0 + c0 + c;
            if (n < g.N) out[(long long)n * g.ld + m] = __uint_as_float(v[c]);
          }
        }
      } else if (vec_ok && n0 + c0 + 32 <= g.N) {
        // A lane holds 32 columns of ONE row; stored directly, every 16-byte store of a warp would hit 32
        // different rows and fill half a 32-byte sector each (measured: 197 MB of store traffic for the 98 MB
        // weight gradient of fc 32768->750).  The 32x32 block goes through shared memory instead (the stage
        // ring is idle once the accumulator is final) so that eight lanes cover one row's 128 bytes.
        float* sbuf = reinterpret_cast<float*>(smem) + (warp - 2) * (32 * 33);
#pragma unroll
        for (int c = 0; c < 32; ++c) sbuf[lane * 33 + c] = __uint_as_float(v[c]);
        __syncwarp();
        const int m_base = m - lane;  // first row of this warp's quadrant
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = i * 4 + (lane >> 3), c4 = (lane & 7) * 4;
          const float* s = sbuf + r * 33 + c4;
          if (m_base + r < g.M)
            *reinterpret_cast<float4*>(out + (long long)(m_base + r) * g.ld + n0 + c0 + c4) =
                make_float4(s[0], s[1], s[2], s[3]);
        }
        __syncwarp();
      } else if (m < g.M) {
        float* row = out + (long long)m * g.ld + n0 + c0;
#pragma unroll
        for (int c = 0; c < 32; ++c)
          if (n0 + c0 + c < g.N) row[c] = __uint_as_float(v[c]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
  // neither CTA may exit while the other can still multicast data or barrier arrivals into it
  if (CL > 1) cluster_sync_all();
}

// fp32 row-major src[R][K] (leading dimension ld) -> split row planes.  Optional `mask` (same shape and
// ld as src): elements whose mask value is <= 0 are packed as zero (ReLU backward fused into the pack).
// Block = 256 rows x 8 chunks (64 K elements); lane = row, so a warp's stores to one plane are 512
// contiguous bytes, and a thread walks 256 contiguous source bytes (its later loads hit the lines the
// first ones brought into L1).  Only the real region is written — the padding of a plane tensor is zero
// from its allocation and stays zero — plus zeros for the rows [R, next multiple of 32): a reduction
// over rows (MN-major use) runs in stages of 32 rows, and the buffer may hold a larger, earlier batch
// beyond R.
constexpr int kPackChunks = 8;  // chunks per block for large matrices; small ones use 1 (more blocks)
__global__ void __launch_bounds__(256)
pack_rows_split_kernel(const float* __restrict__ src, const float* __restrict__ mask, long long ld, int R, int K,
                       uint4* __restrict__ dst, long long rows_pad, long long set_stride, int vec, int cpb) {
  pdl_sync();
  const int r = blockIdx.y * 256 + threadIdx.x;
  const int Rz = (R + kGemmKS - 1) / kGemmKS * kGemmKS;
  if (r >= Rz) return;
  const int c_begin = blockIdx.x * cpb;
  if (vec && !mask && r < R && cpb == kPackChunks && (c_begin + kPackChunks) * 8 <= K) {
    // fast path (whole 64-element group inside the matrix): all sixteen loads are issued before the
    // first conversion, so a thread makes one global round trip instead of eight
    const float4* p4 = reinterpret_cast<const float4*>(src + (long long)r * ld + c_begin * 8);
    float4 v[2 * kPackChunks];
#pragma unroll
    for (int i = 0; i < 2 * kPackChunks; ++i) v[i] = p4[i];
#pragma unroll
    for (int cc = 0; cc < kPackChunks; ++cc) {
      const float x[8] = {v[2 * cc].x, v[2 * cc].y, v[2 * cc].z, v[2 * cc].w,
                          v[2 * cc + 1].x, v[2 * cc + 1].y, v[2 * cc + 1].z, v[2 * cc + 1].w};
      float hi[8], lo[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) split_bf16(x[e], hi[e], lo[e]);
      const long long o = (long long)(c_begin + cc) * rows_pad + r;
      dst[o] = make_uint4(pack_bf16x2(hi[0], hi[1]), pack_bf16x2(hi[2], hi[3]), pack_bf16x2(hi[4], hi[5]),
                          pack_bf16x2(hi[6], hi[7]));
      dst[set_stride + o] = make_uint4(pack_bf16x2(lo[0], lo[1]), pack_bf16x2(lo[2], lo[3]),
                                       pack_bf16x2(lo[4], lo[5]), pack_bf16x2(lo[6], lo[7]));
    }
    return;
  }
#pragma unroll 4
  for (int cc = 0; cc < cpb; ++cc) {
    const int c = c_begin + cc;
    if (c * 8 >= K) break;
    const long long o = (long long)c * rows_pad + r;
    if (r >= R) {
      dst[o] = make_uint4(0u, 0u, 0u, 0u);
      dst[set_stride + o] = make_uint4(0u, 0u, 0u, 0u);
      continue;
    }
    float x[8];
    const float* p = src + (long long)r * ld + c * 8;
    if (vec && c * 8 + 8 <= K) {
      const float4 u = *reinterpret_cast<const float4*>(p), w = *reinterpret_cast<const float4*>(p + 4);
      x[0] = u.x; x[1] = u.y; x[2] = u.z; x[3] = u.w; x[4] = w.x; x[5] = w.y; x[6] = w.z; x[7] = w.w;
      if (mask) {
        const float* mp = mask + (long long)r * ld + c * 8;
        const float4 a = *reinterpret_cast<const float4*>(mp), b = *reinterpret_cast<const float4*>(mp + 4);
        const float mv[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) x[e] = mv[e] > 0.f ? x[e] : 0.f;
      }
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const bool in = c * 8 + e < K;
        x[e] = in ? p[e] : 0.f;
        if (in && mask && !(mask[(long long)r * ld + c * 8 + e] > 0.f)) x[e] = 0.f;
      }
    }
    float hi[8], lo[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) split_bf16(x[e], hi[e], lo[e]);
    dst[o] = make_uint4(pack_bf16x2(hi[0], hi[1]), pack_bf16x2(hi[2], hi[3]), pack_bf16x2(hi[4], hi[5]),
                        pack_bf16x2(hi[6], hi[7]));
    dst[set_stride + o] = make_uint4(pack_bf16x2(lo[0], lo[1]), pack_bf16x2(lo[2], lo[3]),
                                     pack_bf16x2(lo[4], lo[5]), pack_bf16x2(lo[6], lo[7]));
  }
}

// out[r][c] = act( sum_s partial[s*split_stride + r*ld + c] + bias[c] ), fixed summation order.
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ partial, int nsplit, long long split_stride, int R, int C,
                     long long ld, const float* __restrict__ bias, int relu, float* __restrict__ out,
                     long long ld_out) {
  pdl_sync();
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)R * C) return;
  const int r = (int)(idx / C), c = (int)(idx - (long long)r * C);
  const float* p = partial + (long long)r * ld + c;
  float s = 0.f;
  for (int k = 0; k < nsplit; ++k) s += p[(long long)k * split_stride];
  if (bias) s += bias[c];
  if (relu) s = fmaxf(s, 0.f);
  out[(long long)r * ld_out + c] = s;
}

// out[c] = sum_r src[r][c] (mask: only where mask[r][c] > 0) — a linear layer's bias gradient.
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ src, const float* __restrict__ mask, int R_all, int C, long long ld,
              float* __restrict__ out, int rows_per_split) {
  // block = 32 columns x 8 row groups over the rows of split blockIdx.y; fixed order: each group sums its
  // rows r = g, g+8, ..., then groups 0..7.  Split y writes out[y*C + c].
  __shared__ float sm[8][33];
  pdl_sync();
  const int lane = threadIdx.x & 31, grp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const long long row0 = (long long)blockIdx.y * rows_per_split;
  const int R = (int)min((long long)rows_per_split, (long long)R_all - row0);
  src += row0 * ld;
  if (mask) mask += row0 * ld;
  out += (long long)blockIdx.y * C;
  float s = 0.f;
  if (c < C) {
    // four independent chains keep the loads in flight; the order of the additions is fixed
    float a[4] = {0.f, 0.f, 0.f, 0.f};
    int r = grp;
    for (; r + 24 < R; r += 32) {
      float v[4], m[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        v[u] = src[(long long)(r + 8 * u) * ld + c];
        m[u] = mask ? mask[(long long)(r + 8 * u) * ld + c] : 1.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) a[u] += m[u] > 0.f ? v[u] : 0.f;
    }
    for (; r < R; r += 8) {
      const float v = src[(long long)r * ld + c];
      a[0] += (!mask || mask[(long long)r * ld + c] > 0.f) ? v : 0.f;
    }
    s = (a[0] + a[1]) + (a[2] + a[3]);
  }
  sm[grp][lane] = s;
  __syncthreads();
  if (grp == 0 && c < C) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sm[k][lane];
    out[c] = t;
  }
}

// ------------------------------------------------------------------------------------------
// Fused (bias +) ReLU + 2x2/stride-2 max pool over NCHW fp32, forward and backward.
// relu(max(.)) == max(relu(.)), so one pass produces what nn.ReLU -> nn.MaxPool2d(2) produce; `code`
// keeps, per output, which of the four inputs won (0..3, first maximum in row-major order like ATen)
// or 4 when the pooled value is <= 0 (no gradient).  Thread = two horizontally adjacent outputs.
__global__ void __launch_bounds__(256)
relu_maxpool2_fwd_kernel(const float* __restrict__ in, const float* __restrict__ bias, float* __restrict__ out,
                         uint8_t* __restrict__ code, long long nplanes, int C, int H, int W) {
  pdl_sync();
  const int OW2 = W / 4, OH = H / 2;  // output pairs per row, output rows
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= nplanes * OH * OW2) return;
  const int xp = (int)(idx % OW2);
  const int oy = (int)((idx / OW2) % OH);
  const long long pl = idx / ((long long)OW2 * OH);
  const float b = bias ? bias[(int)(pl % C)] : 0.f;
  const float* p = in + (pl * H + 2 * oy) * W + 4 * xp;
  const float4 r0 = *reinterpret_cast<const float4*>(p), r1 = *reinterpret_cast<const float4*>(p + W);
  const float v[2][4] = {{r0.x, r0.y, r1.x, r1.y}, {r0.z, r0.w, r1.z, r1.w}};
  float o[2];
  uint8_t cd[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    float best = v[j][0];
    int arg = 0;
#pragma unroll
    for (int e = 1; e < 4; ++e)
      if (v[j][e] > best) { best = v[j][e]; arg = e; }
    best += b;
    o[j] = fmaxf(best, 0.f);
    cd[j] = best > 0.f ? (uint8_t)arg : (uint8_t)4;
  }
  const long long oo = (pl * OH + oy) * (W / 2) + 2 * xp;
  *reinterpret_cast<float2*>(out + oo) = make_float2(o[0], o[1]);
  *reinterpret_cast<uchar2*>(code + oo) = make_uchar2(cd[0], cd[1]);
}

__global__ void __launch_bounds__(256)
relu_maxpool2_bwd_kernel(const float* __restrict__ dout, const uint8_t* __restrict__ code,
                         float* __restrict__ din, long long nplanes, int H, int W) {
  pdl_sync();
  const int OW2 = W / 4, OH = H / 2;
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= nplanes * OH * OW2) return;
  const int xp = (int)(idx % OW2);
  const int oy = (int)((idx / OW2) % OH);
  const long long pl = idx / ((long long)OW2 * OH);
  const long long oo = (pl * OH + oy) * (W / 2) + 2 * xp;
  const float2 d = *reinterpret_cast<const float2*>(dout + oo);
  const uchar2 cd = *reinterpret_cast<const uchar2*>(code + oo);
  float4 r0, r1;
  r0.x = cd.x == 0 ? d.x : 0.f; r0.y = cd.x == 1 ? d.x : 0.f;
  r1.x = cd.x == 2 ? d.x : 0.f; r1.y = cd.x == 3 ? d.x : 0.f;
  r0.z = cd.y == 0 ? d.y : 0.f; r0.w = cd.y == 1 ? d.y : 0.f;
  r1.z = cd.y == 2 ? d.y : 0.f; r1.w = cd.y == 3 ? d.y : 0.f;
  float* p = din + (pl * H + 2 * oy) * W + 4 * xp;
  *reinterpret_cast<float4*>(p) = r0;
  *reinterpret_cast<float4*>(p + W) = r1;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime (no link dependency on libcuda); NULL when unavailable or
// switched off with DDPB_GEMM_TMA=0.
static EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    const char* e = getenv("DDPB_GEMM_TMA");
    if (!(e && e[0] == '0')) {
      void* p = nullptr;
      cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
      if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
          q == cudaDriverEntryPointSuccess)
        fn = reinterpret_cast<EncodeTiledFn>(p);
      else
        cudaGetLastError();
    }
  }
  return fn;
}

// Tensor map over a packed tensor for tiles of T elements along M/N read in the given major.
// (A box dimension holds at most 256 elements, i.e. 512-byte runs here.  Describing the same bytes in 8-byte
// elements gives 2 KB runs for K-major tiles; measured, it changes nothing (forward GEMM of fc 32768->750:
// 46.2 vs 46.4 us): these GEMMs are bound by L2 throughput — ~300 MB of operand tiles in 46 us — not by the
// TMA unit.  What would lower the traffic without giving up the second CTA per SM is multicasting the shared
// B tile across a cluster of CTAs that differ only in their M tile.)
static int encode_operand_map(EncodeTiledFn enc, CUtensorMap* m, const GemmOperand& op, bool mn, int T,
                              int box_sets = 2) {
  const cuuint64_t planes = (cuuint64_t)(op.set_stride / op.rows);
  cuuint64_t dims[4] = {256, (cuuint64_t)op.rows / 32, planes, 2};
  cuuint64_t strides[3] = {512, (cuuint64_t)op.rows * 16, (cuuint64_t)op.set_stride * 16};
  cuuint32_t box[4] = {256, mn ? 1u : (cuuint32_t)T / 32, mn ? (cuuint32_t)T / 8 : (cuuint32_t)kGemmKS / 8,
                       (cuuint32_t)box_sets};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<uint4*>(op.base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : DDPB_ERR_UNSUPPORTED;
}

// Launch as clusters of two CTAs along x (plain stream order: no programmatic dependent launch attribute).
template <typename K>
static cudaError_t launch_cluster2(K kern, dim3 grid, size_t smem, cudaStream_t st, const GemmArgs& g,
                                   const CUtensorMap& ma, const CUtensorMap& mb) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, g, ma, mb);
}

// Clusters of two CTAs sharing the B tile: opt-in with DDPB_GEMM_CLUSTER=1.  Measured on B200, fc 32768->750 at
// batch 256 (us, independent CTAs / clusters): forward 45.8 / 47.6, data gradient 36.7 / 41.2, weight gradient
// 65.0 / 72.8 — halving the B traffic does not pay for coupling the two CTAs' pipelines (each waits for the
// slower one at every stage), so independent CTAs stay the default.
static bool gemm_cluster_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("DDPB_GEMM_CLUSTER");
    on = (e && e[0] == '1') ? 1 : 0;
  }
  return on == 1;
}

template <bool A_MN, bool B_MN, int BN, int MT>
static int launch_gemm(const GemmArgs& g, dim3 grid, cudaStream_t st) {
  // MT = 1: ~96 KB of stages, two CTAs per SM (one's epilogue under the other's main loop);
  // MT = 2: three 64 KB stages, one CTA per SM
  constexpr int STAGES = MT == 2 ? 3 : BN == 256 ? 2 : BN == 128 ? 3 : 4;
  using S = GemmSmem<BN, STAGES, MT>;
  alignas(64) CUtensorMap ma, mb;
  memset(&ma, 0, sizeof(ma));
  memset(&mb, 0, sizeof(mb));
  EncodeTiledFn enc = encode_tiled();
  if constexpr (MT == 1 && BN == 256) {
    // two or more M tiles: pairs of CTAs share each B tile (the packed rows / planes are padded to 256, so
    // an odd tile count is rounded up inside the padding)
    if (enc && gemm_cluster_enabled() && grid.x >= 2 &&
        !encode_operand_map(enc, &ma, g.a, A_MN, kGemmBM) && !encode_operand_map(enc, &mb, g.b, B_MN, BN, 1)) {
      auto kern = gemm_split_tc_kernel<A_MN, B_MN, BN, STAGES, true, 1, 2>;
      DDPB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::total));
      dim3 cgrid((grid.x + 1) / 2 * 2, grid.y, grid.z);
      DDPB_CUDA_OK(launch_cluster2(kern, cgrid, S::total, st, g, ma, mb));
      return 0;
    }
  }
  if (enc && (encode_operand_map(enc, &ma, g.a, A_MN, MT * kGemmBM) || encode_operand_map(enc, &mb, g.b, B_MN, BN)))
    enc = nullptr;  // fall back to the 1-D copies
  if constexpr (MT == 1 && BN == 256) {
    // At most one CTA per SM anyway (a split-K forward pass): spend the whole shared memory on a deeper ring —
    // with two 48 KB stages a lone CTA keeps only ~96 KB of loads in flight, which bounds it by latency.
    if (enc && (long long)grid.x * grid.y * grid.z <= ddpb_num_sms()) {
      using S4 = GemmSmem<BN, 4, MT>;
      auto kern = gemm_split_tc_kernel<A_MN, B_MN, BN, 4, true, MT, 1>;
      DDPB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S4::total));
      DDPB_CUDA_OK(launch_k(kern, grid, dim3(kGemmThreads), S4::total, st, g, ma, mb));
      return 0;
    }
  }
  if (enc) {
    auto kern = gemm_split_tc_kernel<A_MN, B_MN, BN, STAGES, true, MT, 1>;
    DDPB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::total));
    DDPB_CUDA_OK(launch_k(kern, grid, dim3(kGemmThreads), S::total, st, g, ma, mb));
  } else {
    auto kern = gemm_split_tc_kernel<A_MN, B_MN, BN, STAGES, false, MT, 1>;
    DDPB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::total));
    DDPB_CUDA_OK(launch_k(kern, grid, dim3(kGemmThreads), S::total, st, g, ma, mb));
  }
  return 0;
}

template <int BN, int MT>
static int launch_gemm_major(int a_mn, int b_mn, const GemmArgs& g, dim3 grid, cudaStream_t st) {
  if (!a_mn && !b_mn) return launch_gemm<false, false, BN, MT>(g, grid, st);
  if (a_mn && !b_mn) return launch_gemm<true, false, BN, MT>(g, grid, st);
  if (a_mn && b_mn) return launch_gemm<true, true, BN, MT>(g, grid, st);
  return DDPB_ERR_SHAPE;  // (K-major A, MN-major B) is not needed by a linear layer
}

// N tile and M tiles per CTA for an (M, N) output.  One M tile per CTA by default; DDPB_GEMM_MT=2 selects
// two from 256 rows up at the full N tile.  Measured on B200 for fc 32768->750 at batch 256 (us, one / two M
// tiles): forward 46.0 / 42.5, data gradient 36.6 / 41.0, weight gradient 65.0 / 64.7 — the third less
// operand traffic is paid for with the lost overlap of one CTA's epilogue with the other's main loop, so the
// simpler configuration stays the default.
static void gemm_tiling(int M, int N, int* bn, int* mt) {
  *bn = N <= 64 ? 64 : N <= 128 ? 128 : 256;
  static int mt_max = 0;
  if (!mt_max) {
    const char* e = getenv("DDPB_GEMM_MT");
    mt_max = (e && e[0] == '2') ? 2 : 1;
  }
  *mt = (*bn == 256 && M > kGemmBM && mt_max == 2) ? 2 : 1;
}

}  // namespace ddpb

using namespace ddpb;

extern "C" {

int ddpb_gemm_pad() { return kGemmPad; }

// Bytes of a packed (split row planes) tensor for a matrix of `rows` x `k` elements.
size_t ddpb_pack_rows_split_bytes(int rows, int k) {
  const long long rp = ((long long)rows + kGemmPad - 1) / kGemmPad * kGemmPad;
  const long long kp = ((long long)k + kGemmPad - 1) / kGemmPad * kGemmPad;
  return (size_t)(2 * rp * kp * 2);
}

// src fp32 [rows][k] (leading dimension ld) -> packed; `packed` must have been zero-initialised once
// (the padding is never written).  mask (optional, same shape/ld): values with mask <= 0 pack as zero.
int ddpb_pack_rows_split(const float* src, const float* mask, long long ld, int rows, int k, void* packed,
                         void* stream) {
  if (rows <= 0 || k <= 0 || ld < k) return DDPB_ERR_SHAPE;
  const long long rp = ((long long)rows + kGemmPad - 1) / kGemmPad * kGemmPad;
  const long long kp = ((long long)k + kGemmPad - 1) / kGemmPad * kGemmPad;
  const int nchunks = (k + 7) / 8;
  const int vec = (ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
                  (!mask || (reinterpret_cast<uintptr_t>(mask) & 15) == 0);
  const int row_blocks = ((rows + kGemmKS - 1) / kGemmKS * kGemmKS + 255) / 256;
  const int cpb = (long long)((nchunks + kPackChunks - 1) / kPackChunks) * row_blocks >= 2 * ddpb_num_sms()
                      ? kPackChunks : 1;
  dim3 grid((nchunks + cpb - 1) / cpb, row_blocks);
  DDPB_CUDA_OK(launch_k(pack_rows_split_kernel, grid, dim3(256), 0, (cudaStream_t)stream, src, mask, ld, rows, k,
                        reinterpret_cast<uint4*>(packed), rp, rp * (kp / 8), vec, cpb));
  return 0;
}

// out (+ split-K partial images) = A x B^T over packed operands (see the top of this file).
//   a_rows/a_k, b_rows/b_k: the REAL shape each packed tensor was made from (rows x K elements);
//   a_mn / b_mn = 1 reads the operand MN-major (its K elements index M/N, its rows are reduced over);
//   M, N, K: the GEMM extents (masking / reduction length); out fp32 with leading dimension ld,
//   transposed = 1 stores out[n*ld + m]; k_splits > 1 writes k_splits partial images split_stride
//   elements apart (sum them with ddpb_splitk_reduce).
int ddpb_gemm_split(const void* a, int a_rows, int a_k, int a_mn, const void* b, int b_rows, int b_k, int b_mn,
                    int M, int N, int K, float* out, long long ld, int transposed, int k_splits,
                    long long split_stride, void* stream) {
  auto pad = [](long long v) { return (v + kGemmPad - 1) / kGemmPad * kGemmPad; };
  if (M <= 0 || N <= 0 || K <= 0 || k_splits < 1 || !a || !b || !out) return DDPB_ERR_SHAPE;
  GemmArgs g;
  g.a.base = reinterpret_cast<const uint4*>(a);
  g.a.rows = pad(a_rows);
  g.a.set_stride = pad(a_rows) * (pad(a_k) / 8);
  g.b.base = reinterpret_cast<const uint4*>(b);
  g.b.rows = pad(b_rows);
  g.b.set_stride = pad(b_rows) * (pad(b_k) / 8);
  // the operand's extent along M/N and along the reduction must cover the GEMM's
  const int a_mn_ext = a_mn ? a_k : a_rows, a_red = a_mn ? a_rows : a_k;
  const int b_mn_ext = b_mn ? b_k : b_rows, b_red = b_mn ? b_rows : b_k;
  if (a_mn_ext < M || b_mn_ext < N || a_red < K || b_red < K) return DDPB_ERR_SHAPE;
  g.out = out;
  g.ld = ld;
  g.split_stride = split_stride;
  g.M = M;
  g.N = N;
  g.k_stages = (K + kGemmKS - 1) / kGemmKS;
  g.transposed = transposed;
  if (k_splits > g.k_stages) k_splits = g.k_stages;
  int bn, mt;
  gemm_tiling(M, N, &bn, &mt);
  dim3 grid((M + mt * kGemmBM - 1) / (mt * kGemmBM), (N + bn - 1) / bn, k_splits);
  if (grid.y > 65535 || grid.z > 65535) return DDPB_ERR_LIMIT;
  cudaStream_t st = (cudaStream_t)stream;
  if (bn == 64) return launch_gemm_major<64, 1>(a_mn, b_mn, g, grid, st);
  if (bn == 128) return launch_gemm_major<128, 1>(a_mn, b_mn, g, grid, st);
  if (mt == 2) return launch_gemm_major<256, 2>(a_mn, b_mn, g, grid, st);
  return launch_gemm_major<256, 1>(a_mn, b_mn, g, grid, st);
}

// The K-split count ddpb_gemm_split would be best called with for an (M, N, K) problem: enough CTAs
// to fill the GPU once, never fewer than 4 stages of work per split.
int ddpb_gemm_split_suggest_splits(int M, int N, int K) {
  int bn, mt;
  gemm_tiling(M, N, &bn, &mt);
  const long long tiles = (long long)((M + mt * kGemmBM - 1) / (mt * kGemmBM)) * ((N + bn - 1) / bn);
  const int stages = (K + kGemmKS - 1) / kGemmKS;
  const int sms = ddpb_num_sms();
  long long s = tiles >= sms ? 1 : sms / tiles;
  if (s > stages / 4) s = stages / 4;
  return s < 1 ? 1 : (int)s;
}

int ddpb_splitk_reduce(const float* partial, int nsplit, long long split_stride, int rows, int cols, long long ld,
                       const float* bias, int relu, float* out, long long ld_out, void* stream) {
  if (rows <= 0 || cols <= 0 || nsplit < 1) return DDPB_ERR_SHAPE;
  const long long n = (long long)rows * cols;
  DDPB_CUDA_OK(launch_k(splitk_reduce_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0,
                        (cudaStream_t)stream, partial, nsplit, split_stride, rows, cols, ld, bias, relu, out,
                        ld_out));
  return 0;
}

// row_splits > 1: out receives row_splits partial sums [row_splits][cols] over consecutive row ranges of
// ceil(rows / row_splits) rows (sum them with a second call) — for tall matrices.
int ddpb_colsum(const float* src, const float* mask, int rows, int cols, long long ld, float* out, int row_splits,
                void* stream) {
  if (rows <= 0 || cols <= 0 || row_splits < 1 || row_splits > 65535) return DDPB_ERR_SHAPE;
  const int rps = (rows + row_splits - 1) / row_splits;
  if ((long long)rps * (row_splits - 1) >= rows) return DDPB_ERR_SHAPE;  // every split must own a row
  DDPB_CUDA_OK(launch_k(colsum_kernel, dim3((cols + 31) / 32, row_splits), dim3(256), 0, (cudaStream_t)stream, src,
                        mask, rows, cols, ld, out, rps));
  return 0;
}

// out[b][c][H/2][W/2] = relu(maxpool2x2(in[b][c][H][W]) + bias[c]) (bias optional), code = winner index.
int ddpb_relu_maxpool2_fwd(const float* in, const float* bias, float* out, unsigned char* code, int batch, int C,
                           int H, int W, void* stream) {
  if (batch <= 0 || C <= 0 || H < 2 || W < 4 || (H & 1) || (W & 3)) return DDPB_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(in) & 15) || (reinterpret_cast<uintptr_t>(out) & 7)) return DDPB_ERR_SHAPE;
  const long long n = (long long)batch * C * (H / 2) * (W / 4);
  DDPB_CUDA_OK(launch_k(relu_maxpool2_fwd_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0,
                        (cudaStream_t)stream, in, bias, out, code, (long long)batch * C, C, H, W));
  return 0;
}

// din[b][c][H][W] = gradient of the above wrt `in` given dout[b][c][H/2][W/2] and the forward's code.
int ddpb_relu_maxpool2_bwd(const float* dout, const unsigned char* code, float* din, int batch, int C, int H, int W,
                           void* stream) {
  if (batch <= 0 || C <= 0 || H < 2 || W < 4 || (H & 1) || (W & 3)) return DDPB_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(din) & 15) || (reinterpret_cast<uintptr_t>(dout) & 7)) return DDPB_ERR_SHAPE;
  const long long n = (long long)batch * C * (H / 2) * (W / 4);
  DDPB_CUDA_OK(launch_k(relu_maxpool2_bwd_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0,
                        (cudaStream_t)stream, dout, code, din, (long long)batch * C, H, W));
  return 0;
}

}  // extern "C"
