// Measurement-only kernels (tools/mma_bench.py): the issue cost of tcgen05.mma for the operand
// shapes, majors and start-address alignments the conv kernels use.  Not on any product path.
#include "ptx.cuh"
#include "api_internal.cuh"

namespace ddpb {

// 32 MMAs fully unrolled per outer iteration, accumulators rotate over NACC column blocks.
template <int M, int N, int AMN, int BMN, int NACC>
__global__ void __launch_bounds__(128, 1)
mma_microbench_kernel(int a_shift16, int b_shift16, int outer, uint32_t lbo_a, uint32_t sbo_a,
                      uint32_t lbo_b, uint32_t sbo_b, unsigned long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) unsigned long long bar;
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_mbar_init();
  }
  if (threadIdx.x < 32) {
    tmem_alloc(smem_u32(&s_tmem), 512);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(M, N, AMN, BMN);
    const uint32_t sa = smem_u32(smem) + a_shift16 * 16;
    const uint32_t sb = smem_u32(smem) + 64 * 1024 + b_shift16 * 16;
    const uint64_t ad = umma_smem_desc(sa, lbo_a, sbo_a);
    const uint64_t bd = umma_smem_desc(sb, lbo_b, sbo_b);
    umma_bf16(tmem, ad, bd, idesc, 0);
    umma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    const long long t0 = clock64();
    for (int o = 0; o < outer; ++o) {
#pragma unroll
      for (int i = 0; i < 32; ++i) umma_bf16(tmem + (i % NACC) * N, ad + 2 * (i & 3), bd, idesc, 1);
    }
    const long long t1 = clock64();
    umma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 1);
    const long long t2 = clock64();
    out[0] = t1 - t0;
    out[1] = t2 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
}

template <int M, int N, int AMN, int BMN, int NACC>
static int run_mb(int ash, int bsh, int outer, unsigned la, unsigned sa, unsigned lb, unsigned sb,
                  unsigned long long* out, cudaStream_t st) {
  auto k = mma_microbench_kernel<M, N, AMN, BMN, NACC>;
  DDPB_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  k<<<1, 128, 200 * 1024, st>>>(ash, bsh, outer, la, sa, lb, sb, out);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace ddpb

using namespace ddpb;

// config: 0..: see table in tools/mma_bench.py
extern "C" int ddpb_mma_microbench(int config, int a_shift16, int b_shift16, int outer, unsigned lbo_a,
                                   unsigned sbo_a, unsigned lbo_b, unsigned sbo_b,
                                   unsigned long long* out_dev, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define MB(c, M, N, A, B, NACC) \
  case c: return run_mb<M, N, A, B, NACC>(a_shift16, b_shift16, outer, lbo_a, sbo_a, lbo_b, sbo_b, out_dev, st)
  switch (config) {
    MB(0, 128, 256, 0, 0, 2);
    MB(1, 128, 128, 0, 0, 4);
    MB(2, 128, 64, 0, 0, 4);
    MB(3, 128, 32, 0, 0, 8);
    MB(4, 128, 16, 0, 0, 8);
    MB(5, 128, 64, 0, 0, 1);
    MB(6, 64, 32, 1, 1, 9);
    MB(7, 64, 96, 1, 1, 3);
    MB(8, 64, 64, 1, 1, 4);
    MB(9, 64, 8, 1, 1, 8);
    MB(10, 128, 32, 1, 1, 9);
    MB(11, 64, 32, 1, 1, 1);
    MB(12, 64, 32, 0, 0, 9);
    MB(13, 64, 256, 0, 0, 2);
    default: return DDPB_ERR_SHAPE;
  }
#undef MB
}
