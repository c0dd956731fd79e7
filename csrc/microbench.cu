// Measurement-only kernels (tools/mma_bench.py): the issue cost of tcgen05.mma for the operand
// shapes, majors and start-address alignments the conv kernels use.  Not on any product path.
#include "ptx.cuh"
#include "api_internal.cuh"

namespace ddpb {

__global__ void __launch_bounds__(128, 1)
mma_microbench_kernel(int M, int N, int a_mn, int b_mn, int a_shift16, int b_shift16, int nacc,
                      int iters, uint32_t lbo_a, uint32_t sbo_a, uint32_t lbo_b, uint32_t sbo_b,
                      unsigned long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) unsigned long long bar;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
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
    const uint32_t idesc = umma_idesc_bf16(M, N, a_mn, b_mn);
    const uint32_t sa = smem_u32(smem) + a_shift16 * 16;
    const uint32_t sb = smem_u32(smem) + 48 * 1024 + b_shift16 * 16;
    const uint64_t ad = umma_smem_desc(sa, lbo_a, sbo_a);
    const uint64_t bd = umma_smem_desc(sb, lbo_b, sbo_b);
    // warm-up
    umma_bf16(tmem, ad, bd, idesc, 0);
    umma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) umma_bf16(tmem + (i % nacc) * N, ad, bd, idesc, 1);
    const long long t1 = clock64();
    umma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 1);
    const long long t2 = clock64();
    out[0] = t1 - t0;  // issue time
    out[1] = t2 - t0;  // until all MMAs completed
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
}

}  // namespace ddpb

extern "C" int ddpb_mma_microbench(int M, int N, int a_mn, int b_mn, int a_shift16, int b_shift16,
                                   int nacc, int iters, unsigned lbo_a, unsigned sbo_a,
                                   unsigned lbo_b, unsigned sbo_b, unsigned long long* out_dev,
                                   void* stream) {
  static bool configured = false;
  if (!configured) {
    DDPB_CUDA_OK(cudaFuncSetAttribute(ddpb::mma_microbench_kernel,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    configured = true;
  }
  if (nacc * N > 512) return DDPB_ERR_SHAPE;
  ddpb::mma_microbench_kernel<<<1, 128, 100 * 1024, static_cast<cudaStream_t>(stream)>>>(
      M, N, a_mn, b_mn, a_shift16, b_shift16, nacc, iters, lbo_a, sbo_a, lbo_b, sbo_b, out_dev);
  return static_cast<int>(cudaGetLastError());
}
