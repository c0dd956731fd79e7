// Layout probe for tcgen05.mma kind::tf32 operands in the no-swizzle canonical layout.
// The operand under test is filled with its own word index (exact in tf32 below 2048) and multiplied
// by a K-major identity, so the accumulator shows which shared-memory word the tensor core read for
// every (row, k).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -I csrc tools/umma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "ptx.cuh"
using namespace ddpb;

// mode 0: A MN-major under test (M=64), B = identity K-major (N=8).  D[m][n] = A(m, k=n)
// mode 1: B MN-major under test (N=32), A = identity K-major (M=64).  D[m][n] = B(n, k=m) for m<8
__global__ void probe(int mode, uint32_t lbo, uint32_t sbo, int words, float* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tm;
  float* T = reinterpret_cast<float*>(smem);            // operand under test: `words` floats
  float* I = reinterpret_cast<float*>(smem + 32768);    // identity operand, K-major: [2 kgroups][rows][4]
  for (int i = threadIdx.x; i < words; i += blockDim.x) T[i] = (float)i;
  for (int i = threadIdx.x; i < 2 * 64 * 4; i += blockDim.x) {
    const int kg = i / 256, r = (i % 256) / 4, kk = i % 4;
    I[i] = (r == kg * 4 + kk) ? 1.f : 0.f;  // row r has a one at k = r (r < 8)
  }
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tm), 32); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tm;
  if (threadIdx.x == 0) {
    const uint32_t sT = smem_u32(T), sI = smem_u32(I);
    if (mode == 0) {
      const uint32_t idesc = umma_idesc_tf32(64, 8, 1, 0);
      umma_tf32(tb, umma_smem_desc(sT, lbo, sbo), umma_smem_desc(sI, 64 * 16, 128), idesc, 0);
    } else {
      const uint32_t idesc = umma_idesc_tf32(64, 32, 0, 1);
      umma_tf32(tb, umma_smem_desc(sI, 64 * 16, 128), umma_smem_desc(sT, lbo, sbo), idesc, 0);
    }
    umma_commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  tc_fence_after();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t v[32];
  tmem_ld32(tb + ((uint32_t)(warp * 32) << 16), v);
  tmem_ld_wait();
  if (lane < 16) {  // M=64: row m = warp*16 + lane
    const int m = warp * 16 + lane;
    for (int c = 0; c < 32; ++c) out[m * 32 + c] = __uint_as_float(v[c]);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tb, 32);
}

int main() {
  float* d;
  cudaMalloc(&d, 64 * 32 * 4);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  std::vector<float> h(64 * 32);
  struct Cfg { int mode; uint32_t lbo, sbo; int words; const char* what; };
  const Cfg cfgs[] = {
      {0, 128, 512, 2048, "A MN-major: lbo=128 sbo=512"},
      {0, 512, 128, 2048, "A MN-major: lbo=512 sbo=128"},
      {1, 128, 512, 2048, "B MN-major: lbo=128 sbo=512"},
      {1, 512, 128, 2048, "B MN-major: lbo=512 sbo=128"},
  };
  for (const Cfg& c : cfgs) {
    cudaMemset(d, 0xff, 64 * 32 * 4);
    probe<<<1, 128, 65536>>>(c.mode, c.lbo, c.sbo, c.words, d);
    cudaError_t e = cudaDeviceSynchronize();
    printf("== %s : %s\n", c.what, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    cudaMemcpy(h.data(), d, 64 * 32 * 4, cudaMemcpyDeviceToHost);
    if (c.mode == 0) {
      for (int m = 0; m < 64; m += (m < 9 ? 1 : 11)) {
        printf("  m=%2d words(k=0..7):", m);
        for (int k = 0; k < 8; ++k) printf(" %5.0f", h[m * 32 + k]);
        printf("\n");
      }
    } else {
      for (int n = 0; n < 32; n += (n < 9 ? 1 : 5)) {
        printf("  n=%2d words(k=0..7):", n);
        for (int k = 0; k < 8; ++k) printf(" %5.0f", h[k * 32 + n]);
        printf("\n");
      }
    }
  }
  return 0;
}
