// Kernel launch helper: every kernel of the training step is launched through launch_k so that
// programmatic dependent launch (PDL) can be switched on for the whole step at once.  With PDL a
// kernel's CTAs may become resident while the previous kernel of the stream is still running; they
// run their private prologue (barrier init, TMEM allocation, constant staging) and then block in
// griddepcontrol.wait until the previous kernel has completed and flushed.  Each kernel calls
// pdl_launch_dependents() right AFTER its own pdl_wait(), so at most two kernels overlap and
// completion stays transitive along the stream.  DDPB_PDL=0 disables it (plain launches).
#pragma once
#include <cuda_runtime.h>

namespace ddpb {

int pdl_enabled();  // simt_ops.cu

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                            cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  if (pdl_enabled()) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace ddpb
