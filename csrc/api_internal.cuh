// Internal helpers shared by the translation units behind include/ddp_b200.h.
#pragma once
#include <cuda_runtime.h>

#define DDPB_ERR_SHAPE 9001   // unsupported shape / argument
#define DDPB_ERR_STATE 9002   // call made in the wrong state (e.g. communicator not connected)
#define DDPB_ERR_LIMIT 9003   // size beyond a documented fixed limit
#define DDPB_ERR_UNSUPPORTED 9004  // feature not available on this driver / system (e.g. multicast)

#define DDPB_CUDA_OK(expr)                              \
  do {                                                  \
    cudaError_t _e = (expr);                            \
    if (_e != cudaSuccess) return static_cast<int>(_e); \
  } while (0)

extern "C" int ddpb_num_sms();
