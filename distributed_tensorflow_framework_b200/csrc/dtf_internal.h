// Internal shared definitions for the dtf_b200 CUDA library (error codes, helpers).
#pragma once
#include <cuda_runtime.h>

#define DTF_OK 0
#define DTF_ERR_ARG 1
#define DTF_ERR_ALIGN 2
#define DTF_ERR_CUDA 3
#define DTF_ERR_DRIVER 4
#define DTF_ERR_TENSORMAP 5
#define DTF_ERR_TOO_LARGE 6
#define DTF_ERR_UNSUPPORTED 7

#define DTF_CHECK_LAUNCH() (cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA)

static inline cudaStream_t dtf_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
