// Thread-local error reporting behind the C ABI (no exception crosses the boundary).
#pragma once
#include <cuda_runtime.h>
#include "../../include/bgb.h"

namespace bgb {
int set_error(int status, const char* fmt, ...);          // stores the message, returns status
int check_cuda(cudaError_t e, const char* what);
cudaError_t note_launch();                                // counts one kernel launch, returns cudaGetLastError()          // BGB_OK or BGB_CUDA_ERROR / BGB_OUT_OF_MEMORY with message
int device_sm_count();                                    // SMs of the current device (cached per device; 148 on B200)
}  // namespace bgb
