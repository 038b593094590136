// Library-private stream-ordered memory pool used when the caller passes no bgb_allocator.
// It keeps its pages across calls: the CUDA default pool returns them to the driver at every stream synchronise, which
// costs hundreds of milliseconds for the tens-of-GB group tables of the high-cardinality configs.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace bgb {
// nullptr on failure (error state cleared). transient: scratch that is freed again before the call returns (record buffers):
// ANY cached block of the stream that is large enough is taken, however much larger -- it is back in the cache within the call
void* pool_alloc_async(size_t bytes, cudaStream_t s, bool transient = false);
void pool_free_async(void* p, cudaStream_t s);
void pool_trim();  // hand every cached block of the current device back to the driver
void pool_release_stream(cudaStream_t s);  // hand back the blocks cached for stream `s` (call before destroying the stream)
}  // namespace bgb
