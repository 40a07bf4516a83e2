// pdl.cuh -- programmatic dependent launch: a kernel launched with launch_dependent() may be set up
// and scheduled while the kernel before it on the stream is still running; it must not touch anything
// that kernel writes before it has passed griddep_wait(), which returns once that kernel has finished
// and its writes are visible.  What is gained is the launch latency of the second kernel (a few
// microseconds per boundary, which is a tenth of state.New for text in the indented form).
#pragma once
#include <cuda_runtime.h>

namespace tk {

__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <class... KArgs, class... Args>
inline cudaError_t launch_dependent(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace tk
