// tools/graph_timeline.py: one-thread kernel that writes %globaltimer (ns) into a slot -- launched before and after every
// C-ABI call of a captured training step, so that a graph replay leaves a timeline of its branches behind.
#include <cuda_runtime.h>
__global__ void stamp_kernel(unsigned long long* slot) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    *slot = t;
}
extern "C" int nwb_tool_stamp(unsigned long long* slots, int index, void* stream) {
    stamp_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(slots + index);
    return static_cast<int>(cudaGetLastError());
}
