// Test helper: a kernel that keeps one CTA busy for a given number of nanoseconds.  The stress tests of the
// cross-GPU flag protocols put it in front of the publish / pull / all-reduce kernels with random lengths to shake
// the order in which the two sides reach their flags.  Not part of libdtf_b200.so.
#include <cuda_runtime.h>
#include <stdint.h>

__global__ void debug_spin_kernel(unsigned long long ns) {
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    do {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    } while (t - t0 < ns);
}

extern "C" int dtf_debug_spin(unsigned long long ns, cudaStream_t stream) {
    if (ns > 50000000ull) ns = 50000000ull;      // never more than 50 ms
    debug_spin_kernel<<<1, 1, 0, stream>>>(ns);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}
