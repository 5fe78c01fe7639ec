// Debug aid: fills the shared memory and the tensor memory of every SM with a chosen bit pattern, so that a kernel which
// reads on-chip memory it has not written itself gives a different answer afterwards.
//   nvcc -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC -o tools/libdirty.so tools/dirty_chip.cu
#include <cuda_runtime.h>
#include <stdint.h>
__global__ void __launch_bounds__(128, 1) dirty_kernel(uint32_t pat, int nwords)
{
    extern __shared__ uint32_t sm[];
    __shared__ uint32_t tbase;
    for (int i = threadIdx.x; i < nwords; i += blockDim.x) sm[i] = pat;
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tbase)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t t = tbase + (((threadIdx.x >> 5) * 32u) << 16);
    for (int c = 0; c < 512; c += 8)
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(t + c), "r"(pat) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(512u) : "memory");
}
extern "C" int dirty_chip(uint32_t pat, int sms)
{
    const int bytes = 220 * 1024;
    cudaFuncSetAttribute(dirty_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    dirty_kernel<<<sms, 128, bytes>>>(pat, bytes / 4);
    return (int)cudaDeviceSynchronize();
}
