// tmem_probe.cu -- two hardware questions behind the design of conv_tc.cu (run on a B200):
//   1. how many bytes per clock can tcgen05.ld read from tensor memory, SM-wide, with 4 / 8 / 16 reading warps?
//   2. how does tcgen05.mma kind::f16 lay out a float16 accumulator (D format f16) in tensor memory?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/tmem_probe tools/tmem_probe.cu
#include "../explainn_b200/csrc/tc_common.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>

__global__ void __launch_bounds__(512, 1) ld_bw_kernel(int iters, int ncols, long long *clk_out, float *sink)
{
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tc_alloc(&tmem_base_s, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t taddr = tmem_base_s + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * ncols);
    float acc = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        for (int c = 0; c < ncols; c += 32) {
            uint32_t r[32];
            tc_ld16_nowait(taddr + c, r);
            tc_ld16_nowait(taddr + c + 16, r + 16);
            tc_wait_ld32(r);
            acc += __uint_as_float(r[0] ^ r[17] ^ r[31]);
        }
    }
    const long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) clk_out[blockIdx.x] = t1 - t0;
    if (acc == 123.456f) sink[0] = acc;
    __syncthreads();
    if (warp == 0) tc_dealloc(tmem_base_s, 512);
}

// D (128 x 64) = A (128 x 16) * B (64 x 16)^T with A[r][0] = r + 1, B[n][0] = n + 1, everything else 0.
__global__ void __launch_bounds__(128, 1) f16_acc_kernel(int d_is_f16, uint32_t *out /* (128 lanes, 64 words) */)
{
    __shared__ __align__(128) __half sA[128 * 16];
    __shared__ __align__(128) __half sB[64 * 16];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // K-major core-matrix layout, K = 16: element (r, k) at (r/8)*128 + (k/8)*64 + (r%8)*8 + k%8   (halfs)
    for (int i = threadIdx.x; i < 128 * 16; i += 128) {
        const int r = i / 16, k = i % 16;
        sA[(r >> 3) * 128 + (k >> 3) * 64 + (r & 7) * 8 + (k & 7)] = __float2half(k == 0 ? (float)(r + 1) : 0.f);
    }
    for (int i = threadIdx.x; i < 64 * 16; i += 128) {
        const int r = i / 16, k = i % 16;
        sB[(r >> 3) * 128 + (k >> 3) * 64 + (r & 7) * 8 + (k & 7)] = __float2half(k == 0 ? (float)(r + 1) : 0.f);
    }
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) tc_alloc(&tmem_base_s, 128);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tmem_base_s;
    if (tc_warp_uniform() == 0) {
        // LBO = distance between K core matrices (128 B), SBO = distance between 8-row groups (256 B)
        const uint64_t dA = make_smem_desc(smem_u32(sA), 128, 256), dB = make_smem_desc(smem_u32(sB), 128, 256);
        uint32_t idesc = ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);      // A, B f16, K-major
        if (!d_is_f16) idesc |= 1u << 4;
        if (tc_elect_one()) {
            tc_mma_f16(tb, dA, dB, idesc, 0u);
            tc_commit(&bar);
        }
        __syncwarp();
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    float v[32];
    const uint32_t taddr = tb + ((uint32_t)(32 * warp) << 16);
    tc_ld32(taddr, v);
    for (int i = 0; i < 32; ++i) out[(32 * warp + lane) * 64 + i] = __float_as_uint(v[i]);
    tc_ld32(taddr + 32, v);
    for (int i = 0; i < 32; ++i) out[(32 * warp + lane) * 64 + 32 + i] = __float_as_uint(v[i]);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tc_dealloc(tb, 128);
}

static float h2f(uint16_t h)
{
    __half_raw r; r.x = h;
    return __half2float(__half(r));
}

int main()
{
    int dev_clk = 0;
    cudaDeviceGetAttribute(&dev_clk, cudaDevAttrClockRate, 0);
    long long *clk; float *sink; uint32_t *out;
    cudaMalloc(&clk, 8 * 148); cudaMalloc(&sink, 4); cudaMalloc(&out, 128 * 64 * 4);
    printf("== tcgen05.ld bandwidth (one CTA per SM on all SMs; 32x32b.x16 pairs, bytes = warps*32 lanes*cols*4) ==\n");
    for (int warps : {4, 8, 16}) {
        const int ncols = 512 / (warps / 4), iters = 2000;
        ld_bw_kernel<<<148, warps * 32, 0>>>(iters, ncols, clk, sink);
        cudaDeviceSynchronize();
        ld_bw_kernel<<<148, warps * 32, 0>>>(iters, ncols, clk, sink);
        cudaError_t e = cudaDeviceSynchronize();
        long long h = 0;
        cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
        const double bytes = (double)warps * 32 * ncols * 4 * iters;
        printf("warps %2d cols/warp %3d: %lld clk, %.1f B/clk/SM (%s)\n", warps, ncols, h, bytes / (double)h, cudaGetErrorString(e));
    }
    std::vector<uint32_t> host(128 * 64);
    for (int f16 = 0; f16 < 2; ++f16) {
        cudaMemset(out, 0xFF, 128 * 64 * 4);
        f16_acc_kernel<<<1, 128>>>(f16, out);
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpy(host.data(), out, host.size() * 4, cudaMemcpyDeviceToHost);
        printf("== accumulator format %s (%s): lane 1 (expected D[1][n] = 2(n+1)) ==\n", f16 ? "f16" : "f32", cudaGetErrorString(e));
        for (int c = 0; c < 64; ++c) {
            const uint32_t w = host[1 * 64 + c];
            if (f16) printf("c%02d %08x lo=%g hi=%g\n", c, w, h2f((uint16_t)(w & 0xFFFF)), h2f((uint16_t)(w >> 16)));
            else { float f; memcpy(&f, &w, 4); printf("c%02d %g\n", c, f); }
            if (c == 7 && !f16) break;
        }
    }
    return 0;
}
