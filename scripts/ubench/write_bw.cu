// Store-bandwidth microbenchmark: what can a pure write stream reach on this GPU?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o write_bw write_bw.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__device__ __forceinline__ void st16(void* p, unsigned v)
{
    if (MODE == 0) asm volatile("st.global.v4.u32 [%0], {%1,%1,%1,%1};" :: "l"(p), "r"(v) : "memory");
    if (MODE == 1) asm volatile("st.global.cs.v4.u32 [%0], {%1,%1,%1,%1};" :: "l"(p), "r"(v) : "memory");
    if (MODE == 2) asm volatile("st.global.wt.v4.u32 [%0], {%1,%1,%1,%1};" :: "l"(p), "r"(v) : "memory");
    if (MODE == 3) asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%1,%1,%1};" :: "l"(p), "r"(v) : "memory");
}

// each warp writes contiguous `span` bytes per iteration (span = 512 * UNR), warps grid-stride
template <int MODE, int UNR>
__global__ void k_fill(char* dst, size_t bytes, unsigned v)
{
    const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((size_t)gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t span = 512 * UNR;
    for (size_t off = warp * span; off + span <= bytes; off += nw * span) {
#pragma unroll
        for (int u = 0; u < UNR; u++) st16<MODE>(dst + off + u * 512 + lane * 16, v);
    }
}

// 32-byte stores (sm_100 256-bit)
template <int UNR>
__global__ void k_fill32(char* dst, size_t bytes, unsigned v)
{
    const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((size_t)gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t span = 1024 * UNR;
    for (size_t off = warp * span; off + span <= bytes; off += nw * span) {
#pragma unroll
        for (int u = 0; u < UNR; u++)
            asm volatile("st.global.v8.u32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" :: "l"(dst + off + u * 1024 + lane * 32), "r"(v) : "memory");
    }
}

template <typename F> float timeit(F f)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 2; i++) f();
    float best = 1e9;
    for (int i = 0; i < 5; i++) { cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms; }
    return best;
}

int main()
{
    const size_t bytes = (size_t)3 << 30;
    char* d; cudaMalloc(&d, bytes);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    printf("cudaMemset %.0f GB/s\n", bytes / timeit([&] { cudaMemsetAsync(d, 0, bytes); }) / 1e6);
    for (int cps : {2, 4, 8}) {
        const int g = sms * cps;
        printf("ctas/sm %d: st %.0f  st.cs %.0f  st.wt %.0f  no_alloc %.0f | unr8: st %.0f st.cs %.0f | v8 %.0f v8x4 %.0f GB/s\n", cps,
               bytes / timeit([&] { k_fill<0, 1><<<g, 256>>>(d, bytes, 1); }) / 1e6,
               bytes / timeit([&] { k_fill<1, 1><<<g, 256>>>(d, bytes, 1); }) / 1e6,
               bytes / timeit([&] { k_fill<2, 1><<<g, 256>>>(d, bytes, 1); }) / 1e6,
               bytes / timeit([&] { k_fill<3, 1><<<g, 256>>>(d, bytes, 1); }) / 1e6,
               bytes / timeit([&] { k_fill<0, 8><<<g, 256>>>(d, bytes, 1); }) / 1e6,
               bytes / timeit([&] { k_fill<1, 8><<<g, 256>>>(d, bytes, 1); }) / 1e6,
               bytes / timeit([&] { k_fill32<1><<<g, 256>>>(d, bytes, 1); }) / 1e6,
               bytes / timeit([&] { k_fill32<4><<<g, 256>>>(d, bytes, 1); }) / 1e6);
    }
    // two streams interleaved like depth (2 B/pos) + prefix (4 B/pos)
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
