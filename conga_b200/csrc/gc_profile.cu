// SONIC's GC profile of one chromosome from its bases.
//
// Replaces the character loop of sonic_write_gc_profile (sonic/sonic_reference.c:30-106) for one FASTA
// record whose white space has been removed:
//   * char_count starts at 1 at the '>' line (:82) and is incremented before the test (:90), so the first
//     byte closes after 99 bases and every later one after 100: base r (0-based) belongs to window
//     (r + 1) / 100, i.e. window 0 = bases [0, 99), window k = [100k - 1, 100k + 99);
//   * a byte is written only when its window is complete (:95): floor((n_bases + 1) / 100) bytes;
//   * the byte is (char)(100.0 * gc / 100) (:96) = the number of G/C/g/c characters in the window
//     (the double product and quotient are exact for 0..100); every other character — N included — counts
//     as a base and not as GC (:88-93).
// HBM-bound byte work: n_bases bytes read once with 16-byte loads, n_bases / 100 bytes written.
#include "cg_internal.h"

namespace {

constexpr int GC_WINDOWS = 128;          // windows per CTA: 12.8 kB of bases, 128 bytes out
constexpr int GC_THREADS = 256;

// bit 7 of each byte of the result is set where the byte of w is one of "GCgc"
__device__ __forceinline__ unsigned gc_flags(unsigned w)
{
    // clearing bit 5 maps only 'c','g' onto 'C','G' (what toupper does for them); 'C' ^ 'G' == 0x04
    const unsigned y = ((w & 0xDFDFDFDFu) ^ 0x43434343u) & 0xFBFBFBFBu;
    const unsigned t = ((y & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y;      // bit 7 set where the byte is non-zero
    return ~t & 0x80808080u;
}

// bits 7,15,23,31 -> bits 0..3
__device__ __forceinline__ unsigned nibble(unsigned z) { return (((z >> 7) & 0x01010101u) * 0x01020408u) >> 24; }

__global__ void __launch_bounds__(GC_THREADS) k_gc_profile(const char* __restrict__ bases, long long n_bases,
                                                           signed char* __restrict__ out, int n_windows)
{
    __shared__ int bins[GC_WINDOWS];
    for (int w_first = blockIdx.x * GC_WINDOWS; w_first < n_windows; w_first += gridDim.x * GC_WINDOWS) {
        const int n_w = min(GC_WINDOWS, n_windows - w_first);
        if (threadIdx.x < GC_WINDOWS) bins[threadIdx.x] = 0;
        __syncthreads();
        const long long lo = w_first == 0 ? 0 : 100ll * w_first - 1;           // first base of the first window
        const long long hi = 100ll * (w_first + n_w) - 1;                      // one past the last base of the last
        // 16-byte blocks of the ADDRESS space (the caller's pointer need not be aligned); the bytes outside
        // [lo, hi) that such a block drags in lie in the same 16-byte granule as a valid one
        const unsigned long long addr0 = reinterpret_cast<unsigned long long>(bases);
        const long long first_blk = (long long)((addr0 + (unsigned long long)lo) & ~15ull) - (long long)addr0;
        for (long long o = first_blk + 16ll * threadIdx.x; o < hi; o += 16ll * GC_THREADS) {
            const uint4 v = __ldcs(reinterpret_cast<const uint4*>(bases + o));
            const unsigned m = nibble(gc_flags(v.x)) | (nibble(gc_flags(v.y)) << 4) | (nibble(gc_flags(v.z)) << 8) |
                               (nibble(gc_flags(v.w)) << 12);                  // bit j: base o + j is G or C
            const long long r0 = max(o, lo), r1 = min(o + 16, hi);
            const int w0 = (int)((r0 + 1) / 100);
            const long long split = min(r1, 100ll * (w0 + 1) - 1);             // first base of window w0 + 1
            const int a = (int)(r0 - o), b = (int)(split - o), e = (int)(r1 - o);
            const int c0 = __popc(m & ((1u << b) - 1u) & ~((1u << a) - 1u));
            const int c1 = __popc(m & ((1u << e) - 1u) & ~((1u << b) - 1u));
            if (c0) atomicAdd(&bins[w0 - w_first], c0);
            if (c1) atomicAdd(&bins[w0 + 1 - w_first], c1);
        }
        __syncthreads();
        if (threadIdx.x < n_w) out[w_first + threadIdx.x] = (signed char)bins[threadIdx.x];
        __syncthreads();
    }
}

}  // namespace

cudaError_t cgk_gc_profile(cudaStream_t s, const char* bases, long long n_bases, int8_t* out, int n_windows, int n_sms)
{
    if (n_windows <= 0) return cudaSuccess;
    const int chunks = (n_windows + GC_WINDOWS - 1) / GC_WINDOWS;
    const int blocks = chunks < n_sms * 16 ? chunks : n_sms * 16;
    k_gc_profile<<<blocks, GC_THREADS, 0, s>>>(bases, n_bases, reinterpret_cast<signed char*>(out), n_windows);
    return cudaGetLastError();
}

cudaError_t cgk_configure_gc_profile(void)
{
    return cudaFuncSetAttribute(k_gc_profile, cudaFuncAttributePreferredSharedMemoryCarveout, cgk_carveout());
}
