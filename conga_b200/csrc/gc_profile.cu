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
// HBM-bound byte work: n_bases bytes read once, n_bases / 100 bytes written.
#include "cg_internal.h"

namespace {

// bit 7 of each byte of the result is set where the byte of w is one of "GCgc"
__device__ __forceinline__ unsigned gc_flags(unsigned w)
{
    // clearing bit 5 maps only 'c','g' onto 'C','G' (what toupper does for them); 'C' ^ 'G' == 0x04
    const unsigned y = ((w & 0xDFDFDFDFu) ^ 0x43434343u) & 0xFBFBFBFBu;
    const unsigned t = ((y & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y;      // bit 7 set where the byte is non-zero
    return ~t & 0x80808080u;
}

// Chunks of GT_WINDOWS windows are staged in shared memory by 1-D TMA bulk copies (cp.async.bulk + mbarrier,
// GT_STAGES deep, issued by thread 0) and one thread sums one window: 26 conflict-free word loads (stride 25
// words between neighbouring threads), 4 logic ops per word to flag its G/C bytes, one IDP4A to count them.
// Measured r01g (B200): 6.15 TB/s = 0.94 of the copy peak on 1.5 G bases; 56 us for chr1 (249 M bases, 0.68:
// about 16 us of launch and first-wave latency do not scale).  A first version with direct 16-byte loads and
// shared-memory atomics per 16-byte block reached 0.55 on chr1: twice the instructions per byte.
constexpr int GT_PAD     = 16;                          // bytes in front of a stage: window 0 starts at byte -1
template <int GT_WINDOWS, int GT_STAGES> struct gt_cfg {
    static constexpr int STAGE_BYTES = GT_PAD + 100 * GT_WINDOWS + 32;      // + the 16-byte blocks at both ends
    static constexpr int SMEM = GT_STAGES * STAGE_BYTES + GT_STAGES * 8;
};

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}

struct gc_chunk { long long first_blk; int head, bytes, n_w, b0; };

template <int GT_WINDOWS>
__device__ __forceinline__ gc_chunk gc_chunk_of(const char* bases, int chunk, int n_windows)
{
    gc_chunk c;
    const int w_first = chunk * GT_WINDOWS;
    c.n_w = min(GT_WINDOWS, n_windows - w_first);
    c.b0 = w_first == 0;
    const long long lo = c.b0 ? 0 : 100ll * w_first - 1;
    const long long hi = 100ll * (w_first + c.n_w) - 1;
    const unsigned long long addr0 = reinterpret_cast<unsigned long long>(bases);
    c.first_blk = (long long)((addr0 + (unsigned long long)lo) & ~15ull) - (long long)addr0;
    c.head = (int)(lo - c.first_blk);
    c.bytes = ((int)(hi - c.first_blk) + 15) & ~15;
    return c;
}

template <int GT_WINDOWS, int GT_STAGES>
__global__ void __launch_bounds__(GT_WINDOWS) k_gc_profile(const char* __restrict__ bases, signed char* __restrict__ out, int n_windows)
{
    constexpr int GT_STAGE_BYTES = gt_cfg<GT_WINDOWS, GT_STAGES>::STAGE_BYTES;
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(smem + GT_STAGES * GT_STAGE_BYTES);
    const int n_chunks = (n_windows + GT_WINDOWS - 1) / GT_WINDOWS;
    const int mine = (int)blockIdx.x < n_chunks ? (n_chunks - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

    auto issue = [&](int k) {                           // one thread: chunk k of this CTA into stage k % STAGES
        const gc_chunk c = gc_chunk_of<GT_WINDOWS>(bases, blockIdx.x + k * gridDim.x, n_windows);
        const unsigned bar = smem_u32(&bars[k % GT_STAGES]);
        const unsigned dst = smem_u32(smem + (k % GT_STAGES) * GT_STAGE_BYTES + GT_PAD);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(c.bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(bases + c.first_blk), "r"(c.bytes), "r"(bar) : "memory");
    };

    if (threadIdx.x == 0) {
        for (int s = 0; s < GT_STAGES; s++)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bars[s])), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < GT_STAGES * 4)                    // the pad in front of each stage: never a G or C
        reinterpret_cast<unsigned*>(smem + (threadIdx.x >> 2) * GT_STAGE_BYTES)[threadIdx.x & 3] = 0;
    __syncthreads();
    if (threadIdx.x == 0)
        for (int k = 0; k < GT_STAGES && k < mine; k++) issue(k);

    for (int k = 0; k < mine; k++) {
        const int chunk = blockIdx.x + k * gridDim.x;
        const gc_chunk c = gc_chunk_of<GT_WINDOWS>(bases, chunk, n_windows);
        mbar_wait(smem_u32(&bars[k % GT_STAGES]), (unsigned)(k / GT_STAGES) & 1u);
        if ((int)threadIdx.x < c.n_w) {
            // window t of the chunk: stage bytes [s, s + 100), less its first byte for window 0 of the chromosome
            const int s = GT_PAD + c.head - c.b0 + 100 * (int)threadIdx.x;
            const unsigned* w = reinterpret_cast<const unsigned*>(smem + (k % GT_STAGES) * GT_STAGE_BYTES) + (s >> 2);
            const int sh = s & 3;
            const int skip = sh + ((c.b0 && threadIdx.x == 0) ? 1 : 0);          // 0..4 low bytes of word 0 dropped
            const unsigned m_first = skip >= 4 ? 0u : 0xFFFFFFFFu << (8 * skip);
            const unsigned m_last = sh == 0 ? 0u : 0xFFFFFFFFu >> (8 * (4 - sh));  // word 25: its low `sh` bytes
            unsigned acc0 = 0, acc1 = 0;
#pragma unroll
            for (int j = 0; j < 26; j++) {
                unsigned z = gc_flags(w[j]);
                if (j == 0) z &= m_first;
                if (j == 25) z &= m_last;
                if (j & 1) acc1 = __dp4a(z, 0x01010101u, acc1);                 // 128 per G/C byte
                else acc0 = __dp4a(z, 0x01010101u, acc0);
            }
            out[chunk * GT_WINDOWS + threadIdx.x] = (signed char)((acc0 + acc1) >> 7);
        }
        __syncthreads();                                                        // the stage is free again
        if (threadIdx.x == 0 && k + GT_STAGES < mine) issue(k + GT_STAGES);
    }
}

}  // namespace

constexpr int GT_W = 256, GT_S = 3, GT_PER_SM = 2;     // (256,4) (256,2) (128,4) (512,2) measured within 3 % of this

cudaError_t cgk_gc_profile(cudaStream_t s, const char* bases, long long n_bases, int8_t* out, int n_windows, int n_sms)
{
    (void)n_bases;
    if (n_windows <= 0) return cudaSuccess;
    const int chunks = (n_windows + GT_W - 1) / GT_W;
    const int blocks = chunks < n_sms * GT_PER_SM ? chunks : n_sms * GT_PER_SM;
    k_gc_profile<GT_W, GT_S><<<blocks, GT_W, gt_cfg<GT_W, GT_S>::SMEM, s>>>(bases, reinterpret_cast<signed char*>(out), n_windows);
    return cudaGetLastError();
}

template <int W, int S> static cudaError_t gt_configure()
{
    cudaError_t e = cudaFuncSetAttribute(k_gc_profile<W, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, gt_cfg<W, S>::SMEM);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_gc_profile<W, S>, cudaFuncAttributePreferredSharedMemoryCarveout, cgk_carveout());
}

cudaError_t cgk_configure_gc_profile(void) { return gt_configure<GT_W, GT_S>(); }
