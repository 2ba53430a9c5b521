/* no_alloc.cu — cg_depth_stats_noalloc: the depth side of one chromosome (k_mark -> k_depth_pass -> k_finalize)
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.
 *
 * The kernels are the ones every other entry point launches.  What differs is where their scratch lives: the chunk
 * bounds, the three levels of depth sums, the per-CTA partial rows, the control words, k_finalize's scratch row and
 * the int16 depth array itself sit in ONE file-scope __device__ array of fixed size (g_scratch below, 17.4 MB, held by
 * every GPU that loads the library until the process exits) instead of a context's slot.  Consequences:
 *   - chrom_len is limited to CG_NOALLOC_MAX_LEN; larger sizes are rejected, never truncated;
 *   - every call uses the same array, which keeps the previous call's contents: the words that must start at zero are
 *     cleared at the head of every call, on the caller's stream;
 *   - two calls that overlap on one GPU (different streams, or different threads) race on the array.  Calls queued on
 *     one stream are ordered and safe.
 * The host never uses g_scratch's name as an address (in host code it is a stand-in symbol): the device address is
 * looked up once per device with cudaGetSymbolAddress, with that device current. */
#include "cg_internal.h"
#include <mutex>

#define NA_SPACE      (4 * CG_SUPER_CHUNKS)                      /* chunks of the chunk space: 4 supers */
static_assert((long long)(NA_SPACE - 1) * CG_CHUNK == CG_NOALLOC_MAX_LEN, "CG_NOALLOC_MAX_LEN (conga_gpu.h) = the chunk space minus one chunk");
#define NA_PASS_CTAS  640                                        /* rows of the partial table */
#define NA_MAX_DEV    64

/* layout of g_scratch, in bytes; every region starts on a multiple of 256 */
#define NA_OFF_CTL      0                                        /* [0, 256): job control words (8), slot words (4), sticky, flags */
#define NA_OFF_FIN      256                                      /* k_finalize: int64[256] scratch row + counter */
#define NA_OFF_COUNTS   (NA_OFF_FIN + 256 * 8 + 256)
#define NA_COUNT_WORDS  (2 * NA_SPACE + NA_SPACE / 32 + 64)      /* >= cgk_counts_words(NA_SPACE), checked at set-up */
#define NA_OFF_ZERO_END (NA_OFF_COUNTS + NA_COUNT_WORDS * 4)     /* everything in front of this is cleared per call */
#define NA_OFF_ROW      NA_OFF_ZERO_END                          /* int64[204] statistics row (+ padding) */
#define NA_OFF_EXPECTED (NA_OFF_ROW + 2048)                      /* float[128] */
#define NA_OFF_PARTIAL  (NA_OFF_EXPECTED + 512)
#define NA_OFF_DEPTH    (NA_OFF_PARTIAL + ((NA_PASS_CTAS * CG_PART_WORDS * 4 + 255) / 256) * 256)
#define NA_BYTES        ((size_t)NA_OFF_DEPTH + (size_t)NA_SPACE * CG_CHUNK * 2)
static_assert(NA_OFF_ZERO_END % 256 == 0 && NA_OFF_DEPTH % 256 == 0, "regions are 256-byte aligned");

static __device__ __align__(256) unsigned char g_scratch[NA_BYTES];

struct na_device { bool ready = false; unsigned char* base = nullptr; int n_sms = 0; };
static na_device g_dev[NA_MAX_DEV];
static std::mutex g_mu;

/* once per device, with that device current */
static int na_setup(int device, na_device** out)
{
    std::lock_guard<std::mutex> lock(g_mu);
    na_device& d = g_dev[device];
    if (!d.ready) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
        if (e != cudaSuccess || !p) return cg_fail(nullptr, CG_ECUDA, "cg_depth_stats_noalloc: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(e));
        int major = 0, sms = 0;
        if ((e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device)) != cudaSuccess
            || (e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess)
            return cg_fail(nullptr, CG_ECUDA, "cg_depth_stats_noalloc: device %d: %s", device, cudaGetErrorString(e));
        if (major != 10) return cg_fail(nullptr, CG_ECUDA, "cg_depth_stats_noalloc: device %d is not sm_100 (compute capability %d.x)", device, major);
        if ((e = cgk_configure_depth()) != cudaSuccess)
            return cg_fail(nullptr, CG_ECUDA, "cg_depth_stats_noalloc: kernel attributes on device %d: %s", device, cudaGetErrorString(e));
        if (cgk_counts_words(NA_SPACE) > (size_t)NA_COUNT_WORDS) return cg_fail(nullptr, CG_ECUDA, "cg_depth_stats_noalloc: scratch layout too small for the count arrays");
        d.base = static_cast<unsigned char*>(p); d.n_sms = sms; d.ready = true;
    }
    *out = &d;
    return CG_OK;
}

extern "C" int cg_depth_stats_noalloc(int device, void* stream, const int32_t* pos, const uint8_t* mapq, size_t n_reads,
                                      int mq_threshold, int32_t chrom_len, const int8_t* gc_profile, int32_t n_windows,
                                      int16_t* depth_out, cg_chrom_stats* stats_out, float* expected_out, uint32_t* flags_out)
{
    if (device < 0 || device >= NA_MAX_DEV) return cg_fail(nullptr, CG_EINVAL, "cg_depth_stats_noalloc: device %d outside 0..%d", device, NA_MAX_DEV - 1);
    if (chrom_len <= 0) return cg_fail(nullptr, CG_EINVAL, "cg_depth_stats_noalloc: chrom_len %d must be positive", chrom_len);
    if (chrom_len > CG_NOALLOC_MAX_LEN)
        return cg_fail(nullptr, CG_EINVAL, "cg_depth_stats_noalloc: chrom_len %d exceeds CG_NOALLOC_MAX_LEN = %d (the fixed scratch array); use cg_chrom_begin ... cg_chrom_finish",
                       chrom_len, CG_NOALLOC_MAX_LEN);
    const int need_windows = (chrom_len + 99) / 100;
    if (!gc_profile || n_windows < need_windows)
        return cg_fail(nullptr, CG_EINVAL, "cg_depth_stats_noalloc: gc_profile needs >= %d windows for length %d (got %d)", need_windows, chrom_len, n_windows);
    if (n_reads && !pos) return cg_fail(nullptr, CG_EINVAL, "cg_depth_stats_noalloc: null pos");
    if (n_reads >= 0x7fffffffull) return cg_fail(nullptr, CG_EINVAL, "cg_depth_stats_noalloc: more than 2^31-1 records");
    if (!mapq && mq_threshold >= 0) return cg_fail(nullptr, CG_EINVAL, "cg_depth_stats_noalloc: a NULL mapq plane needs a negative mq_threshold (got %d)", mq_threshold);

    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess || (e = cudaSetDevice(device)) != cudaSuccess)
        return cg_fail(nullptr, CG_ECUDA, "cg_depth_stats_noalloc: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    struct restore { int d; ~restore() { if (d >= 0) cudaSetDevice(d); } } back{prev};

    na_device* D = nullptr;
    int rc = na_setup(device, &D);
    if (rc) return rc;
    unsigned char* base = D->base;                                /* device address of g_scratch on this GPU */
    cudaStream_t st = static_cast<cudaStream_t>(stream);
#define NA_CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return cg_fail(nullptr, CG_ECUDA, "cg_depth_stats_noalloc: %s: %s", #x, cudaGetErrorString(e_)); } while (0)

    /* the array keeps the last call's contents: control words, finalize scratch and the count arrays start at zero */
    NA_CU(cudaMemsetAsync(base, 0, NA_OFF_ZERO_END, st));

    unsigned int* ctl = reinterpret_cast<unsigned int*>(base + NA_OFF_CTL);
    const int n_chunks = (chrom_len + CG_CHUNK - 1) / CG_CHUNK;
    cg_dev_job J;
    memset(&J, 0, sizeof J);
    J.rec_pos = pos; J.rec_mapq = mapq;
    J.n_reads = (long long)n_reads; J.mark_first = 0; J.mark_n = (long long)n_reads; J.mark_blk0 = 0;
    J.mq_threshold = mq_threshold; J.chrom_len = chrom_len; J.n_chunks = n_chunks; J.chunk0 = 0; J.work0 = 0;
    J.gc = gc_profile;
    J.stats_row = reinterpret_cast<long long*>(base + NA_OFF_ROW);
    J.row_out = reinterpret_cast<long long*>(stats_out);          /* k_finalize writes the caller's row itself (NULL: not wanted) */
    J.expected = expected_out ? expected_out : reinterpret_cast<float*>(base + NA_OFF_EXPECTED);
    J.ctl = ctl;

    cg_sub S;
    memset(&S, 0, sizeof S);
    S.jobs = nullptr; S.n_jobs = 1; S.n_work = n_chunks; S.n_space = NA_SPACE;
    S.n_mark_blocks = cgk_mark_blocks(n_reads);
    S.k = cgk_counts_layout(reinterpret_cast<int32_t*>(base + NA_OFF_COUNTS), NA_SPACE);
    S.depth = reinterpret_cast<int16_t*>(base + NA_OFF_DEPTH);
    S.depth32 = nullptr;                                          /* no out-of-order path here: records must be sorted */
    S.partial = reinterpret_cast<int32_t*>(base + NA_OFF_PARTIAL);
    S.n_pass_ctas = cgk_pass_grid(n_chunks, (long long)n_reads, D->n_sms);
    if (S.n_pass_ctas > NA_PASS_CTAS) S.n_pass_ctas = NA_PASS_CTAS;
    S.sctl = ctl + CG_CTL_WORDS;
    S.fin_acc = reinterpret_cast<long long*>(base + NA_OFF_FIN);
    S.fin_cnt = reinterpret_cast<unsigned int*>(base + NA_OFF_FIN + 256 * 8);
    S.sticky = ctl + CG_CTL_WORDS + 4;
    S.h_flags = flags_out ? flags_out : ctl + CG_CTL_WORDS + 5;   /* k_finalize stores the job's CG_FLAG_* word here */
    S.n_peer_rows = 0;

    NA_CU(cgk_mark(st, S, J));
    NA_CU(cgk_depth_pass(st, S, J));
    NA_CU(cgk_finalize(st, S, J));
    if (depth_out) NA_CU(cudaMemcpyAsync(depth_out, S.depth, (size_t)chrom_len * sizeof(int16_t), cudaMemcpyDeviceToDevice, st));
#undef NA_CU
    return CG_OK;
}
