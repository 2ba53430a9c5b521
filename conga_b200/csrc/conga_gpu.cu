/* conga_gpu.cu — context management and the C-ABI of libconga_gpu.so
 * (include/conga_gpu.h).  Host code only; the kernels live in depth_pass.cu and
 * genotype.cu.  There is no CPU implementation of any stage in this library:
 * without a CUDA device cg_create fails and nothing else can be called.
 *
 * Execution model.  The unit of device work is a SUBMISSION: up to CG_JMAX chromosomes ("jobs") that
 * every kernel of the path handles in one launch (cg_internal.h).  cg_run_chromosomes hands a whole
 * sample over as one submission; the per-chromosome entry points are submissions of one job whose
 * phases (init | record sweep per pushed batch | pass + finalize | genotype) are launched by
 * cg_chrom_begin | cg_push_reads | cg_chrom_finish | cg_genotype.  A context owns CG_NSLOT slots
 * (cg_set_slots puts 2..CG_NSLOT of them in rotation); submission k lives in slot k mod n_slot with
 * its own arrays and streams:
 *   s_copy     host->device copies of GC bytes, records, mappability runs (HOST pointer mode)
 *   s_work     job table upload, init, record sweep, depth pass, finalize
 *   s_geno[2]  mappability index + genotype kernel ([1]: the duplications of a per-chromosome
 *              cg_genotype call, so that both kinds run side by side).  The genotype kernel is a chain
 *              of dependent loads whose latency grows several-fold under a depth pass that saturates
 *              HBM, so it runs on high-priority streams underneath the pass of the next submission.
 *   s_main     cg_stream(): what cg_join() orders behind everything else; the collective
 * No slot depends on another.  Events order the streams per slot; nothing ever waits on the host
 * unless the caller asked for blocking HOST-mode semantics, a slot is reused whose previous
 * submission has not finished yet, or a buffer has to grow.  With profiling on (cg_set_profiling)
 * the slots are chained one behind the other so that the event brackets time each kernel alone.
 */
#include "cg_internal.h"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <mutex>
#include <vector>

#define CG_MAX_CHROMS 4096            /* rows of the statistics table */
#ifndef CG_NSLOT
#define CG_NSLOT 6
#endif
#define CG_NBANK 4
#define CG_PASS_GRID_MAX (160 * 4)    /* rows of a slot's partial table: SMs x CTAs per SM the pass may ask for */

struct cg_evpair { cudaEvent_t a, b; int slot; };

/* everything one submission needs from its first launch to its last genotype kernel */
struct cg_slot {
    /* the submission */
    int n_jobs = 0;
    bool sample = false;                      /* cg_run_chromosomes submission: the job table lives in d_jobs */
    bool check = false;                       /* ... whose per-job flags the host has not looked at yet */
    int ptr_mode = CG_PTR_HOST; bool async = false;      /* modes the submission was made under (re-runs) */
    cg_dev_job hj[CG_JMAX];                   /* host mirror of the job table */
    cg_sub sub;                               /* what the kernels are launched with */
    std::vector<cg_chrom_job> src;            /* the caller's jobs (a flagged job is re-run from here) */
    cg_dev_job* d_jobs = nullptr;             /* [CG_JMAX] */
    cg_dev_job* h_jobs = nullptr;             /* page-locked staging of the table upload */
    unsigned int* h_flags = nullptr;          /* page-locked [CG_JMAX]: flag word of every job, written by k_finalize */
    cudaEvent_t ev_table = nullptr;           /* s_work: the table upload has read h_jobs */
    /* arrays over the slot's chunk space */
    int16_t* d_depth = nullptr;       size_t pos_cap = 0;
    int32_t* d_counts = nullptr;      size_t counts_cap = 0;      /* chunk bounds + three count levels (cg_counts) */
    uint32_t* d_depth32 = nullptr;    size_t depth32_cap = 0;     /* out-of-order path (per-chromosome entry points); zero at rest */
    int32_t* d_partial = nullptr;             /* [CG_PASS_GRID_MAX][CG_JMAX][CG_PART_WORDS] */
    float* d_expected = nullptr;              /* [CG_JMAX][128] */
    unsigned int* d_ctl = nullptr;            /* [CG_JMAX][CG_CTL_WORDS] + slot words */
    int32_t* d_run_idx = nullptr;     size_t run_idx_cap = 0;     /* first run per 16384 positions, all jobs */
    /* staging of the caller's host arrays */
    cg_read* d_reads = nullptr;       size_t reads_cap = 0;       /* per-chromosome entry points: records appended batch by batch */
    int32_t* d_pos = nullptr;         size_t rpos_cap = 0;
    uint8_t* d_mapq = nullptr;        size_t rmapq_cap = 0;
    int8_t* d_gc = nullptr;           size_t gc_cap = 0;
    cg_maprun* d_runs = nullptr;      size_t runs_cap = 0;
    char* d_in = nullptr;             size_t in_cap = 0;          /* sample submissions: one arena for records, GC bytes, runs */
    int32_t* d_dec = nullptr;         size_t dec_cap = 0;         /* DEVICE mode: position planes decoded from delta streams */
    char* d_arena = nullptr;          size_t arena_cap = 0, arena_used = 0;   /* CNV + call staging, blocking HOST mode */
    int rec_fmt = 0;                          /* per-chromosome pushes: 0 none yet, 1 cg_read, 2 planes, 3 planes without mapq */
    /* ordering */
    cudaStream_t s_copy = nullptr, s_work = nullptr, s_geno[2] = {nullptr, nullptr};
    cudaEvent_t ev_copy = nullptr;            /* s_copy: last upload for this slot */
    cudaEvent_t ev_pass = nullptr;            /* s_work: the pass and the finalize of this slot are done */
    cudaEvent_t ev_passk = nullptr;           /* s_work: the pass kernel alone is done (the next submission's pass is chained on it) */
    cudaEvent_t ev_geno[2] = {nullptr, nullptr};   /* s_geno[t]: last genotype work queued for this slot */
    cudaEvent_t ev_runs = nullptr;            /* s_geno[0]: mappability runs uploaded and indexed */
    bool used = false, geno_queued[2] = {false, false}, runs_uploaded = false;
    bool open = false;                        /* a chromosome was begun here and not finished: the control words are not at rest */
    long long* d_fin = nullptr;               /* [CG_JMAX][256] scratch rows + CG_JMAX counters of k_finalize (zero at rest) */
    bool dirty_work = false, dirty_geno[2] = {false, false};     /* work queued since the last cg_join */
};

/* NCCL, resolved at run time (cg_comm_*): only the five calls the one collective needs */
typedef struct cg_nccl_comm* cg_nccl_comm_t;
struct cg_nccl_id { char internal[CG_COMM_ID_BYTES]; };      /* ncclUniqueId: 128 bytes, passed by value */
struct cg_nccl_api {
    void* handle = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(cg_nccl_comm_t*, int, cg_nccl_id, int) = nullptr;
    int (*CommInitAll)(cg_nccl_comm_t*, int, const int*) = nullptr;
    int (*CommDestroy)(cg_nccl_comm_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, cg_nccl_comm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GroupStart)(void) = nullptr;
    int (*GroupEnd)(void) = nullptr;
};

struct cg_ctx {
    int device = 0;
    int n_sms = 148;
    cudaStream_t s_main = nullptr;
    cudaEvent_t ev_join = nullptr;
    int ptr_mode = CG_PTR_HOST;
    bool async = false;
    char err[512] = {0};

    cg_slot slot[CG_NSLOT];
    int n_slot = CG_NSLOT;                    /* slots in rotation (cg_set_slots); the others stay idle */
    int cur = 0;                              /* slot of the current chromosome (per-chromosome entry points) */
    long long n_begun = 0;

    /* current chromosome (per-chromosome entry points) */
    bool in_chrom = false, finished = false;
    int n_rows = 0;

    /* statistics tables: [bank][CG_MAX_CHROMS][204]; d_stats_all only with a communicator */
    long long* d_stats = nullptr;
    long long* d_stats_all = nullptr;
    int bank = 0;
    cudaEvent_t ev_bank_read[CG_NBANK] = {};                     /* s_main: the collective that reads this bank */
    bool bank_read[CG_NBANK] = {};
    cg_nccl_comm_t comm = nullptr;
    int comm_ranks = 0;
    /* peer exchange (cg_peer_*): every rank's reduced tables and arrival words, mapped into this context's address space */
    int n_peers = 0, peer_rank = 0;
    long long* peer_base[CG_MAX_PEERS] = {};
    bool peer_ipc[CG_MAX_PEERS] = {};         /* opened with cudaIpcOpenMemHandle: closed by cg_peer_disconnect */
    unsigned int epoch[CG_NBANK] = {};

    unsigned int* d_sticky = nullptr;         /* [0] CG_FLAG_RANGE seen since the last report */

    /* pinned staging */
    cg_chrom_stats* h_stats = nullptr;
    unsigned int* h_flags = nullptr;

    /* split reads (--rp): one chromosome at a time on its own stream */
    struct {
        cudaStream_t st = nullptr;
        unsigned char* d_ref = nullptr; size_t ref_cap = 0; int ref_len = 0;
        int* d_pos_list = nullptr; size_t idx_cap = 0;      /* positions grouped by k-mer value */
        int* d_cnt = nullptr;                               /* [2^20] positions per value, then slot counters */
        int *d_lo = nullptr, *d_hi = nullptr;
        cg_half* d_halves[2] = {nullptr, nullptr}; size_t halves_cap[2] = {0, 0};      /* two staging sets: batch k+1 is copied while batch k is searched */
        unsigned char* d_bases[2] = {nullptr, nullptr}; size_t bases_cap[2] = {0, 0};
        cudaEvent_t ev_buf[2] = {nullptr, nullptr}; unsigned n_push = 0;
        cg_splitrow* d_rows = nullptr; unsigned rows_cap = 0;
        unsigned* d_n_rows = nullptr; unsigned long long* d_n_map = nullptr;     /* one 16-byte block */
        char* d_count = nullptr; size_t count_cap = 0;
        bool have_ref = false;
        int chrom_len = 0, mq_threshold = -1;
    } sr;

    /* largest sizes any slot has needed, per buffer kind (see grow()) */
    size_t hw_gc = 0, hw_reads = 0, hw_rpos = 0, hw_rmapq = 0, hw_runs = 0, hw_run_idx = 0, hw_pos = 0, hw_in = 0, hw_dec = 0;

    /* cg_gc_profile staging (HOST pointer mode) */
    char* d_gc_bases = nullptr; size_t gc_bases_cap = 0;
    int8_t* d_gc_out = nullptr; size_t gc_out_cap = 0;

    /* profiling */
    bool profiling = false;
    std::vector<cg_evpair> ev_live, ev_free;
    double ms_acc[CG_NSLOTS] = {0, 0, 0, 0};
    long long launches_acc[CG_NSLOTS] = {0, 0, 0, 0};
    long long launch_count = 0;
};

/* what cg_last_error(NULL) returns: per thread, because contexts are driven from one thread each (one per GPU) */
static thread_local char g_err[512] = "no error";

int cg_fail(cg_ctx* c, int code, const char* fmt, ...)
{
    char* dst = c ? c->err : g_err;
    va_list ap; va_start(ap, fmt);
    vsnprintf(dst, 512, fmt, ap);
    va_end(ap);
    if (c) memcpy(g_err, c->err, 512);
    return code;
}
#define fail cg_fail

#define CU(c, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return fail((c), e_ == cudaErrorMemoryAllocation ? CG_ENOMEM : CG_ECUDA, "%s: %s", #call, cudaGetErrorString(e_)); } while (0)

static inline bool blocking(const cg_ctx* c) { return c->ptr_mode == CG_PTR_HOST && !c->async; }

static int sync_all(cg_ctx* c)
{
    CU(c, cudaStreamSynchronize(c->s_main));
    if (c->sr.st) CU(c, cudaStreamSynchronize(c->sr.st));
    for (int i = 0; i < CG_NSLOT; i++) {
        cg_slot& s = c->slot[i];
        CU(c, cudaStreamSynchronize(s.s_copy));
        CU(c, cudaStreamSynchronize(s.s_work));
        for (int t = 0; t < 2; t++) CU(c, cudaStreamSynchronize(s.s_geno[t]));
    }
    return CG_OK;
}

/* Replace a device buffer by a larger one.  Everything queued is drained first (the old
 * buffer may still be read by any stream); growth is rare and amortised. */
template <typename T>
static int grow(cg_ctx* c, T** p, size_t* cap, size_t need, size_t keep_elems = 0, size_t* high_water = nullptr)
{
    if (need <= *cap) return CG_OK;
    /* submissions rotate through the slots, so a slot that has to grow goes straight to the largest size any
     * slot has needed: every slot reallocates at most once per buffer after the largest submission was seen */
    size_t ncap = std::max(need, *cap + *cap / 2);
    if (high_water) { ncap = std::max(ncap, *high_water); *high_water = ncap; }
    int rc = sync_all(c);
    if (rc) return rc;
    T* np = nullptr;
    CU(c, cudaMalloc(&np, ncap * sizeof(T)));
    if (keep_elems && *p) CU(c, cudaMemcpy(np, *p, keep_elems * sizeof(T), cudaMemcpyDeviceToDevice));
    if (*p) cudaFree(*p);
    *p = np; *cap = ncap;
    return CG_OK;
}

/* profiling brackets */
struct prof_scope {
    cg_ctx* c; cudaStream_t s; cg_evpair ev; bool on;
    prof_scope(cg_ctx* c_, cudaStream_t s_, int slot, int n_launches) : c(c_), s(s_), on(c_->profiling) {
        c->launch_count += n_launches;
        if (!on) return;
        if (!c->ev_free.empty()) { ev = c->ev_free.back(); c->ev_free.pop_back(); }
        else { cudaEventCreate(&ev.a); cudaEventCreate(&ev.b); }
        ev.slot = slot;
        c->launches_acc[slot] += n_launches;
        cudaEventRecord(ev.a, s);
    }
    ~prof_scope() { if (on) { cudaEventRecord(ev.b, s); c->ev_live.push_back(ev); } }
};

static void prof_collect(cg_ctx* c)
{
    for (auto& ev : c->ev_live) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ev.a, ev.b) == cudaSuccess) c->ms_acc[ev.slot] += ms;
        c->ev_free.push_back(ev);
    }
    c->ev_live.clear();
}

/* CG_FLAG_RANGE raised on the device since the last report -> CG_ERANGE (caller has synchronised) */
static int report_sticky(cg_ctx* c, const char* what)
{
    CU(c, cudaMemcpyAsync(c->h_flags, c->d_sticky, 4, cudaMemcpyDeviceToHost, c->s_main));
    CU(c, cudaStreamSynchronize(c->s_main));
    if (c->h_flags[0] & CG_FLAG_RANGE) {
        CU(c, cudaMemsetAsync(c->d_sticky, 0, 4, c->s_main));
        CU(c, cudaStreamSynchronize(c->s_main));
        return fail(c, CG_ERANGE, "%s: a record or interval lies outside [0, chrom_len); it was clipped", what);
    }
    return CG_OK;
}

#define SR_ROWS_CAP (8u << 20)

template <typename T>
static int sr_grow(cg_ctx* c, T** p, size_t* cap, size_t need)
{
    if (need <= *cap) return CG_OK;
    CU(c, cudaStreamSynchronize(c->sr.st));
    if (*p) cudaFree(*p);
    *p = nullptr; *cap = 0;
    const size_t ncap = need + need / 4 + 1024;
    CU(c, cudaMalloc(p, ncap * sizeof(T)));
    *cap = ncap;
    return CG_OK;
}

static int settle_samples(cg_ctx* c);
static int settle_slot(cg_ctx* c, cg_slot& s);

extern "C" {

int cg_abi_version(void) { return CG_ABI_VERSION; }

const char* cg_last_error(cg_ctx* ctx) { return ctx ? ctx->err : g_err; }

cg_ctx* cg_create(int device)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) { fail(nullptr, CG_ECUDA, "no CUDA device: %s", cudaGetErrorString(e)); return nullptr; }
    if (device < 0 || device >= n) { fail(nullptr, CG_EINVAL, "device %d out of range (0..%d)", device, n - 1); return nullptr; }
    if (cudaSetDevice(device) != cudaSuccess) { fail(nullptr, CG_ECUDA, "cudaSetDevice(%d) failed", device); return nullptr; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { fail(nullptr, CG_ECUDA, "cudaGetDeviceProperties failed"); return nullptr; }
    if (prop.major != 10) {
        fail(nullptr, CG_ECUDA, "device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor);
        return nullptr;
    }
    if ((e = cgk_configure_depth()) != cudaSuccess || (e = cgk_configure_genotype()) != cudaSuccess
        || (e = cgk_configure_split()) != cudaSuccess || (e = cgk_configure_gc_profile()) != cudaSuccess) {
        fail(nullptr, CG_ECUDA, "kernel configuration failed (no sm_100a image for this device?): %s", cudaGetErrorString(e));
        return nullptr;
    }
    cg_ctx* c = new cg_ctx();
    c->device = device;
    c->n_sms = prop.multiProcessorCount;
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);     /* hi is the numerically smaller one */
    const size_t table_bytes = sizeof(long long) * CG_STATS_WORDS * CG_MAX_CHROMS * CG_NBANK;
    /* cg_stream carries the collective: high priority, so that its kernel is placed in the first gap the passes leave */
    bool ok = cudaStreamCreateWithPriority(&c->s_main, cudaStreamNonBlocking, prio_hi) == cudaSuccess
        && cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) == cudaSuccess
        && cudaMalloc(&c->d_stats, table_bytes) == cudaSuccess
        && cudaMemset(c->d_stats, 0, table_bytes) == cudaSuccess
        && cudaMalloc(&c->d_sticky, 16) == cudaSuccess
        && cudaMemset(c->d_sticky, 0, 16) == cudaSuccess
        && cudaMallocHost(&c->h_stats, sizeof(cg_chrom_stats)) == cudaSuccess
        && cudaMallocHost(&c->h_flags, 16) == cudaSuccess;
    for (int b = 0; ok && b < CG_NBANK; b++) ok = cudaEventCreateWithFlags(&c->ev_bank_read[b], cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < CG_NSLOT; i++) {
        cg_slot& s = c->slot[i];
        ok = cudaStreamCreateWithPriority(&s.s_copy, cudaStreamNonBlocking, prio_lo) == cudaSuccess
            && cudaStreamCreateWithPriority(&s.s_work, cudaStreamNonBlocking, prio_lo) == cudaSuccess
            /* the genotype kernels are small and latency-bound: let their CTAs be placed first */
            && cudaStreamCreateWithPriority(&s.s_geno[0], cudaStreamNonBlocking, prio_hi) == cudaSuccess
            && cudaStreamCreateWithPriority(&s.s_geno[1], cudaStreamNonBlocking, prio_hi) == cudaSuccess
            && cudaMalloc(&s.d_expected, sizeof(float) * 128 * CG_JMAX) == cudaSuccess
            && cudaMemset(s.d_expected, 0, sizeof(float) * 128 * CG_JMAX) == cudaSuccess
            && cudaMalloc(&s.d_ctl, sizeof(unsigned) * (CG_CTL_WORDS * CG_JMAX + 8)) == cudaSuccess
            && cudaMemset(s.d_ctl, 0, sizeof(unsigned) * (CG_CTL_WORDS * CG_JMAX + 8)) == cudaSuccess
            && cudaMalloc(&s.d_partial, sizeof(int32_t) * (size_t)CG_PASS_GRID_MAX * CG_JMAX * CG_PART_WORDS) == cudaSuccess
            && cudaMalloc(&s.d_jobs, sizeof(cg_dev_job) * CG_JMAX) == cudaSuccess
            && cudaMalloc(&s.d_fin, sizeof(long long) * (CG_JMAX * 256 + CG_JMAX)) == cudaSuccess
            && cudaMemset(s.d_fin, 0, sizeof(long long) * (CG_JMAX * 256 + CG_JMAX)) == cudaSuccess
            && cudaMallocHost(&s.h_jobs, sizeof(cg_dev_job) * CG_JMAX) == cudaSuccess
            && cudaMallocHost(&s.h_flags, sizeof(unsigned) * CG_JMAX) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_table, cudaEventDisableTiming) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_copy, cudaEventDisableTiming) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_pass, cudaEventDisableTiming) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_passk, cudaEventDisableTiming) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_geno[0], cudaEventDisableTiming) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_geno[1], cudaEventDisableTiming) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_runs, cudaEventDisableTiming) == cudaSuccess;
        if (ok) memset(s.h_flags, 0, sizeof(unsigned) * CG_JMAX);
    }
    if (!ok) {
        fail(nullptr, CG_ECUDA, "context allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
        cg_destroy(c);
        return nullptr;
    }
    if (c->n_sms * 4 > CG_PASS_GRID_MAX) { fail(nullptr, CG_ECUDA, "device has %d SMs; the library is sized for <= %d", c->n_sms, CG_PASS_GRID_MAX / 4); cg_destroy(c); return nullptr; }
    snprintf(c->err, sizeof c->err, "no error");
    return c;
}

void cg_destroy(cg_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    cg_peer_disconnect(c);
    cg_comm_destroy(c);
    for (auto& ev : c->ev_live) { cudaEventDestroy(ev.a); cudaEventDestroy(ev.b); }
    for (auto& ev : c->ev_free) { cudaEventDestroy(ev.a); cudaEventDestroy(ev.b); }
    for (int i = 0; i < CG_NSLOT; i++) {
        cg_slot& s = c->slot[i];
        cudaFree(s.d_reads); cudaFree(s.d_pos); cudaFree(s.d_mapq); cudaFree(s.d_gc); cudaFree(s.d_counts);
        cudaFree(s.d_depth); cudaFree(s.d_expected); cudaFree(s.d_ctl); cudaFree(s.d_partial); cudaFree(s.d_jobs); cudaFree(s.d_fin);
        cudaFree(s.d_runs); cudaFree(s.d_run_idx); cudaFree(s.d_arena); cudaFree(s.d_depth32); cudaFree(s.d_in); cudaFree(s.d_dec);
        if (s.h_jobs) cudaFreeHost(s.h_jobs);
        if (s.h_flags) cudaFreeHost(s.h_flags);
        cudaEvent_t evs[] = {s.ev_table, s.ev_copy, s.ev_pass, s.ev_passk, s.ev_geno[0], s.ev_geno[1], s.ev_runs};
        for (cudaEvent_t ev : evs) if (ev) cudaEventDestroy(ev);
        cudaStream_t sts[] = {s.s_copy, s.s_work, s.s_geno[0], s.s_geno[1]};
        for (cudaStream_t st : sts) if (st) cudaStreamDestroy(st);
    }
    cudaFree(c->d_stats); cudaFree(c->d_stats_all); cudaFree(c->d_sticky);
    if (c->sr.st) cudaStreamDestroy(c->sr.st);
    cudaFree(c->sr.d_ref); cudaFree(c->sr.d_pos_list); cudaFree(c->sr.d_cnt);
    cudaFree(c->sr.d_lo); cudaFree(c->sr.d_hi);
    for (int b = 0; b < 2; b++) { cudaFree(c->sr.d_halves[b]); cudaFree(c->sr.d_bases[b]); if (c->sr.ev_buf[b]) cudaEventDestroy(c->sr.ev_buf[b]); }
    cudaFree(c->sr.d_rows); cudaFree(c->sr.d_n_rows); cudaFree(c->sr.d_count);
    cudaFree(c->d_gc_bases); cudaFree(c->d_gc_out);
    if (c->h_stats) cudaFreeHost(c->h_stats);
    if (c->h_flags) cudaFreeHost(c->h_flags);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    for (int b = 0; b < CG_NBANK; b++) if (c->ev_bank_read[b]) cudaEventDestroy(c->ev_bank_read[b]);
    if (c->s_main) cudaStreamDestroy(c->s_main);
    delete c;
}

int cg_set_pointer_mode(cg_ctx* c, int mode)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (mode != CG_PTR_HOST && mode != CG_PTR_DEVICE) return fail(c, CG_EINVAL, "bad pointer mode %d", mode);
    c->ptr_mode = mode;
    return CG_OK;
}

int cg_set_async(cg_ctx* c, int on)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    c->async = on != 0;
    return CG_OK;
}

void* cg_alloc_pinned(cg_ctx* c, size_t bytes)
{
    if (!c) { fail(nullptr, CG_EINVAL, "null context"); return nullptr; }
    void* p = nullptr;
    cudaSetDevice(c->device);
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { fail(c, CG_ENOMEM, "cudaMallocHost(%zu) failed", bytes); return nullptr; }
    return p;
}

void cg_free_pinned(cg_ctx* c, void* p) { (void)c; if (p) cudaFreeHost(p); }

void* cg_stream(cg_ctx* c) { return c ? (void*)c->s_main : nullptr; }
int64_t cg_launch_count(cg_ctx* c) { return c ? c->launch_count : 0; }

int cg_set_slots(cg_ctx* c, int n)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (n < 2 || n > CG_NSLOT) return fail(c, CG_EINVAL, "%d slots (2..%d)", n, CG_NSLOT);
    if (c->in_chrom && !c->finished) return fail(c, CG_EINVAL, "cg_set_slots inside a chromosome");
    CU(c, cudaSetDevice(c->device));
    int rc = sync_all(c);                      /* slots that leave the rotation must be idle: cg_join no longer visits them */
    if (rc) return rc;
    if ((rc = settle_samples(c)) != CG_OK) return rc;
    c->n_slot = n;
    /* the chromosome that was current may sit in a slot that has just left the rotation: it is closed, the next
     * call that needs a chromosome has to follow a cg_chrom_begin */
    c->in_chrom = false; c->finished = false;
    return CG_OK;
}

/* make the main stream (cg_stream) wait for everything queued on the others since the last join */
int cg_join(cg_ctx* c)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    CU(c, cudaSetDevice(c->device));
    /* every upload on a copy stream is waited for by a kernel on its slot's work or genotype stream */
    for (int i = 0; i < CG_NSLOT; i++) {
        cg_slot& s = c->slot[i];
        if (s.dirty_work) {
            CU(c, cudaEventRecord(c->ev_join, s.s_work));
            CU(c, cudaStreamWaitEvent(c->s_main, c->ev_join, 0));
            s.dirty_work = false;
        }
        for (int t = 0; t < 2; t++) if (s.dirty_geno[t]) {
            CU(c, cudaEventRecord(c->ev_join, s.s_geno[t]));
            CU(c, cudaStreamWaitEvent(c->s_main, c->ev_join, 0));
            s.dirty_geno[t] = false;
        }
    }
    return CG_OK;
}

int cg_sync(cg_ctx* c)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    CU(c, cudaSetDevice(c->device));
    int rc = sync_all(c);
    if (rc) return rc;
    if ((rc = settle_samples(c)) != CG_OK) return rc;
    prof_collect(c);
    if (c->n_peers) {                          /* a peer's arrival word that never came (k_peer_exchange gave up) */
        CU(c, cudaMemcpyAsync(c->h_flags, c->d_sticky, 4, cudaMemcpyDeviceToHost, c->s_main));
        CU(c, cudaStreamSynchronize(c->s_main));
        if (c->h_flags[0] & CG_FLAG_PEER_TIMEOUT) return fail(c, CG_ECUDA, "cg_sync: the peer exchange timed out: a rank did not deliver its rows");
    }
    if (c->ptr_mode == CG_PTR_HOST) return report_sticky(c, "cg_sync");
    return CG_OK;
}

int cg_set_profiling(cg_ctx* c, int on)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    c->profiling = on != 0;
    return CG_OK;
}

int cg_kernel_times(cg_ctx* c, double ms[CG_NSLOTS], int64_t launches[CG_NSLOTS])
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    int rc = sync_all(c);
    if (rc) return rc;
    prof_collect(c);
    for (int i = 0; i < CG_NSLOTS; i++) {
        if (ms) ms[i] = c->ms_acc[i];
        if (launches) launches[i] = c->launches_acc[i];
        c->ms_acc[i] = 0; c->launches_acc[i] = 0;
    }
    return CG_OK;
}

} /* extern "C" */

/* ------------------------------------------------------------------------- */
/* slots                                                                     */

/* The next slot of the rotation, ordered behind its previous tenant: the streams that will write what the
 * tenant's kernels may still read wait for its events.  HOST mode writes the slot's upload buffers from the
 * copy stream; DEVICE mode writes none of them and needs only the work and genotype streams ordered. */
static int take_slot(cg_ctx* c, int* si_out)
{
    int si;
    for (;;) {
        si = (int)(c->n_begun % (unsigned long long)c->n_slot);
        if (!c->slot[si].check) break;
        int rc = settle_slot(c, c->slot[si]);          /* may re-run a chromosome, which takes slots itself */
        if (rc) return rc;
    }
    cg_slot& s = c->slot[si];
    const bool uploads = c->ptr_mode == CG_PTR_HOST;
    if (s.used && uploads) CU(c, cudaStreamWaitEvent(s.s_copy, s.ev_pass, 0));
    for (int t = 0; t < 2; t++) {
        if (!s.geno_queued[t]) continue;
        if (uploads) CU(c, cudaStreamWaitEvent(s.s_copy, s.ev_geno[t], 0));
        CU(c, cudaStreamWaitEvent(s.s_work, s.ev_geno[t], 0));
        CU(c, cudaStreamWaitEvent(s.s_geno[t ^ 1], s.ev_geno[t], 0));
    }
    if (c->profiling && !getenv("CG_PROF_NOCHAIN")) {   /* time kernels alone: chain this slot behind the one before it */
        cg_slot& o = c->slot[(si + c->n_slot - 1) % c->n_slot];
        if (o.used) CU(c, cudaStreamWaitEvent(s.s_work, o.ev_pass, 0));
        for (int t = 0; t < 2; t++) if (o.geno_queued[t]) CU(c, cudaStreamWaitEvent(s.s_work, o.ev_geno[t], 0));
    }
    if (s.open) {       /* begun, records pushed, never finished: no finalize has put the control words back to rest */
        CU(c, cudaMemsetAsync(s.d_ctl, 0, sizeof(unsigned) * (CG_CTL_WORDS * CG_JMAX + 8), s.s_work));
        s.open = false;
    }
    s.used = true; s.geno_queued[0] = s.geno_queued[1] = false; s.runs_uploaded = false;
    s.arena_used = 0; s.rec_fmt = 0; s.sample = false; s.check = false; s.n_jobs = 0;
    c->n_begun++;
    *si_out = si;
    return CG_OK;
}

/* depth, counts (and, for the per-chromosome entry points, the out-of-order scratch) for a chunk space */
static int ensure_space(cg_ctx* c, cg_slot& s, int n_space, bool want_depth32)
{
    const size_t n_pos = (size_t)n_space * CG_CHUNK;
    if (n_pos > c->hw_pos) c->hw_pos = n_pos;
    const size_t a_pos = c->hw_pos;                       /* the largest chunk space any slot has seen (see grow()) */
    const size_t a_words = cgk_counts_words((int)(a_pos / CG_CHUNK));
    if (cgk_counts_words(n_space) > s.counts_cap) {
        int rc = sync_all(c);
        if (rc) return rc;
        cudaFree(s.d_counts); s.d_counts = nullptr; s.counts_cap = 0;
        CU(c, cudaMalloc(&s.d_counts, a_words * sizeof(int32_t)));
        s.counts_cap = a_words;
    }
    if (n_pos > s.pos_cap) {
        int rc = sync_all(c);
        if (rc) return rc;
        cudaFree(s.d_depth); s.d_depth = nullptr; s.pos_cap = 0;
        CU(c, cudaMalloc(&s.d_depth, a_pos * sizeof(int16_t)));
        s.pos_cap = a_pos;
    }
    if (want_depth32 && n_pos > s.depth32_cap) {
        int rc = sync_all(c);
        if (rc) return rc;
        cudaFree(s.d_depth32); s.d_depth32 = nullptr; s.depth32_cap = 0;
        CU(c, cudaMalloc(&s.d_depth32, n_pos * sizeof(uint32_t)));
        /* the out-of-order path keeps these counts zero at rest (the pass clears what it consumes) */
        CU(c, cudaMemsetAsync(s.d_depth32, 0, n_pos * sizeof(uint32_t), s.s_work));
        s.depth32_cap = n_pos;
    }
    return CG_OK;
}

static void fill_sub(cg_ctx* c, cg_slot& s, int n_jobs, int n_work, int n_space, long long n_records)
{
    cg_sub& S = s.sub;
    memset(&S, 0, sizeof S);
    S.n_jobs = n_jobs; S.n_work = n_work; S.n_space = n_space;
    S.k = cgk_counts_layout(s.d_counts, n_space);
    S.depth = s.d_depth;
    S.partial = s.d_partial;
    S.n_pass_ctas = cgk_pass_grid(n_work, n_records, c->n_sms);
    S.sctl = s.d_ctl + CG_CTL_WORDS * CG_JMAX;
    S.fin_acc = s.d_fin;
    S.fin_cnt = reinterpret_cast<unsigned int*>(s.d_fin + CG_JMAX * 256);
    S.n_peer_rows = c->n_peers;
    for (int r = 0; r < c->n_peers; r++) S.peer_rows[r] = c->peer_base[r] + (size_t)c->bank * CG_MAX_CHROMS * CG_STATS_WORDS;
    S.sticky = c->d_sticky;
}

/* The pass of a submission starts when the pass of the submission before it has ended (CG_PASS_OVERLAP=1 lets
 * them overlap).  Two passes at once double the concurrent write streams and cost HBM efficiency, and the
 * genotype kernel and record sweep of the neighbours already fill the pass's ramp and tail; measured r02b on C3:
 * the pass alone 1.10 ms per sample, the pipelined step with free overlap 1.63 ms. */
static int chain_pass(cg_ctx* c, int si)
{
    static const bool overlap = getenv("CG_PASS_OVERLAP") && atoi(getenv("CG_PASS_OVERLAP")) != 0;
    cg_slot& s = c->slot[si];
    cg_slot& o = c->slot[(si + c->n_slot - 1) % c->n_slot];
    if (!overlap && o.used && &o != &s) CU(c, cudaStreamWaitEvent(s.s_work, o.ev_passk, 0));
    return CG_OK;
}

static long long* stats_row_of(cg_ctx* c, int chr_index)
{
    return c->d_stats + ((size_t)c->bank * CG_MAX_CHROMS + (size_t)chr_index) * CG_STATS_WORDS;
}

extern "C" {

/* ------------------------------------------------------------------------- */
int cg_chrom_begin(cg_ctx* c, int chr_index, int32_t chrom_len, const int8_t* gc_profile,
                   int32_t n_windows, int mq_threshold)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (chr_index < 0 || chr_index >= CG_MAX_CHROMS) return fail(c, CG_EINVAL, "chr_index %d outside 0..%d", chr_index, CG_MAX_CHROMS - 1);
    if (chrom_len <= 0) return fail(c, CG_EINVAL, "chrom_len %d must be positive", chrom_len);
    if (chrom_len > INT32_MAX - 2 * CG_CHUNK * CG_SUPER_CHUNKS) return fail(c, CG_EINVAL, "chrom_len %d too large", chrom_len);
    const int need_windows = (chrom_len + 99) / 100;
    if (!gc_profile || n_windows < need_windows)
        return fail(c, CG_EINVAL, "gc_profile needs >= %d windows for length %d (got %d)", need_windows, chrom_len, n_windows);
    CU(c, cudaSetDevice(c->device));

    int si, rc;
    if ((rc = take_slot(c, &si)) != CG_OK) return rc;
    cg_slot& s = c->slot[si];
    const int n_chunks = (chrom_len + CG_CHUNK - 1) / CG_CHUNK;
    const int n_space = cgk_job_space(n_chunks);
    if ((rc = ensure_space(c, s, n_space, true)) != CG_OK) return rc;
    cg_dev_job& J = s.hj[0];
    memset(&J, 0, sizeof J);
    if (c->ptr_mode == CG_PTR_HOST) {
        if ((rc = grow(c, &s.d_gc, &s.gc_cap, (size_t)need_windows, 0, &c->hw_gc)) != CG_OK) return rc;
        CU(c, cudaMemcpyAsync(s.d_gc, gc_profile, (size_t)need_windows, cudaMemcpyHostToDevice, s.s_copy));
        CU(c, cudaEventRecord(s.ev_copy, s.s_copy));
        J.gc = s.d_gc;
    } else {
        J.gc = gc_profile;
    }
    J.chrom_len = chrom_len; J.n_chunks = n_chunks; J.mq_threshold = mq_threshold; J.chr_index = chr_index;
    J.stats_row = stats_row_of(c, chr_index);
    J.expected = s.d_expected; J.ctl = s.d_ctl; J.run_idx = nullptr;
    s.n_jobs = 1;
    fill_sub(c, s, 1, n_chunks, n_space, 0);
    s.sub.depth32 = s.d_depth32;
    c->cur = si;
    c->in_chrom = true; c->finished = false;
    if (chr_index + 1 > c->n_rows) c->n_rows = chr_index + 1;
    s.open = true;                            /* until cg_chrom_finish: its finalize leaves the control words zero again */
    s.dirty_work = true;
    return CG_OK;
}

/* one batch of records of the current chromosome, either format: staged (HOST), used in place (first DEVICE
 * batch) or gathered into the slot's own buffers (later DEVICE batches), then looked at by k_mark */
static int push_records(cg_ctx* c, const cg_read* reads, const int32_t* pos, const uint8_t* mapq, size_t n, const char* who)
{
    if (!c->in_chrom || c->finished) return fail(c, CG_EINVAL, "%s outside cg_chrom_begin/cg_chrom_finish", who);
    if (n == 0) return CG_OK;
    cg_slot& s = c->slot[c->cur];
    cg_dev_job& J = s.hj[0];
    const int fmt = reads ? 1 : (mapq ? 2 : 3);
    if (fmt == 3 && J.mq_threshold >= 0) return fail(c, CG_EINVAL, "%s: a NULL mapq plane needs a negative mq_threshold (got %d)", who, J.mq_threshold);
    if (s.rec_fmt && s.rec_fmt != fmt) return fail(c, CG_EINVAL, "%s: a chromosome is pushed in one record format only", who);
    if ((unsigned long long)J.n_reads + n >= 0x7fffffffull) return fail(c, CG_EINVAL, "more than 2^31-1 records on one chromosome");
    CU(c, cudaSetDevice(c->device));
    const size_t first = (size_t)J.n_reads;
    const cudaMemcpyKind kind = c->ptr_mode == CG_PTR_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
    const bool in_place = c->ptr_mode == CG_PTR_DEVICE && first == 0;
    if (in_place) {
        J.reads = reads; J.rec_pos = pos; J.rec_mapq = mapq;       /* single device batch: used where it lies */
    } else {
        /* the copies run on the copy stream in HOST mode (they overlap the previous slot's kernels) and on the
         * work stream in DEVICE mode (device-to-device gathers of a caller that pushes in several batches) */
        cudaStream_t st = c->ptr_mode == CG_PTR_HOST ? s.s_copy : s.s_work;
        int rc;
        if (fmt == 1) {
            const bool external = first && J.reads != s.d_reads;
            if ((rc = grow(c, &s.d_reads, &s.reads_cap, first + n, external ? 0 : first, &c->hw_reads)) != CG_OK) return rc;
            if (external) CU(c, cudaMemcpyAsync(s.d_reads, J.reads, first * sizeof(cg_read), cudaMemcpyDeviceToDevice, st));
            CU(c, cudaMemcpyAsync(s.d_reads + first, reads, n * sizeof(cg_read), kind, st));
            J.reads = s.d_reads;
        } else {
            const bool external = first && J.rec_pos != s.d_pos;
            if ((rc = grow(c, &s.d_pos, &s.rpos_cap, first + n, external ? 0 : first, &c->hw_rpos)) != CG_OK) return rc;
            if (external) CU(c, cudaMemcpyAsync(s.d_pos, J.rec_pos, first * 4, cudaMemcpyDeviceToDevice, st));
            CU(c, cudaMemcpyAsync(s.d_pos + first, pos, n * 4, kind, st));
            if (fmt == 2) {
                if ((rc = grow(c, &s.d_mapq, &s.rmapq_cap, first + n, external ? 0 : first, &c->hw_rmapq)) != CG_OK) return rc;
                if (external) CU(c, cudaMemcpyAsync(s.d_mapq, J.rec_mapq, first, cudaMemcpyDeviceToDevice, st));
                CU(c, cudaMemcpyAsync(s.d_mapq + first, mapq, n, kind, st));
                J.rec_mapq = s.d_mapq;
            }
            J.rec_pos = s.d_pos;
        }
        if (c->ptr_mode == CG_PTR_HOST) {
            CU(c, cudaEventRecord(s.ev_copy, s.s_copy));
            CU(c, cudaStreamWaitEvent(s.s_work, s.ev_copy, 0));
        }
    }
    s.rec_fmt = fmt;
    J.n_reads = (long long)(first + n);
    J.mark_first = (long long)first; J.mark_n = (long long)n; J.mark_blk0 = 0;
    s.sub.n_mark_blocks = cgk_mark_blocks(n);
    {
        prof_scope ps(c, s.s_work, 0, 1);
        CU(c, cgk_mark(s.s_work, s.sub, J));
    }
    return CG_OK;
}

int cg_push_reads(cg_ctx* c, const cg_read* reads, size_t n)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (n && !reads) return fail(c, CG_EINVAL, "null reads");
    return push_records(c, reads, nullptr, nullptr, n, "cg_push_reads");
}

int cg_push_reads_compact(cg_ctx* c, const int32_t* pos, const uint8_t* mapq, size_t n)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (n && !pos) return fail(c, CG_EINVAL, "null pos");
    return push_records(c, nullptr, pos, mapq, n, "cg_push_reads_compact");
}

int cg_chrom_finish(cg_ctx* c, cg_chrom_stats* out)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom || c->finished) return fail(c, CG_EINVAL, "cg_chrom_finish without cg_chrom_begin");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    cg_dev_job& J = s.hj[0];
    if (c->ptr_mode == CG_PTR_HOST) CU(c, cudaStreamWaitEvent(s.s_work, s.ev_copy, 0));   /* GC bytes */
    {
        prof_scope ps(c, s.s_work, 0, 1);
        CU(c, cgk_rare_scatter(s.s_work, s.sub, J, c->n_sms));
    }
    s.sub.n_pass_ctas = cgk_pass_grid(s.sub.n_work, J.n_reads, c->n_sms);
    { int rc = chain_pass(c, c->cur); if (rc) return rc; }
    {
        prof_scope ps(c, s.s_work, 1, 1);
        CU(c, cgk_depth_pass(s.s_work, s.sub, J));
    }
    CU(c, cudaEventRecord(s.ev_passk, s.s_work));
    /* a collective that still reads this bank must be through before the row is rewritten */
    if (c->bank_read[c->bank]) CU(c, cudaStreamWaitEvent(s.s_work, c->ev_bank_read[c->bank], 0));
    /* asynchronous HOST mode hands over a page-locked pointer, DEVICE mode a device pointer: the kernel writes the row there itself */
    J.row_out = (out && !blocking(c)) ? reinterpret_cast<long long*>(out) : nullptr;
    {
        prof_scope ps(c, s.s_work, 0, 1);
        CU(c, cgk_finalize(s.s_work, s.sub, J));
    }
    CU(c, cudaEventRecord(s.ev_pass, s.s_work));
    s.dirty_work = true;
    s.open = false;
    c->finished = true;
    if (!blocking(c)) return CG_OK;
    CU(c, cudaMemcpyAsync(c->h_stats, J.stats_row, sizeof(cg_chrom_stats), cudaMemcpyDeviceToHost, s.s_work));
    CU(c, cudaStreamSynchronize(s.s_work));
    if (out) memcpy(out, c->h_stats, sizeof(cg_chrom_stats));
    return report_sticky(c, "cg_chrom_finish");
}

int cg_get_expected(cg_ctx* c, float expected[CG_GC_BINS])
{
    if (!c || !expected) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_get_expected before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    CU(c, cudaMemcpyAsync(expected, s.hj[0].expected, sizeof(float) * CG_GC_BINS,
                          c->ptr_mode == CG_PTR_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, s.s_work));
    if (blocking(c)) CU(c, cudaStreamSynchronize(s.s_work));
    return CG_OK;
}

int cg_set_expected(cg_ctx* c, const float expected[CG_GC_BINS])
{
    if (!c || !expected) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_set_expected before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    CU(c, cudaMemcpyAsync(s.hj[0].expected, expected, sizeof(float) * CG_GC_BINS,
                          c->ptr_mode == CG_PTR_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, s.s_work));
    CU(c, cudaEventRecord(s.ev_pass, s.s_work));        /* the genotype stream waits for this */
    return CG_OK;
}

/* ------------------------------------------------------------------------- */
int cg_set_mappability_runs(cg_ctx* c, const cg_maprun* runs, size_t n, int enabled)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom) return fail(c, CG_EINVAL, "cg_set_mappability outside a chromosome");
    if (n > 0x7ffffff0ull) return fail(c, CG_EINVAL, "too many mappability runs");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    cg_dev_job& J = s.hj[0];
    J.has_map = enabled ? 1 : 0;
    J.n_runs = (int)n;
    if (n == 0) { J.runs = nullptr; return CG_OK; }
    if (!runs) return fail(c, CG_EINVAL, "null runs");
    int rc = grow(c, &s.d_run_idx, &s.run_idx_cap, (size_t)cgk_run_index_words(J.chrom_len), 0, &c->hw_run_idx);
    if (rc) return rc;
    J.run_idx = s.d_run_idx;
    if (c->ptr_mode == CG_PTR_DEVICE) {
        J.runs = runs;
    } else {
        rc = grow(c, &s.d_runs, &s.runs_cap, n, 0, &c->hw_runs);
        if (rc) return rc;
        /* on a genotype stream: behind the slot's previous tenant, which may still read d_runs */
        CU(c, cudaMemcpyAsync(s.d_runs, runs, n * sizeof(cg_maprun), cudaMemcpyHostToDevice, s.s_geno[0]));
        J.runs = s.d_runs;
    }
    J.run_blk0 = 0;
    cg_sub S = s.sub;
    S.n_run_blocks = (int)((n + 1 + 255) / 256);
    {
        prof_scope ps(c, s.s_geno[0], 2, 1);
        CU(c, cgk_run_index(s.s_geno[0], S, J));
    }
    CU(c, cudaEventRecord(s.ev_runs, s.s_geno[0]));
    s.runs_uploaded = true;
    s.dirty_geno[0] = true;
    return CG_OK;
}

int cg_set_mappability(cg_ctx* c, const cg_mapiv* rows, size_t n)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom) return fail(c, CG_EINVAL, "cg_set_mappability outside a chromosome");
    if (n && !rows) return fail(c, CG_EINVAL, "null rows");
    std::vector<cg_maprun> v;
    cg_flatten_rows(rows, n, c->slot[c->cur].hj[0].chrom_len, v);
    const int saved = c->ptr_mode;
    c->ptr_mode = CG_PTR_HOST;                                 /* rows are always host memory */
    int rc = cg_set_mappability_runs(c, v.data(), v.size(), n > 0);
    if (rc == CG_OK && !v.empty()) {
        cudaError_t e = cudaStreamSynchronize(c->slot[c->cur].s_geno[0]);   /* v dies at return */
        if (e != cudaSuccess) rc = fail(c, CG_ECUDA, "cudaStreamSynchronize: %s", cudaGetErrorString(e));
    }
    c->ptr_mode = saved;
    return rc;
}

} /* extern "C" */

/* HOST-mode staging for one cg_genotype call: room for n CNVs and n calls in the slot's arena.
 * The arena is rewound when the slot is taken; when it has to grow, everything queued is drained
 * first, so earlier calls of this chromosome have already delivered their results. */
static int arena_take(cg_ctx* c, cg_slot& s, size_t n, cg_cnv** d_in, cg_call** d_out)
{
    const size_t in_b = (n * sizeof(cg_cnv) + 255) & ~(size_t)255, out_b = (n * sizeof(cg_call) + 255) & ~(size_t)255;
    if (s.arena_used + in_b + out_b > s.arena_cap) {
        int rc = sync_all(c);
        if (rc) return rc;
        cudaFree(s.d_arena); s.d_arena = nullptr; s.arena_cap = 0; s.arena_used = 0;
        const size_t cap = std::max<size_t>(2 * (in_b + out_b), (size_t)1 << 20);
        CU(c, cudaMalloc(&s.d_arena, cap));
        s.arena_cap = cap;
    }
    *d_in = reinterpret_cast<cg_cnv*>(s.d_arena + s.arena_used);
    *d_out = reinterpret_cast<cg_call*>(s.d_arena + s.arena_used + in_b);
    s.arena_used += in_b + out_b;
    return CG_OK;
}

extern "C" {

int cg_genotype(cg_ctx* c, const cg_cnv* cnvs, size_t n, int is_dup, cg_call* out)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_genotype before cg_chrom_finish");
    if (n == 0) return CG_OK;
    if (!cnvs || !out) return fail(c, CG_EINVAL, "null argument");
    if (n > 0x3fffffffull) return fail(c, CG_EINVAL, "too many CNVs in one call");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    const int t = is_dup ? 1 : 0;
    cudaStream_t st = s.s_geno[t];            /* deletions and duplications of one chromosome run side by side */
    const cg_cnv* d_in = cnvs; cg_call* d_out = out;
    /* Asynchronous HOST mode: both arrays are page-locked (cg_set_async), i.e. addressable from the device; the
     * kernel reads 8 bytes and writes 56 per CNV across PCIe itself, which is cheaper than two copy calls per launch. */
    const bool staged = blocking(c);
    if (staged) {
        cg_cnv* a_in = nullptr;
        int rc = arena_take(c, s, n, &a_in, &d_out);
        if (rc) return rc;
        CU(c, cudaMemcpyAsync(a_in, cnvs, n * sizeof(cg_cnv), cudaMemcpyHostToDevice, st));
        d_in = a_in;
    }
    CU(c, cudaStreamWaitEvent(st, s.ev_pass, 0));              /* depth, chunk sums and expected table of this slot */
    if (t == 1 && s.runs_uploaded) CU(c, cudaStreamWaitEvent(st, s.ev_runs, 0));
    cg_dev_job J = s.hj[0];                    /* this launch's view of the job: one kind of CNV */
    J.cnv[t] = d_in; J.calls[t] = d_out; J.n_cnv[t] = (int)n; J.n_cnv[t ^ 1] = 0;
    cg_sub S = s.sub;
    S.geno_g = cgk_geno_g(n, c->n_sms);
    S.n_cnv = cgk_geno_groups(n, S.geno_g);
    J.cnv0[0] = 0; J.cnv0[1] = t ? 0 : S.n_cnv;
    {
        prof_scope ps(c, st, 2, 1);
        CU(c, cgk_genotype(st, S, J, c->n_sms));
    }
    if (staged) CU(c, cudaMemcpyAsync(out, d_out, n * sizeof(cg_call), cudaMemcpyDeviceToHost, st));
    CU(c, cudaEventRecord(s.ev_geno[t], st));
    s.geno_queued[t] = true;
    s.dirty_geno[t] = true;
    if (blocking(c)) {
        CU(c, cudaStreamSynchronize(st));
        return report_sticky(c, "cg_genotype");
    }
    return CG_OK;
}

} /* extern "C" */

/* ------------------------------------------------------------------------- */
/* a whole job list as one submission                                         */

static int validate_job(cg_ctx* c, const cg_chrom_job& j, size_t k)
{
    if (j.chr_index < 0 || j.chr_index >= CG_MAX_CHROMS) return fail(c, CG_EINVAL, "job %zu: chr_index %d outside 0..%d", k, j.chr_index, CG_MAX_CHROMS - 1);
    if (j.chrom_len <= 0 || j.chrom_len > INT32_MAX - 2 * CG_CHUNK * CG_SUPER_CHUNKS) return fail(c, CG_EINVAL, "job %zu: chrom_len %d", k, j.chrom_len);
    if (!j.gc || j.n_windows < (j.chrom_len + 99) / 100) return fail(c, CG_EINVAL, "job %zu: gc_profile needs >= %d windows (got %d)", k, (j.chrom_len + 99) / 100, j.n_windows);
    if (j.n_reads >= 0x7fffffffull) return fail(c, CG_EINVAL, "job %zu: more than 2^31-1 records on one chromosome", k);
    if (j.n_reads && !j.reads && !j.reads_pos && !j.reads_delta) return fail(c, CG_EINVAL, "job %zu: null records", k);
    if (j.n_reads && !j.reads && !j.reads_pos && (!j.reads_anchor || (j.n_big && !j.reads_big) || j.n_big > j.n_reads))
        return fail(c, CG_EINVAL, "job %zu: a delta-coded stream needs its anchors and its side array", k);
    if (j.n_reads && !j.reads && !j.reads_mapq && j.mq_threshold >= 0) return fail(c, CG_EINVAL, "job %zu: a NULL mapq plane needs a negative mq_threshold", k);
    if (j.n_runs > 0x7ffffff0ull || (j.n_runs && j.use_mappability && !j.runs)) return fail(c, CG_EINVAL, "job %zu: bad mappability runs", k);
    if (j.n_del > 0x3fffffffull || j.n_dup > 0x3fffffffull) return fail(c, CG_EINVAL, "job %zu: too many CNVs", k);
    if ((j.n_del && (!j.dels || !j.del_calls)) || (j.n_dup && (!j.dups || !j.dup_calls))) return fail(c, CG_EINVAL, "job %zu: null CNV argument", k);
    return CG_OK;
}

static inline size_t al256(size_t b) { return (b + 255) & ~(size_t)255; }

/* n <= CG_JMAX jobs, DEVICE or asynchronous HOST pointer mode: one launch per kernel for all of them */
static int submit_sample(cg_ctx* c, const cg_chrom_job* jobs, int n)
{
    int rc;
    /* geometry of the submission */
    long long space = 0, work = 0, mark_blocks = 0, run_blocks = 0, n_cnv = 0;
    unsigned long long all_cnvs = 0;
    unsigned long long all_records = 0;
    for (int k = 0; k < n; k++) { all_cnvs += jobs[k].n_del + jobs[k].n_dup; all_records += jobs[k].n_reads; }
    const int G = cgk_geno_g(all_cnvs, c->n_sms);
    size_t in_bytes = 0, ridx_words = 0, dec_words = 0;
    const bool host = c->ptr_mode == CG_PTR_HOST;
    for (int k = 0; k < n; k++) {
        const cg_chrom_job& j = jobs[k];
        if ((rc = validate_job(c, j, (size_t)k)) != CG_OK) return rc;
        const int n_chunks = (j.chrom_len + CG_CHUNK - 1) / CG_CHUNK;
        space += cgk_job_space(n_chunks); work += n_chunks;
        mark_blocks += (long long)cgk_mark_blocks(j.n_reads);
        if (j.use_mappability && j.n_runs) { run_blocks += (long long)((j.n_runs + 1 + 255) / 256); ridx_words += (size_t)cgk_run_index_words(j.chrom_len); }
        n_cnv += (long long)cgk_geno_groups(j.n_del, G) + cgk_geno_groups(j.n_dup, G);
        if (!host && !j.reads && !j.reads_pos) dec_words += (size_t)((j.n_reads + 63) & ~(uint64_t)63);   /* DEVICE mode: the plane a delta stream decodes into */
        if (host) {
            in_bytes += j.reads ? al256(j.n_reads * sizeof(cg_read)) : al256(j.n_reads * 4) + (j.reads_mapq ? al256(j.n_reads) : 0);
            if (!j.reads && !j.reads_pos)          /* delta stream: the bytes, the side array, the anchors; the plane above is what they decode into */
                in_bytes += al256(j.n_reads) + al256(j.n_big * 4) + al256((size_t)cgk_mark_blocks(j.n_reads) * sizeof(cg_delta_anchor));
            if (!(j.flags & CG_JOB_ANNOTATION_RESIDENT)) {
                in_bytes += al256((size_t)(j.chrom_len + 99) / 100);
                if (j.use_mappability) in_bytes += al256(j.n_runs * sizeof(cg_maprun));
            }
        }
    }
    if (space * CG_CHUNK >= ((long long)1 << 40) || space > 0x7fffffffll || mark_blocks > 0x7fffffffll || run_blocks > 0x7fffffffll || n_cnv > 0x7fffffffll)
        return fail(c, CG_EINVAL, "submission too large for one launch");

    int si;
    if ((rc = take_slot(c, &si)) != CG_OK) return rc;
    cg_slot& s = c->slot[si];
    if ((rc = ensure_space(c, s, (int)space, false)) != CG_OK) return rc;
    if (ridx_words && (rc = grow(c, &s.d_run_idx, &s.run_idx_cap, ridx_words, 0, &c->hw_run_idx)) != CG_OK) return rc;
    if (host && (rc = grow(c, &s.d_in, &s.in_cap, in_bytes, 0, &c->hw_in)) != CG_OK) return rc;
    if (!host && dec_words && (rc = grow(c, &s.d_dec, &s.dec_cap, dec_words, 0, &c->hw_dec)) != CG_OK) return rc;
    /* the staging copy of the table is rewritten below: its previous upload must have read it */
    CU(c, cudaEventSynchronize(s.ev_table));

    s.sample = true; s.n_jobs = n; s.ptr_mode = c->ptr_mode; s.async = c->async;
    s.src.assign(jobs, jobs + n);
    fill_sub(c, s, n, (int)work, (int)space, (long long)all_records);
    s.sub.jobs = s.d_jobs;
    s.sub.n_mark_blocks = (int)mark_blocks; s.sub.n_run_blocks = (int)run_blocks; s.sub.n_cnv = (int)n_cnv;
    s.sub.h_flags = s.h_flags;
    s.sub.geno_g = G;

    int chunk0 = 0, work0 = 0, blk0 = 0, rblk0 = 0, cnv0 = 0;
    size_t in_off = 0, ridx_off = 0;
    bool any_runs = false, any_delta = false;
    size_t dec_off = 0;
    auto stage = [&](const void* src, size_t bytes) -> void* {           /* upload into the slot's arena (HOST mode) */
        char* dst = s.d_in + in_off;
        in_off += al256(bytes);
        if (bytes && src) cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s.s_copy);     /* no source: room only */
        return dst;
    };
    for (int k = 0; k < n; k++) {
        const cg_chrom_job& j = jobs[k];
        cg_dev_job& J = s.hj[k];
        memset(&J, 0, sizeof J);
        const int n_chunks = (j.chrom_len + CG_CHUNK - 1) / CG_CHUNK;
        J.chrom_len = j.chrom_len; J.n_chunks = n_chunks; J.chunk0 = chunk0; J.work0 = work0;
        J.mq_threshold = j.mq_threshold; J.chr_index = j.chr_index;
        J.n_reads = (long long)j.n_reads; J.mark_first = 0; J.mark_n = (long long)j.n_reads; J.mark_blk0 = blk0;
        const bool annot_dev = !host || (j.flags & CG_JOB_ANNOTATION_RESIDENT);
        const bool delta = !j.reads && !j.reads_pos && j.n_reads;
        if (host) {
            J.gc = annot_dev ? j.gc : static_cast<const int8_t*>(stage(j.gc, (size_t)(j.chrom_len + 99) / 100));
            if (j.reads) J.reads = static_cast<const cg_read*>(stage(j.reads, j.n_reads * sizeof(cg_read)));
            else if (!delta) J.rec_pos = static_cast<const int32_t*>(stage(j.reads_pos, j.n_reads * 4));
            else {
                /* the delta bytes, the side array and the anchors are uploaded; the position plane is only reserved:
                 * k_decode_delta fills it on the device */
                J.rec_delta = static_cast<const uint8_t*>(stage(j.reads_delta, j.n_reads));
                J.rec_big = static_cast<const int32_t*>(stage(j.reads_big, j.n_big * 4));
                J.rec_anchor = static_cast<const cg_delta_anchor*>(stage(j.reads_anchor, (size_t)cgk_mark_blocks(j.n_reads) * sizeof(cg_delta_anchor)));
                J.rec_pos_out = static_cast<int32_t*>(stage(nullptr, j.n_reads * 4));
                J.rec_pos = J.rec_pos_out;
                any_delta = true;
            }
            if (!j.reads && j.reads_mapq) J.rec_mapq = static_cast<const uint8_t*>(stage(j.reads_mapq, j.n_reads));
        } else {
            J.gc = j.gc;
            if (j.reads) J.reads = j.reads;
            else if (!delta) { J.rec_pos = j.reads_pos; J.rec_mapq = j.reads_mapq; }
            else {
                J.rec_delta = j.reads_delta; J.rec_big = j.reads_big; J.rec_anchor = j.reads_anchor; J.rec_mapq = j.reads_mapq;
                J.rec_pos_out = s.d_dec + dec_off; dec_off += (size_t)((j.n_reads + 63) & ~(uint64_t)63);
                J.rec_pos = J.rec_pos_out;
                any_delta = true;
            }
        }
        J.stats_row = stats_row_of(c, j.chr_index);
        J.row_out = reinterpret_cast<long long*>(j.stats);          /* page-locked (asynchronous HOST) or device memory */
        J.expected = s.d_expected + 128 * k;
        J.ctl = s.d_ctl + CG_CTL_WORDS * k;
        J.has_map = j.use_mappability ? 1 : 0;
        if (j.use_mappability && j.n_runs) {
            J.runs = annot_dev ? j.runs : static_cast<const cg_maprun*>(stage(j.runs, j.n_runs * sizeof(cg_maprun)));
            J.n_runs = (int)j.n_runs;
            J.run_idx = s.d_run_idx + ridx_off; ridx_off += (size_t)cgk_run_index_words(j.chrom_len);
            J.run_blk0 = rblk0; rblk0 += (int)((j.n_runs + 1 + 255) / 256);
            any_runs = true;
        } else {
            J.run_blk0 = rblk0;
        }
        /* CNVs and calls: device memory, or the caller's page-locked arrays read and written across PCIe by the kernel */
        J.cnv[0] = j.dels; J.calls[0] = j.del_calls; J.n_cnv[0] = (int)j.n_del; J.cnv0[0] = cnv0; cnv0 += cgk_geno_groups(j.n_del, G);
        J.cnv[1] = j.dups; J.calls[1] = j.dup_calls; J.n_cnv[1] = (int)j.n_dup; J.cnv0[1] = cnv0; cnv0 += cgk_geno_groups(j.n_dup, G);
        chunk0 += cgk_job_space(n_chunks); work0 += n_chunks; blk0 += cgk_mark_blocks(j.n_reads);
        if (j.chr_index + 1 > c->n_rows) c->n_rows = j.chr_index + 1;
    }
    if (host) {
        CU(c, cudaGetLastError());
        CU(c, cudaEventRecord(s.ev_copy, s.s_copy));
    }
    memcpy(s.h_jobs, s.hj, sizeof(cg_dev_job) * (size_t)n);
    CU(c, cudaMemcpyAsync(s.d_jobs, s.h_jobs, sizeof(cg_dev_job) * (size_t)n, cudaMemcpyHostToDevice, s.s_work));
    CU(c, cudaEventRecord(s.ev_table, s.s_work));

    const cg_dev_job& J0 = s.hj[0];            /* ignored by the kernels: the table is used */
    if (host) CU(c, cudaStreamWaitEvent(s.s_work, s.ev_copy, 0));
    if (mark_blocks && any_delta) {
        prof_scope ps(c, s.s_work, 0, 1);
        CU(c, cgk_decode_delta(s.s_work, s.sub, J0));
    }
    if (mark_blocks) {
        prof_scope ps(c, s.s_work, 0, 1);
        CU(c, cgk_mark(s.s_work, s.sub, J0));
    }
    if ((rc = chain_pass(c, si)) != CG_OK) return rc;
    {
        prof_scope ps(c, s.s_work, 1, 1);
        CU(c, cgk_depth_pass(s.s_work, s.sub, J0));
    }
    CU(c, cudaEventRecord(s.ev_passk, s.s_work));
    s.dirty_work = true;
    s.check = true;

    /* finalize, mappability index and genotype kernel on the slot's high-priority stream: when the pass ends, the pass
     * of the next submission (chained on ev_passk, low priority) and this finalize become runnable together, and the
     * small kernel should be placed first */
    cudaStream_t st = s.s_geno[0];
    CU(c, cudaStreamWaitEvent(st, s.ev_table, 0));
    if (n_cnv && any_runs) {
        if (host) CU(c, cudaStreamWaitEvent(st, s.ev_copy, 0));
        prof_scope ps(c, st, 2, 1);
        CU(c, cgk_run_index(st, s.sub, J0));
    }
    CU(c, cudaStreamWaitEvent(st, s.ev_passk, 0));
    /* a collective that still reads this bank must be through before the rows are rewritten */
    if (c->bank_read[c->bank]) CU(c, cudaStreamWaitEvent(st, c->ev_bank_read[c->bank], 0));
    {
        prof_scope ps(c, st, 0, 1);
        CU(c, cgk_finalize(st, s.sub, J0));
    }
    CU(c, cudaEventRecord(s.ev_pass, st));
    if (n_cnv) {
        prof_scope ps(c, st, 2, 1);
        CU(c, cgk_genotype(st, s.sub, J0, c->n_sms));
    }
    CU(c, cudaEventRecord(s.ev_geno[0], st));
    s.geno_queued[0] = true;
    s.dirty_geno[0] = true;
    c->in_chrom = false; c->finished = false;       /* no current chromosome: the per-chromosome calls need a cg_chrom_begin */
    return CG_OK;
}

/* one job through the per-chromosome entry points (blocking HOST mode, and re-runs of flagged jobs) */
static int run_one_job(cg_ctx* c, const cg_chrom_job& j)
{
    int rc;
    if (j.n_reads && !j.reads && !j.reads_pos)             /* before anything is begun */
        return fail(c, CG_EINVAL, "a delta-coded record stream is taken by submissions only (DEVICE or asynchronous HOST mode)");
    if ((rc = cg_chrom_begin(c, j.chr_index, j.chrom_len, j.gc, j.n_windows, j.mq_threshold)) != CG_OK) return rc;
    if (j.n_reads) {
        rc = j.reads ? cg_push_reads(c, j.reads, (size_t)j.n_reads) : cg_push_reads_compact(c, j.reads_pos, j.reads_mapq, (size_t)j.n_reads);
        if (rc != CG_OK) return rc;
    }
    if ((rc = cg_chrom_finish(c, j.stats)) != CG_OK) return rc;
    if (j.use_mappability && (rc = cg_set_mappability_runs(c, j.runs, (size_t)j.n_runs, 1)) != CG_OK) return rc;
    if (j.n_del && (rc = cg_genotype(c, j.dels, (size_t)j.n_del, 0, j.del_calls)) != CG_OK) return rc;
    if (j.n_dup && (rc = cg_genotype(c, j.dups, (size_t)j.n_dup, 1, j.dup_calls)) != CG_OK) return rc;
    return CG_OK;
}

/* A sample submission runs the sorted-input path unconditionally.  When it is complete the host looks at the
 * flag words k_finalize left in page-locked memory: a job whose records were out of order (the reference asks
 * for coordinate-sorted input, README.md:93, but its result does not depend on the order) is run again through
 * the per-chromosome entry points, whose out-of-order path gives the reference's result. */
static int settle_slot(cg_ctx* c, cg_slot& s)
{
    if (!s.check) return CG_OK;
    s.check = false;
    CU(c, cudaStreamSynchronize(s.s_work));
    for (int t = 0; t < 2; t++) CU(c, cudaStreamSynchronize(s.s_geno[t]));
    std::vector<cg_chrom_job> redo;
    for (int k = 0; k < s.n_jobs; k++) if (s.h_flags[k] & CG_FLAG_UNSORTED) redo.push_back(s.src[(size_t)k]);
    if (redo.empty()) return CG_OK;
    const int mode = c->ptr_mode; const bool async = c->async;
    const bool in_chrom = c->in_chrom, finished = c->finished; const int cur = c->cur;
    c->ptr_mode = s.ptr_mode; c->async = s.async;
    int rc = CG_OK;
    for (const cg_chrom_job& j : redo) if ((rc = run_one_job(c, j)) != CG_OK) break;
    if (rc == CG_OK) rc = sync_all(c);
    c->ptr_mode = mode; c->async = async;
    c->in_chrom = in_chrom; c->finished = finished; c->cur = cur;
    return rc;
}

static int settle_samples(cg_ctx* c)
{
    for (int i = 0; i < CG_NSLOT; i++) {
        int rc = settle_slot(c, c->slot[i]);
        if (rc) return rc;
    }
    return CG_OK;
}

extern "C" {

/* host-side input preparation: positions -> delta bytes + side array + anchors (include/conga_gpu.h) */
int cg_encode_positions_delta(const int32_t* pos, size_t n, uint8_t* delta, int32_t* big, size_t big_cap, size_t* n_big,
                              cg_delta_anchor* anchors)
{
    if ((n && (!pos || !delta || !anchors)) || !n_big) return fail(nullptr, CG_EINVAL, "null argument");
    size_t nb = 0;
    int32_t prev = 0;
    bool fits = true;
    for (size_t i = 0; i < n; i++) {
        if ((i % CG_DELTA_BLOCK) == 0) {
            anchors[i / CG_DELTA_BLOCK].pos_before = prev;
            anchors[i / CG_DELTA_BLOCK].big_before = (uint32_t)nb;
        }
        const int32_t p = pos[i];
        if (p < prev) return fail(nullptr, CG_EINVAL, "record %zu: position %d behind %d — a delta-coded stream needs non-negative, non-decreasing positions", i, p, prev);
        const int64_t d = (int64_t)p - (int64_t)prev;
        if (d >= 255) {
            delta[i] = 255;
            if (nb < big_cap && big) big[nb] = (int32_t)d; else fits = false;
            nb++;
        } else {
            delta[i] = (uint8_t)d;
        }
        prev = p;
    }
    *n_big = nb;
    if (!fits) return fail(nullptr, CG_ERANGE, "the side array needs %zu entries (%zu given)", nb, big_cap);
    return CG_OK;
}

int cg_run_chromosomes(cg_ctx* c, const cg_chrom_job* jobs, size_t n_jobs)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (n_jobs && !jobs) return fail(c, CG_EINVAL, "null jobs");
    if (c->in_chrom && !c->finished) return fail(c, CG_EINVAL, "cg_run_chromosomes inside a chromosome");
    CU(c, cudaSetDevice(c->device));
    int rc;
    if (blocking(c)) {
        for (size_t k = 0; k < n_jobs; k++) if ((rc = run_one_job(c, jobs[k])) != CG_OK) return rc;
        return CG_OK;
    }
    for (size_t k = 0; k < n_jobs; k += CG_JMAX)
        if ((rc = submit_sample(c, jobs + k, (int)std::min<size_t>(CG_JMAX, n_jobs - k))) != CG_OK) return rc;
    return CG_OK;
}

/* ------------------------------------------------------------------------- */
int cg_download_depth(cg_ctx* c, int16_t* out)
{
    if (!c || !out) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_download_depth before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    CU(c, cudaMemcpyAsync(out, s.d_depth, (size_t)s.hj[0].chrom_len * sizeof(int16_t), cudaMemcpyDeviceToHost, s.s_work));
    CU(c, cudaStreamSynchronize(s.s_work));
    return CG_OK;
}

int cg_download_prefix(cg_ctx* c, int32_t* out)
{
    if (!c || !out) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_download_prefix before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    cudaStream_t st = s.s_work;
    /* no prefix array exists on the product path: rebuilt here from the depth array and the chunk sums */
    int32_t* d_prefix = nullptr;
    CU(c, cudaMalloc(&d_prefix, (size_t)s.hj[0].n_chunks * CG_CHUNK * sizeof(int32_t)));
    cudaError_t e = cgk_prefix_from_depth(st, s.d_depth, s.sub.k, 0, s.hj[0].n_chunks, d_prefix);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_prefix, (size_t)s.hj[0].chrom_len * sizeof(int32_t), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d_prefix);
    CU(c, e);
    return CG_OK;
}

int cg_device_arrays(cg_ctx* c, const int16_t** depth, const float** expected)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    cg_slot& s = c->slot[c->cur];
    if (depth) *depth = s.d_depth;
    if (expected) *expected = s.d_expected;
    return CG_OK;
}

/* ---- statistics tables and the collective -------------------------------------------------- */
static long long* table_of(cg_ctx* c, int bank)
{
    return ((c->comm || c->n_peers) ? c->d_stats_all : c->d_stats) + (size_t)bank * CG_MAX_CHROMS * CG_STATS_WORDS;
}

int cg_stats_table(cg_ctx* c, void** device_ptr, int32_t* n_rows)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (device_ptr) *device_ptr = table_of(c, c->bank);
    if (n_rows) *n_rows = c->n_rows;
    return CG_OK;
}

int cg_stats_reset(cg_ctx* c)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    CU(c, cudaSetDevice(c->device));
    int rc = cg_join(c);                      /* rows may still be written or copied by queued work */
    if (rc) return rc;
    const size_t bytes = sizeof(long long) * CG_STATS_WORDS * CG_MAX_CHROMS, off = (size_t)c->bank * CG_MAX_CHROMS * CG_STATS_WORDS;
    CU(c, cudaMemsetAsync(c->d_stats + off, 0, bytes, c->s_main));
    if (c->d_stats_all) CU(c, cudaMemsetAsync(c->d_stats_all + off, 0, bytes, c->s_main));
    /* chromosomes that write into this bank from now on wait for the clear, like for a collective that reads it */
    CU(c, cudaEventRecord(c->ev_bank_read[c->bank], c->s_main));
    c->bank_read[c->bank] = true;
    c->n_rows = 0;
    return CG_OK;
}

int cg_stats_bank(cg_ctx* c, int bank)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (bank < 0 || bank >= CG_NBANK) return fail(c, CG_EINVAL, "bank %d must be 0..%d", bank, CG_NBANK - 1);
    if (c->in_chrom && !c->finished) return fail(c, CG_EINVAL, "cg_stats_bank inside a chromosome");
    c->bank = bank;
    return CG_OK;
}

int cg_stats_row(cg_ctx* c, int chr_index, cg_chrom_stats* out)
{
    if (!c || !out) return fail(c, CG_EINVAL, "null argument");
    if (chr_index < 0 || chr_index >= CG_MAX_CHROMS) return fail(c, CG_EINVAL, "chr_index %d out of range", chr_index);
    CU(c, cudaSetDevice(c->device));
    int rc = sync_all(c);
    if (rc) return rc;
    if ((rc = settle_samples(c)) != CG_OK) return rc;
    CU(c, cudaMemcpyAsync(c->h_stats, table_of(c, c->bank) + (size_t)chr_index * CG_STATS_WORDS, sizeof(cg_chrom_stats),
                          cudaMemcpyDeviceToHost, c->s_main));
    CU(c, cudaStreamSynchronize(c->s_main));
    memcpy(out, c->h_stats, sizeof(cg_chrom_stats));
    return CG_OK;
}

} /* extern "C" */

/* NCCL is looked up at run time: libconga_gpu.so has no link-time dependency on it, and in a process that
 * already carries one (PyTorch's) the same library is used. */
static cg_nccl_api g_nccl;
static std::mutex g_nccl_mutex;

static int nccl_load(cg_ctx* c)
{
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    if (g_nccl.handle) return CG_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(c, CG_ECUDA, "NCCL not found: %s", dlerror());
    cg_nccl_api a;
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
    a.CommInitAll = reinterpret_cast<decltype(a.CommInitAll)>(dlsym(h, "ncclCommInitAll"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(h, "ncclAllReduce"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
    a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(dlsym(h, "ncclGroupStart"));
    a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(dlsym(h, "ncclGroupEnd"));
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommInitAll || !a.CommDestroy || !a.AllReduce || !a.GetErrorString || !a.GroupStart || !a.GroupEnd)
        return fail(c, CG_ECUDA, "NCCL library lacks a required symbol");
    a.handle = h;
    g_nccl = a;
    return CG_OK;
}

#define CG_TABLE_WORDS ((size_t)CG_STATS_WORDS * CG_MAX_CHROMS * CG_NBANK)       /* int64 words of the tables of all banks */
static inline unsigned int* arrival_words(long long* tables) { return reinterpret_cast<unsigned int*>(tables + CG_TABLE_WORDS); }

static int comm_tables(cg_ctx* c)
{
    if (c->d_stats_all) return CG_OK;
    /* the reduced tables, and behind them the peers' arrival words [bank][rank] of the peer exchange */
    const size_t bytes = sizeof(long long) * CG_TABLE_WORDS + sizeof(unsigned int) * CG_NBANK * CG_MAX_PEERS;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMalloc(&c->d_stats_all, bytes));
    CU(c, cudaMemset(c->d_stats_all, 0, bytes));
    return CG_OK;
}

extern "C" {

int cg_peer_export(cg_ctx* c, void* handle)
{
    if (!c || !handle) return fail(c, CG_EINVAL, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) <= CG_PEER_HANDLE_BYTES, "IPC handle size");
    int rc = comm_tables(c);
    if (rc) return rc;
    cudaIpcMemHandle_t h;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaIpcGetMemHandle(&h, c->d_stats_all));
    memset(handle, 0, CG_PEER_HANDLE_BYTES);
    memcpy(handle, &h, sizeof h);
    return CG_OK;
}

static int peer_ready(cg_ctx* c, int n_ranks, int rank)
{
    if (n_ranks < 1 || n_ranks > CG_MAX_PEERS || rank < 0 || rank >= n_ranks) return fail(c, CG_EINVAL, "rank %d of %d (at most %d)", rank, n_ranks, CG_MAX_PEERS);
    if (c->n_peers || c->comm) return fail(c, CG_EINVAL, "the context already has peers or a communicator");
    if (c->in_chrom && !c->finished) return fail(c, CG_EINVAL, "cg_peer_connect inside a chromosome");
    CU(c, cudaSetDevice(c->device));
    int rc = sync_all(c);
    if (rc) return rc;
    return comm_tables(c);
}

int cg_peer_connect(cg_ctx* c, int n_ranks, int rank, const void* handles)
{
    if (!c || !handles) return fail(c, CG_EINVAL, "null argument");
    int rc = peer_ready(c, n_ranks, rank);
    if (rc) return rc;
    for (int r = 0; r < n_ranks; r++) {
        if (r == rank) { c->peer_base[r] = c->d_stats_all; c->peer_ipc[r] = false; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const char*>(handles) + (size_t)r * CG_PEER_HANDLE_BYTES, sizeof h);
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            for (int q = 0; q < r; q++) if (c->peer_ipc[q]) { cudaIpcCloseMemHandle(c->peer_base[q]); c->peer_ipc[q] = false; }
            return fail(c, CG_ECUDA, "cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
        }
        c->peer_base[r] = static_cast<long long*>(p); c->peer_ipc[r] = true;
    }
    c->n_peers = n_ranks; c->peer_rank = rank;
    return CG_OK;
}

int cg_peer_connect_local(cg_ctx** ctxs, int n)
{
    if (!ctxs || n < 1 || n > CG_MAX_PEERS) return fail(nullptr, CG_EINVAL, "bad argument");
    int rc;
    for (int i = 0; i < n; i++) {
        if (!ctxs[i]) return fail(nullptr, CG_EINVAL, "context %d is null", i);
        if ((rc = peer_ready(ctxs[i], n, i)) != CG_OK) return rc;
    }
    for (int i = 0; i < n; i++) {
        cudaSetDevice(ctxs[i]->device);
        for (int r = 0; r < n; r++) {
            if (ctxs[r]->device != ctxs[i]->device) {
                const cudaError_t e = cudaDeviceEnablePeerAccess(ctxs[r]->device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(ctxs[i], CG_ECUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", ctxs[i]->device, ctxs[r]->device, cudaGetErrorString(e));
                cudaGetLastError();
            }
            ctxs[i]->peer_base[r] = ctxs[r]->d_stats_all; ctxs[i]->peer_ipc[r] = false;
        }
        ctxs[i]->n_peers = n; ctxs[i]->peer_rank = i;
    }
    return CG_OK;
}

int cg_peer_disconnect(cg_ctx* c)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->n_peers) return CG_OK;
    cudaSetDevice(c->device);
    sync_all(c);
    for (int r = 0; r < c->n_peers; r++) if (c->peer_ipc[r]) { cudaIpcCloseMemHandle(c->peer_base[r]); c->peer_ipc[r] = false; }
    c->n_peers = 0;
    return CG_OK;
}

int cg_comm_unique_id(void* id)
{
    if (!id) return fail(nullptr, CG_EINVAL, "null id");
    int rc = nccl_load(nullptr);
    if (rc) return rc;
    const int e = g_nccl.GetUniqueId(id);
    if (e) return fail(nullptr, CG_ECUDA, "ncclGetUniqueId: %s", g_nccl.GetErrorString(e));
    return CG_OK;
}

int cg_comm_init_rank(cg_ctx* c, int n_ranks, int rank, const void* id)
{
    if (!c || !id) return fail(c, CG_EINVAL, "null argument");
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(c, CG_EINVAL, "rank %d of %d", rank, n_ranks);
    if (c->comm) return fail(c, CG_EINVAL, "the context already has a communicator");
    int rc = nccl_load(c);
    if (rc) return rc;
    CU(c, cudaSetDevice(c->device));
    if ((rc = sync_all(c)) != CG_OK) return rc;
    if ((rc = comm_tables(c)) != CG_OK) return rc;
    cg_nccl_id uid;
    memcpy(&uid, id, sizeof uid);
    const int e = g_nccl.CommInitRank(&c->comm, n_ranks, uid, rank);
    if (e) { c->comm = nullptr; return fail(c, CG_ECUDA, "ncclCommInitRank: %s", g_nccl.GetErrorString(e)); }
    c->comm_ranks = n_ranks;
    return CG_OK;
}

int cg_comm_init_all(cg_ctx** ctxs, int n)
{
    if (!ctxs || n < 1) return fail(nullptr, CG_EINVAL, "bad argument");
    int rc = nccl_load(ctxs[0]);
    if (rc) return rc;
    std::vector<int> devs((size_t)n);
    std::vector<cg_nccl_comm_t> comms((size_t)n, nullptr);
    for (int i = 0; i < n; i++) {
        if (!ctxs[i] || ctxs[i]->comm) return fail(ctxs[i], CG_EINVAL, "context %d is null or already has a communicator", i);
        devs[(size_t)i] = ctxs[i]->device;
        if ((rc = comm_tables(ctxs[i])) != CG_OK) return rc;
    }
    const int e = g_nccl.CommInitAll(comms.data(), n, devs.data());
    if (e) return fail(ctxs[0], CG_ECUDA, "ncclCommInitAll: %s", g_nccl.GetErrorString(e));
    for (int i = 0; i < n; i++) { ctxs[i]->comm = comms[(size_t)i]; ctxs[i]->comm_ranks = n; }
    return CG_OK;
}

int cg_comm_destroy(cg_ctx* c)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (c->comm && g_nccl.CommDestroy) { cudaSetDevice(c->device); g_nccl.CommDestroy(c->comm); }
    c->comm = nullptr; c->comm_ranks = 0;
    return CG_OK;
}

static int allreduce_one(cg_ctx* c, int32_t n_rows)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->comm && !c->n_peers) return fail(c, CG_EINVAL, "cg_stats_allreduce without a communicator (cg_comm_init_rank) or peers (cg_peer_connect)");
    if (n_rows < 0 || n_rows > CG_MAX_CHROMS) return fail(c, CG_EINVAL, "n_rows %d", n_rows);
    if (c->in_chrom && !c->finished) return fail(c, CG_EINVAL, "cg_stats_allreduce inside a chromosome");
    CU(c, cudaSetDevice(c->device));
    int rc = cg_join(c);                       /* cg_stream now waits for every row written so far */
    if (rc) return rc;
    if (c->n_peers) {
        /* the rows are already in every rank's table (k_finalize stored them there): tell the peers, wait for theirs */
        cg_peer_flags f;
        memset(&f, 0, sizeof f);
        for (int r = 0; r < c->n_peers; r++) f.theirs[r] = arrival_words(c->peer_base[r]) + (size_t)c->bank * CG_MAX_PEERS;
        f.mine = arrival_words(c->d_stats_all) + (size_t)c->bank * CG_MAX_PEERS;
        f.n = c->n_peers; f.rank = c->peer_rank; f.epoch = ++c->epoch[c->bank]; f.sticky = c->d_sticky;
        CU(c, cgk_peer_exchange(c->s_main, f));
        CU(c, cudaEventRecord(c->ev_bank_read[c->bank], c->s_main));
        c->bank_read[c->bank] = true;
        c->launch_count += 1;
        return CG_OK;
    }
    if (n_rows == 0) return CG_OK;
    const size_t off = (size_t)c->bank * CG_MAX_CHROMS * CG_STATS_WORDS;
    const int e = g_nccl.AllReduce(c->d_stats + off, c->d_stats_all + off, (size_t)n_rows * CG_STATS_WORDS, /* ncclInt64 */ 4, /* ncclSum */ 0,
                                   c->comm, c->s_main);
    if (e) return fail(c, CG_ECUDA, "ncclAllReduce: %s", g_nccl.GetErrorString(e));
    CU(c, cudaEventRecord(c->ev_bank_read[c->bank], c->s_main));
    c->bank_read[c->bank] = true;
    c->launch_count += 1;
    return CG_OK;
}

int cg_stats_allreduce(cg_ctx* c, int32_t n_rows) { return allreduce_one(c, n_rows); }

int cg_stats_allreduce_all(cg_ctx** ctxs, int n, int32_t n_rows)
{
    if (!ctxs || n < 1) return fail(nullptr, CG_EINVAL, "bad argument");
    int rc = nccl_load(ctxs[0]);
    if (rc) return rc;
    g_nccl.GroupStart();                       /* one thread issuing for several communicators */
    for (int i = 0; i < n && rc == CG_OK; i++) rc = allreduce_one(ctxs[i], n_rows);
    const int e = g_nccl.GroupEnd();
    if (rc == CG_OK && e) rc = fail(ctxs[0], CG_ECUDA, "ncclGroupEnd: %s", g_nccl.GetErrorString(e));
    return rc;
}

/* ---- split reads (--rp) ----------------------------------------------------------------- */
int cg_sr_reference(cg_ctx* c, const char* ref_seq, int32_t ref_len)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom) return fail(c, CG_EINVAL, "cg_sr_reference outside a chromosome");
    if (!ref_seq || ref_len < 0) return fail(c, CG_EINVAL, "bad reference sequence");
    CU(c, cudaSetDevice(c->device));
    auto& r = c->sr;
    /* each piece on its own test: a call that failed half way (CG_ENOMEM) is completed by the next one */
    if (!r.st) CU(c, cudaStreamCreateWithFlags(&r.st, cudaStreamNonBlocking));
    if (!r.d_lo) CU(c, cudaMalloc(&r.d_lo, sizeof(int) << 20));
    if (!r.d_hi) CU(c, cudaMalloc(&r.d_hi, sizeof(int) << 20));
    if (!r.d_n_rows) {
        CU(c, cudaMalloc(&r.d_n_rows, 16));
        r.d_n_map = reinterpret_cast<unsigned long long*>(r.d_n_rows + 2);
    }
    if (!r.d_rows) {
        CU(c, cudaMalloc(&r.d_rows, sizeof(cg_splitrow) * (size_t)SR_ROWS_CAP));
        r.rows_cap = SR_ROWS_CAP;
    }
    int rc;
    if ((rc = sr_grow(c, &r.d_ref, &r.ref_cap, (size_t)ref_len + 1))) return rc;
    if (!r.d_cnt) CU(c, cudaMalloc(&r.d_cnt, sizeof(int) << 20));
    if ((size_t)ref_len > r.idx_cap) {
        CU(c, cudaStreamSynchronize(r.st));
        cudaFree(r.d_pos_list); r.d_pos_list = nullptr; r.idx_cap = 0;
        const size_t n = (size_t)ref_len + 1024;
        CU(c, cudaMalloc(&r.d_pos_list, n * 4));
        r.idx_cap = n;
    }
    CU(c, cudaMemcpyAsync(r.d_ref, ref_seq, (size_t)ref_len, cudaMemcpyHostToDevice, r.st));
    CU(c, cudaMemsetAsync(r.d_ref + ref_len, 0, 1, r.st));
    CU(c, cudaMemsetAsync(r.d_n_rows, 0, 16, r.st));
    {
        prof_scope ps(c, r.st, 3, 3);
        CU(c, cgk_sr_index(r.st, r.d_ref, ref_len, r.d_cnt, r.d_pos_list, r.d_lo, r.d_hi, c->n_sms));
    }
    CU(c, cudaStreamSynchronize(r.st));            /* ref_seq may be pageable and may die at return */
    r.ref_len = ref_len;
    r.have_ref = true;
    r.chrom_len = c->slot[c->cur].hj[0].chrom_len;
    r.mq_threshold = c->slot[c->cur].hj[0].mq_threshold;
    return CG_OK;
}

int cg_sr_push(cg_ctx* c, const cg_half* halves, size_t n, const char* bases, size_t n_bases)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->sr.have_ref) return fail(c, CG_EINVAL, "cg_sr_push before cg_sr_reference");
    if (n == 0) return CG_OK;
    if (!halves || !bases) return fail(c, CG_EINVAL, "null argument");
    if (n_bases > 0xffffffffull) return fail(c, CG_EINVAL, "more than 2^32 bases in one cg_sr_push");
    CU(c, cudaSetDevice(c->device));
    auto& r = c->sr;
    int rc;
    /* two staging sets in rotation: this batch's copies and search are queued behind the previous batch's without the
     * host waiting for either; a set is written again only when the search that read it is done (a pageable host
     * buffer is consumed by cudaMemcpyAsync before it returns, so the caller may reuse it at once) */
    const int b = (int)(r.n_push++ & 1u);
    if (!r.ev_buf[b]) CU(c, cudaEventCreateWithFlags(&r.ev_buf[b], cudaEventDisableTiming));
    CU(c, cudaEventSynchronize(r.ev_buf[b]));
    if ((rc = sr_grow(c, &r.d_halves[b], &r.halves_cap[b], n))) return rc;
    if ((rc = sr_grow(c, &r.d_bases[b], &r.bases_cap[b], n_bases + 1))) return rc;
    CU(c, cudaMemcpyAsync(r.d_halves[b], halves, n * sizeof(cg_half), cudaMemcpyHostToDevice, r.st));
    CU(c, cudaMemcpyAsync(r.d_bases[b], bases, n_bases, cudaMemcpyHostToDevice, r.st));
    {
        prof_scope ps(c, r.st, 3, 1);
        CU(c, cgk_sr_search(r.st, r.d_ref, r.ref_len, r.d_pos_list, r.d_lo, r.d_hi, r.d_halves[b], (long long)n, r.d_bases[b], r.chrom_len,
                            r.mq_threshold, r.d_rows, r.d_n_rows, r.rows_cap, r.d_n_map, c->n_sms));
    }
    CU(c, cudaEventRecord(r.ev_buf[b], r.st));
    return CG_OK;
}

int cg_sr_rows(cg_ctx* c, cg_splitrow* out, size_t cap, size_t* n_rows, uint64_t* n_mappings)
{
    if (!c || !n_rows) return fail(c, CG_EINVAL, "null argument");
    if (!c->sr.have_ref) return fail(c, CG_EINVAL, "cg_sr_rows before cg_sr_reference");
    CU(c, cudaSetDevice(c->device));
    auto& r = c->sr;
    unsigned long long head[2] = {0, 0};
    CU(c, cudaMemcpyAsync(head, r.d_n_rows, 16, cudaMemcpyDeviceToHost, r.st));
    CU(c, cudaStreamSynchronize(r.st));
    const unsigned n = (unsigned)(head[0] & 0xffffffffull);
    if (n > r.rows_cap) return fail(c, CG_ENOMEM, "%u split-read rows exceed the device buffer (%u)", n, r.rows_cap);
    *n_rows = n;
    if (n_mappings) *n_mappings = head[1];
    const size_t take = n < cap ? n : cap;
    if (take && !out) return fail(c, CG_EINVAL, "null rows buffer");
    if (take) {
        CU(c, cudaMemcpyAsync(out, r.d_rows, take * sizeof(cg_splitrow), cudaMemcpyDeviceToHost, r.st));
        CU(c, cudaStreamSynchronize(r.st));
    }
    return CG_OK;
}

int cg_sr_count(cg_ctx* c, const cg_splitrow* rows, size_t n, const cg_cnv* dels, size_t n_del,
                const cg_cnv* dups, size_t n_dup, int32_t* border_rp, int32_t* rp)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if ((n && !rows) || (n_del && (!dels || !border_rp)) || (n_dup && (!dups || !rp))) return fail(c, CG_EINVAL, "null argument");
    if (n > 0x7fffffffull || n_del > 0x3fffffffull || n_dup > 0x3fffffffull) return fail(c, CG_EINVAL, "too many rows or CNVs");
    CU(c, cudaSetDevice(c->device));
    auto& r = c->sr;
    if (!r.st) CU(c, cudaStreamCreateWithFlags(&r.st, cudaStreamNonBlocking));
    const size_t b_rows = (n * sizeof(cg_splitrow) + 255) & ~(size_t)255, b_del = (n_del * sizeof(cg_cnv) + 255) & ~(size_t)255,
                 b_dup = (n_dup * sizeof(cg_cnv) + 255) & ~(size_t)255, b_cd = (n_del * 4 + 255) & ~(size_t)255, b_cu = (n_dup * 4 + 255) & ~(size_t)255;
    int rc;
    if ((rc = sr_grow(c, &r.d_count, &r.count_cap, b_rows + b_del + b_dup + b_cd + b_cu + 256))) return rc;
    char* p = r.d_count;
    cg_splitrow* d_rows = reinterpret_cast<cg_splitrow*>(p); p += b_rows;
    cg_cnv* d_del = reinterpret_cast<cg_cnv*>(p); p += b_del;
    cg_cnv* d_dup = reinterpret_cast<cg_cnv*>(p); p += b_dup;
    int* d_cd = reinterpret_cast<int*>(p); p += b_cd;
    int* d_cu = reinterpret_cast<int*>(p);
    if (n) CU(c, cudaMemcpyAsync(d_rows, rows, n * sizeof(cg_splitrow), cudaMemcpyHostToDevice, r.st));
    if (n_del) CU(c, cudaMemcpyAsync(d_del, dels, n_del * sizeof(cg_cnv), cudaMemcpyHostToDevice, r.st));
    if (n_dup) CU(c, cudaMemcpyAsync(d_dup, dups, n_dup * sizeof(cg_cnv), cudaMemcpyHostToDevice, r.st));
    CU(c, cudaMemsetAsync(d_cd, 0, b_cd + b_cu, r.st));
    {
        prof_scope ps(c, r.st, 3, 1);
        CU(c, cgk_sr_count(r.st, d_rows, (int)n, d_del, (int)n_del, d_dup, (int)n_dup, d_cd, d_cu, c->n_sms));
    }
    if (n_del) CU(c, cudaMemcpyAsync(border_rp, d_cd, n_del * 4, cudaMemcpyDeviceToHost, r.st));
    if (n_dup) CU(c, cudaMemcpyAsync(rp, d_cu, n_dup * 4, cudaMemcpyDeviceToHost, r.st));
    CU(c, cudaStreamSynchronize(r.st));
    return CG_OK;
}

int cg_gc_profile(cg_ctx* c, const char* bases, int64_t n_bases, int8_t* out_gc, int32_t n_windows)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (n_bases < 0 || n_bases > 0x7fffffffll) return fail(c, CG_EINVAL, "chromosome of %lld bases", (long long)n_bases);
    const int32_t need = (int32_t)((n_bases + 1) / 100);           /* complete windows only: sonic_reference.c:95 */
    if (n_windows < need) return fail(c, CG_EINVAL, "out_gc holds %d bytes, %d needed", n_windows, need);
    if (need == 0) return CG_OK;
    if (!bases || !out_gc) return fail(c, CG_EINVAL, "null argument");
    CU(c, cudaSetDevice(c->device));
    cudaStream_t st = c->s_main;
    if (c->ptr_mode == CG_PTR_DEVICE) {
        prof_scope ps(c, st, 0, 1);
        CU(c, cgk_gc_profile(st, bases, n_bases, out_gc, need, c->n_sms));
        return CG_OK;
    }
    if ((size_t)n_bases + 16 > c->gc_bases_cap) {
        CU(c, cudaStreamSynchronize(st));
        cudaFree(c->d_gc_bases); c->d_gc_bases = nullptr; c->gc_bases_cap = 0;
        const size_t cap = (size_t)n_bases + (size_t)n_bases / 8 + 4096;
        CU(c, cudaMalloc(&c->d_gc_bases, cap));
        c->gc_bases_cap = cap;
    }
    if ((size_t)need > c->gc_out_cap) {
        CU(c, cudaStreamSynchronize(st));
        cudaFree(c->d_gc_out); c->d_gc_out = nullptr; c->gc_out_cap = 0;
        const size_t cap = (size_t)need + (size_t)need / 8 + 4096;
        CU(c, cudaMalloc(&c->d_gc_out, cap));
        c->gc_out_cap = cap;
    }
    CU(c, cudaMemcpyAsync(c->d_gc_bases, bases, (size_t)n_bases, cudaMemcpyHostToDevice, st));
    {
        prof_scope ps(c, st, 0, 1);
        CU(c, cgk_gc_profile(st, c->d_gc_bases, n_bases, c->d_gc_out, need, c->n_sms));
    }
    CU(c, cudaMemcpyAsync(out_gc, c->d_gc_out, (size_t)need, cudaMemcpyDeviceToHost, st));
    CU(c, cudaStreamSynchronize(st));                  /* bases / out_gc may be pageable */
    return CG_OK;
}

} /* extern "C" */
