/* conga_gpu.cu — context management and the C-ABI of libconga_gpu.so
 * (include/conga_gpu.h).  Host code only; the kernels live in depth_pass.cu and
 * genotype.cu.  There is no CPU implementation of any stage in this library:
 * without a CUDA device cg_create fails and nothing else can be called.
 *
 * Execution model.  A context owns CG_NSLOT chromosome slots (cg_set_slots puts 2..CG_NSLOT of them in rotation)
 * and four streams per slot:
 *   s_copy[k]  host->device copies of GC bytes and read batches of slot k (HOST pointer mode); one stream per slot,
 *              so that the few-megabyte copies of consecutive chromosomes overlap their start-up gaps on the copy engines
 *   s_slot[k]  one per slot: chromosome init, batch marking, rare paths, the fused depth pass
 *   s_geno[2]  per slot, deletions / duplications: mappability + CNV uploads, the genotype kernel,
 *              downloads.  The genotype kernel is a chain of dependent loads whose latency grows
 *              several-fold under a depth pass that saturates HBM, so the genotype kernels of
 *              up to CG_NSLOT chromosomes are allowed to run side by side.
 *   s_main     cg_stream(): statistics reset, and what cg_join() orders behind everything else
 * Chromosome k lives in slot k mod n_slot (its own records, depth, GC, expected-table, count and
 * staging buffers) and no slot depends on another.  So the preparation kernels and the
 * ramp-up of chromosome k+1's depth pass fill the SMs that chromosome k's pass is draining from,
 * the latency-bound genotype kernels of k run underneath the bandwidth-bound pass of k+1, and the
 * copies of k+1 run underneath the pass of k.  Events order the streams per slot; nothing ever
 * waits on the host unless the caller asked for blocking HOST-mode semantics.  With profiling on
 * (cg_set_profiling) the slots are chained one behind the other so that the event brackets time
 * each kernel alone.
 */
#include "cg_internal.h"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CG_MAX_CHROMS 4096            /* rows of the statistics table */
#ifndef CG_NSLOT
#define CG_NSLOT 6                    /* measured r01h, C2 resident: 3 slots 1.48 ms per sample, 4 slots 1.22, 6 slots 1.09 (host-issue bound) */
#endif

struct cg_evpair { cudaEvent_t a, b; int slot; };

/* everything one chromosome needs from cg_chrom_begin to its last cg_genotype */
struct cg_slot {
    /* geometry */
    int chr_index = -1;
    int32_t chrom_len = 0;
    int n_chunks = 0;
    long long* d_row = nullptr;               /* this chromosome's int64[204] row in the statistics bank it was begun under */
    /* depth side */
    cg_read* d_reads = nullptr;       size_t reads_cap = 0;
    int8_t* d_gc = nullptr;           size_t gc_cap = 0;
    int32_t* d_counts = nullptr;      size_t counts_cap = 0;      /* chunk bounds + three count levels (cg_counts) */
    int16_t* d_depth = nullptr;       size_t pos_cap = 0;
    float* d_expected = nullptr;              /* [101] (128 allocated) */
    unsigned int* d_ctl = nullptr;            /* CG_W_* control words */
    const cg_read* reads_ptr = nullptr;       /* own buffer or the caller's (DEVICE mode) */
    const int8_t* gc_ptr = nullptr;
    /* genotype side */
    cg_maprun* d_runs = nullptr;      size_t runs_cap = 0;
    int32_t* d_run_idx = nullptr;     size_t run_idx_cap = 0;    /* first run per 16384 positions */
    const cg_maprun* runs_ptr = nullptr; int n_runs = 0; int has_map = 0;
    char* d_arena = nullptr;          size_t arena_cap = 0, arena_used = 0;   /* CNV + call staging, HOST mode */
    /* ordering */
    cudaEvent_t ev_copy = nullptr;            /* s_copy: last upload for this slot */
    cudaEvent_t ev_pass = nullptr;            /* slot stream: the pass (and the table) of this slot is done */
    cudaStream_t s_geno[2] = {nullptr, nullptr};   /* [0] deletions (and uploads), [1] duplications */
    cudaEvent_t ev_geno[2] = {nullptr, nullptr};   /* s_geno[t]: last genotype work queued for this slot */
    cudaEvent_t ev_runs = nullptr;            /* s_geno[0]: mappability runs uploaded */
    bool used = false, geno_queued[2] = {false, false}, runs_uploaded = false, wait_reset = false;
    uint32_t* d_depth32 = nullptr;    size_t depth32_cap = 0;    /* out-of-order path only; zero at rest */
};

struct cg_ctx {
    int device = 0;
    int n_sms = 148;
    cudaStream_t s_copy[CG_NSLOT] = {}, s_main = nullptr;
    cudaStream_t s_slot[CG_NSLOT] = {};
    cudaStream_t s_bank = nullptr;            /* clears a statistics bank that is selected again */
    cudaEvent_t ev_reset = nullptr;           /* the current statistics bank has been cleared */
    cudaEvent_t ev_left[2] = {nullptr, nullptr};   /* s_main: everything queued when the bank was left (its collective) */
    bool bank_left[2] = {false, false};
    int bank = 0;
    bool dirty = false;                       /* work queued since the last cg_join */
    cudaEvent_t ev_join = nullptr;
    int ptr_mode = CG_PTR_HOST;
    bool async = false;
    char err[512] = {0};

    cg_slot slot[CG_NSLOT];
    int n_slot = CG_NSLOT;                    /* slots in rotation (cg_set_slots); the others stay idle */
    int cur = 0;                              /* slot of the current chromosome */
    long long n_begun = 0;

    /* current chromosome */
    bool in_chrom = false, finished = false;
    int mq_threshold = -1;
    long long n_reads = 0;
    int n_rows = 0;

    long long* d_stats = nullptr;             /* [2 banks][CG_MAX_CHROMS][204] */
    unsigned int* d_sticky = nullptr;         /* [0] CG_FLAG_RANGE seen since the last report */

    /* pinned staging */
    cg_chrom_stats* h_stats = nullptr;
    unsigned int* h_flags = nullptr;

    /* split reads (--rp): one chromosome at a time on its own stream */
    struct {
        cudaStream_t st = nullptr;
        unsigned char* d_ref = nullptr; size_t ref_cap = 0; int ref_len = 0;
        unsigned *d_keys_a = nullptr, *d_keys_b = nullptr; int *d_vals_a = nullptr, *d_vals_b = nullptr; size_t idx_cap = 0;
        void* d_temp = nullptr; size_t temp_cap = 0;
        int *d_lo = nullptr, *d_hi = nullptr;
        cg_half* d_halves = nullptr; size_t halves_cap = 0;
        unsigned char* d_bases = nullptr; size_t bases_cap = 0;
        cg_splitrow* d_rows = nullptr; unsigned rows_cap = 0;
        unsigned* d_n_rows = nullptr; unsigned long long* d_n_map = nullptr;     /* one 16-byte block */
        char* d_count = nullptr; size_t count_cap = 0;
        bool have_ref = false;
        int chrom_len = 0, mq_threshold = -1;
    } sr;

    /* largest sizes any slot has needed, per buffer kind (see grow()) */
    size_t hw_gc = 0, hw_reads = 0, hw_runs = 0, hw_run_idx = 0, hw_chunks = 0;

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
    for (int i = 0; i < CG_NSLOT; i++) CU(c, cudaStreamSynchronize(c->s_copy[i]));
    CU(c, cudaStreamSynchronize(c->s_main));
    CU(c, cudaStreamSynchronize(c->s_bank));
    if (c->sr.st) CU(c, cudaStreamSynchronize(c->sr.st));
    for (int i = 0; i < CG_NSLOT; i++) CU(c, cudaStreamSynchronize(c->s_slot[i]));
    for (int i = 0; i < CG_NSLOT; i++) for (int t = 0; t < 2; t++) CU(c, cudaStreamSynchronize(c->slot[i].s_geno[t]));
    return CG_OK;
}

/* Replace a device buffer by a larger one.  Everything queued is drained first (the old
 * buffer may still be read by any stream); growth is rare and amortised. */
template <typename T>
static int grow(cg_ctx* c, T** p, size_t* cap, size_t need, size_t keep_elems = 0, size_t* high_water = nullptr)
{
    if (need <= *cap) return CG_OK;
    /* chromosomes rotate through the slots, so a slot that has to grow goes straight to the largest size any
     * slot has needed: every slot reallocates at most once per buffer after the largest chromosome was seen */
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
        c->dirty = true;
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
    bool ok = true;
    for (int i = 0; ok && i < CG_NSLOT; i++) ok = cudaStreamCreateWithPriority(&c->s_copy[i], cudaStreamNonBlocking, prio_lo) == cudaSuccess;
    ok = ok && cudaStreamCreateWithPriority(&c->s_main, cudaStreamNonBlocking, prio_lo) == cudaSuccess
        && cudaStreamCreateWithPriority(&c->s_bank, cudaStreamNonBlocking, prio_lo) == cudaSuccess
        && cudaEventCreateWithFlags(&c->ev_reset, cudaEventDisableTiming) == cudaSuccess
        && cudaEventCreateWithFlags(&c->ev_left[0], cudaEventDisableTiming) == cudaSuccess
        && cudaEventCreateWithFlags(&c->ev_left[1], cudaEventDisableTiming) == cudaSuccess
        && cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) == cudaSuccess
        && cudaMalloc(&c->d_stats, sizeof(long long) * CG_STATS_WORDS * CG_MAX_CHROMS * 2) == cudaSuccess
        && cudaMemset(c->d_stats, 0, sizeof(long long) * CG_STATS_WORDS * CG_MAX_CHROMS * 2) == cudaSuccess
        && cudaMalloc(&c->d_sticky, 16) == cudaSuccess
        && cudaMemset(c->d_sticky, 0, 16) == cudaSuccess
        && cudaMallocHost(&c->h_stats, sizeof(cg_chrom_stats)) == cudaSuccess
        && cudaMallocHost(&c->h_flags, 16) == cudaSuccess;
    for (int i = 0; ok && i < CG_NSLOT; i++) {
        cg_slot& s = c->slot[i];
        ok = cudaStreamCreateWithPriority(&c->s_slot[i], cudaStreamNonBlocking, prio_lo) == cudaSuccess
            /* the genotype kernels are small and latency-bound: let their CTAs be placed first */
            && cudaStreamCreateWithPriority(&s.s_geno[0], cudaStreamNonBlocking, prio_hi) == cudaSuccess
            && cudaStreamCreateWithPriority(&s.s_geno[1], cudaStreamNonBlocking, prio_hi) == cudaSuccess
            && cudaMalloc(&s.d_expected, sizeof(float) * 128) == cudaSuccess
            && cudaMemset(s.d_expected, 0, sizeof(float) * 128) == cudaSuccess
            && cudaMalloc(&s.d_ctl, 16) == cudaSuccess
            && cudaMemset(s.d_ctl, 0, 16) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_copy, cudaEventDisableTiming) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_pass, cudaEventDisableTiming) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_geno[0], cudaEventDisableTiming) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_geno[1], cudaEventDisableTiming) == cudaSuccess
            && cudaEventCreateWithFlags(&s.ev_runs, cudaEventDisableTiming) == cudaSuccess;
    }
    if (!ok) {
        fail(nullptr, CG_ECUDA, "context allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
        cg_destroy(c);
        return nullptr;
    }
    snprintf(c->err, sizeof c->err, "no error");
    return c;
}

void cg_destroy(cg_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    for (int i = 0; i < CG_NSLOT; i++) if (c->s_copy[i]) cudaStreamSynchronize(c->s_copy[i]);
    if (c->s_main) cudaStreamSynchronize(c->s_main);
    if (c->s_bank) cudaStreamSynchronize(c->s_bank);
    for (int t = 0; t < CG_NSLOT; t++) if (c->s_slot[t]) cudaStreamSynchronize(c->s_slot[t]);
    for (int i = 0; i < CG_NSLOT; i++) for (int t = 0; t < 2; t++) if (c->slot[i].s_geno[t]) cudaStreamSynchronize(c->slot[i].s_geno[t]);
    for (auto& ev : c->ev_live) { cudaEventDestroy(ev.a); cudaEventDestroy(ev.b); }
    for (auto& ev : c->ev_free) { cudaEventDestroy(ev.a); cudaEventDestroy(ev.b); }
    for (int i = 0; i < CG_NSLOT; i++) {
        cg_slot& s = c->slot[i];
        cudaFree(s.d_reads); cudaFree(s.d_gc); cudaFree(s.d_counts);
        cudaFree(s.d_depth); cudaFree(s.d_expected); cudaFree(s.d_ctl);
        cudaFree(s.d_runs); cudaFree(s.d_run_idx); cudaFree(s.d_arena); cudaFree(s.d_depth32);
        if (s.ev_copy) cudaEventDestroy(s.ev_copy);
        if (s.ev_pass) cudaEventDestroy(s.ev_pass);
        for (int t = 0; t < 2; t++) if (s.ev_geno[t]) cudaEventDestroy(s.ev_geno[t]);
        if (s.ev_runs) cudaEventDestroy(s.ev_runs);
    }
    cudaFree(c->d_stats); cudaFree(c->d_sticky);
    if (c->sr.st) { cudaStreamSynchronize(c->sr.st); cudaStreamDestroy(c->sr.st); }
    cudaFree(c->sr.d_ref); cudaFree(c->sr.d_keys_a); cudaFree(c->sr.d_keys_b); cudaFree(c->sr.d_vals_a); cudaFree(c->sr.d_vals_b);
    cudaFree(c->sr.d_temp); cudaFree(c->sr.d_lo); cudaFree(c->sr.d_hi); cudaFree(c->sr.d_halves); cudaFree(c->sr.d_bases);
    cudaFree(c->sr.d_rows); cudaFree(c->sr.d_n_rows); cudaFree(c->sr.d_count);
    cudaFree(c->d_gc_bases); cudaFree(c->d_gc_out);
    if (c->h_stats) cudaFreeHost(c->h_stats);
    if (c->h_flags) cudaFreeHost(c->h_flags);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    for (int i = 0; i < CG_NSLOT; i++) if (c->s_copy[i]) cudaStreamDestroy(c->s_copy[i]);
    if (c->s_main) cudaStreamDestroy(c->s_main);
    if (c->s_bank) cudaStreamDestroy(c->s_bank);
    for (int b = 0; b < 2; b++) if (c->ev_left[b]) cudaEventDestroy(c->ev_left[b]);
    for (int t = 0; t < CG_NSLOT; t++) if (c->s_slot[t]) cudaStreamDestroy(c->s_slot[t]);
    if (c->ev_reset) cudaEventDestroy(c->ev_reset);
    for (int i = 0; i < CG_NSLOT; i++) for (int t = 0; t < 2; t++) if (c->slot[i].s_geno[t]) cudaStreamDestroy(c->slot[i].s_geno[t]);
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
    c->n_slot = n;
    /* the chromosome that was current may sit in a slot that has just left the rotation: it is closed, the next
     * call that needs a chromosome has to follow a cg_chrom_begin */
    c->in_chrom = false; c->finished = false;
    return CG_OK;
}

/* make the main stream (cg_stream) wait for everything queued on the others */
int cg_join(cg_ctx* c)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    CU(c, cudaSetDevice(c->device));
    /* every upload on a copy stream is waited for by a kernel on its slot's stream (cg_push_reads, cg_chrom_finish),
     * so the slot and genotype streams cover the copy streams */
    cudaStream_t all[3 * CG_NSLOT];
    for (int i = 0; i < c->n_slot; i++) { all[3 * i] = c->s_slot[i]; all[1 + 3 * i] = c->slot[i].s_geno[0]; all[2 + 3 * i] = c->slot[i].s_geno[1]; }
    for (int t = 0; t < 3 * c->n_slot; t++) {
        CU(c, cudaEventRecord(c->ev_join, all[t]));
        CU(c, cudaStreamWaitEvent(c->s_main, c->ev_join, 0));
    }
    c->dirty = false;
    return CG_OK;
}

int cg_sync(cg_ctx* c)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    CU(c, cudaSetDevice(c->device));
    int rc = sync_all(c);
    if (rc) return rc;
    prof_collect(c);
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

/* ------------------------------------------------------------------------- */
int cg_chrom_begin(cg_ctx* c, int chr_index, int32_t chrom_len, const int8_t* gc_profile,
                   int32_t n_windows, int mq_threshold)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (chr_index < 0 || chr_index >= CG_MAX_CHROMS) return fail(c, CG_EINVAL, "chr_index %d outside 0..%d", chr_index, CG_MAX_CHROMS - 1);
    if (chrom_len <= 0) return fail(c, CG_EINVAL, "chrom_len %d must be positive", chrom_len);
    if (chrom_len > INT32_MAX - CG_CHUNK) return fail(c, CG_EINVAL, "chrom_len %d too large", chrom_len);
    const int need_windows = (chrom_len + 99) / 100;
    if (!gc_profile || n_windows < need_windows)
        return fail(c, CG_EINVAL, "gc_profile needs >= %d windows for length %d (got %d)", need_windows, chrom_len, n_windows);
    CU(c, cudaSetDevice(c->device));

    const int si = (int)(c->n_begun % (unsigned long long)c->n_slot);
    cg_slot& s = c->slot[si];
    const int n_chunks = (chrom_len + CG_CHUNK - 1) / CG_CHUNK;
    const size_t n_pos = (size_t)n_chunks * CG_CHUNK;
    if ((size_t)n_chunks > c->hw_chunks) c->hw_chunks = (size_t)n_chunks;
    const int a_chunks = (int)c->hw_chunks;               /* what a slot that has to grow is sized for: the largest */
    const size_t a_pos = (size_t)a_chunks * CG_CHUNK;      /* chromosome any slot has seen (see grow()) */
    if (cgk_counts_words(n_chunks) > s.counts_cap) {
        int rc = sync_all(c);
        if (rc) return rc;
        cudaFree(s.d_counts); s.d_counts = nullptr; s.counts_cap = 0;
        CU(c, cudaMalloc(&s.d_counts, cgk_counts_words(a_chunks) * sizeof(int32_t)));
        s.counts_cap = cgk_counts_words(a_chunks);
    }
    if (n_pos > s.pos_cap) {
        int rc = sync_all(c);
        if (rc) return rc;
        cudaFree(s.d_depth);
        s.d_depth = nullptr; s.pos_cap = 0;
        CU(c, cudaMalloc(&s.d_depth, a_pos * sizeof(int16_t)));
        s.pos_cap = a_pos;
    }
    cudaStream_t ss = c->s_slot[si];
    if (n_pos > s.depth32_cap) {
        int rc = sync_all(c);
        if (rc) return rc;
        cudaFree(s.d_depth32); s.d_depth32 = nullptr; s.depth32_cap = 0;
        CU(c, cudaMalloc(&s.d_depth32, a_pos * sizeof(uint32_t)));
        /* the out-of-order path keeps these counts zero at rest (the pass clears what it consumes) */
        CU(c, cudaMemsetAsync(s.d_depth32, 0, a_pos * sizeof(uint32_t), ss));
        s.depth32_cap = a_pos;
    }
    if (c->ptr_mode == CG_PTR_HOST) {
        int rc = grow(c, &s.d_gc, &s.gc_cap, (size_t)need_windows, 0, &c->hw_gc);
        if (rc) return rc;
    }
    /* the slot's previous tenant (two chromosomes ago) must be done with what we overwrite */
    /* HOST mode writes the slot's own upload buffers (d_reads, d_gc from the copy stream; d_runs and the staging
     * arena from both genotype streams); DEVICE mode writes none of them and needs only the slot stream ordered
     * behind the previous tenant's genotype kernels — 3 waits per chromosome instead of 7 on a host that is the
     * limit of the resident leg */
    const bool uploads = c->ptr_mode == CG_PTR_HOST;
    if (s.used && uploads) CU(c, cudaStreamWaitEvent(c->s_copy[si], s.ev_pass, 0));   /* d_reads */
    for (int t = 0; t < 2; t++) {
        if (!s.geno_queued[t]) continue;
        if (uploads) CU(c, cudaStreamWaitEvent(c->s_copy[si], s.ev_geno[t], 0)); /* d_gc */
        CU(c, cudaStreamWaitEvent(ss, s.ev_geno[t], 0));                         /* d_depth, d_counts, d_expected */
        /* d_run_idx is rebuilt on s_geno[0] in either mode: behind the previous tenant's duplication kernel */
        if (uploads || t == 1) CU(c, cudaStreamWaitEvent(s.s_geno[t ^ 1], s.ev_geno[t], 0));
    }
    if (s.wait_reset) { CU(c, cudaStreamWaitEvent(ss, c->ev_reset, 0)); s.wait_reset = false; }
    if (c->profiling && !getenv("CG_PROF_NOCHAIN")) {   /* time kernels alone: chain this slot behind the other one */
        cg_slot& o = c->slot[(si + c->n_slot - 1) % c->n_slot];
        if (o.used) CU(c, cudaStreamWaitEvent(ss, o.ev_pass, 0));
        for (int t = 0; t < 2; t++) if (o.geno_queued[t]) CU(c, cudaStreamWaitEvent(ss, o.ev_geno[t], 0));
    }
    if (c->ptr_mode == CG_PTR_HOST) {
        CU(c, cudaMemcpyAsync(s.d_gc, gc_profile, (size_t)need_windows, cudaMemcpyHostToDevice, c->s_copy[si]));
        CU(c, cudaEventRecord(s.ev_copy, c->s_copy[si]));
        s.gc_ptr = s.d_gc;
    } else {
        s.gc_ptr = gc_profile;
    }
    s.chr_index = chr_index; s.chrom_len = chrom_len; s.n_chunks = n_chunks;
    s.d_row = c->d_stats + ((size_t)c->bank * CG_MAX_CHROMS + (size_t)chr_index) * CG_STATS_WORDS;
    c->dirty = true;
    s.reads_ptr = s.d_reads;
    s.n_runs = 0; s.has_map = 0; s.runs_ptr = nullptr;
    s.arena_used = 0;
    s.used = true; s.geno_queued[0] = s.geno_queued[1] = false; s.runs_uploaded = false;
    c->cur = si; c->n_begun++;
    c->mq_threshold = mq_threshold;
    c->n_reads = 0;
    c->in_chrom = true; c->finished = false;
    if (chr_index + 1 > c->n_rows) c->n_rows = chr_index + 1;
    {
        prof_scope ps(c, c->s_slot[c->cur], 0, 1);
        CU(c, cgk_chrom_init(c->s_slot[c->cur], s.d_counts, n_chunks, s.d_row, s.d_ctl));
    }
    return CG_OK;
}

int cg_push_reads(cg_ctx* c, const cg_read* reads, size_t n)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom || c->finished) return fail(c, CG_EINVAL, "cg_push_reads outside cg_chrom_begin/cg_chrom_finish");
    if (n == 0) return CG_OK;
    if (!reads) return fail(c, CG_EINVAL, "null reads");
    if ((unsigned long long)c->n_reads + n >= 0x7fffffffull) return fail(c, CG_EINVAL, "more than 2^31-1 records on one chromosome");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    const long long first = c->n_reads;
    if (c->ptr_mode == CG_PTR_HOST) {
        int rc = grow(c, &s.d_reads, &s.reads_cap, (size_t)first + n, (size_t)first, &c->hw_reads);
        if (rc) return rc;
        CU(c, cudaMemcpyAsync(s.d_reads + first, reads, n * sizeof(cg_read), cudaMemcpyHostToDevice, c->s_copy[c->cur]));
        CU(c, cudaEventRecord(s.ev_copy, c->s_copy[c->cur]));
        CU(c, cudaStreamWaitEvent(c->s_slot[c->cur], s.ev_copy, 0));
        s.reads_ptr = s.d_reads;
    } else if (first == 0) {
        s.reads_ptr = reads;                  /* single device batch: used in place */
    } else {
        const bool external = s.reads_ptr != s.d_reads;
        int rc = grow(c, &s.d_reads, &s.reads_cap, (size_t)first + n, external ? 0 : (size_t)first, &c->hw_reads);
        if (rc) return rc;
        if (external) CU(c, cudaMemcpyAsync(s.d_reads, s.reads_ptr, (size_t)first * sizeof(cg_read), cudaMemcpyDeviceToDevice, c->s_slot[c->cur]));
        CU(c, cudaMemcpyAsync(s.d_reads + first, reads, n * sizeof(cg_read), cudaMemcpyDeviceToDevice, c->s_slot[c->cur]));
        s.reads_ptr = s.d_reads;
    }
    c->n_reads = first + (long long)n;
    {
        prof_scope ps(c, c->s_slot[c->cur], 0, 1);
        CU(c, cgk_mark_batch(c->s_slot[c->cur], s.reads_ptr, first, (long long)n, cgk_counts_layout(s.d_counts, s.n_chunks),
                             s.n_chunks, s.chrom_len, c->mq_threshold, s.d_ctl, s.d_row));
    }
    return CG_OK;
}

int cg_chrom_finish(cg_ctx* c, cg_chrom_stats* out)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom || c->finished) return fail(c, CG_EINVAL, "cg_chrom_finish without cg_chrom_begin");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    long long* row = s.d_row;
    if (c->ptr_mode == CG_PTR_HOST) CU(c, cudaStreamWaitEvent(c->s_slot[c->cur], s.ev_copy, 0));   /* GC bytes */
    {
        prof_scope ps(c, c->s_slot[c->cur], 0, 2);
        CU(c, cgk_rare_paths(c->s_slot[c->cur], s.reads_ptr, c->n_reads, s.d_depth32, cgk_counts_layout(s.d_counts, s.n_chunks),
                             s.n_chunks, s.chrom_len, c->mq_threshold, s.d_ctl, c->n_sms));
    }
    {
        /* Passes of consecutive chromosomes may run side by side (CG_PASS_OVERLAP=0 chains them).  At 6 bytes per
         * position (rounds r01a-g) two passes doubled the concurrent write streams and cost HBM efficiency, so
         * they were chained; at 2 bytes per position the overlap hides each launch's ramp and tail instead:
         * C2 sample 1.25 -> 1.20 ms (measured r01h). */
        cg_slot& o = c->slot[(c->cur + c->n_slot - 1) % c->n_slot];
        static const bool overlap = !(getenv("CG_PASS_OVERLAP") && atoi(getenv("CG_PASS_OVERLAP")) == 0);
        if (o.used && !overlap) CU(c, cudaStreamWaitEvent(c->s_slot[c->cur], o.ev_pass, 0));
    }
    {
        prof_scope ps(c, c->s_slot[c->cur], 1, 1);
        cg_pass_args a;
        a.reads = s.reads_ptr; a.n_reads = c->n_reads; a.k = cgk_counts_layout(s.d_counts, s.n_chunks);
        a.gc = s.gc_ptr;
        a.chrom_len = s.chrom_len; a.n_chunks = s.n_chunks; a.mq_threshold = c->mq_threshold;
        a.depth = s.d_depth; a.depth32 = s.d_depth32; a.ctl = s.d_ctl;
        a.stats = row; a.expected = s.d_expected; a.sticky = c->d_sticky;
        /* asynchronous HOST mode hands over page-locked pointers: the kernel writes the row there itself */
        a.host_row = (c->ptr_mode == CG_PTR_HOST && c->async && out) ? reinterpret_cast<long long*>(out) : nullptr;
        CU(c, cgk_depth_pass(c->s_slot[c->cur], a, c->n_sms));
    }
    CU(c, cudaEventRecord(s.ev_pass, c->s_slot[c->cur]));
    c->finished = true;
    if (c->ptr_mode == CG_PTR_DEVICE) {
        if (out) CU(c, cudaMemcpyAsync(out, row, sizeof(cg_chrom_stats), cudaMemcpyDeviceToDevice, c->s_slot[c->cur]));
        return CG_OK;
    }
    if (c->async) return CG_OK;           /* `out` is page-locked, written by the pass, valid after cg_sync */
    CU(c, cudaMemcpyAsync(c->h_stats, row, sizeof(cg_chrom_stats), cudaMemcpyDeviceToHost, c->s_slot[c->cur]));
    CU(c, cudaStreamSynchronize(c->s_slot[c->cur]));
    if (out) memcpy(out, c->h_stats, sizeof(cg_chrom_stats));
    return report_sticky(c, "cg_chrom_finish");
}

int cg_get_expected(cg_ctx* c, float expected[CG_GC_BINS])
{
    if (!c || !expected) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_get_expected before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    CU(c, cudaMemcpyAsync(expected, s.d_expected, sizeof(float) * CG_GC_BINS,
                          c->ptr_mode == CG_PTR_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, c->s_slot[c->cur]));
    if (blocking(c)) CU(c, cudaStreamSynchronize(c->s_slot[c->cur]));
    return CG_OK;
}

int cg_set_expected(cg_ctx* c, const float expected[CG_GC_BINS])
{
    if (!c || !expected) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_set_expected before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    CU(c, cudaMemcpyAsync(s.d_expected, expected, sizeof(float) * CG_GC_BINS,
                          c->ptr_mode == CG_PTR_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, c->s_slot[c->cur]));
    CU(c, cudaEventRecord(s.ev_pass, c->s_slot[c->cur]));        /* the genotype stream waits for this */
    return CG_OK;
}

/* ------------------------------------------------------------------------- */
int cg_set_mappability_runs(cg_ctx* c, const cg_maprun* runs, size_t n, int enabled)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom) return fail(c, CG_EINVAL, "cg_set_mappability outside a chromosome");
    if (n > 0x7fffffffull) return fail(c, CG_EINVAL, "too many mappability runs");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    s.has_map = enabled ? 1 : 0;
    s.n_runs = (int)n;
    if (n == 0) { s.runs_ptr = nullptr; return CG_OK; }
    if (!runs) return fail(c, CG_EINVAL, "null runs");
    int rc = grow(c, &s.d_run_idx, &s.run_idx_cap, (size_t)cgk_run_index_words(s.chrom_len), 0, &c->hw_run_idx);
    if (rc) return rc;
    if (c->ptr_mode == CG_PTR_DEVICE) {
        s.runs_ptr = runs;
    } else {
        rc = grow(c, &s.d_runs, &s.runs_cap, n, 0, &c->hw_runs);
        if (rc) return rc;
        /* on a genotype stream: behind the slot's previous tenant, which may still read d_runs */
        CU(c, cudaMemcpyAsync(s.d_runs, runs, n * sizeof(cg_maprun), cudaMemcpyHostToDevice, s.s_geno[0]));
        s.runs_ptr = s.d_runs;
    }
    {
        prof_scope ps(c, s.s_geno[0], 2, 1);
        CU(c, cgk_run_index(s.s_geno[0], s.runs_ptr, (int)n, s.d_run_idx, s.chrom_len));
    }
    CU(c, cudaEventRecord(s.ev_runs, s.s_geno[0]));
    s.runs_uploaded = true;
    return CG_OK;
}

int cg_set_mappability(cg_ctx* c, const cg_mapiv* rows, size_t n)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom) return fail(c, CG_EINVAL, "cg_set_mappability outside a chromosome");
    if (n && !rows) return fail(c, CG_EINVAL, "null rows");
    std::vector<cg_maprun> v;
    cg_flatten_rows(rows, n, c->slot[c->cur].chrom_len, v);
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

/* HOST-mode staging for one cg_genotype call: room for n CNVs and n calls in the slot's arena.
 * The arena is rewound by cg_chrom_begin; when it has to grow, everything queued is drained
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
     * kernel reads 8 bytes and writes 56 per CNV across PCIe itself, which is cheaper than two copy calls per
     * launch on a host that issues a chromosome every 50 us. */
    const bool staged = c->ptr_mode == CG_PTR_HOST && !c->async;
    if (staged) {
        cg_cnv* a_in = nullptr;
        int rc = arena_take(c, s, n, &a_in, &d_out);
        if (rc) return rc;
        CU(c, cudaMemcpyAsync(a_in, cnvs, n * sizeof(cg_cnv), cudaMemcpyHostToDevice, st));
        d_in = a_in;
    }
    CU(c, cudaStreamWaitEvent(st, s.ev_pass, 0));              /* depth, chunk sums and expected table of this slot */
    if (t == 1 && s.runs_uploaded) CU(c, cudaStreamWaitEvent(st, s.ev_runs, 0));
    {
        prof_scope ps(c, st, 2, 1);
        CU(c, cgk_genotype(st, d_in, (int)n, t, s.d_depth, cgk_counts_layout(s.d_counts, s.n_chunks), s.gc_ptr, s.d_expected, s.runs_ptr, s.d_run_idx, s.n_runs, s.has_map,
                           s.chrom_len, c->d_sticky, d_out, c->n_sms));
    }
    if (staged) CU(c, cudaMemcpyAsync(out, d_out, n * sizeof(cg_call), cudaMemcpyDeviceToHost, st));
    CU(c, cudaEventRecord(s.ev_geno[t], st));
    s.geno_queued[t] = true;
    if (blocking(c)) {
        CU(c, cudaStreamSynchronize(st));
        return report_sticky(c, "cg_genotype");
    }
    return CG_OK;
}

int cg_run_chromosomes(cg_ctx* c, const cg_chrom_job* jobs, size_t n_jobs)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (n_jobs && !jobs) return fail(c, CG_EINVAL, "null jobs");
    const size_t BATCH = (size_t)1 << 22;
    int rc;
    for (size_t k = 0; k < n_jobs; k++) {
        const cg_chrom_job& j = jobs[k];
        if ((rc = cg_chrom_begin(c, j.chr_index, j.chrom_len, j.gc, j.n_windows, j.mq_threshold)) != CG_OK) return rc;
        for (size_t i = 0; i < j.n_reads; i += BATCH)
            if ((rc = cg_push_reads(c, j.reads + i, j.n_reads - i < BATCH ? (size_t)(j.n_reads - i) : BATCH)) != CG_OK) return rc;
        if ((rc = cg_chrom_finish(c, j.stats)) != CG_OK) return rc;
        if (j.use_mappability && (rc = cg_set_mappability_runs(c, j.runs, (size_t)j.n_runs, 1)) != CG_OK) return rc;
        if (j.n_del && (rc = cg_genotype(c, j.dels, (size_t)j.n_del, 0, j.del_calls)) != CG_OK) return rc;
        if (j.n_dup && (rc = cg_genotype(c, j.dups, (size_t)j.n_dup, 1, j.dup_calls)) != CG_OK) return rc;
    }
    return CG_OK;
}

/* ------------------------------------------------------------------------- */
int cg_download_depth(cg_ctx* c, int16_t* out)
{
    if (!c || !out) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_download_depth before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    CU(c, cudaMemcpyAsync(out, s.d_depth, (size_t)s.chrom_len * sizeof(int16_t), cudaMemcpyDeviceToHost, c->s_slot[c->cur]));
    CU(c, cudaStreamSynchronize(c->s_slot[c->cur]));
    return CG_OK;
}

int cg_download_prefix(cg_ctx* c, int32_t* out)
{
    if (!c || !out) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_download_prefix before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    cg_slot& s = c->slot[c->cur];
    cudaStream_t st = c->s_slot[c->cur];
    /* no prefix array exists on the product path: rebuilt here from the depth array and the chunk sums */
    int32_t* d_prefix = nullptr;
    CU(c, cudaMalloc(&d_prefix, (size_t)s.n_chunks * CG_CHUNK * sizeof(int32_t)));
    cudaError_t e = cgk_prefix_from_depth(st, s.d_depth, cgk_counts_layout(s.d_counts, s.n_chunks), s.n_chunks, d_prefix);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_prefix, (size_t)s.chrom_len * sizeof(int32_t), cudaMemcpyDeviceToHost, st);
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

int cg_stats_table(cg_ctx* c, void** device_ptr, int32_t* n_rows)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (device_ptr) *device_ptr = c->d_stats + (size_t)c->bank * CG_MAX_CHROMS * CG_STATS_WORDS;
    if (n_rows) *n_rows = c->n_rows;
    return CG_OK;
}

int cg_stats_reset(cg_ctx* c)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    CU(c, cudaSetDevice(c->device));
    int rc = cg_join(c);                      /* rows may still be written or copied by queued work */
    if (rc) return rc;
    /* the whole bank: an in-place collective also fills the rows of chromosomes this context never saw */
    CU(c, cudaMemsetAsync(c->d_stats + (size_t)c->bank * CG_MAX_CHROMS * CG_STATS_WORDS, 0,
                          sizeof(long long) * CG_STATS_WORDS * CG_MAX_CHROMS, c->s_main));
    CU(c, cudaEventRecord(c->ev_reset, c->s_main));
    for (int i = 0; i < CG_NSLOT; i++) c->slot[i].wait_reset = true;
    c->n_rows = 0;
    return CG_OK;
}

/* Two statistics tables so that the collective over sample k's table (queued on cg_stream after
 * cg_join) can be in flight while the chromosomes of sample k+1 fill the other one.  Selecting a
 * bank clears it on a side stream, ordered behind everything that was queued on cg_stream when the
 * bank was left (the collective that read it); the chromosomes begun afterwards wait for that
 * clear only, never for the other bank's collective. */
int cg_stats_bank(cg_ctx* c, int bank)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (bank != 0 && bank != 1) return fail(c, CG_EINVAL, "bank %d must be 0 or 1", bank);
    if (c->in_chrom && !c->finished) return fail(c, CG_EINVAL, "cg_stats_bank inside a chromosome");
    if (bank == c->bank) return CG_OK;
    CU(c, cudaSetDevice(c->device));
    if (c->dirty) { int rc = cg_join(c); if (rc) return rc; }      /* the table being left is complete behind this point */
    CU(c, cudaEventRecord(c->ev_left[c->bank], c->s_main));
    c->bank_left[c->bank] = true;
    if (c->bank_left[bank]) CU(c, cudaStreamWaitEvent(c->s_bank, c->ev_left[bank], 0));
    if (c->bank_left[bank])
        CU(c, cudaMemsetAsync(c->d_stats + (size_t)bank * CG_MAX_CHROMS * CG_STATS_WORDS, 0,
                              sizeof(long long) * CG_STATS_WORDS * CG_MAX_CHROMS, c->s_bank));
    CU(c, cudaEventRecord(c->ev_reset, c->s_bank));
    for (int i = 0; i < CG_NSLOT; i++) c->slot[i].wait_reset = true;
    c->bank = bank; c->n_rows = 0;
    return CG_OK;
}

int cg_stats_row(cg_ctx* c, int chr_index, cg_chrom_stats* out)
{
    if (!c || !out) return fail(c, CG_EINVAL, "null argument");
    if (chr_index < 0 || chr_index >= CG_MAX_CHROMS) return fail(c, CG_EINVAL, "chr_index %d out of range", chr_index);
    CU(c, cudaSetDevice(c->device));
    int rc = sync_all(c);
    if (rc) return rc;
    CU(c, cudaMemcpyAsync(c->h_stats, c->d_stats + ((size_t)c->bank * CG_MAX_CHROMS + (size_t)chr_index) * CG_STATS_WORDS, sizeof(cg_chrom_stats),
                          cudaMemcpyDeviceToHost, c->s_main));
    CU(c, cudaStreamSynchronize(c->s_main));
    memcpy(out, c->h_stats, sizeof(cg_chrom_stats));
    return CG_OK;
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
    if ((size_t)ref_len > r.idx_cap) {
        CU(c, cudaStreamSynchronize(r.st));
        cudaFree(r.d_keys_a); cudaFree(r.d_keys_b); cudaFree(r.d_vals_a); cudaFree(r.d_vals_b);
        r.d_keys_a = r.d_keys_b = nullptr; r.d_vals_a = r.d_vals_b = nullptr; r.idx_cap = 0;
        const size_t n = (size_t)ref_len + 1024;
        CU(c, cudaMalloc(&r.d_keys_a, n * 4)); CU(c, cudaMalloc(&r.d_keys_b, n * 4));
        CU(c, cudaMalloc(&r.d_vals_a, n * 4)); CU(c, cudaMalloc(&r.d_vals_b, n * 4));
        r.idx_cap = n;
    }
    const size_t temp = cgk_sr_sort_temp_bytes(ref_len > 0 ? ref_len : 1);
    if (temp > r.temp_cap) {
        CU(c, cudaStreamSynchronize(r.st));
        cudaFree(r.d_temp); r.d_temp = nullptr; r.temp_cap = 0;
        CU(c, cudaMalloc(&r.d_temp, temp + 1024));
        r.temp_cap = temp + 1024;
    }
    CU(c, cudaMemcpyAsync(r.d_ref, ref_seq, (size_t)ref_len, cudaMemcpyHostToDevice, r.st));
    CU(c, cudaMemsetAsync(r.d_ref + ref_len, 0, 1, r.st));
    CU(c, cudaMemsetAsync(r.d_n_rows, 0, 16, r.st));
    {
        prof_scope ps(c, r.st, 3, 2);
        CU(c, cgk_sr_index(r.st, r.d_ref, ref_len, r.d_keys_a, r.d_keys_b, r.d_vals_a, r.d_vals_b, r.d_temp, r.temp_cap, r.d_lo, r.d_hi, c->n_sms));
    }
    CU(c, cudaStreamSynchronize(r.st));            /* ref_seq may be pageable and may die at return */
    r.ref_len = ref_len;
    r.have_ref = true;
    r.chrom_len = c->slot[c->cur].chrom_len;
    r.mq_threshold = c->mq_threshold;
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
    if ((rc = sr_grow(c, &r.d_halves, &r.halves_cap, n))) return rc;
    if ((rc = sr_grow(c, &r.d_bases, &r.bases_cap, n_bases + 1))) return rc;
    CU(c, cudaMemcpyAsync(r.d_halves, halves, n * sizeof(cg_half), cudaMemcpyHostToDevice, r.st));
    CU(c, cudaMemcpyAsync(r.d_bases, bases, n_bases, cudaMemcpyHostToDevice, r.st));
    {
        prof_scope ps(c, r.st, 3, 1);
        CU(c, cgk_sr_search(r.st, r.d_ref, r.ref_len, r.d_vals_b, r.d_lo, r.d_hi, r.d_halves, (long long)n, r.d_bases, r.chrom_len,
                            r.mq_threshold, r.d_rows, r.d_n_rows, r.rows_cap, r.d_n_map, c->n_sms));
    }
    CU(c, cudaStreamSynchronize(r.st));            /* the staging buffers are reused by the next batch */
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
