/* conga_gpu.cu — context management and the C-ABI of libconga_gpu.so
 * (include/conga_gpu.h).  Host code only; the kernels live in depth_pass.cu and
 * genotype.cu.  There is no CPU implementation of any stage in this library:
 * without a CUDA device cg_create fails and nothing else can be called. */
#include "cg_internal.h"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <queue>
#include <vector>

#define CG_MAX_CHROMS 4096            /* rows of the statistics table */


struct cg_evpair { cudaEvent_t a, b; int slot; };

struct cg_ctx {
    int device = 0;
    int n_sms = 148;
    cudaStream_t stream = nullptr;
    int ptr_mode = CG_PTR_HOST;
    char err[512] = {0};

    /* current chromosome */
    bool in_chrom = false, finished = false;
    int chr_index = -1;
    int32_t chrom_len = 0;
    int n_tiles = 0;
    int mq_threshold = -1;
    long long n_reads = 0;
    const cg_read* reads_ptr = nullptr;       /* device: own buffer or the caller's (DEVICE mode) */
    const int8_t* gc_ptr = nullptr;
    int n_rows = 0;

    /* grow-only device buffers */
    cg_read* d_reads = nullptr;       size_t reads_cap = 0;      /* records */
    int8_t* d_gc = nullptr;           size_t gc_cap = 0;
    int32_t* d_tile_lo = nullptr;     int32_t* d_chunk_cnt = nullptr; uint32_t* d_tile_carry = nullptr; size_t tiles_cap = 0;
    int16_t* d_depth = nullptr;       int32_t* d_prefix = nullptr; uint32_t* d_depth32 = nullptr; size_t pos_cap = 0;
    cg_maprun* d_runs = nullptr;      size_t runs_cap = 0;
    cg_cnv* d_cnvs = nullptr;         cg_call* d_calls = nullptr; size_t cnv_cap = 0;
    const cg_maprun* runs_ptr = nullptr; int n_runs = 0; int has_map = 0;

    long long* d_stats = nullptr;             /* [CG_MAX_CHROMS][204] */
    float* d_expected = nullptr;              /* [101] */
    unsigned int* d_flags = nullptr;          /* control words CG_W_*: flags, tile counter, done counter */
    unsigned int* d_work = nullptr;      /* monotonic work counter of the genotype kernel */
    unsigned int work_base = 0;

    /* pinned staging */
    cg_chrom_stats* h_stats = nullptr;
    unsigned int* h_flags = nullptr;

    /* profiling */
    bool profiling = false;
    std::vector<cg_evpair> ev_live, ev_free;
    double ms_acc[CG_NSLOTS] = {0, 0, 0, 0};
    long long launches_acc[CG_NSLOTS] = {0, 0, 0, 0};
    long long launch_count = 0;
};

static char g_err[512] = "no error";

static int fail(cg_ctx* c, int code, const char* fmt, ...)
{
    char* dst = c ? c->err : g_err;
    va_list ap; va_start(ap, fmt);
    vsnprintf(dst, 512, fmt, ap);
    va_end(ap);
    if (c) memcpy(g_err, c->err, 512);
    return code;
}

#define CU(c, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return fail((c), e_ == cudaErrorMemoryAllocation ? CG_ENOMEM : CG_ECUDA, "%s: %s", #call, cudaGetErrorString(e_)); } while (0)

template <typename T>
static int grow(cg_ctx* c, T** p, size_t* cap, size_t need, size_t keep_elems = 0)
{
    if (need <= *cap) return CG_OK;
    size_t ncap = std::max(need, *cap + *cap / 2);
    T* np = nullptr;
    CU(c, cudaMalloc(&np, ncap * sizeof(T)));
    if (keep_elems && *p) CU(c, cudaMemcpyAsync(np, *p, keep_elems * sizeof(T), cudaMemcpyDeviceToDevice, c->stream));
    if (*p) { CU(c, cudaStreamSynchronize(c->stream)); cudaFree(*p); }
    *p = np; *cap = ncap;
    return CG_OK;
}

/* profiling brackets */
struct prof_scope {
    cg_ctx* c; cg_evpair ev; bool on; int n;
    prof_scope(cg_ctx* c_, int slot, int n_launches) : c(c_), on(c_->profiling), n(n_launches) {
        c->launch_count += n_launches;
        if (!on) return;
        if (!c->ev_free.empty()) { ev = c->ev_free.back(); c->ev_free.pop_back(); }
        else { cudaEventCreate(&ev.a); cudaEventCreate(&ev.b); }
        ev.slot = slot;
        c->launches_acc[slot] += n_launches;
        cudaEventRecord(ev.a, c->stream);
    }
    ~prof_scope() { if (on) { cudaEventRecord(ev.b, c->stream); c->ev_live.push_back(ev); } }
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
    cg_ctx* c = new cg_ctx();
    c->device = device;
    c->n_sms = prop.multiProcessorCount;
    bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess
        && cudaMalloc(&c->d_stats, sizeof(long long) * CG_STATS_WORDS * CG_MAX_CHROMS) == cudaSuccess
        && cudaMemset(c->d_stats, 0, sizeof(long long) * CG_STATS_WORDS * CG_MAX_CHROMS) == cudaSuccess
        && cudaMalloc(&c->d_expected, sizeof(float) * 128) == cudaSuccess
        && cudaMemset(c->d_expected, 0, sizeof(float) * 128) == cudaSuccess
        && cudaMalloc(&c->d_flags, 16) == cudaSuccess
        && cudaMemset(c->d_flags, 0, 16) == cudaSuccess
        && cudaMalloc(&c->d_work, 16) == cudaSuccess
        && cudaMemset(c->d_work, 0, 16) == cudaSuccess
        && cudaMallocHost(&c->h_stats, sizeof(cg_chrom_stats)) == cudaSuccess
        && cudaMallocHost(&c->h_flags, 16) == cudaSuccess;
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
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (auto& ev : c->ev_live) { cudaEventDestroy(ev.a); cudaEventDestroy(ev.b); }
    for (auto& ev : c->ev_free) { cudaEventDestroy(ev.a); cudaEventDestroy(ev.b); }
    cudaFree(c->d_reads); cudaFree(c->d_gc); cudaFree(c->d_tile_lo); cudaFree(c->d_chunk_cnt); cudaFree(c->d_tile_carry);
    cudaFree(c->d_depth); cudaFree(c->d_prefix); cudaFree(c->d_depth32); cudaFree(c->d_runs);
    cudaFree(c->d_cnvs); cudaFree(c->d_calls); cudaFree(c->d_stats); cudaFree(c->d_expected);
    cudaFree(c->d_flags); cudaFree(c->d_work);
    if (c->h_stats) cudaFreeHost(c->h_stats);
    if (c->h_flags) cudaFreeHost(c->h_flags);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int cg_set_pointer_mode(cg_ctx* c, int mode)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (mode != CG_PTR_HOST && mode != CG_PTR_DEVICE) return fail(c, CG_EINVAL, "bad pointer mode %d", mode);
    c->ptr_mode = mode;
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

void* cg_stream(cg_ctx* c) { return c ? (void*)c->stream : nullptr; }
int64_t cg_launch_count(cg_ctx* c) { return c ? c->launch_count : 0; }

int cg_sync(cg_ctx* c)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    CU(c, cudaStreamSynchronize(c->stream));
    prof_collect(c);
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
    CU(c, cudaStreamSynchronize(c->stream));
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
    if (chrom_len > INT32_MAX - CG_TILE) return fail(c, CG_EINVAL, "chrom_len %d too large", chrom_len);
    const int need_windows = (chrom_len + 99) / 100;
    if (!gc_profile || n_windows < need_windows)
        return fail(c, CG_EINVAL, "gc_profile needs >= %d windows for length %d (got %d)", need_windows, chrom_len, n_windows);
    CU(c, cudaSetDevice(c->device));

    const int n_tiles = (chrom_len + CG_TILE - 1) / CG_TILE;
    const size_t n_pos = (size_t)n_tiles * CG_TILE;
    if ((size_t)n_tiles + 1 > c->tiles_cap) {
        CU(c, cudaStreamSynchronize(c->stream));
        cudaFree(c->d_tile_lo); cudaFree(c->d_chunk_cnt); cudaFree(c->d_tile_carry);
        c->d_tile_lo = nullptr; c->d_chunk_cnt = nullptr; c->d_tile_carry = nullptr; c->tiles_cap = 0;
        CU(c, cudaMalloc(&c->d_tile_lo, ((size_t)n_tiles + 1) * sizeof(int32_t)));
        CU(c, cudaMalloc(&c->d_chunk_cnt, (size_t)n_tiles * CG_CHUNKS_PER_TILE * sizeof(int32_t)));
        CU(c, cudaMalloc(&c->d_tile_carry, ((size_t)n_tiles + 1) * sizeof(uint32_t)));
        c->tiles_cap = (size_t)n_tiles + 1;
    }
    if (n_pos > c->pos_cap) {
        CU(c, cudaStreamSynchronize(c->stream));
        cudaFree(c->d_depth); cudaFree(c->d_prefix); cudaFree(c->d_depth32);
        c->d_depth = nullptr; c->d_prefix = nullptr; c->d_depth32 = nullptr; c->pos_cap = 0;
        CU(c, cudaMalloc(&c->d_depth, n_pos * sizeof(int16_t)));
        CU(c, cudaMalloc(&c->d_prefix, n_pos * sizeof(int32_t)));
        CU(c, cudaMalloc(&c->d_depth32, n_pos * sizeof(uint32_t)));
        /* the out-of-order path keeps these counts zero at rest (the pass clears what it consumes) */
        CU(c, cudaMemsetAsync(c->d_depth32, 0, n_pos * sizeof(uint32_t), c->stream));
        c->pos_cap = n_pos;
    }
    if (c->ptr_mode == CG_PTR_HOST) {
        int rc = grow(c, &c->d_gc, &c->gc_cap, (size_t)need_windows);
        if (rc) return rc;
        CU(c, cudaMemcpyAsync(c->d_gc, gc_profile, (size_t)need_windows, cudaMemcpyHostToDevice, c->stream));
        c->gc_ptr = c->d_gc;
    } else {
        c->gc_ptr = gc_profile;
    }
    c->chr_index = chr_index; c->chrom_len = chrom_len; c->n_tiles = n_tiles;
    c->mq_threshold = mq_threshold;
    c->n_reads = 0; c->reads_ptr = c->d_reads;
    c->in_chrom = true; c->finished = false;
    c->n_runs = 0; c->has_map = 0; c->runs_ptr = nullptr;
    if (chr_index + 1 > c->n_rows) c->n_rows = chr_index + 1;
    {
        prof_scope ps(c, 0, 1);
        CU(c, cgk_chrom_init(c->stream, c->d_chunk_cnt, n_tiles * CG_CHUNKS_PER_TILE,
                             c->d_stats + (size_t)chr_index * CG_STATS_WORDS, c->d_flags));
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
    const long long first = c->n_reads;
    if (c->ptr_mode == CG_PTR_HOST) {
        int rc = grow(c, &c->d_reads, &c->reads_cap, (size_t)first + n, (size_t)first);
        if (rc) return rc;
        CU(c, cudaMemcpyAsync(c->d_reads + first, reads, n * sizeof(cg_read), cudaMemcpyHostToDevice, c->stream));
        c->reads_ptr = c->d_reads;
    } else if (first == 0) {
        c->reads_ptr = reads;                 /* single device batch: used in place */
    } else {
        const bool external = c->reads_ptr != c->d_reads;
        int rc = grow(c, &c->d_reads, &c->reads_cap, (size_t)first + n, external ? 0 : (size_t)first);
        if (rc) return rc;
        if (external) CU(c, cudaMemcpyAsync(c->d_reads, c->reads_ptr, (size_t)first * sizeof(cg_read), cudaMemcpyDeviceToDevice, c->stream));
        CU(c, cudaMemcpyAsync(c->d_reads + first, reads, n * sizeof(cg_read), cudaMemcpyDeviceToDevice, c->stream));
        c->reads_ptr = c->d_reads;
    }
    c->n_reads = first + (long long)n;
    {
        prof_scope ps(c, 0, 1);
        CU(c, cgk_mark_batch(c->stream, c->reads_ptr, first, (long long)n, c->d_tile_lo, c->d_chunk_cnt, c->n_tiles,
                             c->chrom_len, c->mq_threshold, c->d_flags,
                             c->d_stats + (size_t)c->chr_index * CG_STATS_WORDS));
    }
    return CG_OK;
}

static int check_flags_host(cg_ctx* c, const char* what)
{
    const unsigned f = c->h_flags[0];
    if (f & CG_FLAG_RANGE) return fail(c, CG_ERANGE, "%s: a record or interval lies outside [0, %d); it was clipped", what, c->chrom_len);
    return CG_OK;
}

int cg_chrom_finish(cg_ctx* c, cg_chrom_stats* out)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom || c->finished) return fail(c, CG_EINVAL, "cg_chrom_finish without cg_chrom_begin");
    CU(c, cudaSetDevice(c->device));
    long long* row = c->d_stats + (size_t)c->chr_index * CG_STATS_WORDS;
    {
        prof_scope ps(c, 0, 3);
        CU(c, cgk_rare_paths(c->stream, c->reads_ptr, c->n_reads, c->d_depth32, c->d_chunk_cnt, c->n_tiles,
                             c->chrom_len, c->mq_threshold, c->d_flags, c->n_sms));
        CU(c, cgk_tile_scan(c->stream, c->reads_ptr, c->n_reads, c->d_tile_lo, c->d_chunk_cnt, c->d_tile_carry,
                            c->n_tiles, row));
    }
    {
        prof_scope ps(c, 1, 1);
        cg_pass_args a;
        a.reads = c->reads_ptr; a.n_reads = c->n_reads; a.tile_lo = c->d_tile_lo; a.chunk_cnt = c->d_chunk_cnt;
        a.tile_carry = c->d_tile_carry; a.gc = c->gc_ptr;
        a.chrom_len = c->chrom_len; a.n_tiles = c->n_tiles; a.mq_threshold = c->mq_threshold;
        a.depth = c->d_depth; a.prefix = c->d_prefix; a.depth32 = c->d_depth32; a.ctl = c->d_flags;
        a.stats = row; a.expected = c->d_expected;
        CU(c, cgk_depth_pass(c->stream, a, c->n_sms));
    }
    c->finished = true;
    if (c->ptr_mode == CG_PTR_HOST) {
        CU(c, cudaMemcpyAsync(c->h_stats, row, sizeof(cg_chrom_stats), cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaMemcpyAsync(c->h_flags, c->d_flags, 4, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
        if (out) memcpy(out, c->h_stats, sizeof(cg_chrom_stats));
        return check_flags_host(c, "cg_chrom_finish");
    }
    if (out) CU(c, cudaMemcpyAsync(out, row, sizeof(cg_chrom_stats), cudaMemcpyDeviceToDevice, c->stream));
    return CG_OK;
}

int cg_get_expected(cg_ctx* c, float expected[CG_GC_BINS])
{
    if (!c || !expected) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_get_expected before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(expected, c->d_expected, sizeof(float) * CG_GC_BINS,
                          c->ptr_mode == CG_PTR_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, c->stream));
    if (c->ptr_mode == CG_PTR_HOST) CU(c, cudaStreamSynchronize(c->stream));
    return CG_OK;
}

int cg_set_expected(cg_ctx* c, const float expected[CG_GC_BINS])
{
    if (!c || !expected) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_set_expected before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(c->d_expected, expected, sizeof(float) * CG_GC_BINS,
                          c->ptr_mode == CG_PTR_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, c->stream));
    return CG_OK;
}

/* ------------------------------------------------------------------------- */
/* Mappability rows -> disjoint runs.  svs.c:350-371 paints rows in file order,
 * [start, end] inclusive, later rows overwriting earlier ones; the run list is
 * what mappability[] would hold afterwards (zero gaps omitted), clipped to the
 * chromosome.  Input preparation, like BED parsing: done on the host once. */
static void flatten_rows(const cg_mapiv* rows, size_t n, int32_t chrom_len, std::vector<cg_maprun>& out)
{
    out.clear();
    if (n == 0) return;
    auto emit = [&](int32_t s, int32_t e, float v) {           /* [s,e) */
        if (s < 0) s = 0;
        if (e > chrom_len) e = chrom_len;
        if (e <= s) return;
        if (!out.empty() && out.back().end == s && memcmp(&out.back().val, &v, 4) == 0) { out.back().end = e; return; }
        out.push_back(cg_maprun{s, e, v, 0});
    };
    /* fast path: bedGraph order (starts strictly increasing, ends non-decreasing) */
    bool ordered = true;
    for (size_t i = 1; i < n && ordered; i++)
        ordered = rows[i].start > rows[i - 1].start && rows[i].end >= rows[i - 1].end;
    if (ordered) {
        for (size_t i = 0; i < n; i++) {
            if (rows[i].end < rows[i].start) { ordered = false; break; }
        }
    }
    if (ordered) {
        out.reserve(n);
        for (size_t i = 0; i < n; i++) {
            long long e = (long long)rows[i].end + 1;
            if (i + 1 < n && rows[i + 1].start < e) e = rows[i + 1].start;
            emit(rows[i].start, (int32_t)std::min<long long>(e, chrom_len), rows[i].val);
        }
        return;
    }
    /* general case: sweep over elementary segments; the winner is the latest row covering it */
    std::vector<uint32_t> order;
    order.reserve(n);
    std::vector<long long> cuts;
    cuts.reserve(2 * n);
    for (size_t i = 0; i < n; i++) {
        if (rows[i].end < rows[i].start) continue;
        order.push_back((uint32_t)i);
        cuts.push_back(rows[i].start);
        cuts.push_back((long long)rows[i].end + 1);
    }
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
        return rows[a].start != rows[b].start ? rows[a].start < rows[b].start : a < b; });
    std::sort(cuts.begin(), cuts.end());
    cuts.erase(std::unique(cuts.begin(), cuts.end()), cuts.end());
    std::priority_queue<uint32_t> active;                      /* largest file index on top */
    size_t next = 0;
    for (size_t k = 0; k + 1 < cuts.size(); k++) {
        const long long a = cuts[k], b = cuts[k + 1];
        while (next < order.size() && rows[order[next]].start <= a) active.push(order[next++]);
        while (!active.empty() && (long long)rows[active.top()].end < a) active.pop();
        if (active.empty()) continue;
        const long long lo = std::max<long long>(a, 0), hi = std::min<long long>(b, chrom_len);
        if (hi > lo) emit((int32_t)lo, (int32_t)hi, rows[active.top()].val);
    }
}

int cg_flatten_mappability(const cg_mapiv* rows, size_t n, int32_t chrom_len, cg_maprun** runs_out, size_t* n_out)
{
    if ((!rows && n) || !runs_out || !n_out) return fail(nullptr, CG_EINVAL, "null argument");
    std::vector<cg_maprun> v;
    flatten_rows(rows, n, chrom_len, v);
    cg_maprun* p = (cg_maprun*)malloc((v.size() ? v.size() : 1) * sizeof(cg_maprun));
    if (!p) return fail(nullptr, CG_ENOMEM, "out of host memory");
    if (!v.empty()) memcpy(p, v.data(), v.size() * sizeof(cg_maprun));
    *runs_out = p; *n_out = v.size();
    return CG_OK;
}

void cg_free_runs(cg_maprun* runs) { free(runs); }

int cg_set_mappability_runs(cg_ctx* c, const cg_maprun* runs, size_t n, int enabled)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom) return fail(c, CG_EINVAL, "cg_set_mappability outside a chromosome");
    if (n > 0x7fffffffull) return fail(c, CG_EINVAL, "too many mappability runs");
    CU(c, cudaSetDevice(c->device));
    c->has_map = enabled ? 1 : 0;
    c->n_runs = (int)n;
    if (n == 0) { c->runs_ptr = nullptr; return CG_OK; }
    if (!runs) return fail(c, CG_EINVAL, "null runs");
    if (c->ptr_mode == CG_PTR_DEVICE) { c->runs_ptr = runs; return CG_OK; }
    int rc = grow(c, &c->d_runs, &c->runs_cap, n);
    if (rc) return rc;
    CU(c, cudaMemcpyAsync(c->d_runs, runs, n * sizeof(cg_maprun), cudaMemcpyHostToDevice, c->stream));
    c->runs_ptr = c->d_runs;
    return CG_OK;
}

int cg_set_mappability(cg_ctx* c, const cg_mapiv* rows, size_t n)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->in_chrom) return fail(c, CG_EINVAL, "cg_set_mappability outside a chromosome");
    if (n && !rows) return fail(c, CG_EINVAL, "null rows");
    std::vector<cg_maprun> v;
    flatten_rows(rows, n, c->chrom_len, v);
    const int saved = c->ptr_mode;
    c->ptr_mode = CG_PTR_HOST;                                 /* rows are always host memory */
    int rc = cg_set_mappability_runs(c, v.data(), v.size(), n > 0);
    if (rc == CG_OK && !v.empty()) {
        cudaError_t e = cudaStreamSynchronize(c->stream);      /* v dies at return */
        if (e != cudaSuccess) rc = fail(c, CG_ECUDA, "cudaStreamSynchronize: %s", cudaGetErrorString(e));
    }
    c->ptr_mode = saved;
    return rc;
}

int cg_genotype(cg_ctx* c, const cg_cnv* cnvs, size_t n, int is_dup, cg_call* out)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_genotype before cg_chrom_finish");
    if (n == 0) return CG_OK;
    if (!cnvs || !out) return fail(c, CG_EINVAL, "null argument");
    if (n > 0x3fffffffull) return fail(c, CG_EINVAL, "too many CNVs in one call");
    CU(c, cudaSetDevice(c->device));
    const cg_cnv* d_in = cnvs; cg_call* d_out = out;
    if (c->ptr_mode == CG_PTR_HOST) {
        if (n > c->cnv_cap) {
            CU(c, cudaStreamSynchronize(c->stream));
            cudaFree(c->d_cnvs); cudaFree(c->d_calls); c->d_cnvs = nullptr; c->d_calls = nullptr; c->cnv_cap = 0;
            const size_t cap = std::max<size_t>(n, 4096);
            CU(c, cudaMalloc(&c->d_cnvs, cap * sizeof(cg_cnv)));
            CU(c, cudaMalloc(&c->d_calls, cap * sizeof(cg_call)));
            c->cnv_cap = cap;
        }
        CU(c, cudaMemcpyAsync(c->d_cnvs, cnvs, n * sizeof(cg_cnv), cudaMemcpyHostToDevice, c->stream));
        d_in = c->d_cnvs; d_out = c->d_calls;
    }
    {
        prof_scope ps(c, 2, 1);
        CU(c, cgk_genotype(c->stream, d_in, (int)n, is_dup ? 1 : 0, c->d_prefix, c->gc_ptr, c->d_expected,
                           c->runs_ptr, c->n_runs, c->has_map, c->chrom_len, c->d_flags, d_out, c->d_work, &c->work_base, c->n_sms));
    }
    if (c->ptr_mode == CG_PTR_HOST) {
        CU(c, cudaMemcpyAsync(out, c->d_calls, n * sizeof(cg_call), cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaMemcpyAsync(c->h_flags, c->d_flags, 4, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
        return check_flags_host(c, "cg_genotype");
    }
    return CG_OK;
}

/* ------------------------------------------------------------------------- */
int cg_download_depth(cg_ctx* c, int16_t* out)
{
    if (!c || !out) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_download_depth before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(out, c->d_depth, (size_t)c->chrom_len * sizeof(int16_t), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return CG_OK;
}

int cg_download_prefix(cg_ctx* c, int32_t* out)
{
    if (!c || !out) return fail(c, CG_EINVAL, "null argument");
    if (!c->finished) return fail(c, CG_EINVAL, "cg_download_prefix before cg_chrom_finish");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(out, c->d_prefix, (size_t)c->chrom_len * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return CG_OK;
}

int cg_device_arrays(cg_ctx* c, const int16_t** depth, const int32_t** prefix, const float** expected)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (depth) *depth = c->d_depth;
    if (prefix) *prefix = c->d_prefix;
    if (expected) *expected = c->d_expected;
    return CG_OK;
}

int cg_stats_table(cg_ctx* c, void** device_ptr, int32_t* n_rows)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    if (device_ptr) *device_ptr = c->d_stats;
    if (n_rows) *n_rows = c->n_rows;
    return CG_OK;
}

int cg_stats_reset(cg_ctx* c)
{
    if (!c) return fail(nullptr, CG_EINVAL, "null context");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemsetAsync(c->d_stats, 0, sizeof(long long) * CG_STATS_WORDS * (size_t)CG_MAX_CHROMS, c->stream));
    c->n_rows = 0;
    return CG_OK;
}

int cg_stats_row(cg_ctx* c, int chr_index, cg_chrom_stats* out)
{
    if (!c || !out) return fail(c, CG_EINVAL, "null argument");
    if (chr_index < 0 || chr_index >= CG_MAX_CHROMS) return fail(c, CG_EINVAL, "chr_index %d out of range", chr_index);
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(c->h_stats, c->d_stats + (size_t)chr_index * CG_STATS_WORDS, sizeof(cg_chrom_stats),
                          cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    memcpy(out, c->h_stats, sizeof(cg_chrom_stats));
    return CG_OK;
}

} /* extern "C" */
