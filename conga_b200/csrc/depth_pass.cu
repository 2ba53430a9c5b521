/* depth_pass.cu — the fused read-depth pass on sm_100a, over a SUBMISSION of up to CG_JMAX chromosomes.
 *
 * Replaces, in ONE sweep over the positions of every chromosome of the submission (no intermediate HBM
 * round trip):
 *   K1  read_depth[pos]++ per passing record        bam_data.c:201-217, read_distribution.c:12-18
 *   K2  chromosome sum + 101 GC-bin sums/counts     read_distribution.c:27-46, :49-73
 *   K3  per-CNV depth sums in O(1): NOT a prefix array — the chunk / group / super depth sums left by the
 *       record sweep plus the depth array itself serve likelihood.c:115-121 (genotype.cu, depth_sum_before)
 *
 * Kernels, one launch each per submission (a whole sample through cg_run_chromosomes, or one chromosome
 * through cg_chrom_begin ... cg_chrom_finish):
 *   k_mark         one sweep over the records: sortedness / range flags, first record of every chunk,
 *                  passing records per chunk / group / super and per chromosome
 *   k_depth_pass   persistent warps pull chunks of ALL chromosomes from one counter (no per-chromosome ramp
 *                  and tail); shared-memory histogram -> int16 depth vectors; GC-bin sums per CTA and job
 *   k_finalize     per-CTA partial rows -> cg_chrom_stats row, expected_read_depth[101], flags for the host
 *   k_rare_scatter out-of-order input (per-chromosome entry points only; a sample submission that meets
 *                  unsorted records is re-run through them by the host side, conga_gpu.cu)
 *
 * Data layout in HBM (per slot; job j owns chunks [chunk0, chunk0 + n_chunks] of the slot's chunk space):
 *   reads      cg_read[N] or pos int32[N] + mapq u8[N]   coordinate sorted
 *   chunk_lo   int32[n_space]        first record of every 2048-position chunk
 *   cnt1/2/3   int32[...]            sum of the int16 depths per chunk / 32 chunks / 1024 chunks
 *   depth      int16[n_space*2048]   the reference's `short read_depth[]` (wraps mod 2^16)
 *   partial    int32[CTAs][jobs][202]  what each CTA of the pass saw of each chromosome's GC bins
 *
 * A chunk's depth sum is the number of passing records that fall in it (every read adds exactly 1) unless
 * a position collects >= 32768 reads and the int16 goes negative.  The pass notices that while it bins the
 * records, takes the sums of such a chunk from a scan of its bins instead, and corrects the three count
 * levels itself.  The pass is embarrassingly parallel over chunks: no look-back, no spinning, no block
 * barrier inside the loop, no separate scan kernel, no global atomics except the work counter.
 *
 * Algorithmic bytes per chromosome: 8·N (records; 5·N or 4·N in the compact formats) + 2·L (depth, written
 * once) + L/100 (GC bytes); the pass is HBM-write bound.
 */
#include "cg_internal.h"
#include <cstdlib>

#define FULL 0xffffffffu
/* How a finished chunk leaves shared memory.  Default: ONE bulk copy shared -> global of the warp's 4 KB region
 * (cp.async.bulk.global.shared::cta, SASS UBLKCP.G.S), issued by one lane, with an L2 evict-first policy — the depth
 * array is a 5.8 GB write stream nobody reads before the genotype kernel.  The warp's eight LDS.128 + eight STG.128 per
 * lane and chunk become one instruction, and the copy engine writes whole 4 KB lines.  r02k on B200, pass alone /
 * step, ms per sample:   per-lane st.global.cs.v4 (-DPASS_NO_BULK_STORE): C3 0.993 / 1.193, C2 0.911 / 0.999, C4 1.561 / 2.055;
 * bulk copy without the policy: C3 0.976 / 1.154, C2 0.915 / 0.994, C4 1.529 / 1.997;  with it: C3 0.961 / 1.145, C2 0.899 / 0.977.
 * Two regions per warp (-DPASS_BULK_BUFS=2: bin one while the other is read) need 64 KB per CTA, only two CTAs fit an SM,
 * and C3 falls to 1.168 / 1.378: with one region the warp waits for the copy engine to have READ the region (not for
 * the write) before it clears it, which costs nothing measurable beside 23 other warps. */
#ifndef PASS_NO_BULK_STORE
#define PASS_BULK_STORE
#define PASS_BULK_HINT
#endif
#define NT   CG_PASS_THREADS
#define NWARP (NT / 32)
#define ROWS (CG_CHUNK / 256)         /* rows of 256 positions per chunk */
#define MAXWIN 24                     /* GC windows that can overlap one chunk: 2048/100 + 2 */
#define LVL (CG_GROUP_SHIFT - CG_CHUNK_SHIFT)      /* 5: 32 children per level */

static_assert(CG_SUPER_SHIFT - CG_GROUP_SHIFT == LVL && LVL == 5, "32 children per level, one per lane");
static_assert((1 << CG_CHUNK_SHIFT) == CG_CHUNK && ROWS == 8, "chunk geometry");
static_assert(CG_SUPER_CHUNKS == 1 << (2 * LVL), "a super is 1024 chunks");
static_assert(CG_JMAX <= 32, "one lane / one thread per job in the set-up code");

__device__ __forceinline__ int sext16(unsigned v) { return (int)(short)(v & 0xffffu); }

__device__ __forceinline__ int warp_incl_scan(int v, int lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int o = __shfl_up_sync(FULL, v, d);
        if (lane >= d) v += o;
    }
    return v;
}

/* streaming stores: depth is written once and not re-read by this kernel */
__device__ __forceinline__ void st_cs_v4(void* p, unsigned a, unsigned b, unsigned c, unsigned d)
{
    asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

/* ------------------------------------------------------------------------- */
static inline size_t pad4(size_t n) { return (n + 3) & ~(size_t)3; }

/* chunk space of one job: its chunks plus the bound behind the last one, rounded up to whole supers */
int cgk_job_space(int n_chunks) { return (n_chunks + 1 + CG_SUPER_CHUNKS - 1) / CG_SUPER_CHUNKS * CG_SUPER_CHUNKS; }

size_t cgk_counts_words(int n_space)
{
    return pad4((size_t)n_space) * 2 + pad4((size_t)n_space >> LVL) + pad4((size_t)n_space >> (2 * LVL));
}

cg_counts cgk_counts_layout(int32_t* base, int n_space)
{
    cg_counts k;
    k.cnt1 = base;                                            /* the three zeroed arrays first, chunk_lo last */
    k.cnt2 = k.cnt1 + pad4((size_t)n_space);
    k.cnt3 = k.cnt2 + pad4((size_t)n_space >> LVL);
    k.chunk_lo = k.cnt3 + pad4((size_t)n_space >> (2 * LVL));
    return k;
}

/* the records of a job, whichever format they came in */
struct rec_src { const uint2* aos; const int* pos; const unsigned char* mapq; };

__device__ __forceinline__ void load_rec(const rec_src& r, long long i, int& pos, int& mapq)
{
    if (r.aos) {
        const uint2 v = __ldg(r.aos + i);
        pos = (int)v.x; mapq = (int)((v.y >> 16) & 0xffu);
    } else {
        pos = __ldg(r.pos + i);
        mapq = r.mapq ? (int)__ldg(r.mapq + i) : 255;        /* no mapq plane: the threshold is negative, everything passes */
    }
}
/* The same in two steps for software pipelining: load_raw only ISSUES the loads (nothing reads the result, so an
 * in-order warp does not wait), unpack_raw is called where the record is used.  A load whose result is shifted or
 * masked on the spot is not a prefetch: the warp stalls at the shift until the data is there. */
__device__ __forceinline__ uint2 load_raw(const rec_src& r, long long i)
{
    if (r.aos) return __ldg(r.aos + i);
    uint2 v;
    v.x = (unsigned)__ldg(r.pos + i);
    v.y = r.mapq ? (unsigned)__ldg(r.mapq + i) : 255u;
    return v;
}
__device__ __forceinline__ void unpack_raw(bool aos, uint2 v, int& pos, int& mapq)
{
    pos = (int)v.x;
    mapq = aos ? (int)((v.y >> 16) & 0xffu) : (int)v.y;
}

__device__ __forceinline__ int load_pos(const rec_src& r, long long i) { return r.aos ? (int)__ldg(&r.aos[i].x) : __ldg(r.pos + i); }

/* No kernel prepares a submission.  Every array the kernels accumulate into is ZERO AT REST: the work counters
 * and the per-job control words are cleared by k_finalize when it has read them, its scratch rows likewise; the
 * chunk depth sums are written for every chunk by the pass; the chunk bounds are written for every chunk by the
 * record sweep (the chromosome's last record also closes the chunks behind it) and a chromosome without records
 * has its bounds clamped to [0, 0] by the pass. */

/* ------------------------------------------------------------------------- */
/* one sweep over the records of the submission: sortedness and range checks,  */
/* first record of every chunk, records that pass the mapq threshold           */
__device__ __forceinline__ int chunk_of(int pos, int n_chunks)
{
    int t = pos >> CG_CHUNK_SHIFT;            /* arithmetic shift: negative pos -> negative chunk */
    t = t < 0 ? 0 : t;
    return t >= n_chunks ? n_chunks - 1 : t;
}

/* first lane of a run of equal keys in the warp, and the run's length */
__device__ __forceinline__ bool run_leader(int key, int lane, int& cnt)
{
    const int prev = __shfl_up_sync(FULL, key, 1);
    const bool leader = lane == 0 || key != prev;
    const unsigned lm = __ballot_sync(FULL, leader);
    const unsigned above = lane == 31 ? 0u : (lm & ~((2u << lane) - 1u));
    cnt = (above ? __ffs(above) - 1 : 32) - lane;
    return leader;
}

#define MARK_U 4                      /* records per thread and block visit: MARK_U independent loads in flight */
#define MARK_BLOCK (256 * MARK_U)     /* records per block visit = the unit of cg_dev_job::mark_blk0 */

__global__ void __launch_bounds__(256)
k_mark(const __grid_constant__ cg_sub S, const __grid_constant__ cg_dev_job J1)
{
    const cg_dev_job* jobs = S.jobs ? S.jobs : &J1;
    __shared__ int s_blk0[CG_JMAX + 1];
    if (threadIdx.x <= S.n_jobs) s_blk0[threadIdx.x] = threadIdx.x < S.n_jobs ? jobs[threadIdx.x].mark_blk0 : S.n_mark_blocks;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    int j = 0, cur = -1;
    /* what the current job says, in registers */
    rec_src R = {nullptr, nullptr, nullptr};
    long long first = 0, n = 0;
    int n_chunks = 1, chrom_len = 0, mq = -1, chunk0 = 0;
    unsigned int* ctl = nullptr;
    unsigned f = 0;
    int n_pass = 0;
    for (int b = blockIdx.x; b < S.n_mark_blocks; b += gridDim.x) {
        while (b >= s_blk0[j + 1]) j++;                       /* blocks are visited in ascending order */
        if (j != cur) {
            if (cur >= 0) {                                   /* leave the previous job: flags and passing records */
                if (f) atomicOr(&ctl[CG_W_FLAGS], f);
                n_pass = __reduce_add_sync(FULL, n_pass);
                if (lane == 0 && n_pass) atomicAdd(reinterpret_cast<unsigned long long*>(ctl + CG_W_PASSED), (unsigned long long)n_pass);
            }
            const cg_dev_job& J = jobs[j];
            R.aos = reinterpret_cast<const uint2*>(J.reads); R.pos = J.rec_pos; R.mapq = J.rec_mapq;
            first = J.mark_first; n = J.mark_n;
            n_chunks = J.n_chunks; chrom_len = J.chrom_len; mq = J.mq_threshold; chunk0 = J.chunk0; ctl = J.ctl;
            f = 0; n_pass = 0; cur = j;
        }
        /* The sweep is a chain load -> compare -> store per record, so its speed is the number of loads in flight and
         * the instructions per record.  A thread takes MARK_U CONSECUTIVE records: all their loads are issued before any
         * is looked at (load_raw), and every record's predecessor sits in the same thread except the first one's, which
         * comes from the lane below (one extra load for lane 0).  Only a record that opens a new chunk does more than a
         * dozen instructions.  (r02d: one record per thread and step, 100 instructions per 32 records, 0.2 ms per 43 M.) */
        const int k0 = (b - s_blk0[j]) * MARK_BLOCK + (int)threadIdx.x * MARK_U;      /* batches hold < 2^31 records */
        const int n32 = (int)n;
        const bool aos = R.aos != nullptr;
        uint2 raw[MARK_U];
        int prev0 = 0;
#pragma unroll
        for (int u = 0; u < MARK_U; u++) {
            raw[u] = make_uint2(0u, 0u);
            if (k0 + u < n32) raw[u] = load_raw(R, first + k0 + u);
        }
        if (lane == 0 && k0 < n32 && first + k0 > 0) prev0 = load_pos(R, first + k0 - 1);
        int p[MARK_U], cp[MARK_U], mapq[MARK_U];
#pragma unroll
        for (int u = 0; u < MARK_U; u++) { unpack_raw(aos, raw[u], p[u], mapq[u]); cp[u] = chunk_of(p[u], n_chunks); CG_CHECK(cp[u] >= 0 && cp[u] < n_chunks); }
        CG_CHECK(chunk0 >= 0 && chunk0 + n_chunks < S.n_space);
        int q = __shfl_up_sync(FULL, p[MARK_U - 1], 1), cq = __shfl_up_sync(FULL, cp[MARK_U - 1], 1);
        if (lane == 0) { q = prev0; cq = chunk_of(prev0, n_chunks); }
        if (first + k0 == 0) cq = -1;                         /* the chromosome's first record opens every chunk up to its own */
        int32_t* lo = S.k.chunk_lo + chunk0;
#pragma unroll
        for (int u = 0; u < MARK_U; u++) {
            if (k0 + u < n32) {
                if (p[u] < q && first + k0 + u > 0) f |= CG_FLAG_UNSORTED;
                if ((unsigned)p[u] >= (unsigned)chrom_len) f |= CG_FLAG_RANGE;
                n_pass += mapq[u] > mq;                                  /* bam_data.c:205, :214 */
                if (cp[u] > cq) {                                        /* this record opens chunk cp (and the empty ones before it) */
                    const int i = (int)first + k0 + u;
                    lo[cp[u]] = i;
#pragma unroll 1
                    for (int t = cq + 1; t < cp[u]; t++) lo[t] = i;
                }
            }
            q = p[u]; cq = cp[u];
        }
        /* the last record pushed so far closes every chunk behind its own, the bound behind the last chunk included
         * (a later batch's first record rewrites the ones it opens); the warp that holds it does that together */
        const int last_u = n32 - 1 - k0;
        const unsigned last_m = __ballot_sync(FULL, last_u >= 0 && last_u < MARK_U);
        if (last_m) {
            int mine = 0;
#pragma unroll
            for (int u = 0; u < MARK_U; u++) if (u == last_u) mine = cp[u];
            const int cpl = __shfl_sync(FULL, mine, __ffs(last_m) - 1);
            const int end = (int)first + n32;
            for (int t = cpl + 1 + lane; t <= n_chunks; t += 32) lo[t] = end;
        }
    }
    if (cur >= 0) {
        if (f) atomicOr(&ctl[CG_W_FLAGS], f);
        n_pass = __reduce_add_sync(FULL, n_pass);
        if (lane == 0 && n_pass) atomicAdd(reinterpret_cast<unsigned long long*>(ctl + CG_W_PASSED), (unsigned long long)n_pass);
    }
}

int cgk_mark_blocks(unsigned long long n_records) { return (int)((n_records + MARK_BLOCK - 1) / MARK_BLOCK); }

cudaError_t cgk_mark(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1)
{
    if (S.n_mark_blocks <= 0) return cudaSuccess;
    int blocks = S.n_mark_blocks;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_mark<<<blocks, 256, 0, s>>>(S, J1);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* delta-coded positions -> int32 position plane (include/conga_gpu.h, cg_encode_positions_delta).  One CTA per block
 * of CG_DELTA_BLOCK = MARK_BLOCK records (the block partition of k_mark), four records per thread: the four delta
 * bytes are one 32-bit load; escapes (255) take their gap from the side array at anchor.big_before + (escapes in
 * front inside the block); positions are anchor.pos_before + inclusive block scan of the gaps.  ~1 byte read and 4
 * written per record: HBM time a twentieth of what the bytes saved on PCIe cost. */
static_assert(MARK_BLOCK == CG_DELTA_BLOCK && MARK_U == 4, "k_decode_delta shares k_mark's block partition");
__global__ void __launch_bounds__(256)
k_decode_delta(const __grid_constant__ cg_sub S, const __grid_constant__ cg_dev_job J1)
{
    const cg_dev_job* jobs = S.jobs ? S.jobs : &J1;
    __shared__ int s_blk0[CG_JMAX + 1];
    __shared__ int s_wsum[8], s_wesc[8];
    if (threadIdx.x <= S.n_jobs) s_blk0[threadIdx.x] = threadIdx.x < S.n_jobs ? jobs[threadIdx.x].mark_blk0 : S.n_mark_blocks;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int j = 0;
    for (int b = blockIdx.x; b < S.n_mark_blocks; b += gridDim.x) {
        while (b >= s_blk0[j + 1]) j++;
        const cg_dev_job& J = jobs[j];
        if (!J.rec_delta) continue;                           /* block-uniform */
        const int lb = b - s_blk0[j], n = (int)J.n_reads;
        const int k0 = lb * CG_DELTA_BLOCK + (int)threadIdx.x * 4;
        const cg_delta_anchor an = J.rec_anchor[lb];
        unsigned d4 = 0u;
        if (k0 + 3 < n && (reinterpret_cast<uintptr_t>(J.rec_delta) & 3u) == 0u)
            d4 = __ldg(reinterpret_cast<const unsigned*>(J.rec_delta + k0));                    /* k0 is a multiple of 4 */
        else for (int u = 0; u < 4; u++) if (k0 + u < n) d4 |= (unsigned)__ldg(J.rec_delta + k0 + u) << (8 * u);
        int g[4], esc = 0;
#pragma unroll
        for (int u = 0; u < 4; u++) { g[u] = (int)((d4 >> (8 * u)) & 0xffu); esc += (k0 + u < n && g[u] == 255); }
        /* escapes in front of this thread inside the block */
        int e_incl = esc;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(FULL, e_incl, d); if (lane >= d) e_incl += o; }
        if (lane == 31) s_wesc[warp] = e_incl;
        __syncthreads();
        int e_before = e_incl - esc;
        for (int w = 0; w < warp; w++) e_before += s_wesc[w];
        int bi = (int)an.big_before + e_before;
#pragma unroll
        for (int u = 0; u < 4; u++) if (k0 + u < n && g[u] == 255) g[u] = __ldg(J.rec_big + bi++);
        /* positions: inclusive scan of the gaps */
        int tsum = 0;
#pragma unroll
        for (int u = 0; u < 4; u++) { if (k0 + u >= n) g[u] = 0; tsum += g[u]; }
        int s_incl = tsum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(FULL, s_incl, d); if (lane >= d) s_incl += o; }
        if (lane == 31) s_wsum[warp] = s_incl;
        __syncthreads();
        int p = an.pos_before + s_incl - tsum;
        for (int w = 0; w < warp; w++) p += s_wsum[w];
        int out[4];
#pragma unroll
        for (int u = 0; u < 4; u++) { p += g[u]; out[u] = p; }
        if (k0 + 3 < n) *reinterpret_cast<int4*>(J.rec_pos_out + k0) = make_int4(out[0], out[1], out[2], out[3]);
        else for (int u = 0; u < 4; u++) if (k0 + u < n) J.rec_pos_out[k0 + u] = out[u];
        __syncthreads();                                      /* s_wsum / s_wesc are reused by the next block */
    }
}

cudaError_t cgk_decode_delta(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1)
{
    if (S.n_mark_blocks <= 0) return cudaSuccess;
    int blocks = S.n_mark_blocks;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_decode_delta<<<blocks, 256, 0, s>>>(S, J1);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* out-of-order input (per-chromosome entry points): global-atomic histogram into u32 words (zero at
 * rest: the pass clears every word it consumes).  Returns at once for a sorted stream. */
__global__ void k_rare_scatter(const __grid_constant__ cg_sub S, const __grid_constant__ cg_dev_job J1)
{
    const cg_dev_job& J = S.jobs ? S.jobs[0] : J1;
    if (!(J.ctl[CG_W_FLAGS] & CG_FLAG_UNSORTED)) return;
    const rec_src R = {reinterpret_cast<const uint2*>(J.reads), J.rec_pos, J.rec_mapq};
    uint32_t* d32 = S.depth32 + ((size_t)J.chunk0 << CG_CHUNK_SHIFT);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < J.n_reads; i += stride) {
        int pos, mapq;
        load_rec(R, i, pos, mapq);
        if (mapq > J.mq_threshold && pos >= 0 && pos < J.chrom_len) atomicAdd(&d32[pos], 1u);
    }
}

cudaError_t cgk_rare_scatter(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1, int n_sms)
{
    if (!S.depth32) return cudaSuccess;
    k_rare_scatter<<<n_sms * 4, 256, 0, s>>>(S, J1);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* the fused pass                                                            */
struct pass_job {                             /* what the pass needs of a job, staged in shared memory */
    rec_src R;
    const signed char* gc;
    int n_reads, chrom_len, mq, chunk0, work0, unsorted;
    int pad[2];
};

struct __align__(16) pass_smem {
#ifdef PASS_BULK_STORE
#ifndef PASS_BULK_BUFS
#define PASS_BULK_BUFS 1                      /* 2: one region is binned while the bulk store of the other is read (only two CTAs fit an SM then) */
#endif
    uint4 bins4[NWARP][PASS_BULK_BUFS * (CG_CHUNK / 8)];
#else
    uint4 bins4[NWARP][CG_CHUNK / 8];         /* 8 x 4 KB: eight packed u16 counts per vector, one region per warp */
#endif
    int win_end[NWARP][MAXWIN];               /* per GC window of the warp's chunk: records (or, wrapped chunks, depth sum) */
    pass_job job[CG_JMAX];
    /* followed by int gcs[n_jobs][CG_PART_WORDS]: per-CTA rd_per_gc / window_per_gc of every job.  32-bit is
     * exact here: a CTA's share of the depth sum is bounded by the record count (< 2^31) and it sweeps < 2^32
     * positions; native 32-bit shared atomics instead of the CAS loop a 64-bit shared add compiles to */
};

#define PASS_MAX_CTAS_PER_SM 4        /* register budget the kernel is compiled for: 64 per thread, so that three CTAs leave
                                         a quarter of the register file to the record sweep and the genotype kernel of the
                                         neighbouring submissions */
/* CTAs per SM one launch asks for.  Measured on B200 with one launch per sample (r02b, pass alone): C3 1 -> 2.38 ms,
 * 2 -> 1.40, 3 -> 1.12 (0.85 of the copy peak); C2 2 -> 1.06, 3 -> 0.99 (0.91): the loop waits on record loads that
 * are asked for one chunk ahead, and under a saturated HBM a chunk-time of 16 warps is shorter than that latency. */
#define PASS_GRID_CTAS_PER_SM 3       /* upper value; cgk_pass_grid picks 2 or 3 by record density */
#ifndef PASS_BATCH
#define PASS_BATCH 8                  /* chunks per visit to the work counter.  r02j, pass alone / step, ms per sample:
                                       * 4: C3 1.002 / 1.198, C2 1.175 / 1.113;  6: 0.989 / 1.189, 0.899 / 0.977;  8: 0.985 / 1.185,
                                       * 0.902 / 0.987 (C4 and a rank's eighth of C3 unchanged): with a batch a warp writes 32 KB of
                                       * consecutive depth before it moves on, and the sparse stream of C2 is all writes */
#endif
#ifndef PASS_PRE
#define PASS_PRE 2                    /* 32-record groups of a chunk asked for one chunk ahead */
#endif
static_assert(PASS_BATCH >= 3 && PASS_BATCH <= 32, "the bounds of batch n+1 are asked for at chunk 1 and used after chunk PASS_BATCH-1");

/* expected_read_depth[101], read_distribution.c:75-83:
 *   expected[0] = 0; expected[i] = (float)rd_per_gc[i] / window_per_gc[i]; NaN / Inf -> 0.
 * long->float and int->float round to nearest like x86's cvtsi2ss, and the quotient is one
 * IEEE fp32 division. */
__device__ __forceinline__ float expected_of(long long rd, long long cnt, int i)
{
    float e = 0.0f;
    if (i > 0) {
        e = __fdiv_rn(__ll2float_rn(rd), __int2float_rn((int)cnt));
        if (isnan(e) || isinf(e)) e = 0.0f;
    }
    return e;
}

/* sum of the two sign-extended 16-bit halves of w, added to acc (one IDP.2A) */
__device__ __forceinline__ int add_halves(unsigned w, int acc) { return __dp2a_lo((int)w, 0x0101, acc); }

/* 32 records (one per lane) into the warp's packed bins and window counters.  Warp aggregation: the stream
 * is sorted, so equal positions sit in adjacent lanes and one lane adds for all of them. */
__device__ __forceinline__ void bin32(bool in, int pos, int mapq, int base, int w0, int mq, int chrom_len,
                                      unsigned* bins, int* wend, int lane, bool& big)
{
    const unsigned off = (unsigned)(pos - base);
    const bool ok = in && mapq > mq                                              /* bam_data.c:205 */
                    && off < (unsigned)CG_CHUNK && pos < chrom_len;
    int cnt;
    const bool leader = run_leader(ok ? (int)off : -1 - lane, lane, cnt);
    if (ok && leader) {
        unsigned now;                                                            /* the position's count so far, unwrapped */
        if (off & 1u) {
            now = (atomicAdd(&bins[off >> 1], (unsigned)cnt << 16) >> 16) + (unsigned)cnt;   /* bam_data.c:213; carry-out = wrap */
        } else {
            now = (atomicAdd(&bins[off >> 1], (unsigned)cnt) & 0xffffu) + (unsigned)cnt;
            /* the low half wrapped into its neighbour's count: take the carry back out */
            if (now > 0xffffu) atomicSub(&bins[off >> 1], 0x10000u);
        }
        big |= now >= 0x8000u;
        CG_CHECK((off >> 1) < (unsigned)(CG_CHUNK / 2) && pos / 100 - w0 >= 0 && pos / 100 - w0 < MAXWIN);
        atomicAdd(&wend[pos / 100 - w0], cnt);
    }
}

__global__ void __launch_bounds__(NT, PASS_MAX_CTAS_PER_SM)
k_depth_pass(const __grid_constant__ cg_sub S, const __grid_constant__ cg_dev_job J1)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    pass_smem& sm = *reinterpret_cast<pass_smem*>(smem_raw);
    int* gcs = reinterpret_cast<int*>(smem_raw + sizeof(pass_smem));
    const cg_dev_job* jobs = S.jobs ? S.jobs : &J1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_jobs = S.n_jobs;
    uint4* my4 = sm.bins4[warp];
    unsigned* bins = reinterpret_cast<unsigned*>(my4);                 /* word = positions 2k (low), 2k+1 (high) */
#ifdef PASS_BULK_STORE
    int bulk_buf = 0;
#ifdef PASS_BULK_HINT
    unsigned long long bulk_policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(bulk_policy));
#endif
#endif
    int* wend = sm.win_end[warp];

    if (tid < n_jobs) {
        const cg_dev_job& J = jobs[tid];
        pass_job& P = sm.job[tid];
        P.R.aos = reinterpret_cast<const uint2*>(J.reads); P.R.pos = J.rec_pos; P.R.mapq = J.rec_mapq;
        P.gc = reinterpret_cast<const signed char*>(J.gc);
        P.n_reads = (int)J.n_reads; P.chrom_len = J.chrom_len; P.mq = J.mq_threshold; P.chunk0 = J.chunk0; P.work0 = J.work0;
        P.unsorted = (S.depth32 != nullptr && (J.ctl[CG_W_FLAGS] & CG_FLAG_UNSORTED)) ? 1 : 0;
    }
    for (int i = lane; i < CG_CHUNK / 8; i += 32) my4[i] = make_uint4(0, 0, 0, 0);
    if (lane < MAXWIN) wend[lane] = 0;
    for (int i = tid; i < n_jobs * CG_PART_WORDS; i += NT) gcs[i] = 0;
    __syncthreads();

    const unsigned n_work = (unsigned)S.n_work;

    /* lanes 0..PASS_BATCH-1: job, chunk within the job and record bounds of the chunks of one batch.  The two bounds
     * are only ASKED for here (clamped to the job's record count where they are used, see load_raw) */
    auto load_batch = [&](unsigned wfirst, int jhint, int& job, int& lc, int& lo, int& hi) {
        job = -1; lc = 0; lo = 0; hi = 0;
        const unsigned wq = wfirst + (unsigned)lane;
        if (lane < PASS_BATCH && wq < n_work) {
            int j = jhint;
            while (j + 1 < n_jobs && wq >= (unsigned)sm.job[j + 1].work0) j++;
            job = j; lc = (int)(wq - (unsigned)sm.job[j].work0);
            const int g = sm.job[j].chunk0 + lc;
            CG_CHECK(g >= 0 && g + 1 < S.n_space);
            lo = __ldg(&S.k.chunk_lo[g]);
            hi = __ldg(&S.k.chunk_lo[g + 1]);
        }
    };
    /* what a chunk needs from HBM, asked for one chunk ahead of use: its first 32*PASS_PRE records
     * and the GC byte of this lane's window */
    auto prefetch = [&](int job, int lc, int lo, int hi, uint2 (&raw)[PASS_PRE], int& gcb) {
#pragma unroll
        for (int u = 0; u < PASS_PRE; u++) raw[u] = make_uint2(0u, 0u);
        gcb = 0;
        if (job < 0) return;
        const pass_job& P = sm.job[job];
        if (!P.unsorted) {
            const rec_src R = P.R;
#pragma unroll
            for (int u = 0; u < PASS_PRE; u++) {
                const int i = lo + 32 * u + lane;
                CG_CHECK(i < hi ? (i >= 0 && i < P.n_reads) : true);
                if (i < hi) raw[u] = load_raw(R, i);           /* issued here, unpacked when the chunk is binned */
            }
        }
        const int base = lc << CG_CHUNK_SHIFT, w0 = base / 100;
        const int nwin = (base + CG_CHUNK - 1) / 100 - w0 + 1;
        const int gw_s = max(base, 100 * (w0 + lane)), gw_e = min(min(base + CG_CHUNK, P.chrom_len), 100 * (w0 + lane) + 100);
        CG_CHECK(!(lane < nwin && gw_e > gw_s) || (w0 + lane) * 100 < P.chrom_len);
        if (lane < nwin && gw_e > gw_s) gcb = (int)P.gc[w0 + lane];               /* sonic_interval.c:263-272 */
    };

    /* Chunks are handed out PASS_BATCH at a time from one counter: at 2 B per position the pass needs
     * 1.5 G chunks/s to saturate HBM, more than one address takes atomics (measured 0.9 G/s, r01g).
     * Every global load of the loop is issued at least one chunk ahead of its use: the counter value of
     * batch n+1 at chunk 0 of batch n, its chunk bounds at chunk 1, the records and GC bytes of a chunk
     * while the chunk before it is binned and stored. */
    unsigned wcur = 0u, wnext = 0u;
    if (lane == 0) wcur = atomicAdd(&S.sctl[CG_SW_TILE], (unsigned)PASS_BATCH);
    wcur = __shfl_sync(FULL, wcur, 0);
    int b_job, b_lc, b_lo, b_hi;
    load_batch(wcur, 0, b_job, b_lc, b_lo, b_hi);
    int n_job = -1, n_lc = 0, n_lo = 0, n_hi = 0;
    uint2 p_raw[PASS_PRE];
    int p_gc;
    {
        const int j0 = __shfl_sync(FULL, b_job, 0), nr0 = j0 >= 0 ? sm.job[j0].n_reads : 0;
        prefetch(j0, __shfl_sync(FULL, b_lc, 0), max(0, min(__shfl_sync(FULL, b_lo, 0), nr0)), max(0, min(__shfl_sync(FULL, b_hi, 0), nr0)), p_raw, p_gc);
    }
    int q = 0;

    while (wcur < n_work) {
        if (q == 0 && lane == 0) wnext = atomicAdd(&S.sctl[CG_SW_TILE], (unsigned)PASS_BATCH);
        if (q == 1) {
            wnext = __shfl_sync(FULL, wnext, 0);
            load_batch(wnext, max(0, __shfl_sync(FULL, b_job, 0)), n_job, n_lc, n_lo, n_hi);
        }
        const int cjob = __shfl_sync(FULL, b_job, q), clc = __shfl_sync(FULL, b_lc, q);
        const int c_nr = cjob >= 0 ? sm.job[cjob].n_reads : 0;
        const int lo = max(0, min(__shfl_sync(FULL, b_lo, q), c_nr)), hi = max(0, min(__shfl_sync(FULL, b_hi, q), c_nr));
        /* the chunk after this one: the next of the batch, or the first of the next batch */
        const bool last = q == PASS_BATCH - 1;
        const int xs = last ? 0 : q + 1;
        const int xjob = __shfl_sync(FULL, last ? n_job : b_job, xs), xlc = __shfl_sync(FULL, last ? n_lc : b_lc, xs);
        const int x_nr = xjob >= 0 ? sm.job[xjob].n_reads : 0;
        const int xlo = max(0, min(__shfl_sync(FULL, last ? n_lo : b_lo, xs), x_nr)), xhi = max(0, min(__shfl_sync(FULL, last ? n_hi : b_hi, xs), x_nr));
        uint2 x_raw[PASS_PRE];
        int x_gc;
        prefetch(xjob, xlc, xlo, xhi, x_raw, x_gc);

        if (cjob >= 0) {
            const pass_job& P = sm.job[cjob];
            const int chrom_len = P.chrom_len, mq = P.mq;
            const int base = clc << CG_CHUNK_SHIFT;
            const int w0 = base / 100;                    /* first GC window overlapping the chunk */
            const int nwin = (base + CG_CHUNK - 1) / 100 - w0 + 1;
            const int gw_s = max(base, 100 * (w0 + lane)), gw_e = min(min(base + CG_CHUNK, chrom_len), 100 * (w0 + lane) + 100);
            const size_t gpos = (size_t)(P.chunk0 + clc) << CG_CHUNK_SHIFT;          /* the chunk in the slot's depth array */
            CG_CHECK(P.chunk0 + clc >= 0 && P.chunk0 + clc < S.n_space && nwin <= MAXWIN && cjob < n_jobs);
            /* A window's depth sum is the number of records that fell into it as long as no int16 of the chunk
             * went negative (32768 or more records on one position): counted while the records are binned.
             * `wrapped` chunks (and the out-of-order path) take the sums from a scan of the bins instead. */
            bool big = false;
#ifdef PASS_BULK_STORE
            /* this chunk's region: the other one may still be read by the bulk store of the chunk before; the store
             * that read THIS region (two chunks ago) has to be through before the region is cleared */
            my4 = sm.bins4[warp] + bulk_buf * (CG_CHUNK / 8);
            bins = reinterpret_cast<unsigned*>(my4);
            if (lane == 0) {
                if (PASS_BULK_BUFS == 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
            __syncwarp();
            for (int i = lane; i < CG_CHUNK / 8; i += 32) my4[i] = make_uint4(0, 0, 0, 0);
            __syncwarp();
#endif

            /* ---- phase A: the chunk's records -> packed shared-memory histogram, records per window ---- */
            if (!P.unsorted) {
                const rec_src R = P.R;
                const bool aos = R.aos != nullptr;
                uint2 c_raw[PASS_PRE];
#pragma unroll
                for (int u = 0; u < PASS_PRE; u++) c_raw[u] = p_raw[u];
                for (int i0 = lo; i0 < hi; i0 += 32 * PASS_PRE) {
                    uint2 f_raw[PASS_PRE];                        /* the group after this one */
#pragma unroll
                    for (int u = 0; u < PASS_PRE; u++) {
                        f_raw[u] = make_uint2(0u, 0u);
                        const int i = i0 + 32 * (PASS_PRE + u) + lane;
                        CG_CHECK(i < hi ? (i >= 0 && i < P.n_reads) : true);
                        if (i < hi) f_raw[u] = load_raw(R, i);
                    }
#pragma unroll
                    for (int u = 0; u < PASS_PRE; u++)
                        if (i0 + 32 * u < hi) {
                            int r_pos, r_mq;
                            unpack_raw(aos, c_raw[u], r_pos, r_mq);
                            bin32(i0 + 32 * u + lane < hi, r_pos, r_mq, base, w0, mq, chrom_len, bins, wend, lane, big);
                        }
#pragma unroll
                    for (int u = 0; u < PASS_PRE; u++) c_raw[u] = f_raw[u];
                }
            } else {
                uint4* src = reinterpret_cast<uint4*>(S.depth32 + gpos);
                for (int v = lane; v < CG_CHUNK / 8; v += 32) {
                    const uint4 x = src[2 * v], y = src[2 * v + 1];
                    my4[v] = make_uint4((x.x & 0xffffu) | (x.y << 16), (x.z & 0xffffu) | (x.w << 16),
                                        (y.x & 0xffffu) | (y.y << 16), (y.z & 0xffffu) | (y.w << 16));
                    src[2 * v] = make_uint4(0, 0, 0, 0);                              /* zero at rest */
                    src[2 * v + 1] = make_uint4(0, 0, 0, 0);
                }
                big = true;
            }
            const bool wrapped = __any_sync(FULL, big);
            __syncwarp();

            if (!wrapped) {
                /* the chunk's depth sum = its records (k_finalize sums the group and super levels from these) */
                const int chunk_total = __reduce_add_sync(FULL, lane < nwin ? wend[lane] : 0);
                if (lane == 0) S.k.cnt1[P.chunk0 + clc] = chunk_total;
                /* ---- rows -> depth stores: the packed bins ARE the int16 output vectors ---- */
#ifdef PASS_BULK_STORE
                /* one bulk copy shared -> global of the whole 4 KB region, issued by one lane */
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
                    const unsigned src = (unsigned)__cvta_generic_to_shared(my4);
                    int16_t* dst = S.depth + gpos;
#ifdef PASS_BULK_HINT
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
                                 :: "l"(dst), "r"(src), "r"(CG_CHUNK * 2), "l"(bulk_policy) : "memory");
#else
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(dst), "r"(src), "r"(CG_CHUNK * 2) : "memory");
#endif
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
                bulk_buf ^= PASS_BULK_BUFS - 1;
#else
                char* dp = reinterpret_cast<char*>(S.depth) + (gpos + (size_t)lane * 8) * 2;
#pragma unroll
                for (int r = 0; r < ROWS; r++) {
                    const uint4 v = my4[r * 32 + lane];
                    my4[r * 32 + lane] = make_uint4(0, 0, 0, 0);          /* clean bins for the next chunk */
                    st_cs_v4(dp + r * 512, v.x, v.y, v.z, v.w);
                }
#endif
            } else {
                if (lane < MAXWIN) wend[lane] = 0;                        /* the scan below writes window-END sums */
                __syncwarp();
                /* ---- eight row scans at once: lane sums, then eight independent shuffle chains ---- */
                int ex[ROWS];                                 /* -> prefix in front of this lane's 8 positions, per row */
                int chunk_total = 0;
                {
                    int tot[ROWS];
#pragma unroll
                    for (int r = 0; r < ROWS; r++) {
                        const uint4 v = my4[r * 32 + lane];
                        tot[r] = add_halves(v.w, add_halves(v.z, add_halves(v.y, add_halves(v.x, 0))));
                        ex[r] = tot[r];
                    }
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
                        for (int r = 0; r < ROWS; r++) {
                            const int o = __shfl_up_sync(FULL, ex[r], d);
                            if (lane >= d) ex[r] += o;
                        }
                    }
                    int run = 0;                                /* sums are relative to the chunk's first position */
#pragma unroll
                    for (int r = 0; r < ROWS; r++) {
                        const int row_total = __shfl_sync(FULL, ex[r], 31);
                        ex[r] = run + ex[r] - tot[r];
                        run += row_total;
                    }
                    chunk_total = run;
                }
                if (lane == 0) S.k.cnt1[P.chunk0 + clc] = chunk_total;     /* depth sum of the chunk, wrapped int16s included */

                /* ---- rows -> depth stores, window-end sums ---- */
                {
                    const int cpos = base + lane * 8;         /* this lane's first position in row 0 */
                    int wq = cpos / 100, wm = cpos - wq * 100;                /* its GC window and offset within it */
                    char* dp = reinterpret_cast<char*>(S.depth) + (gpos + (size_t)lane * 8) * 2;
#pragma unroll
                    for (int r = 0; r < ROWS; r++) {
                        const uint4 v = my4[r * 32 + lane];
                        my4[r * 32 + lane] = make_uint4(0, 0, 0, 0);          /* clean bins for the next chunk */
                        const int b = ex[r];
                        const int s0 = sext16(v.x), s1 = add_halves(v.x, 0);
                        const int s2 = s1 + sext16(v.y), s3 = add_halves(v.y, s1);
                        const int s4 = s3 + sext16(v.z), s5 = add_halves(v.z, s3);
                        const int s6 = s5 + sext16(v.w), s7 = add_halves(v.w, s5);
                        st_cs_v4(dp + r * 512, v.x, v.y, v.z, v.w);
                        /* the lane whose 8 positions contain a window's last position (== 99 mod 100) */
                        if (wm >= 92) {
                            const int k = 99 - wm;
                            const int sk = k == 0 ? s0 : k == 1 ? s1 : k == 2 ? s2 : k == 3 ? s3
                                         : k == 4 ? s4 : k == 5 ? s5 : k == 6 ? s6 : s7;
                            wend[wq - w0] = b + sk;
                        }
                        if (r == ROWS - 1 && lane == 31)                      /* the chunk's last position closes a partial window */
                            wend[(base + CG_CHUNK - 1) / 100 - w0] = b + s7;
                        wm += 56; wq += 2;                                    /* 256 = 2*100 + 56 */
                        if (wm >= 100) { wm -= 100; wq++; }
                    }
                }
                __syncwarp();

                /* window-end sums -> per-window sums */
                const int d = lane < nwin ? wend[lane] - (lane ? wend[lane - 1] : 0) : 0;
                __syncwarp();
                if (lane < nwin) wend[lane] = d;
            }
            __syncwarp();

            /* ---- GC-bin sums from the per-window depth sums (read_distribution.c:63-73) ---- */
            if (lane < nwin && gw_e > gw_s) {
                const int bin = (unsigned)p_gc > 100u ? 0 : p_gc;
                int* acc = gcs + cjob * CG_PART_WORDS;
                const int sum = wend[lane];
                if (sum) atomicAdd(&acc[bin], sum);
                atomicAdd(reinterpret_cast<unsigned*>(&acc[CG_GC_BINS + bin]), (unsigned)(gw_e - gw_s));
            }
            if (lane < MAXWIN) wend[lane] = 0;                            /* clean for the next chunk */
            __syncwarp();
        }

#pragma unroll
        for (int u = 0; u < PASS_PRE; u++) p_raw[u] = x_raw[u];
        p_gc = x_gc;
        if (++q == PASS_BATCH) {
            q = 0;
            wcur = wnext;
            b_job = n_job; b_lc = n_lc; b_lo = n_lo; b_hi = n_hi;
        }
    }

#ifdef PASS_BULK_STORE
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      /* shared memory is read until then */
#endif
    /* ---- CTA epilogue: what this CTA saw of every job's GC bins, plain coalesced stores ---- */
    __syncthreads();
    CG_CHECK((int)blockIdx.x < S.n_pass_ctas);
    int* out = S.partial + (size_t)blockIdx.x * n_jobs * CG_PART_WORDS;
    for (int i = tid; i < n_jobs * CG_PART_WORDS; i += NT) out[i] = gcs[i];
}

static int pass_ctas_per_sm(void)
{
    static int ctas_per_sm = 0;               /* CG_PASS_CTAS: CTAs per SM one launch asks for (tuning) */
    if (!ctas_per_sm) {
        const char* e = getenv("CG_PASS_CTAS");
        ctas_per_sm = e ? atoi(e) : PASS_GRID_CTAS_PER_SM;
        if (ctas_per_sm < 1 || ctas_per_sm > PASS_MAX_CTAS_PER_SM) ctas_per_sm = PASS_MAX_CTAS_PER_SM;
    }
    return ctas_per_sm;
}

int cgk_pass_grid(int n_work, long long n_records, int n_sms)
{
    /* Three CTAs per SM.  Alone, a sparse record stream (C2: 3 per chunk) is a pure write stream and runs faster with
     * two (1.00 against 1.14 ms per sample, r02d) — but in the pipeline of samples two CTAs leave so much of the SM to
     * the neighbours' record sweep and genotype kernel that their traffic breaks the write stream (1.47 ms per step
     * against 1.09); the dense stream (C3: 29 per chunk) needs the third CTA for its record loads either way. */
    (void)n_records;
    int per_sm = pass_ctas_per_sm();
    int grid = n_sms * per_sm;
    const int want = (n_work + NWARP * PASS_BATCH - 1) / (NWARP * PASS_BATCH);
    if (grid > want) grid = want;
    return grid < 1 ? 1 : grid;
}

static size_t pass_smem_bytes(int n_jobs) { return sizeof(pass_smem) + (size_t)n_jobs * CG_PART_WORDS * sizeof(int); }

cudaError_t cgk_depth_pass(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1)
{
    k_depth_pass<<<S.n_pass_ctas, NT, pass_smem_bytes(S.n_jobs), s>>>(S, J1);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* per job: the partial rows of the pass -> the cg_chrom_stats row (read_distribution.c:27-73), the
 * expected table (:75-83), the chromosome's flags; and the group / super levels of the depth sums from the
 * chunk level (what depth_sum_before in genotype.cu reads).  grid = (jobs, FIN_Y): the CTAs of a job share
 * its supers, CTA (j, 0) also writes the row. */
/* 16 CTAs of 256 threads per job rather than a few large ones: a CTA of this size fits beside three CTAs of the pass
 * on an SM, so the finalize of submission k does not have to wait for a gap in the pass of submission k+1 */
#define FIN_Y 16
#define FIN_T 256
#define FIN_PARTS (FIN_T / 256)       /* slices of a CTA's share of the pass's CTAs summed side by side */
__global__ void __launch_bounds__(FIN_T)
k_finalize(const __grid_constant__ cg_sub S, const __grid_constant__ cg_dev_job J1)
{
    const int j = blockIdx.x;
    const cg_dev_job& J = S.jobs ? S.jobs[j] : J1;
    __shared__ long long part[FIN_PARTS][CG_PART_WORDS];
    __shared__ long long row[CG_STATS_WORDS];
    __shared__ int s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    /* ---- depth sums per group of 32 chunks and per super of 32 groups, from the chunk level the pass wrote
     *      (chunks behind the chromosome's last one count as zero: nobody ever writes them) ---- */
    {
        const int s0 = J.chunk0 >> (2 * LVL), s1 = (J.chunk0 + J.n_chunks + CG_SUPER_CHUNKS - 1) >> (2 * LVL);
        const int c_end = J.chunk0 + J.n_chunks;
        for (int sp = s0 + (int)blockIdx.y * (FIN_T / 32) + warp; sp < s1; sp += (FIN_T / 32) * FIN_Y) {
            const int cb = sp << (2 * LVL);
            CG_CHECK(cb >= 0 && cb + CG_SUPER_CHUNKS <= S.n_space);
            const int32_t* c1 = S.k.cnt1 + cb;
            int v[32];
#pragma unroll
            for (int g = 0; g < 32; g++) { const int ci = (g << LVL) + lane; v[g] = cb + ci < c_end ? c1[ci] : 0; }   /* 32 independent coalesced loads */
            int mine = 0;                                     /* lane g ends up with the sum of group g */
#pragma unroll
            for (int g = 0; g < 32; g++) {
                const int t = __reduce_add_sync(FULL, v[g]);
                if (lane == g) mine = t;
            }
            S.k.cnt2[((size_t)sp << LVL) + lane] = mine;
            const int tot = __reduce_add_sync(FULL, mine);
            if (lane == 0) S.k.cnt3[sp] = tot;
        }
    }

    /* ---- the statistics row: CTA (j, y) adds word w of every FIN_Y-th pass CTA's partial row into the job's scratch
     *      row (zero at rest); the last of the job's CTAs to arrive finishes the row ---- */
    long long* acc = S.fin_acc + (size_t)j * 256;
    {
        const int w = tid & 255, sl = tid >> 8;
        if (w < CG_PART_WORDS) {
            const int32_t* p = S.partial + (size_t)j * CG_PART_WORDS + w;
            const size_t stride = (size_t)S.n_jobs * CG_PART_WORDS;
            long long s = 0;
            if (w < CG_GC_BINS) {
#pragma unroll 8
                for (int c = (int)blockIdx.y + FIN_Y * sl; c < S.n_pass_ctas; c += FIN_Y * FIN_PARTS) s += (long long)p[c * stride];
            } else {
#pragma unroll 8
                for (int c = (int)blockIdx.y + FIN_Y * sl; c < S.n_pass_ctas; c += FIN_Y * FIN_PARTS) s += (long long)(unsigned)p[c * stride];
            }
            part[sl][w] = s;
        }
    }
    __syncthreads();
    if (tid < CG_PART_WORDS) {
        long long s = 0;
#pragma unroll
        for (int k = 0; k < FIN_PARTS; k++) s += part[k][tid];
        if (s) atomicAdd(reinterpret_cast<unsigned long long*>(acc + tid), (unsigned long long)s);
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(&S.fin_cnt[j], 1u) == FIN_Y - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (tid < CG_PART_WORDS) { row[tid] = __ldcg(acc + tid); acc[tid] = 0; }        /* scratch row zero at rest again */
    if (tid == 0) S.fin_cnt[j] = 0u;
    __syncthreads();
    if (warp == 0) {
        /* calc_mu_per_chr, read_distribution.c:33-37: the sum of every depth = the sum over the GC bins, each
         * position belonging to exactly one bin (:70-71) */
        long long tot = 0;
        for (int b = lane; b < CG_GC_BINS; b += 32) tot += row[b];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) tot += __shfl_xor_sync(FULL, tot, d);
        if (lane == 0) {
            row[2 * CG_GC_BINS] = tot;
            row[2 * CG_GC_BINS + 1] = (long long)*reinterpret_cast<const unsigned long long*>(J.ctl + CG_W_PASSED);
        }
    }
    __syncthreads();
    if (tid < CG_STATS_WORDS) {
        J.stats_row[tid] = row[tid];
        if (J.row_out) J.row_out[tid] = row[tid];
        /* the collective, fused: the finished row goes straight into every rank's table (peer memory over NVLink;
         * rows are disjoint per rank, so there is nothing to reduce) */
        for (int p = 0; p < S.n_peer_rows; p++) S.peer_rows[p][(size_t)J.chr_index * CG_STATS_WORDS + tid] = row[tid];
        if (S.n_peer_rows) __threadfence_system();
    }
    if (tid < CG_GC_BINS) J.expected[tid] = expected_of(row[tid], row[CG_GC_BINS + tid], tid);
    if (tid == 0) {
        const unsigned fl = J.ctl[CG_W_FLAGS];
        if (fl & CG_FLAG_RANGE) atomicOr(S.sticky, CG_FLAG_RANGE);
        if (S.h_flags) S.h_flags[j] = fl;
    }
    __syncthreads();
    /* zero at rest: this job's control words, and (once per submission) the slot's work counters — the record sweep
     * and the pass that used them are behind this kernel in stream order */
    if (tid < CG_CTL_WORDS) J.ctl[tid] = 0u;
    if (j == 0 && tid < 4) S.sctl[tid] = 0u;
}

cudaError_t cgk_finalize(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1)
{
    k_finalize<<<dim3((unsigned)S.n_jobs, FIN_Y), FIN_T, 0, s>>>(S, J1);
    return cudaGetLastError();
}

/* What is left of the collective when k_finalize stores the rows into the peers' tables itself: one warp tells every
 * peer "my rows of this sample are in your table" (flags[bank][my rank] = epoch in the peer's memory) and waits for
 * the same word from every peer.  Runs on cg_stream behind cg_join, i.e. behind every finalize of the sample. */
__global__ void k_peer_exchange(const __grid_constant__ cg_peer_flags f)
{
    const int t = threadIdx.x;
    if (t >= f.n) return;
    __threadfence_system();
    if (t != f.rank) {
        volatile unsigned int* dst = f.theirs[t] + f.rank;
        *dst = f.epoch;
    }
    __threadfence_system();
    if (t != f.rank) {
        volatile unsigned int* src = f.mine + t;
        const long long t0 = clock64();
        while ((int)(*src - f.epoch) < 0) {
            if (clock64() - t0 > 20000000000ll) { atomicOr(f.sticky, CG_FLAG_PEER_TIMEOUT); break; }     /* ~10 s: a peer is gone */
            __nanosleep(200);
        }
    }
    __threadfence_system();
}

cudaError_t cgk_peer_exchange(cudaStream_t s, const cg_peer_flags& f)
{
    k_peer_exchange<<<1, 32, 0, s>>>(f);
    return cudaGetLastError();
}

/* Two kernels share an SM only if they ask for the same L1/shared-memory split; otherwise the
 * CTAs of the second wait until the SM has drained.  Every kernel of the library therefore
 * states the same carve-out, so that the preparation and genotype kernels of one submission
 * really run on the SMs a neighbour's depth pass occupies. */
int cgk_carveout(void)
{
    static int pct = -1;
    if (pct < 0) {
        const char* e = getenv("CG_CARVEOUT");            /* tuning only */
        pct = e ? atoi(e) : 100;
        if (pct < 0 || pct > 100) pct = 100;
    }
    return pct;
}

cudaError_t cgk_configure_depth(void)
{
    const int pct = cgk_carveout();
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_depth_pass, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pass_smem_bytes(CG_JMAX))) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_depth_pass, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_mark, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_decode_delta, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_rare_scatter, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_finalize, cudaFuncAttributePreferredSharedMemoryCarveout, pct)) != cudaSuccess) return e;
    return cudaSuccess;
}

/* Diagnostic (cg_download_prefix): the int32 inclusive prefix of a job's depth array, P[i] = sum of depth[0..i]
 * mod 2^32, rebuilt from the depth array and the chunk / group / super sums.  One warp per chunk. */
__global__ void __launch_bounds__(256)
k_prefix_from_depth(const int16_t* __restrict__ depth, cg_counts k, int chunk0, int n_chunks, int32_t* __restrict__ prefix)
{
    const int lane = threadIdx.x & 31;
    const int n_warps = gridDim.x * (blockDim.x >> 5);
    for (int lc = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); lc < n_chunks; lc += n_warps) {
        const int c = chunk0 + lc;
        int part = 0;
        const int g = c >> LVL, sp = c >> (2 * LVL);
        const int i1 = (g << LVL) + lane, i2 = (sp << LVL) + lane;
        if (i1 < c) part += k.cnt1[i1];
        if (i2 < g) part += k.cnt2[i2];
        for (int j = (chunk0 >> (2 * LVL)) + lane; j < sp; j += 32) part += k.cnt3[j];
        int run = __reduce_add_sync(FULL, part);
        const size_t base = (size_t)c << CG_CHUNK_SHIFT, obase = (size_t)lc << CG_CHUNK_SHIFT;
        const uint4* src = reinterpret_cast<const uint4*>(depth + base);
        for (int r = 0; r < ROWS; r++) {
            const uint4 v = src[r * 32 + lane];
            const int s0 = sext16(v.x), s1 = add_halves(v.x, 0);
            const int s2 = s1 + sext16(v.y), s3 = add_halves(v.y, s1);
            const int s4 = s3 + sext16(v.z), s5 = add_halves(v.z, s3);
            const int s6 = s5 + sext16(v.w), s7 = add_halves(v.w, s5);
            const int incl = warp_incl_scan(s7, lane);
            const int b = run + incl - s7;
            int4* dst = reinterpret_cast<int4*>(prefix + obase + (size_t)(r * 32 + lane) * 8);
            dst[0] = make_int4(b + s0, b + s1, b + s2, b + s3);
            dst[1] = make_int4(b + s4, b + s5, b + s6, b + s7);
            run += __shfl_sync(FULL, incl, 31);
        }
    }
}

cudaError_t cgk_prefix_from_depth(cudaStream_t s, const int16_t* depth, cg_counts k, int chunk0, int n_chunks, int32_t* prefix)
{
    if (n_chunks <= 0) return cudaSuccess;
    int blocks = (n_chunks + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_prefix_from_depth<<<blocks, 256, 0, s>>>(depth, k, chunk0, n_chunks, prefix);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------------- */
/* the table alone (cg_set_expected replays, tests) */
__global__ void k_expected_table(const long long* __restrict__ stats_row, float* expected)
{
    const int i = threadIdx.x;
    if (i < CG_GC_BINS) expected[i] = expected_of(stats_row[i], stats_row[CG_GC_BINS + i], i);
}

cudaError_t cgk_expected_table(cudaStream_t s, const long long* stats_row, float* expected)
{
    k_expected_table<<<1, 128, 0, s>>>(stats_row, expected);
    return cudaGetLastError();
}
