/* conga_gpu.h — C-ABI of libconga_gpu.so: CONGA's read-depth genotyping hot path
 * on one B200 (sm_100a).  Plain C types only; no torch, no C++ in the signatures.
 *
 * CONGA (reference tree, commit 58d4fe4) has no plugin/FFI interface: the hot
 * path sits behind four plain C call sites inside read_bam() (bam_data.c:224-359).
 * Each entry point below names the reference seam it replaces (file:line under
 * the reference tree).  One cg_ctx per GPU; calls on one context must come from
 * one host thread at a time.  Every int-returning function returns CG_OK (0) or a
 * negative CG_E* code; the text is available from cg_last_error().  There is NO
 * CPU fallback anywhere behind this interface: if the device or the kernels are
 * unavailable every call fails.
 *
 * Per-chromosome call order (mirrors bam_data.c:304-338):
 *   cg_chrom_begin      <- init_rd_per_chr            bam_data.c:304, read_distribution.c:12-18
 *   cg_push_reads  (xN) <- count_reads_bam loop body  bam_data.c:201-217
 *   cg_chrom_finish     <- calc_mean_per_chr          bam_data.c:322, read_distribution.c:27-84
 *   cg_set_mappability  <- init_mappability_per_chr + load_mappability_regions
 *                                                     likelihood.c:352-356, svs.c:317-377
 *   cg_genotype    (x2) <- find_depths                likelihood.c:290-308 -> :108-169, :96-105
 */
#ifndef CONGA_GPU_H
#define CONGA_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CG_ABI_VERSION 3

enum {
    CG_OK = 0,
    CG_EINVAL = -1,   /* bad argument / call order */
    CG_ECUDA = -2,    /* CUDA runtime or kernel failure */
    CG_ENOMEM = -3,   /* host or device allocation failed */
    CG_ERANGE = -4    /* a read or interval lies outside [0, chrom_len) */
};

/* Where the caller's buffers live (cuBLAS-style pointer mode).
 * HOST  : every in/out pointer is host memory (pinned from cg_alloc_pinned gives
 *         asynchronous copies); cg_chrom_finish / cg_genotype block until their
 *         outputs are valid.  This is the drop-in mode.
 * DEVICE: every in/out pointer is device memory on the context's GPU and must stay
 *         valid until cg_sync(); nothing is copied and no call blocks. */
enum { CG_PTR_HOST = 0, CG_PTR_DEVICE = 1 };

/* One alignment record, 8 bytes.  Only pos and mapq feed the depth path
 * (bam_data.c:205,213: `if (core.qual > mq_threshold) read_depth[core.pos]++`).
 * len/fl are carried for the --rp gate (bam_data.c:207): fl bit0 secondary,
 * bit1 qc-fail, bit2 duplicate, bit3 supplementary, bit4 unmapped, bit5 reverse. */
typedef struct cg_read {
    int32_t  pos;    /* bam1_core_t.pos, 0-based leftmost */
    uint16_t len;    /* bam1_core_t.l_qseq (saturated at 65535) */
    uint8_t  mapq;   /* bam1_core_t.qual */
    uint8_t  fl;
} cg_read;

/* A known CNV, BED coordinates as parsed by atoi (svs.c:49-53); half-open
 * [start,end) as used by the per-base loop likelihood.c:115. */
typedef struct cg_cnv { int32_t start, end; } cg_cnv;

/* One mappability BED row of the current chromosome in FILE ORDER (svs.c:350-371):
 * paints [start, end] INCLUSIVE with val; later rows overwrite earlier ones. */
typedef struct cg_mapiv { int32_t start, end; float val; uint32_t line; } cg_mapiv;

/* One disjoint run of the flattened mappability track: positions [start, end)
 * (half-open) all carry `val` — what mappability[] holds after svs.c:350-371 has
 * painted every row.  Runs are sorted by start and never overlap; unpainted (zero)
 * stretches are simply absent. */
typedef struct cg_maprun { int32_t start, end; float val; int32_t pad; } cg_maprun;

/* Result per CNV; mirrors the fields of `svs` (svs.h:17-26) that find_depths fills. */
typedef struct cg_call {
    int32_t  observed;      /* observed_rd_sv   likelihood.c:120 */
    float    expected;      /* expected_rd_sv   likelihood.c:119 (ordered fp32 sum) */
    double   mappability;   /* likelihood.c:128; 0 when no mappability was set */
    double   lhomo, lhetero, lnone;   /* likelihood.c:133-135 / :155-157 */
    double   score;         /* likelihood_score likelihood.c:138,160 (int-truncating max) */
    int32_t  copy_number;   /* likelihood.c:146-149 / :164-167 */
    uint32_t margin_flags;  /* bit0: lhomo or lhetero within 1e-9 relative of an integer, bit1: of each other.  For a
                               flagged CNV the (int) truncations in `score` and the lhomo > lhetero test in
                               `copy_number` were taken from a double-double evaluation of the likelihoods */
} cg_call;

#define CG_GC_BINS 101
/* Per-chromosome integer statistics (read_distribution.c:27-84).  As int64[204]
 * this is one row of the table a multi-GPU run all-reduces (cg_stats_table). */
typedef struct cg_chrom_stats {
    int64_t rd_per_gc[CG_GC_BINS];      /* rd_per_gc_unfiltered  read_distribution.c:71 */
    int64_t window_per_gc[CG_GC_BINS];  /* window_per_gc         read_distribution.c:72 (positions, not windows) */
    int64_t rd_cnt;                     /* calc_mu_per_chr       read_distribution.c:35 */
    int64_t reads_passed;               /* total_read_count_unfiltered  bam_data.c:214 */
} cg_chrom_stats;

typedef struct cg_ctx cg_ctx;

/* Context on CUDA device `device`.  NULL on failure (cg_last_error(NULL) says why: the last error of the
 * calling thread). */
cg_ctx*     cg_create(int device);
void        cg_destroy(cg_ctx* ctx);
const char* cg_last_error(cg_ctx* ctx);
int         cg_abi_version(void);
int         cg_set_pointer_mode(cg_ctx* ctx, int mode);
/* HOST pointer mode only.  on != 0: no call blocks any more — every host pointer handed to
 * cg_chrom_begin / cg_push_reads / cg_chrom_finish / cg_set_mappability_runs / cg_genotype must
 * be page-locked (cg_alloc_pinned) and stay untouched until cg_sync(), which is also when the
 * outputs become valid and when CG_ERANGE is reported.  This is how a host that decodes the
 * next chromosome while the GPU works on the current one drives the library.  The small arrays
 * (the statistics row of cg_chrom_finish, the CNVs and calls of cg_genotype) are then read and
 * written by the kernels directly in that page-locked memory, without a copy call. */
int         cg_set_async(cg_ctx* ctx, int on);

/* Chromosomes in flight: chromosome k of a context occupies slot k mod n (its own streams, depth array,
 * count and upload buffers) and a slot is reused when its tenant's genotype kernels are done.  Default and
 * maximum 6 (measured on a 22-chromosome sample: 3 slots 1.48 ms, 4 slots 1.22, 6 slots 1.09); a caller that
 * owns only two or three chromosomes per sample (one rank of an 8-GPU job) loses nothing with 4 and has a
 * third fewer streams to order in every cg_join.  Waits for the context to drain; not between cg_chrom_begin and
 * cg_chrom_finish; the finished chromosome is closed (no further cg_genotype on it). */
int         cg_set_slots(cg_ctx* ctx, int n);

/* Page-locked host memory for read batches / results (library owns it). */
void* cg_alloc_pinned(cg_ctx* ctx, size_t bytes);
void  cg_free_pinned(cg_ctx* ctx, void* p);

/* Start a chromosome: depth array of chrom_len zeros, read count 0
 * (init_rd_per_chr, read_distribution.c:12-18).  gc_profile[w] is SONIC's byte for
 * positions [100w, 100w+100) (sonic_get_gc_content as called from
 * read_distribution.c:70 / likelihood.c:117 reduces to gc_profile[pos/100]);
 * n_windows must be >= ceil(chrom_len/100).  A read passes iff mapq > mq_threshold. */
int cg_chrom_begin(cg_ctx* ctx, int chr_index, int32_t chrom_len,
                   const int8_t* gc_profile, int32_t n_windows, int mq_threshold);

/* Append n records (bam_data.c:201-217).  Coordinate-sorted input (README.md:93)
 * takes the tiled shared-memory path; out-of-order input is detected on the device
 * and handled by the global-atomic path with identical results. */
int cg_push_reads(cg_ctx* ctx, const cg_read* reads, size_t n);

/* The same records as two planes — all that bam_data.c:205,213 reads of a record: pos[i] = core.pos and
 * mapq[i] = core.qual.  5 bytes per record across PCIe instead of 8; mapq may be NULL when the chromosome's
 * mq_threshold is negative (the default, cmdline.c:188-194: every record passes and the plane would never
 * decide anything), which makes it 4.  A chromosome is pushed in one format only. */
int cg_push_reads_compact(cg_ctx* ctx, const int32_t* pos, const uint8_t* mapq, size_t n);

/* Close the chromosome: depth (int16, wraps mod 2^16 like the reference's `short`),
 * chromosome sum and the 101 GC-bin sums/counts, and the
 * expected_read_depth[101] table computed with the reference's exact expression
 * (read_distribution.c:75-83).  `out` may be NULL. */
int cg_chrom_finish(cg_ctx* ctx, cg_chrom_stats* out);

/* expected_read_depth[101] of the current chromosome (bam_info.expected_read_depth,
 * common.h:94).  set overrides what cg_chrom_finish derived (tests, replays). */
int cg_get_expected(cg_ctx* ctx, float expected[CG_GC_BINS]);
int cg_set_expected(cg_ctx* ctx, const float expected[CG_GC_BINS]);

/* Mappability rows of the current chromosome; n == 0 clears (no --mappability).
 * Always a HOST pointer: rows are flattened to disjoint runs on the host (input
 * preparation, like BED parsing) before upload.  Call after cg_chrom_begin. */
int cg_set_mappability(cg_ctx* ctx, const cg_mapiv* rows, size_t n);

/* The same in two steps, for callers that keep tracks resident across samples
 * (batch mode): flatten once on the host (result malloc'ed, free with
 * cg_free_runs), then hand the runs over — a host or device pointer according to
 * the pointer mode.  enabled != 0 with n == 0 means "--mappability given but no
 * row on this chromosome" (mappability 0.0 for every CNV). */
int  cg_flatten_mappability(const cg_mapiv* rows, size_t n, int32_t chrom_len,
                            cg_maprun** runs_out, size_t* n_out);
void cg_free_runs(cg_maprun* runs);
int  cg_set_mappability_runs(cg_ctx* ctx, const cg_maprun* runs, size_t n, int enabled);

/* Genotype n CNVs of the current chromosome (find_depths, likelihood.c:290-308).
 * is_dup = 0: DELETION hypotheses lambda in {0, 0.5E, E}; 1: DUPLICATION {2E, 1.5E, E}.
 * Requires 0 <= start and end <= chrom_len (the reference reads unchecked, A23). */
int cg_genotype(cg_ctx* ctx, const cg_cnv* cnvs, size_t n, int is_dup, cg_call* out);

/* ---- read-pair support from split reads: `--rp` together with `--dups` (split_read.c, bam_data.c:10-154,
 * likelihood.c:41-94).  Per chromosome, after cg_chrom_begin:
 *   cg_sr_reference  <- readReferenceSeq + build_hash_table + create_hash_table   bam_data.c:306-312
 *   cg_sr_push  (xN) <- find_split_reads -> almostPerfect_match_seq_ref           bam_data.c:207-210
 *   cg_sr_rows       <- read_SplitReads + determine_SvType                         bam_data.c:331-335
 *   cg_sr_count      <- count_ReadPairs                                            likelihood.c:344-348
 * All pointers of these four calls are HOST pointers in either pointer mode; cg_sr_rows and cg_sr_count
 * block.  The caller keeps what needs htslib or SONIC: the gate of bam_data.c:207, the average-phred
 * tests of split_read.c:236-246 / :301-311 (which decide whether a half exists at all), and the
 * sonic_is_satellite term of bam_data.c:101-102 on the rows that come back. */

/* One half of a read that passed the gate and its average-phred test (a `splitRead`, 16 bytes). */
typedef struct cg_half {
    int32_t  pos;        /* splitRead.pos: core.pos for the second half of the read (split_read.c:228),
                            core.pos + l_qseq/2 for the first half, the "_read2" element (:291) */
    uint32_t seq_off;    /* offset of the half's bases in the `bases` buffer of the same cg_sr_push call */
    uint16_t len;        /* bases in the half: strlen(str) in almostPerfect_match_seq_ref */
    uint16_t read_len;   /* splitRead.read_length = core.l_qseq */
    uint8_t  mapq;       /* splitRead.qual = core.qual */
    uint8_t  is_read2;   /* readName ends in "_read2" (bam_data.c:40-41) */
    uint16_t pad;
} cg_half;

/* A (half, mapping) pair that passed read_SplitReads' tests other than the satellite one and that
 * determine_SvType classified; coordinates BEFORE the +-SOFTCLIP_WRONGMAP_WINDOW widening of
 * bam_data.c:112-125. */
typedef struct cg_splitrow {
    int32_t pos, pos_map;                                    /* splitRead.pos, posMapSplitRead.posMap */
    int32_t left_start, left_end, right_start, right_end;    /* locMapLeftStart/End, locMapRightStart/End */
    uint8_t sv_type;                                         /* 'D' DELETION, 'E' DUPLICATION (common.h:12-13) */
    uint8_t pad[7];
} cg_splitrow;

/* ref_seq[0..ref_len) is the chromosome as readReferenceSeq leaves it (upper case; common.c:421-458). */
int cg_sr_reference(cg_ctx* ctx, const char* ref_seq, int32_t ref_len);
/* n halves and their bases (one ASCII character per base: A C G T N). */
int cg_sr_push(cg_ctx* ctx, const cg_half* halves, size_t n, const char* bases, size_t n_bases);
/* Rows found since cg_sr_reference, in no particular order.  Copies min(*n_rows, cap) rows; *n_rows is the
 * total.  n_mappings (may be NULL): (half, mapping) pairs listed, reverse-strand ones included. */
int cg_sr_rows(cg_ctx* ctx, cg_splitrow* out, size_t cap, size_t* n_rows, uint64_t* n_mappings);
/* rows: the caller's surviving rows with widened coordinates.  border_rp[n_del], rp[n_dup]: the counts
 * count_ReadPairs adds to svs.border_rp / svs.rp (both start at 0, svs.c:164-167). */
int cg_sr_count(cg_ctx* ctx, const cg_splitrow* rows, size_t n, const cg_cnv* dels, size_t n_del,
                const cg_cnv* dups, size_t n_dup, int32_t* border_rp, int32_t* rp);

/* ---- Delta-coded positions: the record stream at one byte per read ----
 * A coordinate-sorted BAM (README.md:93) hands out non-decreasing positions, a few dozen bases apart at the
 * coverages CONGA is meant for, and all bam_data.c:205,213 reads of a record is pos and qual.  So the host may
 * send, instead of pos[i], delta[i] = min(pos[i] - pos[i-1], 255) as one byte (pos[-1] = 0), the gaps of 255 and
 * more in a side array `big` in stream order, and one anchor per block of CG_DELTA_BLOCK records (the position in
 * front of the block and the number of side-array entries in front of it) so that the device decodes every block
 * on its own: k_decode_delta rebuilds the int32 position plane in HBM with a block scan, and the path continues
 * as for cg_push_reads_compact.  4.0 -> 1.0 bytes per record across PCIe (C3: 201 -> 72 MB per sample).
 * cg_encode_positions_delta is host code for callers that produce positions (it is input preparation, like building
 * cg_read records from bam1_t): returns CG_EINVAL if the positions decrease or are negative (such a stream travels
 * as cg_read records or planes), CG_ERANGE if `big` is too small (*n_big says how many entries are needed). */
#define CG_DELTA_BLOCK 1024
typedef struct cg_delta_anchor { int32_t pos_before; uint32_t big_before; } cg_delta_anchor;
int cg_encode_positions_delta(const int32_t* pos, size_t n, uint8_t* delta /* n */, int32_t* big, size_t big_cap, size_t* n_big,
                              cg_delta_anchor* anchors /* ceil(n / CG_DELTA_BLOCK) */);

/* ---- The loop body of read_bam (bam_data.c:304-338) over several chromosomes as ONE submission ----
 * Job k means exactly: cg_chrom_begin(chr_index, chrom_len, gc, n_windows, mq_threshold); cg_push_reads(reads,
 * n_reads) (or cg_push_reads_compact(reads_pos, reads_mapq, n_reads) when reads == NULL); cg_chrom_finish(stats);
 * cg_set_mappability_runs(runs, n_runs, 1) when use_mappability != 0; cg_genotype(dels, n_del, 0, del_calls);
 * cg_genotype(dups, n_dup, 1, dup_calls) — same pointer mode, same results.
 *
 * In DEVICE pointer mode and in asynchronous HOST mode the jobs are NOT run one after the other: the whole
 * list (in groups of 32) is handed to the device as one job table and each kernel of the path is launched
 * once for all chromosomes of the group — one init, one record sweep, one depth pass whose warps pull the
 * chunks of every chromosome from one counter, one finalize, one mappability index, one genotype launch over
 * all CNVs.  A sample is six launches instead of eight per chromosome, and the pass has one ramp and one tail
 * per sample.  Consecutive calls rotate through the context's slots, so sample k+1's uploads and pass overlap
 * sample k's genotype kernel.  Results are valid after cg_sync() (or after an event recorded on cg_stream
 * behind cg_join()).  A job whose records turn out not to be coordinate-sorted (the reference requires sorted
 * input, README.md:93) is detected on the device and re-run through the per-chromosome entry points inside
 * cg_sync() / when its slot is reused, with identical results; its statistics row is final only then.
 * In blocking HOST mode the call is the plain loop over the per-chromosome entry points. */
typedef struct cg_chrom_job {
    int32_t chr_index, chrom_len;
    const int8_t* gc;          int32_t n_windows; int32_t mq_threshold;
    const cg_read* reads;      uint64_t n_reads;  /* reads == NULL: the compact planes below */
    const cg_maprun* runs;     uint64_t n_runs;   int32_t use_mappability; int32_t flags;   /* CG_JOB_* */
    const cg_cnv* dels;        uint64_t n_del;    cg_call* del_calls;
    const cg_cnv* dups;        uint64_t n_dup;    cg_call* dup_calls;
    cg_chrom_stats* stats;     /* may be NULL */
    const int32_t* reads_pos;  const uint8_t* reads_mapq;    /* see cg_push_reads_compact */
    /* reads == NULL and reads_pos == NULL: the positions as a delta-coded stream (cg_encode_positions_delta);
     * reads_mapq goes with it as with reads_pos.  Submissions only (DEVICE and asynchronous HOST mode). */
    const uint8_t* reads_delta; const int32_t* reads_big; uint64_t n_big; const struct cg_delta_anchor* reads_anchor;
} cg_chrom_job;
/* flags.  CG_JOB_ANNOTATION_RESIDENT: HOST pointer mode only, submissions only (DEVICE and asynchronous HOST
 * mode) — gc, runs, dels and dups are DEVICE pointers although the records and the results are host memory:
 * what a batch driver does with the inputs that are the same for every sample (SONIC's GC bytes, the CNV set,
 * the flattened mappability track).  Only the records cross PCIe. */
#define CG_JOB_ANNOTATION_RESIDENT 1
int cg_run_chromosomes(cg_ctx* ctx, const cg_chrom_job* jobs, size_t n_jobs);

/* ---- SONIC's GC profile from the bases of a chromosome ----
 * Replaces the character loop of sonic_write_gc_profile (sonic/sonic_reference.c:30-106) for ONE FASTA
 * record.  bases: the record's n_bases characters with white space removed (what faidx_fetch_seq returns),
 * either case.  out_gc: floor((n_bases + 1) / 100) bytes, byte w = the number of G/C/g/c among bases
 * [100w - 1, 100w + 99) (window 0: [0, 99) — char_count starts at 1, :82), the value the builder writes
 * as (char)(100.0 * gc / 100) (:96) and sonic_read_gc_profile (:139-159) hands to CONGA.  n_windows: capacity
 * of out_gc.  Pointers follow the pointer mode; in HOST mode the call blocks, in DEVICE mode it is queued
 * on cg_stream.  Needs no chromosome to be open.  DEVICE mode: the kernel's bulk copies start at `bases` rounded
 * down to 16 bytes and end at `bases + n_bases` rounded up, so the buffer needs up to 15 addressable bytes on either
 * side (any cudaMalloc block has them; a pointer at the very end of a mapping has not). */
int cg_gc_profile(cg_ctx* ctx, const char* bases, int64_t n_bases, int8_t* out_gc, int32_t n_windows);

/* Wait for everything queued on the context (all of its streams).  In HOST pointer mode it
 * returns CG_ERANGE if a record or interval was clipped since the last report. */
int cg_sync(cg_ctx* ctx);
/* Order the context's main stream (cg_stream) behind everything queued so far on its
 * internal copy and genotype streams, without blocking the host: work or events queued on
 * cg_stream afterwards (a collective, a timing event) see all results. */
int cg_join(cg_ctx* ctx);

/* Parity/diagnostic downloads of the current chromosome (HOST pointers always). */
int cg_download_depth(cg_ctx* ctx, int16_t* out /* chrom_len */);
/* P[i] = sum of depth[0..i] mod 2^32.  No such array exists on the product path (cg_genotype sums a CNV
 * from per-2048-position depth sums and the depth array); this call rebuilds it on the device for tests. */
int cg_download_prefix(cg_ctx* ctx, int32_t* out /* chrom_len */);
/* Device addresses of the current chromosome's arrays (valid until the next
 * cg_chrom_begin): int16 depth, float expected[101]. */
int cg_device_arrays(cg_ctx* ctx, const int16_t** depth, const float** expected);

/* Device address of the int64[n_rows][204] statistics table of the current bank, one row per
 * chr_index seen since cg_stats_reset (layout = cg_chrom_stats).  With a communicator (cg_comm_init_*)
 * this is the table cg_stats_allreduce fills: every rank holds every chromosome's row. */
int cg_stats_table(cg_ctx* ctx, void** device_ptr, int32_t* n_rows);
int cg_stats_reset(cg_ctx* ctx);
/* Four tables ("banks", 0..3; 0 at creation).  The chromosomes begun after cg_stats_bank write
 * their rows to `bank`, and cg_stats_table / cg_stats_row / cg_stats_allreduce refer to it.  A driver that
 * processes sample after sample rotates through the banks (k mod 4): the collective over sample k's table is then in
 * flight while the chromosomes of the next three samples already fill the other banks.  Selecting a bank clears nothing: a row holds
 * what the last chromosome with that chr_index left there (cg_stats_reset zeroes the current bank); a
 * chromosome that writes into a bank a collective is still reading waits for that collective on the device. */
int cg_stats_bank(cg_ctx* ctx, int bank);
int cg_stats_row(cg_ctx* ctx, int chr_index, cg_chrom_stats* out);

/* ---- the path's one collective: every rank receives every chromosome's statistics row ----
 * One context per GPU; the contexts of a job form a communicator (NCCL over NVLink, loaded with dlopen when
 * the first of these calls is made: a single-GPU user needs no NCCL).  cg_comm_unique_id fills 128 bytes on
 * one rank, which the caller hands to every rank (any out-of-band channel); cg_comm_init_rank is then called
 * by all ranks (one process or thread per GPU).  cg_comm_init_all is the single-process form for n contexts
 * driven by one process.  cg_stats_allreduce orders cg_stream behind everything queued on the context
 * (cg_join) and queues ONE ncclAllReduce(int64, sum) of rows [0, n_rows) of the current bank on it: out of
 * place, from the table this context's chromosomes wrote (rows of other ranks' chromosomes stay zero there)
 * into the table cg_stats_table / cg_stats_row then refer to.  Nothing blocks; every rank must pass the same
 * n_rows.  The reduced rows are bookkeeping (conga.log, read_distribution.c:41): no genotype of any
 * chromosome depends on another chromosome's row (read_distribution.c:49-84 rebuilds the table per chromosome). */
/* ---- the same exchange fused into the path's own kernel: peer stores over NVLink instead of a collective kernel ----
 * The rows are disjoint per rank (a chromosome is owned by one rank), so nothing has to be reduced: with peer
 * memory connected, k_finalize — which has just built a chromosome's row — stores it straight into the table of
 * EVERY rank (its own included), 1632 bytes per peer over NVLink / NVSwitch, and cg_stats_allreduce shrinks to a
 * one-warp kernel on cg_stream that tells every peer "my rows of this sample are in your table" and waits for
 * the same word from all of them.  cg_stats_table / cg_stats_row read the table the peers write.  Measured on
 * B200: the NCCL all-reduce of the 36 KB table costs a step 29 us (it needs an empty SM under a persistent pass);
 * the flag exchange does not.  NCCL stays available (cg_comm_*) and is what the two are checked against.
 *   one process per GPU:  cg_peer_export fills CG_PEER_HANDLE_BYTES (a CUDA IPC handle) on every rank; the
 *                         handles of all ranks, in rank order, go to cg_peer_connect on every rank;
 *   one process, n GPUs:  cg_peer_connect_local(ctxs, n), rank = position in ctxs.
 * A bank's rows are overwritten by a peer as soon as that peer starts the sample that uses the bank again (four
 * samples later): a driver that reads rows on the host does so before it lets any rank go that far. */
#define CG_PEER_HANDLE_BYTES 64
#define CG_MAX_PEERS 16
int cg_peer_export(cg_ctx* ctx, void* handle /* CG_PEER_HANDLE_BYTES */);
int cg_peer_connect(cg_ctx* ctx, int n_ranks, int rank, const void* handles /* n_ranks x CG_PEER_HANDLE_BYTES */);
int cg_peer_connect_local(cg_ctx** ctxs, int n);
int cg_peer_disconnect(cg_ctx* ctx);

#define CG_COMM_ID_BYTES 128
int cg_comm_unique_id(void* id /* CG_COMM_ID_BYTES */);
int cg_comm_init_rank(cg_ctx* ctx, int n_ranks, int rank, const void* id);
int cg_comm_init_all(cg_ctx** ctxs, int n);
int cg_comm_destroy(cg_ctx* ctx);
int cg_stats_allreduce(cg_ctx* ctx, int32_t n_rows);
/* the same for the n contexts of cg_comm_init_all, issued by one thread (inside one NCCL group) */
int cg_stats_allreduce_all(cg_ctx** ctxs, int n, int32_t n_rows);

/* ---- the depth side of one chromosome without any allocation ----
 * No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels and
 * same results as cg_chrom_begin + cg_push_reads_compact + cg_chrom_finish (+ cg_get_expected, cg_download_depth)
 * in DEVICE pointer mode, but needs no context: every scratch array of the kernels, the int16 depth array
 * included, lives in one fixed 17.4 MB __device__ array of the library (held by every GPU that loads the library
 * until the process exits).  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine)
 * and does not block: three kernels, one clearing memset and the copy of the depth array are queued on `stream`
 * (a cudaStream_t; NULL = the default stream) of GPU `device`.
 *   pos, mapq      device: the records' int32 positions, NON-DECREASING, and their uint8 mapq plane (NULL iff
 *                  mq_threshold < 0: every record passes), n_reads of each.  There is no out-of-order path here:
 *                  for unsorted records bit 1 (unsorted) is set in *flags_out and the outputs are undefined.
 *   chrom_len      1 .. CG_NOALLOC_MAX_LEN; anything larger is rejected with CG_EINVAL (use a context then).
 *   gc_profile     device: SONIC GC byte per 100-position window, n_windows >= ceil(chrom_len / 100).
 *   depth_out      device int16[chrom_len] (read_depth[], bam_data.c:205-215) or NULL.
 *   stats_out      device cg_chrom_stats or NULL;  expected_out: device float[CG_GC_BINS] or NULL.
 *   flags_out      device uint32 or NULL: 0, or bit 1 = a record was out of order, bit 2 = a record outside
 *                  [0, chrom_len).
 * All calls share the one array, which keeps the previous call's contents (what must start at zero is cleared at
 * the head of every call): calls queued on ONE stream are ordered and safe; calls that overlap on a GPU — two
 * streams, two threads — race on it.  Errors: cg_last_error(NULL). */
#define CG_NOALLOC_MAX_LEN 8386560          /* 4095 chunks of 2048 positions */
int cg_depth_stats_noalloc(int device, void* stream, const int32_t* pos, const uint8_t* mapq, size_t n_reads,
                           int mq_threshold, int32_t chrom_len, const int8_t* gc_profile, int32_t n_windows,
                           int16_t* depth_out, cg_chrom_stats* stats_out, float* expected_out, uint32_t* flags_out);

/* Device time of this context's kernels, from CUDA events recorded on its stream
 * around every launch group while profiling is on.  cg_kernel_times waits for the
 * stream, returns the milliseconds and launch counts accumulated since the last
 * call, and clears them.  Slots: 0 = submission init + record sweep + out-of-order scatter + finalize,
 * 1 = the fused depth/GC pass, 2 = mappability index + CNV sums + likelihoods (on the genotype stream,
 * overlapping slot 1 of the next submission), 3 = the split-read kernels (k-mer index, re-mapping,
 * overlap counts). */
#define CG_NSLOTS 4
int cg_set_profiling(cg_ctx* ctx, int on);
int cg_kernel_times(cg_ctx* ctx, double ms[CG_NSLOTS], int64_t launches[CG_NSLOTS]);
/* Number of kernels this context has launched since creation. */
int64_t cg_launch_count(cg_ctx* ctx);
/* The context's main CUDA stream (cudaStream_t): the depth pass runs on it; see cg_join. */
void* cg_stream(cg_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* CONGA_GPU_H */
