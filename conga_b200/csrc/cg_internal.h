/* cg_internal.h — shared between the translation units of libconga_gpu.so.
 * Not part of the ABI (include/conga_gpu.h is). */
#ifndef CG_INTERNAL_H
#define CG_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/conga_gpu.h"

/* ---- geometry of the fused depth pass (K1+K2+K3) -------------------------
 * The unit of work is a CHUNK of 2048 positions, owned by one warp (4 KB of packed u16 bins
 * in shared memory).  Depth sums are kept at three levels so that the depth sum in front of a
 * chunk is three short coalesced loads: per chunk, per GROUP of 32 chunks, per SUPER of 32 groups.
 *
 * A SUBMISSION is what one set of kernel launches works on: up to CG_JMAX chromosomes ("jobs").
 * cg_run_chromosomes hands a whole sample over as one submission (one init, one mark, one pass,
 * one finalize, one run-index and one genotype launch for all of its chromosomes); the
 * per-chromosome entry points (cg_chrom_begin ... cg_genotype) are submissions of one job.
 * The jobs of a submission share the slot's chunk space: job j owns the chunks
 * [chunk0, chunk0 + n_chunks] of the slot's depth / count arrays, chunk0 a multiple of 1024 so that
 * groups and supers never straddle two chromosomes. */
#define CG_CHUNK        2048
#define CG_CHUNK_SHIFT  11
#define CG_GROUP_SHIFT  16            /* 32 chunks  = 65,536 positions   */
#define CG_SUPER_SHIFT  21            /* 32 groups  = 2,097,152 positions */
#define CG_SUPER_CHUNKS 1024
#define CG_PASS_THREADS 256
#define CG_STATS_WORDS 204            /* int64 words per chromosome row = sizeof(cg_chrom_stats)/8 */
#define CG_JMAX         32            /* chromosomes one submission carries */
#define CG_PART_WORDS   (2 * CG_GC_BINS)      /* int32 words one CTA of the pass leaves per job: 101 sums, 101 counts */

/* Device-side condition bits, first control word of a job. */
#define CG_FLAG_UNSORTED   1u         /* a pushed record had pos < its predecessor */
#define CG_FLAG_RANGE      2u         /* a record or CNV outside [0, chrom_len) */
#define CG_FLAG_PEER_TIMEOUT 8u       /* a peer's word did not arrive within the exchange kernel's patience */
/* u32 control words of a job (cg_dev_job::ctl) */
#define CG_CTL_WORDS 8
#define CG_W_FLAGS   0
#define CG_W_PASSED  2                /* u64 (words 2,3): records with mapq > threshold, bam_data.c:214 */
/* u32 control words of a slot (cg_sub::sctl) */
#define CG_SW_TILE   0                /* dynamic work counter of the depth pass */
#define CG_SW_GENO   1                /* ... of the genotype kernel */

struct cg_counts {                    /* one allocation per slot, see cgk_counts_layout; indices are chunks of the slot's chunk space */
    int32_t* chunk_lo;                /* [n_space] first record index of each chunk (sorted input) */
    int32_t* cnt1;                    /* [n_space]   sum of the (int16) depths of each chunk */
    int32_t* cnt2;                    /* [n_space/32]   ... of each group of 32 chunks */
    int32_t* cnt3;                    /* [n_space/1024] ... of each super of 32 groups */
};

/* One chromosome of a submission, as the kernels see it. */
struct cg_dev_job {
    /* records: either 8-byte cg_read records or the compact planes (cg_push_reads_compact) */
    const cg_read*  reads;
    const int32_t*  rec_pos;
    const uint8_t*  rec_mapq;         /* may be NULL with rec_pos when mq_threshold < 0: every record passes */
    const uint8_t*  rec_delta;        /* delta-coded positions (cg_encode_positions_delta) that k_decode_delta turns into ... */
    const int32_t*  rec_big;
    const cg_delta_anchor* rec_anchor;
    int32_t*        rec_pos_out;      /* ... this position plane (= rec_pos), or NULL */
    long long       n_reads;          /* records of the chromosome so far */
    long long       mark_first, mark_n;   /* k_mark: the batch this launch looks at */
    int32_t         mark_blk0;        /* first block visit of this job in the k_mark launch (cgk_mark_blocks) */
    int32_t         mq_threshold;
    /* geometry */
    int32_t         chrom_len, n_chunks;
    int32_t         chunk0;           /* first chunk in the slot's chunk space, multiple of CG_SUPER_CHUNKS */
    int32_t         work0;            /* first work index of the pass (prefix of n_chunks over the jobs) */
    const int8_t*   gc;               /* SONIC GC byte per 100-position window */
    /* depth-side results */
    long long*      stats_row;        /* this chromosome's int64[204] row in the statistics bank */
    long long*      row_out;          /* NULL, or where k_finalize also writes the row: page-locked host memory
                                         (asynchronous HOST mode) or the caller's device buffer (DEVICE mode) */
    float*          expected;         /* [128] expected_read_depth[101] */
    unsigned int*   ctl;              /* CG_W_* words of this job */
    /* genotype side */
    const cg_maprun* runs;
    int32_t*        run_idx;          /* first run per 16384 positions */
    int32_t         n_runs, has_map;
    int32_t         run_blk0;         /* first 256-run block of this job in the k_run_index launch */
    int32_t         chr_index;
    const cg_cnv*   cnv[2];           /* [0] deletions, [1] duplications */
    cg_call*        calls[2];
    int32_t         n_cnv[2];
    int32_t         cnv0[2];          /* first group (cgk_geno_groups) of each kind in the k_genotype launch */
};

/* What a set of launches works on.  jobs == NULL: a single job travels as the second kernel parameter. */
struct cg_sub {
    const cg_dev_job* jobs;
    int32_t        n_jobs;
    int32_t        n_work;            /* chunks of all jobs */
    int32_t        n_space;           /* chunks of the slot's chunk space in use (multiple of CG_SUPER_CHUNKS) */
    int32_t        n_mark_blocks, n_run_blocks, n_cnv;   /* n_cnv: CNV groups of the launch */
    int32_t        geno_g;            /* CNVs per group = per warp of k_genotype */
    cg_counts      k;
    int16_t*       depth;             /* [n_space * CG_CHUNK] */
    uint32_t*      depth32;           /* NULL, or the out-of-order path's per-position u32 counts (zero at rest) */
    int32_t*       partial;           /* [pass CTAs][n_jobs][CG_PART_WORDS] */
    int32_t        n_pass_ctas;
    unsigned int*  sctl;              /* CG_SW_* words of the slot (zero at rest) */
    long long*     fin_acc;           /* [CG_JMAX][256] k_finalize's scratch rows (zero at rest) */
    unsigned int*  fin_cnt;           /* [CG_JMAX] CTAs of k_finalize that have added their share (zero at rest) */
    unsigned int*  sticky;            /* context-wide word: CG_FLAG_RANGE is copied here */
    unsigned int*  h_flags;           /* NULL, or page-locked [n_jobs]: each job's flag word for the host */
    int32_t        n_peer_rows;       /* 0, or ranks whose tables k_finalize stores a finished row into (this rank included) */
    long long*     peer_rows[CG_MAX_PEERS];   /* row 0 of every rank's table for the current bank (peer memory over NVLink) */
};

/* Checked build (-DCG_CHECKED, scripts/build_checked.sh): every index the kernels form is asserted against the extent of
 * the array it goes into; a violation traps the kernel and the next CUDA call fails.  compute-sanitizer is not available on
 * the B200 pool this was developed on, so the parity suite is run once against this build instead. */
#ifdef CG_CHECKED
#include <assert.h>
#define CG_CHECK(cond) assert(cond)
#else
#define CG_CHECK(cond) ((void)0)
#endif

#ifdef __cplusplus
#include <vector>
extern "C++" {
/* conga_gpu.cu / flatten.cu */
int  cg_fail(cg_ctx* c, int code, const char* fmt, ...);
void cg_flatten_rows(const cg_mapiv* rows, size_t n, int32_t chrom_len, std::vector<cg_maprun>& out);
/* depth_pass.cu */
size_t      cgk_counts_words(int n_space);                        /* int32 words the counts of a chunk space need */
cg_counts   cgk_counts_layout(int32_t* base, int n_space);
int         cgk_job_space(int n_chunks);                          /* chunks of chunk space a job of n_chunks occupies */
int         cgk_pass_grid(int n_work, long long n_records, int n_sms);   /* CTAs of the pass, = rows of cg_sub::partial */
int         cgk_mark_blocks(unsigned long long n_records);        /* block visits of k_mark for a batch of records */
cudaError_t cgk_mark(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1);
cudaError_t cgk_decode_delta(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1);
cudaError_t cgk_rare_scatter(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1, int n_sms);
cudaError_t cgk_depth_pass(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1);
cudaError_t cgk_finalize(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1);
struct cg_peer_flags { unsigned int* theirs[CG_MAX_PEERS]; unsigned int* mine; int n, rank; unsigned epoch; unsigned int* sticky; };
cudaError_t cgk_peer_exchange(cudaStream_t s, const cg_peer_flags& f);
cudaError_t cgk_prefix_from_depth(cudaStream_t s, const int16_t* depth, cg_counts k, int chunk0, int n_chunks, int32_t* prefix);
cudaError_t cgk_expected_table(cudaStream_t s, const long long* stats_row, float* expected);
int         cgk_carveout(void);                                   /* shared-memory carve-out (percent) every kernel asks for */
cudaError_t cgk_configure_depth(void);                            /* once per process, after the device is selected */
cudaError_t cgk_configure_genotype(void);

/* genotype.cu */
int         cgk_geno_g(unsigned long long n_cnv, int n_sms);      /* CNVs per warp for a launch over n_cnv CNVs */
int         cgk_geno_groups(unsigned long long n_cnv, int g);     /* warps' worth of work n CNVs of one (job, kind) are */
cudaError_t cgk_genotype(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1, int n_sms);
int         cgk_run_index_words(int32_t chrom_len);
cudaError_t cgk_run_index(cudaStream_t s, const cg_sub& S, const cg_dev_job& J1);
/* split_read.cu */
cudaError_t cgk_sr_index(cudaStream_t s, const unsigned char* ref, int n, int* cnt, int* pos_list, int* lo, int* hi, int n_sms);
cudaError_t cgk_sr_search(cudaStream_t s, const unsigned char* ref, int ref_len, const int* pos_list, const int* lo, const int* hi,
                          const cg_half* halves, long long n_halves, const unsigned char* bases, int chrom_len, int mq_threshold,
                          cg_splitrow* rows, unsigned* n_rows, unsigned cap, unsigned long long* n_mappings, int n_sms);
cudaError_t cgk_sr_count(cudaStream_t s, const cg_splitrow* rows, int n_rows, const cg_cnv* dels, int n_del, const cg_cnv* dups, int n_dup,
                         int* border_rp, int* rp, int n_sms);
cudaError_t cgk_configure_split(void);

/* gc_profile.cu — SONIC's GC bytes from the bases of one chromosome (sonic/sonic_reference.c:30-106) */
cudaError_t cgk_gc_profile(cudaStream_t s, const char* bases, long long n_bases, int8_t* out, int n_windows, int n_sms);
cudaError_t cgk_configure_gc_profile(void);
}
#endif

#endif /* CG_INTERNAL_H */
