/* cg_internal.h — shared between the translation units of libconga_gpu.so.
 * Not part of the ABI (include/conga_gpu.h is). */
#ifndef CG_INTERNAL_H
#define CG_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/conga_gpu.h"

/* ---- geometry of the fused depth pass (K1+K2+K3) -------------------------
 * The unit of work is a CHUNK of 2048 positions, owned by one warp (4 KB of packed u16 bins
 * in shared memory).  Record counts are kept at three levels so that a chunk's scan carry is
 * three short coalesced loads: per chunk, per GROUP of 32 chunks, per SUPER of 32 groups. */
#define CG_CHUNK        2048
#define CG_CHUNK_SHIFT  11
#define CG_GROUP_SHIFT  16            /* 32 chunks  = 65,536 positions   */
#define CG_SUPER_SHIFT  21            /* 32 groups  = 2,097,152 positions */
#define CG_PASS_THREADS 256
#define CG_PILEUP_PROBE 16384         /* k_mark_batch looks this many records back for an equal pos */
#define CG_STATS_WORDS 204            /* int64 words per chromosome row = sizeof(cg_chrom_stats)/8 */

/* Device-side error/condition bits, one u32 word per context. */
#define CG_FLAG_UNSORTED   1u         /* a pushed record had pos < its predecessor */
#define CG_FLAG_RANGE      2u         /* a record or CNV outside [0, chrom_len) */
#define CG_FLAG_PILEUP     4u         /* > CG_PILEUP_PROBE records share one pos: int16 wrap possible */
/* words of the per-context u32 control block (d_flags) */
#define CG_W_FLAGS   0
#define CG_W_TILE    1                /* dynamic chunk counter of the depth pass */
#define CG_W_DONE    2                /* CTAs of the depth pass that have flushed their statistics */

struct cg_counts {                    /* one allocation per slot, see cgk_counts_layout */
    int32_t* chunk_lo;                /* [n_chunks+1] first record index of each chunk (sorted input) */
    int32_t* cnt1;                    /* [n_chunks]   sum of the (int16) depths of each chunk */
    int32_t* cnt2;                    /* [n_groups]   ... of each group of 32 chunks */
    int32_t* cnt3;                    /* [n_supers]   ... of each super of 32 groups */
};

struct cg_pass_args {
    const cg_read* reads;             /* all records of the chromosome, device */
    long long      n_reads;
    cg_counts      k;
    const int8_t*  gc;                /* SONIC GC byte per 100-position window */
    int32_t        chrom_len;
    int32_t        n_chunks;
    int            mq_threshold;
    int16_t*       depth;             /* [n_chunks*CG_CHUNK] */
    uint32_t*      depth32;           /* unsorted fallback: per-position u32 counts (zero at rest) */
    unsigned int*  ctl;               /* CG_W_* control words */
    long long*     stats;             /* this chromosome's int64[204] row */
    long long*     host_row;          /* NULL, or page-locked host memory the last CTA copies the finished row to */
    float*         expected;          /* [101] written by the last CTA to finish */
    unsigned int*  sticky;            /* context-wide word: CG_FLAG_RANGE is copied here */
};

#ifdef __cplusplus
#include <vector>
extern "C++" {
/* conga_gpu.cu / flatten.cu */
int  cg_fail(cg_ctx* c, int code, const char* fmt, ...);
void cg_flatten_rows(const cg_mapiv* rows, size_t n, int32_t chrom_len, std::vector<cg_maprun>& out);
/* depth_pass.cu */
size_t      cgk_counts_words(int n_chunks);                       /* int32 words one slot's counts need */
cg_counts   cgk_counts_layout(int32_t* base, int n_chunks);
cudaError_t cgk_chrom_init(cudaStream_t s, int32_t* counts_base, int n_chunks, long long* stats_row, unsigned int* ctl);
cudaError_t cgk_mark_batch(cudaStream_t s, const cg_read* reads, long long first, long long n,
                           cg_counts k, int n_chunks, int32_t chrom_len, int mq_threshold,
                           unsigned int* ctl, long long* stats_row);
cudaError_t cgk_rare_paths(cudaStream_t s, const cg_read* reads, long long n_reads, uint32_t* depth32,
                           cg_counts k, int n_chunks, int32_t chrom_len, int mq_threshold,
                           const unsigned int* ctl, int n_sms);
cudaError_t cgk_depth_pass(cudaStream_t s, const cg_pass_args& a, int n_sms);
cudaError_t cgk_prefix_from_depth(cudaStream_t s, const int16_t* depth, cg_counts k, int n_chunks, int32_t* prefix);
cudaError_t cgk_expected_table(cudaStream_t s, const long long* stats_row, float* expected);
int         cgk_depth_pass_ctas_per_sm(void);
int         cgk_carveout(void);                                   /* shared-memory carve-out (percent) every kernel asks for */
cudaError_t cgk_configure_depth(void);                            /* once per process, after the device is selected */
cudaError_t cgk_configure_genotype(void);

/* genotype.cu */
cudaError_t cgk_genotype(cudaStream_t s, const cg_cnv* cnvs, int n, int is_dup,
                         const int16_t* depth, const cg_counts& counts, const int8_t* gc, const float* expected,
                         const cg_maprun* runs, const int32_t* run_idx, int n_runs, int has_map,
                         int32_t chrom_len, unsigned int* flags, cg_call* out, int n_sms);
/* split_read.cu */
size_t      cgk_sr_sort_temp_bytes(int n);
cudaError_t cgk_sr_index(cudaStream_t s, const unsigned char* ref, int n, unsigned* keys_a, unsigned* keys_b, int* vals_a, int* vals_b,
                         void* temp, size_t temp_bytes, int* lo, int* hi, int n_sms);
cudaError_t cgk_sr_search(cudaStream_t s, const unsigned char* ref, int ref_len, const int* pos_sorted, const int* lo, const int* hi,
                          const cg_half* halves, long long n_halves, const unsigned char* bases, int chrom_len, int mq_threshold,
                          cg_splitrow* rows, unsigned* n_rows, unsigned cap, unsigned long long* n_mappings, int n_sms);
cudaError_t cgk_sr_count(cudaStream_t s, const cg_splitrow* rows, int n_rows, const cg_cnv* dels, int n_del, const cg_cnv* dups, int n_dup,
                         int* border_rp, int* rp, int n_sms);
cudaError_t cgk_configure_split(void);

/* gc_profile.cu — SONIC's GC bytes from the bases of one chromosome (sonic/sonic_reference.c:30-106) */
cudaError_t cgk_gc_profile(cudaStream_t s, const char* bases, long long n_bases, int8_t* out, int n_windows, int n_sms);
cudaError_t cgk_configure_gc_profile(void);
int         cgk_run_index_words(int32_t chrom_len);
cudaError_t cgk_run_index(cudaStream_t s, const cg_maprun* runs, int n_runs, int32_t* run_idx, int32_t chrom_len);
}
#endif

#endif /* CG_INTERNAL_H */
