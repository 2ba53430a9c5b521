/* ref_dump_shim.c — link-time shim around the UNMODIFIED reference objects.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/conga_oracle.c header).  Built only where
 * /root/reference exists (oracle/Makefile target `ref`), into oracle/_ref/.
 *
 * Linked with  -Wl,--wrap=find_SVs -Wl,--wrap=free_SVs  so that the reference's
 * own cross-TU calls land here first:
 *   bam_data.c:338  -> find_SVs : at entry in_bam->read_depth, ->mean,
 *       ->expected_read_depth[101] and ->total_read_count_unfiltered are final.
 *   likelihood.c:369-370 -> free_SVs : at entry the svs array still holds every
 *       per-CNV field calculate_likelihood_CNV (likelihood.c:108-169) wrote.
 * Both are dumped to $CONGA_DUMP_DIR (if set) and the real function is called.
 * No reference source is edited or copied; the reference headers are only
 * #included for the struct layouts.
 *
 * Files written per chromosome index K:
 *   chrK.meta   text: chr_name, chr_index, length, mean(%.9g), total_read_count
 *   chrK.depth  int16[length]
 *   chrK.exp    float[101]
 *   chrK.del / chrK.dup  packed records of struct dump_call (below)
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

#include "common.h"
#include "svs.h"

void __real_find_SVs(bam_info*, parameters*, FILE*, FILE*, FILE*, char*, int);
void __real_free_SVs(svs*, int);

struct dump_call {
    int32_t start, end, observed; float expected;
    int32_t rp, border_rp, copy_number, pad;
    double score, lhomo, lhetero, lnone, mappability;
};

static int g_chr_index = -1;

static FILE* dump_open(const char* suffix)
{
    const char* dir = getenv("CONGA_DUMP_DIR");
    if (!dir) return NULL;
    char path[2048];
    snprintf(path, sizeof path, "%s/chr%d.%s", dir, g_chr_index, suffix);
    FILE* f = fopen(path, "wb");
    if (!f) { perror(path); exit(9); }
    return f;
}

void __wrap_find_SVs(bam_info* in_bam, parameters* params, FILE* fp_del, FILE* fp_dup,
                     FILE* fp_SVs, char* chr_name, int chr_index)
{
    g_chr_index = chr_index;
    int len = params->this_sonic->chromosome_lengths[chr_index];
    FILE* f;
    if ((f = dump_open("meta"))) {
        fprintf(f, "%s\t%d\t%d\t%.9g\t%d\n", chr_name, chr_index, len,
                (double)in_bam->mean, in_bam->total_read_count_unfiltered);
        fclose(f);
    }
    if ((f = dump_open("depth"))) { fwrite(in_bam->read_depth, sizeof(short), len, f); fclose(f); }
    if ((f = dump_open("exp")))   { fwrite(in_bam->expected_read_depth, sizeof(float), 101, f); fclose(f); }
    __real_find_SVs(in_bam, params, fp_del, fp_dup, fp_SVs, chr_name, chr_index);
}

void __wrap_free_SVs(svs* arr, int count)
{
    if (arr != NULL && count > 0) {
        FILE* f = dump_open(arr[0].SV_type == DELETION ? "del" : "dup");
        if (f) {
            for (int i = 0; i < count; i++) {
                struct dump_call d;
                memset(&d, 0, sizeof d);
                d.start = arr[i].start; d.end = arr[i].end;
                d.observed = arr[i].observed_rd_sv; d.expected = arr[i].expected_rd_sv;
                d.rp = arr[i].rp; d.border_rp = arr[i].border_rp; d.copy_number = arr[i].copy_number;
                d.score = arr[i].likelihood_score; d.lhomo = arr[i].lhomo;
                d.lhetero = arr[i].lhetero; d.lnone = arr[i].lnone; d.mappability = arr[i].mappability;
                fwrite(&d, sizeof d, 1, f);
            }
            fclose(f);
        }
    }
    __real_free_SVs(arr, count);
}
