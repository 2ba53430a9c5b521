/* mkbam.c — cg_read record files -> coordinate-sorted BAM + .bai, through libhts.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/conga_oracle.c header).  The image has no
 * samtools/pysam, so fixtures for the reference `conga` binary (and for our own
 * host CLI) are produced with this small writer linked against the libhts.a that
 * oracle/Makefile compiles from the reference's htslib submodule.  Written by us;
 * it contains no reference code.
 *
 * usage: mkbam OUT.bam CHROMS.tsv
 *   CHROMS.tsv: one line per @SQ in BAM header order:  name <TAB> length <TAB> records.bin|- [<TAB> seq.bin]
 *   records.bin holds cg_read records (include/conga_gpu.h) sorted by pos.
 *   seq.bin (optional, for the --rp split-read path): per record, in order, `len` ASCII bases (ACGTN)
 *   followed by `len` raw phred bytes; without it SEQ is all N and QUAL absent.
 * Each record becomes one single-end alignment: QNAME r<n>, FLAG from cg_read.fl
 * (bit0 secondary 0x100, bit1 qc-fail 0x200, bit2 duplicate 0x400, bit3
 * supplementary 0x800, bit4 unmapped 0x4, bit5 reverse 0x10), POS, MAPQ, CIGAR
 * <len>M, SEQ = len x 'N', QUAL = 0xff, no aux.  Header carries
 * "@RG ID:1 SM:synth" so get_sample_name (common.c:325) finds a sample name.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

#include "htslib/sam.h"
#include "htslib/hts.h"
#include "htslib/kstring.h"
#include "conga_gpu.h"

#define MAXCHR 4096

static void die(const char* m, const char* a) { fprintf(stderr, "mkbam: %s %s\n", m, a ? a : ""); exit(2); }

int main(int argc, char** argv)
{
    if (argc != 3) die("usage: mkbam OUT.bam CHROMS.tsv", NULL);
    static char names[MAXCHR][256], paths[MAXCHR][2048], seqs[MAXCHR][2048];
    static long lens[MAXCHR];
    int n_chr = 0;
    FILE* f = fopen(argv[2], "r");
    if (!f) die("cannot open", argv[2]);
    char line[4096];
    while (fgets(line, sizeof line, f)) {
        if (line[0] == '\n' || line[0] == '#') continue;
        if (n_chr >= MAXCHR) die("too many chromosomes", NULL);
        seqs[n_chr][0] = 0;
        if (sscanf(line, "%255s %ld %2047s %2047s", names[n_chr], &lens[n_chr], paths[n_chr], seqs[n_chr]) < 3) die("bad line", line);
        n_chr++;
    }
    fclose(f);

    kstring_t txt = {0, 0, NULL};
    ksprintf(&txt, "@HD\tVN:1.6\tSO:coordinate\n");
    for (int i = 0; i < n_chr; i++) ksprintf(&txt, "@SQ\tSN:%s\tLN:%ld\n", names[i], lens[i]);
    ksprintf(&txt, "@RG\tID:1\tSM:synth\n");
    sam_hdr_t* h = sam_hdr_parse(txt.l, txt.s);
    if (!h) die("sam_hdr_parse failed", NULL);
    h->l_text = txt.l; h->text = txt.s;

    samFile* out = sam_open(argv[1], "wb");
    if (!out) die("cannot create", argv[1]);
    if (sam_hdr_write(out, h) < 0) die("sam_hdr_write failed", NULL);

    bam1_t* b = bam_init1();
    uint64_t serial = 0;
    for (int tid = 0; tid < n_chr; tid++) {
        if (!strcmp(paths[tid], "-")) continue;
        FILE* r = fopen(paths[tid], "rb");
        if (!r) die("cannot open", paths[tid]);
        FILE* sq = seqs[tid][0] ? fopen(seqs[tid], "rb") : NULL;
        if (seqs[tid][0] && !sq) die("cannot open", seqs[tid]);
        static unsigned char bases[65536], quals[65536];
        cg_read q; int32_t last = -1;
        while (fread(&q, sizeof q, 1, r) == 1) {
            if (q.pos < last) die("records not sorted in", paths[tid]);
            last = q.pos;
            char qname[32];
            int lq = snprintf(qname, sizeof qname, "r%llu", (unsigned long long)serial++) + 1;
            int extranul = (4 - (lq & 3)) & 3;
            int l_qseq = q.len;
            int n_cigar = l_qseq > 0 ? 1 : 0;
            size_t need = (size_t)lq + extranul + 4 * n_cigar + (l_qseq + 1) / 2 + l_qseq;
            if (need > b->m_data) {
                b->m_data = (uint32_t)(need + 64);
                b->data = (uint8_t*)realloc(b->data, b->m_data);
                if (!b->data) die("out of memory", NULL);
            }
            uint8_t* p = b->data;
            memcpy(p, qname, lq); p += lq;
            memset(p, 0, extranul); p += extranul;
            if (n_cigar) { uint32_t c = ((uint32_t)l_qseq << BAM_CIGAR_SHIFT) | BAM_CMATCH; memcpy(p, &c, 4); p += 4; }
            if (sq) {
                if (fread(bases, 1, l_qseq, sq) != (size_t)l_qseq || fread(quals, 1, l_qseq, sq) != (size_t)l_qseq) die("short read from", seqs[tid]);
                memset(p, 0, (l_qseq + 1) / 2);
                for (int i = 0; i < l_qseq; i++) {
                    int c = bases[i] == 'A' ? 1 : bases[i] == 'C' ? 2 : bases[i] == 'G' ? 4 : bases[i] == 'T' ? 8 : 15;
                    p[i >> 1] |= (uint8_t)(c << ((~i & 1) << 2));
                }
                p += (l_qseq + 1) / 2;
                memcpy(p, quals, l_qseq); p += l_qseq;
            } else {
                memset(p, 0xff, (l_qseq + 1) / 2); p += (l_qseq + 1) / 2;  /* 'N' = nibble 15 */
                memset(p, 0xff, l_qseq); p += l_qseq;                      /* QUAL absent */
            }
            b->l_data = (int)(p - b->data);
            uint16_t flag = 0;
            if (q.fl & 1)  flag |= BAM_FSECONDARY;
            if (q.fl & 2)  flag |= BAM_FQCFAIL;
            if (q.fl & 4)  flag |= BAM_FDUP;
            if (q.fl & 8)  flag |= BAM_FSUPPLEMENTARY;
            if (q.fl & 16) flag |= BAM_FUNMAP;
            if (q.fl & 32) flag |= BAM_FREVERSE;
            b->core.tid = tid; b->core.pos = q.pos; b->core.qual = q.mapq;
            b->core.l_qname = (uint16_t)(lq + extranul); b->core.l_extranul = (uint8_t)extranul;
            b->core.flag = flag; b->core.n_cigar = n_cigar; b->core.l_qseq = l_qseq;
            b->core.mtid = -1; b->core.mpos = -1; b->core.isize = 0;
            hts_pos_t end = q.pos + (l_qseq > 0 ? l_qseq : 1);
            b->core.bin = hts_reg2bin(q.pos, end, 14, 5);
            if (sam_write1(out, h, b) < 0) die("sam_write1 failed", NULL);
        }
        fclose(r);
        if (sq) fclose(sq);
    }
    bam_destroy1(b);
    if (sam_close(out) < 0) die("sam_close failed", NULL);
    h->text = NULL; h->l_text = 0;      /* owned by txt */
    sam_hdr_destroy(h);
    free(txt.s);
    if (sam_index_build(argv[1], 0) < 0) die("sam_index_build failed", argv[1]);   /* .bai */
    return 0;
}
