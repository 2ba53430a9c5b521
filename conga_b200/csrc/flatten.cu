/* flatten.cu — host-side input preparation for the mappability track (no device code).
 * Part of libconga_gpu.so; declared in include/conga_gpu.h and cg_internal.h. */
#include "cg_internal.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <queue>
#include <vector>

/* ------------------------------------------------------------------------- */
/* Mappability rows -> disjoint runs.  svs.c:350-371 paints rows in file order,
 * [start, end] inclusive, later rows overwriting earlier ones; the run list is
 * what mappability[] would hold afterwards (zero gaps omitted), clipped to the
 * chromosome.  Input preparation, like BED parsing: done on the host once. */
void cg_flatten_rows(const cg_mapiv* rows, size_t n, int32_t chrom_len, std::vector<cg_maprun>& out)
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


extern "C" {

int cg_flatten_mappability(const cg_mapiv* rows, size_t n, int32_t chrom_len, cg_maprun** runs_out, size_t* n_out)
{
    if ((!rows && n) || !runs_out || !n_out) return cg_fail(nullptr, CG_EINVAL, "null argument");
    std::vector<cg_maprun> v;
    cg_flatten_rows(rows, n, chrom_len, v);
    cg_maprun* p = (cg_maprun*)malloc((v.size() ? v.size() : 1) * sizeof(cg_maprun));
    if (!p) return cg_fail(nullptr, CG_ENOMEM, "out of host memory");
    if (!v.empty()) memcpy(p, v.data(), v.size() * sizeof(cg_maprun));
    *runs_out = p; *n_out = v.size();
    return CG_OK;
}

void cg_free_runs(cg_maprun* runs) { free(runs); }

} /* extern "C" */
