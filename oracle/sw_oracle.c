/*
 * TEST INFRASTRUCTURE ONLY — the CPU oracle for the Smith-Waterman / rerank half of the hot path.
 * Nothing in the product (libdrm_cuda.so, pipeline, hnswpq_index) links, loads or calls this file;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg do.
 *
 * Plain-C restatement of the reference algorithms (citations are into /root/reference):
 *   orc_sw_score            <- calc_sw_score           src/utils/metrics.cpp:10-45
 *   orc_partial_sort_desc   <- std::partial_sort call  src/utils/reranker.cpp:35-40 (libstdc++ 13
 *                              heap-select + sort-heap, restated so tie order is reproduced exactly)
 *   orc_find_sequence       <- find_sequence           src/utils/post_processor.cpp:47-66
 *                              + reverse_complement    src/utils/parse_inputs.cpp:5-14,43-53
 *   orc_expand_ids          <- find_sequences + the mapping re-expansion in post_process_sw_*
 *                              src/utils/post_processor.cpp:74-201 (dynamic), :204-336 (static),
 *                              :404-413 / :501-510
 *   orc_sw_rerank_one       <- sw_reranker             src/utils/reranker.cpp:3-51
 *   orc_post_process_sw     <- post_process_sw_static/dynamic  src/utils/post_processor.cpp:357-549
 *
 * Pinned: tests/test_oracle_sw.py checks every function here against oracle/_ref (the unmodified
 * reference compiled in place) and against tests/golden/sw_golden.json (generated from oracle/_ref).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- calc_sw_score: match +1, mismatch -1, linear gap -1, byte equality, max over all cells ---- */
int orc_sw_score(const char *s1, int64_t len1, const char *s2, int64_t len2)
{
    /* The reference keeps the full (len1+1)x(len2+1) matrix; two rolling rows give the same cells. */
    int *prev = (int *)calloc((size_t)len2 + 1, sizeof(int));
    int *cur = (int *)calloc((size_t)len2 + 1, sizeof(int));
    int best = 0;
    for (int64_t i = 1; i <= len1; ++i) {
        cur[0] = 0;
        for (int64_t j = 1; j <= len2; ++j) {
            int m = prev[j - 1] + (s1[i - 1] == s2[j - 1] ? 1 : -1);
            int d = prev[j] - 1;
            int ins = cur[j - 1] - 1;
            int h = 0;
            if (m > h) h = m;
            if (d > h) h = d;
            if (ins > h) h = ins;
            cur[j] = h;
            if (h > best) best = h;
        }
        int *t = prev; prev = cur; cur = t;
    }
    free(prev); free(cur);
    return best;
}

void orc_sw_pairs(const char *a_flat, const int64_t *a_off, const char *b_flat, const int64_t *b_off,
                  int64_t n, int32_t *scores, int threads)
{
    (void)threads;
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
    for (int64_t i = 0; i < n; ++i)
        scores[i] = orc_sw_score(a_flat + a_off[i], a_off[i + 1] - a_off[i], b_flat + b_off[i], b_off[i + 1] - b_off[i]);
}

/* ---- libstdc++ std::partial_sort(idx, idx+k, idx+n, [&](a,b){ return scores[a] > scores[b]; }) ---- */
/* comp(a,b) == "a sorts before b" == scores[a] > scores[b]; the heap keeps the comp-largest (= the
 * lowest score) at the root, exactly as bits/stl_heap.h does. */
#define ORC_COMP(a, b) (scores[(a)] > scores[(b)])

static void orc_push_heap(int64_t *first, int64_t hole, int64_t top, int64_t value, const int32_t *scores)
{
    int64_t parent = (hole - 1) / 2;
    while (hole > top && ORC_COMP(first[parent], value)) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}

static void orc_adjust_heap(int64_t *first, int64_t hole, int64_t len, int64_t value, const int32_t *scores)
{
    const int64_t top = hole;
    int64_t child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (ORC_COMP(first[child], first[child - 1])) child--;
        first[hole] = first[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        first[hole] = first[child - 1];
        hole = child - 1;
    }
    orc_push_heap(first, hole, top, value, scores);
}

/* idx[0..n) is permuted in place; afterwards idx[0..k) is what the reference gathers from. */
void orc_partial_sort_desc(const int32_t *scores, int64_t n, int64_t k, int64_t *idx)
{
    for (int64_t i = 0; i < n; ++i) idx[i] = i; /* std::iota */
    if (k > n) k = n;
    /* __heap_select: __make_heap(first, middle) */
    if (k >= 2) {
        int64_t parent = (k - 2) / 2;
        for (;;) {
            int64_t value = idx[parent];
            orc_adjust_heap(idx, parent, k, value, scores);
            if (parent == 0) break;
            parent--;
        }
    }
    if (k >= 1) {
        for (int64_t i = k; i < n; ++i) {
            if (ORC_COMP(idx[i], idx[0])) { /* __pop_heap(first, middle, i) */
                int64_t value = idx[i];
                idx[i] = idx[0];
                orc_adjust_heap(idx, 0, k, value, scores);
            }
        }
    }
    /* __sort_heap(first, middle) */
    int64_t last = k;
    while (last > 1) {
        --last;
        int64_t value = idx[last];
        idx[last] = idx[0];
        orc_adjust_heap(idx, 0, last, value, scores);
    }
}

/* ---- find_sequence (dynamic window fetch) ---- */
/* parse_inputs.cpp:5-14: only A,T,C,G,N are mapped, every other byte complements to '\0'.  A constant table: the
 * function is called from inside OpenMP loops, so nothing here may be initialised lazily. */
static const unsigned char orc_comp[256] = {['A'] = 'T', ['T'] = 'A', ['C'] = 'G', ['G'] = 'C', ['N'] = 'N'};
static const unsigned char *orc_comp_table(void) { return orc_comp; }

/* Returns the window length written to out (ref_len, or 0 when the window overruns the genome). */
int64_t orc_find_sequence(const char *genome, int64_t glen, uint64_t id, int64_t ref_len, char *out)
{
    uint64_t pos = id / 2;
    int rc = (int)(id % 2);
    if (pos + (uint64_t)ref_len > (uint64_t)glen) return 0;
    if (!rc) {
        memcpy(out, genome + pos, (size_t)ref_len);
    } else {
        const unsigned char *t = orc_comp_table();
        for (int64_t i = 0; i < ref_len; ++i)
            out[i] = (char)t[(unsigned char)genome[pos + (uint64_t)(ref_len - 1 - i)] & 0x7f];
    }
    return ref_len;
}

/* ---- candidate-id expansion: the list sw_reranker finally sees, in order, duplicates kept ----
 * mode 0 = static  (bound = number of dense windows; stride==1 drops ids >= bound)
 * mode 1 = dynamic (bound = genome length in bases; stride==1 keeps every id)
 * `ids` are the search labels already cast long -> size_t (so -1 is 2^64-1, post_processor.cpp:339-355).
 * Returns the number of expanded ids (<= n_ids * (2*stride-1)). */
int64_t orc_expand_ids(int mode, const uint64_t *ids, int64_t n_ids, uint64_t stride, uint64_t bound, uint64_t *out)
{
    int64_t m = 0;
    if (stride == 1) {
        for (int64_t i = 0; i < n_ids; ++i)
            if (mode == 1 || ids[i] < bound) out[m++] = ids[i];
        return m;
    }
    for (int64_t i = 0; i < n_ids; ++i) {
        uint64_t p = ids[i] * stride; /* size_t arithmetic: wraps exactly like the reference */
        if (p >= bound) continue;
        uint64_t start = (p >= stride - 1) ? p - stride + 1 : 0;
        uint64_t end = p + stride < bound ? p + stride : bound;
        for (uint64_t q = start; q < end; ++q) out[m++] = q;
    }
    return m;
}

/* ---- one query of post_process_sw_*: expand, fetch, score, top-k ----
 * ref: mode 0 -> n_ref windows of ref_len bytes (ref_seqs), mode 1 -> genome of n_ref bytes.
 * Returns k on success, -1 when there are fewer than k candidates (sw_reranker throws), 0 when the
 * candidate list is empty (sw_reranker returns empty vectors). */
int64_t orc_sw_rerank_one(int mode, const char *ref, int64_t n_ref, int64_t ref_len,
                          const uint64_t *ids, int64_t n_ids, uint64_t stride,
                          const char *q, int64_t lq, int64_t k,
                          int32_t *out_scores, uint64_t *out_ids, int64_t *n_cands_out)
{
    uint64_t per = stride == 1 ? 1 : 2 * stride - 1;
    uint64_t *cand = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(n_ids * (int64_t)per + 1));
    int64_t nc = orc_expand_ids(mode, ids, n_ids, stride, (uint64_t)n_ref, cand);
    if (n_cands_out) *n_cands_out = nc;
    if (nc == 0 || k == 0) { free(cand); return 0; }
    int32_t *scores = (int32_t *)malloc(sizeof(int32_t) * (size_t)nc);
    char *win = (char *)malloc((size_t)ref_len + 1);
    for (int64_t c = 0; c < nc; ++c) {
        if (mode == 0) {
            scores[c] = orc_sw_score(ref + cand[c] * (uint64_t)ref_len, ref_len, q, lq);
        } else {
            int64_t wl = orc_find_sequence(ref, n_ref, cand[c], ref_len, win);
            scores[c] = orc_sw_score(win, wl, q, lq);
        }
    }
    int64_t ret = -1;
    if (nc >= k) {
        int64_t *idx = (int64_t *)malloc(sizeof(int64_t) * (size_t)nc);
        orc_partial_sort_desc(scores, nc, k, idx);
        for (int64_t i = 0; i < k; ++i) { out_scores[i] = scores[idx[i]]; out_ids[i] = cand[idx[i]]; }
        free(idx);
        ret = k;
    }
    free(win); free(scores); free(cand);
    return ret;
}

/* ---- the whole stage.  neighbors: n x k_in int64 labels.  Outputs n x k.  Returns 0, or -(1+i) for the
 * first query i that has too few candidates; -1000000000 for the "Final k too large" guard. ---- */
int64_t orc_post_process_sw(int mode, const int64_t *neighbors, int64_t n, int64_t k_in,
                            const char *ref, int64_t n_ref, int64_t ref_len,
                            const char *q_flat, const int64_t *q_off,
                            int64_t stride, int64_t k, int64_t k_clusters,
                            int32_t *out_scores, uint64_t *out_ids, int64_t *n_cands, int threads)
{
    (void)threads;
    if ((uint64_t)k > (uint64_t)k_clusters * 2u * (uint64_t)stride) return -1000000000;
    int64_t take = k_clusters < k_in ? k_clusters : k_in;
    int64_t fail = 0;
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads)
    for (int64_t i = 0; i < n; ++i) {
        int64_t nc = 0;
        int64_t r = orc_sw_rerank_one(mode, ref, n_ref, ref_len, (const uint64_t *)(neighbors + i * k_in), take,
                                      (uint64_t)stride, q_flat + q_off[i], q_off[i + 1] - q_off[i], k,
                                      out_scores + i * k, out_ids + i * k, &nc);
        if (n_cands) n_cands[i] = nc;
        if (r < 0 || (r == 0 && k > 0)) {
#pragma omp critical
            { if (fail == 0 || -(1 + i) > fail) fail = -(1 + i); }
        }
    }
    return fail;
}

/* ---- SW with traceback: the "real CIGAR" the reference's write_sam leaves as a TODO (src/utils/utils.cpp:381-384).
 * The reference has no traceback, so this is a definition, not a restatement; it is pinned to the reference through its
 * score: out[0] always equals calc_sw_score(s1, s2) and re-scoring the reported operations gives out[0] back
 * (tests/test_oracle_sw.py).  Conventions: full matrix of calc_sw_score (rows = s1 = window, columns = s2 = read);
 * end cell = first maximum in row-major order; walk back while H > 0 preferring diagonal, then up (D), then left (I).
 * out[0..6] = score, ref_beg, ref_end, q_beg, q_end, n_ops, nm (0-based half-open intervals); ops = (len << 4 | op) in
 * alignment order, op 0 = M, 1 = I, 2 = D; at most max_ops are written (n_ops > max_ops means truncated). */
void orc_sw_align(const char *s1, int64_t len1, const char *s2, int64_t len2, int32_t *out, uint32_t *ops, int max_ops)
{
    const int64_t W = len2 + 1;
    int *H = (int *)calloc((size_t)((len1 + 1) * W), sizeof(int));
    int best = 0;
    int64_t bi = 0, bj = 0;
    for (int64_t i = 1; i <= len1; ++i)
        for (int64_t j = 1; j <= len2; ++j) {
            int d = H[(i - 1) * W + j - 1] + (s1[i - 1] == s2[j - 1] ? 1 : -1);
            int u = H[(i - 1) * W + j] - 1, l = H[i * W + j - 1] - 1;
            int h = d > u ? d : u;
            if (l > h) h = l;
            if (h < 0) h = 0;
            H[i * W + j] = h;
            if (h > best) { best = h; bi = i; bj = j; }
        }
    int64_t i = bi, j = bj;
    int n_ops = 0, nm = 0, cur = -1, cur_len = 0;
    uint32_t *rev = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(len1 + len2 + 2));
    while (i > 0 && j > 0 && H[i * W + j] > 0) {
        const int h = H[i * W + j];
        int op;
        if (h == H[(i - 1) * W + j - 1] + (s1[i - 1] == s2[j - 1] ? 1 : -1)) {
            op = 0;
            if (s1[i - 1] != s2[j - 1]) ++nm;
            --i; --j;
        } else if (h == H[(i - 1) * W + j] - 1) { op = 2; ++nm; --i; }
        else { op = 1; ++nm; --j; }
        if (op == cur) ++cur_len;
        else {
            if (cur >= 0) rev[n_ops++] = ((uint32_t)cur_len << 4) | (uint32_t)cur;
            cur = op; cur_len = 1;
        }
    }
    if (cur >= 0) rev[n_ops++] = ((uint32_t)cur_len << 4) | (uint32_t)cur;
    for (int x = 0; x < n_ops && x < max_ops; ++x) ops[x] = rev[n_ops - 1 - x];
    out[0] = best; out[1] = (int32_t)i; out[2] = (int32_t)bi; out[3] = (int32_t)j; out[4] = (int32_t)bj; out[5] = n_ops; out[6] = nm;
    free(rev);
    free(H);
}
