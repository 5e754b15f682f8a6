/* engine_model.c -- sequential CPU model of the B200 engine's OWN algorithm.
 *
 * TEST INFRASTRUCTURE ONLY.  This is not the reference's algorithm (that is
 * deflate_oracle.c / oracle/_ref) and it is never on the product path: it is
 * the executable specification the CUDA kernels in gzippier_b200/csrc are
 * checked against stage by stage (per-position matches, token stream, code
 * lengths, final bytes must be identical), and the tool used to choose the
 * search-effort table without a GPU.
 *
 * The engine's semantics are defined so that they do NOT depend on how the
 * GPU schedules work:
 *   match(p)   = best (len, dist) over the first `depth` earlier positions, most
 *                recent first, that share p's 6-BYTE hash bucket and lie within
 *                `maxd`; a candidate replaces the best only if strictly longer;
 *                the walk stops at len >= nice.  If that gives less than 6, the
 *                most recent 4-byte twin is tried, and if still nothing, the
 *                most recent 3-byte twin within TOO_FAR.   [cf. longest_match,
 *                lib/zlib/deflate.c:1236-1377]
 *   next(p)    = p+1 (literal) or p+len (match), a pure function of match(p),
 *                match(p+1): greedy for levels 1-3 [deflate_fast :1824-1918],
 *                one-step lazy for 4-9 [deflate_slow :1926-2049].
 *   blocks     = fixed input spans; stored/static/dynamic chosen by size
 *                [_tr_flush_block, lib/zlib/trees.c:911-1008].
 */
#include "engine_model.h"
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

/* ---------------- deflate tables (RFC 1951; cf. lib/zlib/trees.c:62-72) ---------------- */
static const uint8_t  k_len_extra[29]  = {0,0,0,0,0,0,0,0,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,4,5,5,5,5,0};
static const uint16_t k_len_base[29]   = {3,4,5,6,7,8,9,10,11,13,15,17,19,23,27,31,35,43,51,59,67,83,99,115,131,163,195,227,258};
static const uint8_t  k_dist_extra[30] = {0,0,0,0,1,1,2,2,3,3,4,4,5,5,6,6,7,7,8,8,9,9,10,10,11,11,12,12,13,13};
static const uint16_t k_dist_base[30]  = {1,2,3,4,5,7,9,13,17,25,33,49,65,97,129,193,257,385,513,769,1025,1537,2049,3073,4097,6145,8193,12289,16385,24577};
static const uint8_t  k_bl_order[19]   = {16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15};

int gzm_len_code(int len)   /* 3..258 -> 0..28 */
{
    int c = 28;
    if (len < 258) { c = 0; while (c < 28 && k_len_base[c + 1] <= len) c++; }
    return c;
}
int gzm_dist_code(int dist) /* 1..32768 -> 0..29 */
{
    int c = 0; while (c < 29 && k_dist_base[c + 1] <= dist) c++;
    return c;
}

/* ---------------- level table ---------------- */
/* good/lazy/nice are the reference's configuration_table (lib/zlib/deflate.c:134-146);
 * depth is the engine's own candidate budget per position (every position is searched
 * and the chains are keyed on 6 bytes, so far fewer probes are needed than zlib's max_chain
 * for the same ratio -- see DESIGN.md "Level semantics"). */
void gzm_level_params(int level, gzm_params *p)
{
    static const struct { int good, lazy, nice, chain, depth; } t[10] = {
        {0, 0, 0, 0, 0},
        {4, 4, 8, 4, 1}, {4, 5, 16, 8, 2}, {4, 6, 32, 32, 4},
        {4, 4, 16, 16, 4}, {8, 16, 32, 32, 6}, {8, 16, 128, 128, 8},
        {8, 32, 128, 256, 16}, {32, 128, 258, 1024, 32}, {32, 258, 258, 4096, 64}};
    memset(p, 0, sizeof(*p));
    p->level = level;
    p->good = t[level].good; p->max_lazy = t[level].lazy; p->nice = t[level].nice;
    p->ref_chain = t[level].chain; p->depth = t[level].depth;
    p->lazy_parse = level >= 4;
    p->maxd = GZM_MAXD; p->hbits = GZM_HBITS; p->block_span = GZM_BLOCK_SPAN;
    p->chunk = 131072; p->dict = 32768; p->use_h3 = 1; p->use_h4 = 1;
    p->scan = 1;        /* k_scan: incompressible chunks are stored without a search */
    p->good_cut = 0;    /* measured and left off: +0.3..1.1 % bytes for -9 % compare trips per 32-position group */
}

/* ---------------- match finder ---------------- */
static inline uint32_t ld32(const uint8_t *p) { uint32_t v; memcpy(&v, p, 4); return v; }
uint32_t gzm_hash6(uint32_t lo, uint32_t hi16, int hbits) { return (lo * 0x9E3779B1u + hi16 * 0x85EBCA6Bu) >> (32 - hbits); }
uint32_t gzm_hash4(uint32_t v) { return (v * 0x9E3779B1u) >> (32 - GZM_H4BITS); }
uint32_t gzm_hash3(uint32_t v) { return ((v << 8) * 0x9E3779B1u) >> (32 - GZM_H3BITS); }

static int common_len(const uint8_t *a, const uint8_t *b, int maxlen) { int l = 0; while (l < maxlen && a[l] == b[l]) l++; return l; }
#ifdef GZM_STATS
/* cost proxies for the lockstep search of k_lz77 (tools/ experiments only): per 32-position group, rounds = max over its
 * positions, compare trips per round = max over its positions.  [0] positions [1] chain steps [2] measured candidates
 * [3] trips (per position) [4] groups [5] lockstep rounds [6] lockstep trips */
unsigned long long gzm_stats[8];
static int g_rounds[32], g_trips[32][20];
static void stats_group_flush(void)
{
    int mr = 0;
    for (int i = 0; i < 32; i++) if (g_rounds[i] > mr) mr = g_rounds[i];
    gzm_stats[4]++; gzm_stats[5] += (unsigned)mr;
    for (int r = 0; r < mr && r < 20; r++) { int mt = 0; for (int i = 0; i < 32; i++) if (g_trips[i][r] > mt) mt = g_trips[i][r]; gzm_stats[6] += (unsigned)mt; }
    memset(g_rounds, 0, sizeof(g_rounds)); memset(g_trips, 0, sizeof(g_trips));
}
#endif

/* stream = dict||chunk, D = dict length, n = total stream length.  Positions < D are inserted only.
 * bytes beyond the stream end read as zero for hashing (positions n-5..n-1). */
void gzm_find_matches(const gzm_params *P, const uint8_t *s, int D, int n, uint16_t *mlen, uint16_t *mdist)
{
    int hsize = 1 << P->hbits, h3size = 1 << GZM_H3BITS, h4size = 1 << GZM_H4BITS;
    int32_t *head = (int32_t *)malloc(sizeof(int32_t) * hsize);
    int32_t *head3 = (int32_t *)malloc(sizeof(int32_t) * h3size);
    int32_t *head4 = (int32_t *)malloc(sizeof(int32_t) * h4size);
    int32_t *link = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1));
    for (int i = 0; i < hsize; i++) head[i] = -1;
    for (int i = 0; i < h3size; i++) head3[i] = -1;
    for (int i = 0; i < h4size; i++) head4[i] = -1;
    for (int p = 0; p < n; p++) {
        uint8_t tmp[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int k = 0; k < 6 && p + k < n; k++) tmp[k] = s[p + k];
        uint32_t lo = ld32(tmp), hi16 = (uint32_t)tmp[4] | ((uint32_t)tmp[5] << 8);
        uint32_t h = gzm_hash6(lo, hi16, P->hbits), h4 = gzm_hash4(lo), h3 = gzm_hash3(lo);
        int32_t prev = head[h], prev4 = head4[h4], prev3 = head3[h3];
        /* the GPU stores 16-bit deltas: a gap above 32767 ends the chain */
        link[p] = (prev >= 0 && p - prev <= 32767) ? prev : -1;
        head[h] = p; head4[h4] = p; head3[h3] = p;
        if (p < D) continue;
        int maxlen = n - p < 258 ? n - p : 258;
        int best = 2, bdist = 0;
        int limit = p < P->maxd ? p : P->maxd;
        int nice = P->nice < maxlen ? P->nice : maxlen;
        /* the last five positions of a chunk, whose six hashed bytes would run past its end, do not walk a chain
         * (what follows the chunk is the next chunk's business: the engine may or may not have it at hand) */
        int c = maxlen >= 6 ? link[p] : -1, depth = P->depth, cut = 0;
        /* The walk goes in ROUNDS, as the kernel's lanes do: up to GZM_WALK_STEPS chain steps in which a candidate that
         * passes the one-byte filter (against the best length at the start of the round) is only noted -- a second noted
         * candidate ends the round early --, then the noted ones are measured in chain order, strictly longer wins.
         * Without the good_length rule this equals the one-candidate-at-a-time walk.  With it (levels 4-9, cf.
         * `if (s->prev_length >= s->good_match) chain_length >>= 2`, lib/zlib/deflate.c:1254-1257): once the best match
         * has reached good_length at the end of a round, the remaining candidate budget is quartered, once. */
        while (c >= 0 && p - c <= limit && depth > 0) {
            int pend[2], npend = 0;
            const int fbest = best;
#ifdef GZM_STATS
            const int lane_ = (p - D) & 31, r_ = g_rounds[lane_] < 19 ? g_rounds[lane_] : 19; g_rounds[lane_]++;
#endif
            for (int st = 0; st < GZM_WALK_STEPS && npend < 2; st++) {
                if (!(c >= 0 && p - c <= limit && depth > 0)) break;
                if (s[c + fbest] == s[p + fbest]) pend[npend++] = c;
                c = link[c]; depth--;
#ifdef GZM_STATS
                gzm_stats[1]++;
#endif
            }
            int stop = 0;
            for (int i = 0; i < npend && !stop; i++) {
                const int cand = pend[i];
                if (i == 1 && s[cand + best] != s[p + best]) continue;
                int l = common_len(s + cand, s + p, maxlen);
#ifdef GZM_STATS
                gzm_stats[2]++; gzm_stats[3] += (unsigned)(l / 16 + 1); g_trips[lane_][r_] += l / 16 + 1;
#endif
                if (l > best) { best = l; bdist = p - cand; if (l >= nice || l >= maxlen) stop = 1; }
            }
            if (stop) break;
            if (P->good_cut && !cut && best >= P->good) { cut = 1; depth >>= 2; }
        }
        /* nothing of 6+ on the chain: try the most recent 4-byte twin (single-entry table) */
        if (best < 6 && P->use_h4 && prev4 >= 0 && p - prev4 <= (limit < 32767 ? limit : 32767) && maxlen >= 4) {
            int l = common_len(s + prev4, s + p, maxlen);
            if (l >= 4 && l > best) { best = l; bdist = p - prev4; }
        }
        /* still nothing: the most recent 3-byte twin */
        if (best < 3 && P->use_h3 && prev3 >= 0 && p - prev3 <= (limit < GZM_TOO_FAR ? limit : GZM_TOO_FAR) && maxlen >= 3) {
            int l = common_len(s + prev3, s + p, maxlen);
            if (l >= 3) { best = l; bdist = p - prev3; }
        }
        if (best == 3 && bdist > GZM_TOO_FAR) best = 2;       /* TOO_FAR, deflate.c:110-112,1972-1981 */
        if (best < 3) { mlen[p - D] = 0; mdist[p - D] = 0; }
        else { mlen[p - D] = (uint16_t)best; mdist[p - D] = (uint16_t)bdist; }
#ifdef GZM_STATS
        gzm_stats[0]++;
        if (((p - D) & 31) == 31 || p == n - 1) stats_group_flush();
#endif
    }
    free(head); free(head3); free(head4); free(link);
}

/* ---------------- parse ---------------- */
/* token: bit31 = match; match: (len-3) << 16 | (dist-1);  literal: byte */
int gzm_parse(const gzm_params *P, const uint8_t *chunk, int n, const uint16_t *mlen, const uint16_t *mdist,
              uint32_t *tokens, uint32_t *tok_pos)
{
    int nt = 0, p = 0;
    while (p < n) {
        int L = mlen[p];
        int lit = L < 3;
        if (!lit && P->lazy_parse && L < P->max_lazy && p + 1 < n && mlen[p + 1] > L) lit = 1;
        if (tok_pos) tok_pos[nt] = (uint32_t)p;
        if (lit) { tokens[nt++] = chunk[p]; p += 1; }
        else { tokens[nt++] = 0x80000000u | ((uint32_t)(L - 3) << 16) | (uint32_t)(mdist[p] - 1); p += L; }
    }
    return nt;
}

/* ---------------- Huffman ---------------- */
/* In-place minimum-redundancy code lengths (Moffat & Katajainen) over frequencies sorted ascending. */
static void min_redundancy(uint32_t *A, int n)
{
    int root, leaf, next, avbl, used, dpth;
    if (n == 0) return;
    if (n == 1) { A[0] = 1; return; }
    A[0] += A[1]; root = 0; leaf = 2;
    for (next = 1; next < n - 1; next++) {
        if (leaf >= n || A[root] < A[leaf]) { A[next] = A[root]; A[root++] = (uint32_t)next; } else A[next] = A[leaf++];
        if (leaf >= n || (root < next && A[root] < A[leaf])) { A[next] += A[root]; A[root++] = (uint32_t)next; } else A[next] += A[leaf++];
    }
    A[n - 2] = 0;
    for (next = n - 3; next >= 0; next--) A[next] = A[A[next]] + 1;
    avbl = 1; used = dpth = 0; root = n - 2; next = n - 1;
    while (avbl > 0) {
        while (root >= 0 && (int)A[root] == dpth) { used++; root--; }
        while (avbl > used) { A[next--] = (uint32_t)dpth; avbl--; }
        avbl = 2 * used; dpth++; used = 0;
    }
}

/* freq[0..nsym) -> lens[0..nsym), limited to maxbits.  Forces >= 2 coded symbols like
 * build_tree (lib/zlib/trees.c:646-651).  Returns max used symbol index. */
int gzm_huff_lengths(const uint32_t *freq_in, int nsym, int maxbits, uint8_t *lens)
{
    uint32_t freq[288]; uint32_t key[288]; uint32_t A[288]; int n = 0;
    memcpy(freq, freq_in, sizeof(uint32_t) * nsym);
    int used = 0; for (int i = 0; i < nsym; i++) used += freq[i] != 0;
    for (int i = 0; used < 2 && i < nsym; i++) if (!freq[i]) { freq[i] = 1; used++; }
    for (int i = 0; i < nsym; i++) { lens[i] = 0; if (freq[i]) key[n++] = (uint32_t)i; }
    /* sort by (freq, symbol) ascending: insertion sort is fine for the model */
    for (int i = 1; i < n; i++) {
        uint32_t k = key[i]; int j = i - 1;
        while (j >= 0 && (freq[key[j]] > freq[k] || (freq[key[j]] == freq[k] && key[j] > k))) { key[j + 1] = key[j]; j--; }
        key[j + 1] = k;
    }
    for (int i = 0; i < n; i++) A[i] = freq[key[i]];
    min_redundancy(A, n);
    int num[33]; memset(num, 0, sizeof(num));
    for (int i = 0; i < n; i++) num[A[i] > 32 ? 32 : A[i]]++;
    if (n > 1) {
        for (int i = maxbits + 1; i <= 32; i++) { num[maxbits] += num[i]; num[i] = 0; }
        uint32_t total = 0;
        for (int i = maxbits; i > 0; i--) total += (uint32_t)num[i] << (maxbits - i);
        while (total != (1u << maxbits)) {
            num[maxbits]--;
            for (int i = maxbits - 1; i > 0; i--) if (num[i]) { num[i]--; num[i + 1] += 2; break; }
            total--;
        }
    }
    int j = n;   /* most frequent symbols (end of the sorted array) get the shortest codes */
    for (int i = 1; i <= maxbits; i++) for (int l = num[i]; l > 0; l--) lens[key[--j]] = (uint8_t)i;
    int maxsym = 0; for (int i = 0; i < nsym; i++) if (lens[i]) maxsym = i;
    return maxsym;
}

/* canonical codes, bit-reversed for LSB-first emission (gen_codes + bi_reverse, trees.c:572-605,1158-1168) */
void gzm_huff_codes(const uint8_t *lens, int nsym, int maxbits, uint16_t *codes)
{
    int cnt[16] = {0}, next[16];
    for (int i = 0; i < nsym; i++) cnt[lens[i]]++;
    cnt[0] = 0;
    int code = 0;
    for (int b = 1; b <= maxbits; b++) { code = (code + cnt[b - 1]) << 1; next[b] = code; }
    for (int i = 0; i < nsym; i++) {
        int l = lens[i]; if (!l) { codes[i] = 0; continue; }
        unsigned c = (unsigned)next[l]++, r = 0;
        for (int k = 0; k < l; k++) { r = (r << 1) | (c & 1); c >>= 1; }
        codes[i] = (uint16_t)r;
    }
}

/* ---------------- bit writer ---------------- */
typedef struct { uint8_t *out; uint64_t bitpos; } bitw;
static inline void putbits(bitw *w, uint32_t v, int n)
{
    for (int i = 0; i < n; i++, w->bitpos++) {
        if ((w->bitpos & 7) == 0) w->out[w->bitpos >> 3] = 0;
        w->out[w->bitpos >> 3] |= (uint8_t)(((v >> i) & 1u) << (w->bitpos & 7));
    }
}

/* Run-length coding of a code-length sequence with the symbols 16 / 17 / 18 of RFC 1951 3.2.7 (the engine's own policy;
 * zlib's is scan_tree / send_tree, lib/zlib/trees.c:703-793).  The sequence is cut into MAXIMAL runs of equal lengths and
 * every run is coded on its own:
 *   zeros, L of them : as many "18" (138 zeros) as fit; then one "18" (11..137) or "17" (3..10) for the rest, or the
 *                      last one or two zeros as plain lengths
 *   v != 0, L of them: v once, then the other L - 1 as "16" (repeat the previous length 6 times) while 3 or more are
 *                      left -- the last "16" takes 3..6 -- and one or two plain v at the end
 * Returns the number of (symbol, extra-bits value) pairs.  (GPU: k_huffman rle_lengths, one lane per run.) */
static int rle_run_count(int v, int L)
{
    if (v == 0) { int q = L / 138, r = L % 138; return q + (r >= 3 ? 1 : r); }
    int R = L - 1, q = R / 6, r = R % 6;
    return 1 + q + (r >= 3 ? 1 : r);
}
static int rle_run_emit(int v, int L, uint8_t *sym, uint8_t *ext)
{
    int k = 0;
    if (v == 0) {
        while (L >= 138) { sym[k] = 18; ext[k++] = 138 - 11; L -= 138; }
        if (L >= 11) { sym[k] = 18; ext[k++] = (uint8_t)(L - 11); }
        else if (L >= 3) { sym[k] = 17; ext[k++] = (uint8_t)(L - 3); }
        else while (L-- > 0) { sym[k] = 0; ext[k++] = 0; }
        return k;
    }
    sym[k] = (uint8_t)v; ext[k++] = 0;
    int R = L - 1;
    while (R >= 6) { sym[k] = 16; ext[k++] = 6 - 3; R -= 6; }
    if (R >= 3) { sym[k] = 16; ext[k++] = (uint8_t)(R - 3); }
    else while (R-- > 0) { sym[k] = (uint8_t)v; ext[k++] = 0; }
    return k;
}
static int rle_lengths(const uint8_t *lens, int n, uint8_t *sym, uint8_t *ext)
{
    int k = 0;
    for (int i = 0; i < n; ) {
        int j = i + 1;
        while (j < n && lens[j] == lens[i]) j++;
        int c = rle_run_emit(lens[i], j - i, sym + k, ext + k);
        if (c != rle_run_count(lens[i], j - i)) return -1;       /* the two statements of the policy agree */
        k += c; i = j;
    }
    return k;
}

void gzm_static_lengths(uint8_t *ll, uint8_t *dl)
{
    for (int i = 0; i < 288; i++) ll[i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
    for (int i = 0; i < 30; i++) dl[i] = 5;
}

/* Build everything a block needs from its histograms.  (GPU: k_huffman) */
void gzm_plan_block(const uint32_t *lfreq_in, const uint32_t *dfreq, uint32_t in_bytes, gzm_block_plan *B)
{
    uint32_t lfreq[286];
    memcpy(lfreq, lfreq_in, sizeof(lfreq));
    lfreq[256] = 1;                                                /* EOB, init_block trees.c:407-420 */
    memset(B, 0, sizeof(*B));
    int maxl = gzm_huff_lengths(lfreq, 286, 15, B->llen);
    int maxd = gzm_huff_lengths(dfreq, 30, 15, B->dlen);
    if (maxl < 256) maxl = 256;
    B->hlit = maxl + 1; B->hdist = maxd + 1;
    /* code-length alphabet */
    uint8_t rs[320], re[320];
    int k1 = rle_lengths(B->llen, B->hlit, rs, re);
    int k2 = rle_lengths(B->dlen, B->hdist, rs + k1, re + k1);
    int k = k1 + k2;
    uint32_t bfreq[19] = {0};
    for (int i = 0; i < k; i++) bfreq[rs[i]]++;
    gzm_huff_lengths(bfreq, 19, 7, B->blen);
    int hclen = 19; while (hclen > 4 && B->blen[k_bl_order[hclen - 1]] == 0) hclen--;
    B->hclen = hclen;
    /* header bit string */
    uint16_t bcodes[19]; gzm_huff_codes(B->blen, 19, 7, bcodes);
    bitw w = {B->header, 0};
    putbits(&w, (uint32_t)(B->hlit - 257), 5); putbits(&w, (uint32_t)(B->hdist - 1), 5); putbits(&w, (uint32_t)(hclen - 4), 4);
    for (int i = 0; i < hclen; i++) putbits(&w, B->blen[k_bl_order[i]], 3);
    for (int i = 0; i < k; i++) {
        putbits(&w, bcodes[rs[i]], B->blen[rs[i]]);
        if (rs[i] == 16) putbits(&w, re[i], 2); else if (rs[i] == 17) putbits(&w, re[i], 3); else if (rs[i] == 18) putbits(&w, re[i], 7);
    }
    B->header_bits = (uint32_t)w.bitpos;
    /* sizes in bits, header included, 3 block-type bits excluded */
    uint8_t sl[288], sd[30]; gzm_static_lengths(sl, sd);
    uint64_t dyn = B->header_bits, sta = 0;
    for (int i = 0; i < 286; i++) {
        int ex = i >= 257 ? k_len_extra[i - 257] : 0;
        dyn += (uint64_t)lfreq[i] * (uint64_t)(B->llen[i] + ex); sta += (uint64_t)lfreq[i] * (uint64_t)(sl[i] + ex);
    }
    for (int i = 0; i < 30; i++) {
        dyn += (uint64_t)dfreq[i] * (uint64_t)(B->dlen[i] + k_dist_extra[i]); sta += (uint64_t)dfreq[i] * (uint64_t)(5 + k_dist_extra[i]);
    }
    B->dyn_bits = dyn; B->static_bits = sta;
    /* choice (cf. trees.c:945-993): bytes, rounded up, 3 header bits included */
    uint64_t dyn_b = (dyn + 3 + 7) >> 3, sta_b = (sta + 3 + 7) >> 3;
    uint64_t nsub = in_bytes ? (in_bytes + 65534) / 65535 : 1;
    uint64_t sto_b = in_bytes + 5 * nsub;
    uint64_t best = dyn_b < sta_b ? dyn_b : sta_b;
    if (sto_b <= best) B->type = 0;
    else if (sta_b <= dyn_b) B->type = 1;
    else B->type = 2;
    if (B->type == 1) { memcpy(B->llen, sl, 288);   /* all 288: 286 and 287 take part in the code (RFC 1951 3.2.6) */ memcpy(B->dlen, sd, 30); }
    gzm_huff_codes(B->llen, B->type == 1 ? 288 : 286, 15, B->lcode);
    gzm_huff_codes(B->dlen, 30, 15, B->dcode);
}

/* Emit one block at bit position w->bitpos.  (GPU: k_layout computes positions, k_pack writes) */
static void emit_block(bitw *w, const gzm_block_plan *B, const uint8_t *raw, uint32_t in_bytes,
                       const uint32_t *tok, int nt, int bfinal)
{
    if (B->type == 0) {
        uint32_t off = 0;
        uint32_t nsub = in_bytes ? (in_bytes + 65534) / 65535 : 1;
        for (uint32_t sb = 0; sb < nsub; sb++) {
            uint32_t len = in_bytes - off < 65535 ? in_bytes - off : 65535;
            putbits(w, (bfinal && sb == nsub - 1) ? 1 : 0, 1); putbits(w, 0, 2);
            while (w->bitpos & 7) putbits(w, 0, 1);
            putbits(w, len, 16); putbits(w, len ^ 0xffff, 16);
            memcpy(w->out + (w->bitpos >> 3), raw + off, len); w->bitpos += (uint64_t)len * 8; off += len;
        }
        return;
    }
    putbits(w, (uint32_t)bfinal, 1); putbits(w, (uint32_t)B->type, 2);
    if (B->type == 2) for (uint32_t i = 0; i < B->header_bits; i++) putbits(w, (B->header[i >> 3] >> (i & 7)) & 1, 1);
    for (int i = 0; i < nt; i++) {
        uint32_t t = tok[i];
        if (!(t & 0x80000000u)) { putbits(w, B->lcode[t & 0xff], B->llen[t & 0xff]); continue; }
        int len = (int)((t >> 16) & 0xff) + 3, dist = (int)(t & 0xffff) + 1;
        int lc = gzm_len_code(len), dc = gzm_dist_code(dist);
        putbits(w, B->lcode[257 + lc], B->llen[257 + lc]); putbits(w, (uint32_t)(len - k_len_base[lc]), k_len_extra[lc]);
        putbits(w, B->dcode[dc], B->dlen[dc]); putbits(w, (uint32_t)(dist - k_dist_base[dc]), k_dist_extra[dc]);
    }
    putbits(w, B->lcode[256], B->llen[256]);
}

/* ---------------- incompressible chunks (GPU: k_scan) ---------------- */
/* The decision "store this chunk, do not search it" is taken before the match finder from two order-independent
 * integer quantities (gzippier_b200/csrc/k_scan.cu computes the same ones in parallel):
 *   (1) the order-0 entropy, in Q8 fixed point, of one 16-byte word out of every 256 bytes of the chunk must be at
 *       least 8 * 127/128 bits per byte;
 *   (2) of the 8-byte grams at content-defined positions (first byte ending in the bits 101010: one in 64 on near-uniform bytes) of the chunk,
 *       fewer than 1 in 64 may equal a gram seen before in the chunk or in its dictionary.
 * zlib takes the same decision after the search (_tr_flush_block, lib/zlib/trees.c:945-993). */
static const uint8_t k_log2_q8[256] = {
      0,   1,   3,   4,   6,   7,   9,  10,  11,  13,  14,  16,  17,  18,  20,  21,
     22,  24,  25,  26,  28,  29,  30,  32,  33,  34,  36,  37,  38,  40,  41,  42,
     44,  45,  46,  47,  49,  50,  51,  52,  54,  55,  56,  57,  59,  60,  61,  62,
     63,  65,  66,  67,  68,  69,  71,  72,  73,  74,  75,  77,  78,  79,  80,  81,
     82,  84,  85,  86,  87,  88,  89,  90,  92,  93,  94,  95,  96,  97,  98,  99,
    100, 102, 103, 104, 105, 106, 107, 108, 109, 110, 111, 112, 113, 114, 116, 117,
    118, 119, 120, 121, 122, 123, 124, 125, 126, 127, 128, 129, 130, 131, 132, 133,
    134, 135, 136, 137, 138, 139, 140, 141, 142, 143, 144, 145, 146, 147, 148, 149,
    150, 151, 152, 153, 154, 155, 155, 156, 157, 158, 159, 160, 161, 162, 163, 164,
    165, 166, 167, 168, 169, 169, 170, 171, 172, 173, 174, 175, 176, 177, 178, 178,
    179, 180, 181, 182, 183, 184, 185, 185, 186, 187, 188, 189, 190, 191, 192, 192,
    193, 194, 195, 196, 197, 198, 198, 199, 200, 201, 202, 203, 203, 204, 205, 206,
    207, 208, 208, 209, 210, 211, 212, 212, 213, 214, 215, 216, 216, 217, 218, 219,
    220, 220, 221, 222, 223, 224, 224, 225, 226, 227, 228, 228, 229, 230, 231, 231,
    232, 233, 234, 234, 235, 236, 237, 238, 238, 239, 240, 241, 241, 242, 243, 244,
    244, 245, 246, 247, 247, 248, 249, 249, 250, 251, 252, 252, 253, 254, 255, 255};
static uint32_t lg_q8(uint32_t x)      /* 256 * log2(x), x >= 1 */
{
    uint32_t k = 31u - (uint32_t)__builtin_clz(x);
    uint32_t m = k >= 8u ? (x >> (k - 8u)) & 0xFFu : (x << (8u - k)) & 0xFFu;
    return (k << 8) + k_log2_q8[m];
}
int gzm_chunk_is_stored(const uint8_t *s, int D, int n)
{
    if (n < GZM_SCAN_MIN) return 0;
    const uint8_t *c = s + D;
    uint32_t h[256]; memset(h, 0, sizeof(h));
    uint32_t ns = 0;
    for (int o = 0; o + 16 <= n; o += GZM_SCAN_SAMPLE) { for (int k = 0; k < 16; k++) h[c[o + k]]++; ns += 16; }
    uint64_t bits = 0;
    for (int b = 0; b < 256; b++) if (h[b]) bits += (uint64_t)h[b] * (lg_q8(ns) - lg_q8(h[b]));
    if (bits * 128u < (uint64_t)ns * 8u * 256u * 127u) return 0;
    uint32_t *tab = (uint32_t *)calloc(GZM_SCAN_SLOTS, sizeof(uint32_t));
    uint32_t distinct = 0, grams = 0, fresh = 0;
    int P = ((D + n) / 4 - 2) * 4, overflow = 0;            /* positions whose three aligned words lie inside the stream */
    for (int p = 0; p < P && !overflow; p++) {
        if (p >= D || p < (D < P ? D : P)) {
            uint32_t a = ld32(s + p), b = ld32(s + p + 4);
            if ((a & 0x3Fu) != 0x2Au) continue;             /* sampled positions: first byte ends in 101010 */
            uint32_t fp = ((a * 0x85EBCA6Bu) ^ (b * 0xC2B2AE35u)) | 1u, slot = (fp >> 8) & (GZM_SCAN_SLOTS - 1);
            int isnew = 0;
            for (;;) {
                if (tab[slot] == fp) break;
                if (tab[slot] == 0) { tab[slot] = fp; isnew = 1; break; }
                slot = (slot + 1) & (GZM_SCAN_SLOTS - 1);
            }
            if (isnew && ++distinct > GZM_SCAN_MAX_GRAMS) overflow = 1;
            if (p >= D) { grams++; fresh += (uint32_t)isnew; }
        }
    }
    free(tab);
    if (overflow || grams == 0) return 0;
    return (grams - fresh) * 64u < grams;
}

/* ---------------- one chunk ---------------- */
size_t gzm_deflate_chunk(const gzm_params *P, const uint8_t *stream, int D, int n_chunk, int final,
                         uint8_t *out, gzm_chunk_stats *st)
{
    int n = n_chunk;
    const uint8_t *chunk = stream + D;
    if (P->scan && gzm_chunk_is_stored(stream, D, n)) {
        /* stored without a search: blocks are the plain 64 KiB spans of the chunk */
        bitw w = {out, 0};
        gzm_block_plan B; memset(&B, 0, sizeof(B));
        int nblocks = (n + P->block_span - 1) / P->block_span;
        for (int b = 0; b < nblocks; b++) {
            uint32_t lo = (uint32_t)b * (uint32_t)P->block_span, hi = lo + (uint32_t)P->block_span < (uint32_t)n ? lo + (uint32_t)P->block_span : (uint32_t)n;
            emit_block(&w, &B, chunk + lo, hi - lo, NULL, 0, final && b == nblocks - 1);
            if (st) st->blocks[0]++;
        }
        if (final) { while (w.bitpos & 7) putbits(&w, 0, 1); }
        else if (w.bitpos & 7) { putbits(&w, 0, 3); while (w.bitpos & 7) putbits(&w, 0, 1); putbits(&w, 0, 16); putbits(&w, 0xffff, 16); }
        return (size_t)(w.bitpos >> 3);
    }
    uint16_t *mlen = (uint16_t *)malloc(sizeof(uint16_t) * (size_t)(n + 1));
    uint16_t *mdist = (uint16_t *)malloc(sizeof(uint16_t) * (size_t)(n + 1));
    uint32_t *tok = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n + 1));
    uint32_t *tpos = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n + 1));
    gzm_find_matches(P, stream, D, D + n, mlen, mdist);
    int nt = gzm_parse(P, chunk, n, mlen, mdist, tok, tpos);
    bitw w = {out, 0};
    int nblocks = n ? (n + P->block_span - 1) / P->block_span : 1;
    int ti = 0, last_type = -1;
    /* block b holds the tokens that start in [b*span-1, (b+1)*span-1) -- the positions the kernel decides
     * while it parses the tiles of span b (the last byte of a tile waits for its successor's match);
     * its input range runs from its first token to the first token of the next block. */
    uint32_t lo = 0;
    for (int b = 0; b < nblocks; b++) {
        int t0 = ti;
        while (ti < nt && (b == nblocks - 1 || (int)((tpos[ti] + 1) / (uint32_t)P->block_span) <= b)) ti++;
        uint32_t hi = (ti < nt) ? tpos[ti] : (uint32_t)n;
        uint32_t lf[286] = {0}, df[30] = {0};
        for (int i = t0; i < ti; i++) {
            uint32_t t = tok[i];
            if (t & 0x80000000u) { lf[257 + gzm_len_code((int)((t >> 16) & 0xff) + 3)]++; df[gzm_dist_code((int)(t & 0xffff) + 1)]++; }
            else lf[t & 0xff]++;
        }
        gzm_block_plan B; gzm_plan_block(lf, df, hi - lo, &B);
        int bfinal = final && b == nblocks - 1;
        emit_block(&w, &B, chunk + lo, hi - lo, tok + t0, ti - t0, bfinal);
        last_type = B.type;
        if (st) { st->blocks[B.type]++; }
        lo = hi;
    }
    if (final) { while (w.bitpos & 7) putbits(&w, 0, 1); }
    else if (w.bitpos & 7) {          /* byte-align with an empty stored block (cf. Z_SYNC_FLUSH, parallel.c:624-644) */
        putbits(&w, 0, 3); while (w.bitpos & 7) putbits(&w, 0, 1);
        putbits(&w, 0, 16); putbits(&w, 0xffff, 16);
    }
    (void)last_type;
    if (st) { st->tokens += (uint64_t)nt; for (int i = 0; i < nt; i++) st->matches += tok[i] >> 31; }
    free(mlen); free(mdist); free(tok); free(tpos);
    return (size_t)(w.bitpos >> 3);
}

/* ---------------- whole buffer, chunk grid of src/parallel.c ---------------- */
typedef struct { const gzm_params *P; const uint8_t *in; uint64_t n; int halo, final; uint64_t nchunks;
                 uint8_t **frag; size_t *fsize; gzm_chunk_stats *st; int tid, nth; } mjob;

static void *model_worker(void *arg)
{
    mjob *j = (mjob *)arg; const gzm_params *P = j->P;
    for (uint64_t c = (uint64_t)j->tid; c < j->nchunks; c += (uint64_t)j->nth) {
        uint64_t lo = c * (uint64_t)P->chunk, hi = lo + (uint64_t)P->chunk; if (hi > j->n) hi = j->n;
        int D = 0;
        if (P->dict) { uint64_t avail = lo + (uint64_t)j->halo; D = avail < (uint64_t)P->dict ? (int)avail : P->dict; }
        int nc = (int)(hi - lo);
        uint8_t *o = (uint8_t *)malloc((size_t)nc + (size_t)nc / 8 + 1024);
        j->fsize[c] = gzm_deflate_chunk(P, j->in + lo - D, D, nc, j->final && c == j->nchunks - 1, o, &j->st[j->tid]);
        j->frag[c] = o;
    }
    return NULL;
}

int64_t gzm_deflate(const gzm_params *P, const uint8_t *in, uint64_t n, int halo, int final,
                    uint8_t *out, uint64_t cap, int nthreads, gzm_chunk_stats *stats)
{
    uint64_t nchunks = n ? (n + (uint64_t)P->chunk - 1) / (uint64_t)P->chunk : 1;
    uint8_t **frag = (uint8_t **)calloc(nchunks, sizeof(*frag)); size_t *fsize = (size_t *)calloc(nchunks, sizeof(*fsize));
    if (nthreads < 1) nthreads = 1; if (nthreads > 64) nthreads = 64;
    pthread_t th[64]; mjob jobs[64]; gzm_chunk_stats sts[64]; memset(sts, 0, sizeof(sts));
    for (int t = 0; t < nthreads; t++) { jobs[t] = (mjob){P, in, n, halo, final, nchunks, frag, fsize, sts, t, nthreads}; pthread_create(&th[t], NULL, model_worker, &jobs[t]); }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    uint64_t pos = 0; int64_t rc = 0;
    for (uint64_t c = 0; c < nchunks; c++) {
        if (pos + fsize[c] > cap) rc = -1; else memcpy(out + pos, frag[c], fsize[c]);
        pos += fsize[c]; free(frag[c]);
    }
    if (stats) { memset(stats, 0, sizeof(*stats)); for (int t = 0; t < nthreads; t++) { stats->tokens += sts[t].tokens; stats->matches += sts[t].matches; for (int k = 0; k < 3; k++) stats->blocks[k] += sts[t].blocks[k]; } }
    free(frag); free(fsize);
    return rc < 0 ? rc : (int64_t)pos;
}
