/* corpus.c -- deterministic synthetic corpora for the benchmark and the tests.
 *
 * Host-only C.  Not part of the compression path: it only manufactures input
 * bytes (the reference's benchmark inputs are downloads that are not in its
 * tree -- benchmarks/Makefile.am:4-7 -- so every number here is on synthetic
 * data and bench.py says so).
 *
 * Every corpus is defined region by region (1 MiB regions, each seeded from
 * (seed, kind, region index)), so any byte range can be produced independently:
 * GPU shards and host threads fill their own slices and the result does not
 * depend on how the range was cut.
 *
 *   GZB_CORPUS_TEXT    line/word sampler in the style of SURVEY.md App. B:
 *                      50 % whole lines from a fixed pool, 50 % 3-14 words drawn
 *                      by frequency; zlib -6 ratio ~4.6
 *   GZB_CORPUS_ZIPF    Zipfian word stream over an 8192-word vocabulary with
 *                      numbers and punctuation; zlib -6 ratio ~2.75 (natural-English-like)
 *   GZB_CORPUS_BINARY  structured binary: little-endian int32 counters and
 *                      repeated 64-byte records with 10 % random bytes
 *   GZB_CORPUS_RANDOM  uniform random bytes (incompressible)
 *   GZB_CORPUS_MIXED   1 MiB regions cycling TEXT / BINARY / RANDOM (config 3)
 */
#include "gzb_corpus.h"
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <math.h>

#define REGION ((uint64_t)1 << 20)

static inline uint64_t splitmix64(uint64_t *s)
{
    uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline uint32_t rnd32(uint64_t *s) { return (uint32_t)(splitmix64(s) >> 32); }
static inline uint32_t rnd_below(uint64_t *s, uint32_t n) { return (uint32_t)(((uint64_t)rnd32(s) * n) >> 32); }

/* ---- vocabulary (built once, depends on nothing but constants) ---- */
#define VOCAB 32768
#define POOL_LINES 256
#define SAMPLER 65536
typedef struct { uint8_t len; char s[15]; } word_t;
static word_t g_vocab[VOCAB];
static uint16_t g_zipf_sampler[SAMPLER];   /* word id by Zipf(1.5) rank frequency over 8192 words */
static uint16_t g_text_sampler[SAMPLER];   /* steeper distribution over the first 320 words */
static char g_pool[POOL_LINES][128];
static uint8_t g_pool_len[POOL_LINES];
static pthread_once_t g_once = PTHREAD_ONCE_INIT;

static void build_sampler(uint16_t *tab, int nwords, double expo)
{
    double *w = (double *)malloc(sizeof(double) * nwords), tot = 0;
    for (int i = 0; i < nwords; i++) { w[i] = pow((double)i + 2.7, -expo); tot += w[i]; }
    double acc = 0; int k = 0;
    for (int i = 0; i < nwords; i++) {
        acc += w[i] / tot * SAMPLER;
        int hi = (int)(acc + 0.5);
        if (i == nwords - 1) hi = SAMPLER;
        while (k < hi && k < SAMPLER) tab[k++] = (uint16_t)i;
    }
    while (k < SAMPLER) tab[k++] = (uint16_t)(nwords - 1);
    free(w);
}

static void corpus_init(void)
{
    static const char *onset[] = {"", "b", "c", "d", "f", "g", "h", "l", "m", "n", "p", "r", "s", "t", "w", "th", "st", "ch", "sh", "tr", "pr", "br", "gr", "pl", "cl", "v", "k", "qu"};
    static const char *nucleus[] = {"a", "e", "i", "o", "u", "ea", "ou", "ai", "ee", "oo", "io", "ie"};
    static const char *coda[] = {"", "", "n", "r", "s", "t", "l", "m", "d", "ng", "nt", "st", "ck", "ll", "rs", "ss"};
    static const char *common[] = {"the", "of", "and", "to", "a", "in", "is", "that", "it", "for", "was", "on", "with", "as", "be", "at", "by", "this", "have", "from", "or", "had", "not", "but", "what", "all", "were", "when", "we", "there", "can", "an", "your", "which", "their", "said", "if", "do", "will", "each", "about", "how", "up", "out", "them", "then", "she", "many", "some", "so"};
    uint64_t s = 0x5EEDC0DEull;
    int ncommon = (int)(sizeof(common) / sizeof(common[0]));
    for (int i = 0; i < VOCAB; i++) {
        word_t *w = &g_vocab[i];
        if (i < ncommon) { w->len = (uint8_t)strlen(common[i]); memcpy(w->s, common[i], w->len); continue; }
        int syl = 1 + (int)rnd_below(&s, 3) + (i > 2000 ? (int)rnd_below(&s, 2) : 0);
        int n = 0;
        for (int k = 0; k < syl && n < 11; k++) {
            const char *a = onset[rnd_below(&s, 28)], *b = nucleus[rnd_below(&s, 12)], *c = coda[rnd_below(&s, 16)];
            for (; *a && n < 14; a++) w->s[n++] = *a;
            for (; *b && n < 14; b++) w->s[n++] = *b;
            for (; *c && n < 14; c++) w->s[n++] = *c;
        }
        if (rnd_below(&s, 23) == 0) w->s[0] = (char)(w->s[0] - 32);   /* some proper nouns */
        w->len = (uint8_t)n;
    }
    build_sampler(g_zipf_sampler, 8192, 1.5);
    build_sampler(g_text_sampler, 320, 1.15);
    for (int i = 0; i < POOL_LINES; i++) {
        int nw = 4 + (int)rnd_below(&s, 9), n = 0;
        for (int k = 0; k < nw; k++) {
            const word_t *w = &g_vocab[g_text_sampler[rnd32(&s) & (SAMPLER - 1)]];
            if (n + w->len + 2 > 120) break;
            if (k) g_pool[i][n++] = ' ';
            memcpy(&g_pool[i][n], w->s, w->len); n += w->len;
        }
        if (rnd_below(&s, 3) == 0) g_pool[i][n++] = '.';
        g_pool[i][n++] = '\n';
        g_pool_len[i] = (uint8_t)n;
    }
}

/* ---- region generators: fill exactly REGION bytes of `r` (a scratch buffer with slack) ---- */
#define SLACK 512

static void region_text(uint64_t *s, uint8_t *r)
{
    size_t n = 0;
    while (n < REGION) {
        if (rnd32(s) & 1) {
            int i = (int)rnd_below(s, POOL_LINES);
            memcpy(r + n, g_pool[i], g_pool_len[i]); n += g_pool_len[i];
        } else {
            int nw = 3 + (int)rnd_below(s, 12);
            for (int k = 0; k < nw; k++) {
                const word_t *w = &g_vocab[g_text_sampler[rnd32(s) & (SAMPLER - 1)]];
                if (k) r[n++] = ' ';
                memcpy(r + n, w->s, w->len); n += w->len;
            }
            r[n++] = '\n';
        }
    }
}

static void region_zipf(uint64_t *s, uint8_t *r)
{
    size_t n = 0; int col = 0, sentence = 0;
    while (n < REGION) {
        uint32_t x = rnd32(s);
        const word_t *w = &g_vocab[g_zipf_sampler[x & (SAMPLER - 1)]];
        if ((x >> 16) % 41 == 0) {                       /* a number now and then */
            uint32_t v = rnd32(s) % 100000; char tmp[8]; int k = 0;
            do { tmp[k++] = (char)('0' + v % 10); v /= 10; } while (v);
            while (k) r[n++] = (uint8_t)tmp[--k];
            col += 5;
        } else {
            memcpy(r + n, w->s, w->len);
            if (sentence == 0 && r[n] >= 'a' && r[n] <= 'z') r[n] = (uint8_t)(r[n] - 32);
            n += w->len; col += w->len;
        }
        sentence++;
        uint32_t y = (x >> 22) & 63;
        if (sentence > 4 && y < 7) { r[n++] = (y < 5) ? '.' : (y == 5 ? '?' : '!'); sentence = 0; }
        else if (y >= 58) r[n++] = ',';
        if (col > 68) { r[n++] = '\n'; col = 0; if (((x >> 28) & 7) == 0) r[n++] = '\n'; }
        else r[n++] = ' ';
        col++;
    }
}

static void region_binary(uint64_t *s, uint8_t *r, uint64_t region_index)
{
    /* alternating 64 KiB spans: int32 LE counters with a random stride / 64-byte records, 10 % random bytes */
    size_t n = 0; int span = 0;
    while (n < REGION) {
        if ((span++ & 1) == 0) {
            uint32_t v = (uint32_t)(region_index * 2654435761u) + rnd32(s) % 1000, step = 1 + rnd_below(s, 4);
            for (int i = 0; i < 16384; i++, v += step) { memcpy(r + n, &v, 4); n += 4; }
        } else {
            uint8_t rec[64];
            for (int i = 0; i < 64; i++) rec[i] = (uint8_t)rnd32(s);
            rec[0] = 0xCA; rec[1] = 0xFE; rec[62] = 0; rec[63] = 0;
            for (int k = 0; k < 1024; k++) {
                memcpy(r + n, rec, 64);
                for (int j = 0; j < 6; j++) r[n + 4 + rnd_below(s, 56)] = (uint8_t)rnd32(s);   /* ~10 % */
                uint16_t seq = (uint16_t)k; memcpy(r + n + 2, &seq, 2);
                n += 64;
            }
        }
    }
}

static void region_random(uint64_t *s, uint8_t *r)
{
    for (size_t n = 0; n < REGION; n += 8) { uint64_t v = splitmix64(s); memcpy(r + n, &v, 8); }
}

static void gen_region(int kind, uint64_t seed, uint64_t idx, uint8_t *scratch)
{
    uint64_t s = seed * 0xD1342543DE82EF95ull + idx * 0x9E3779B97F4A7C15ull + (uint64_t)kind * 0x632BE59BD9B4E019ull;
    (void)splitmix64(&s);
    int k = kind;
    if (kind == GZB_CORPUS_MIXED) { static const int cyc[3] = {GZB_CORPUS_TEXT, GZB_CORPUS_BINARY, GZB_CORPUS_RANDOM}; k = cyc[idx % 3]; }
    switch (k) {
    case GZB_CORPUS_TEXT:   region_text(&s, scratch); break;
    case GZB_CORPUS_ZIPF:   region_zipf(&s, scratch); break;
    case GZB_CORPUS_BINARY: region_binary(&s, scratch, idx); break;
    default:                region_random(&s, scratch); break;
    }
}

typedef struct { int kind; uint64_t seed, offset, n; uint8_t *dst; int tid, nthreads; } job_t;

static void *fill_worker(void *arg)
{
    job_t *j = (job_t *)arg;
    uint8_t *scratch = (uint8_t *)malloc(REGION + SLACK);
    if (!scratch) return (void *)1;
    uint64_t first = j->offset / REGION, last = (j->offset + j->n - 1) / REGION;
    for (uint64_t r = first + (uint64_t)j->tid; r <= last; r += (uint64_t)j->nthreads) {
        gen_region(j->kind, j->seed, r, scratch);
        uint64_t lo = r * REGION, hi = lo + REGION;
        uint64_t a = lo > j->offset ? lo : j->offset, b = hi < j->offset + j->n ? hi : j->offset + j->n;
        memcpy(j->dst + (a - j->offset), scratch + (a - lo), b - a);
    }
    free(scratch);
    return NULL;
}

int gzb_corpus_fill(int kind, uint64_t seed, uint64_t offset, uint8_t *dst, uint64_t n, int nthreads)
{
    if (kind < GZB_CORPUS_TEXT || kind > GZB_CORPUS_MIXED) return -1;
    if (n == 0) return 0;
    if (!dst) return -1;
    pthread_once(&g_once, corpus_init);
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 64) nthreads = 64;
    uint64_t regions = (offset + n - 1) / REGION - offset / REGION + 1;
    if ((uint64_t)nthreads > regions) nthreads = (int)regions;
    pthread_t th[64]; job_t jobs[64]; int rc = 0;
    for (int t = 0; t < nthreads; t++) {
        jobs[t] = (job_t){kind, seed, offset, n, dst, t, nthreads};
        if (t > 0 && pthread_create(&th[t], NULL, fill_worker, &jobs[t]) != 0) { th[t] = 0; if (fill_worker(&jobs[t])) rc = -1; jobs[t].tid = -1; }
    }
    if (fill_worker(&jobs[0])) rc = -1;
    for (int t = 1; t < nthreads; t++) if (jobs[t].tid >= 0) { void *res; pthread_join(th[t], &res); if (res) rc = -1; }
    return rc;
}
