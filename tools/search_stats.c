/* Exploration: per-position and per-32-group (lockstep) cost of the engine's match search on a corpus.
 *   gcc -O2 -I include -o /tmp/search_stats tools/search_stats.c gzippier_b200/csrc/corpus.c -lpthread -lm && /tmp/search_stats 0 16 8 128
 * Not part of the product or the tests. */
#include "../oracle/engine_model.c"
#include <stdio.h>
int gzb_corpus_fill(int kind, uint64_t seed, uint64_t offset, void *dst, uint64_t n, int threads);
int main(int argc, char **argv)
{
    int kind = argc > 1 ? atoi(argv[1]) : 0, mb = argc > 2 ? atoi(argv[2]) : 16, depth = argc > 3 ? atoi(argv[3]) : 8, nicep = argc > 4 ? atoi(argv[4]) : 128, good = argc > 5 ? atoi(argv[5]) : 0;
    size_t n = (size_t)mb << 20;
    uint8_t *s = malloc(n + 8); memset(s + n, 0, 8);
    gzb_corpus_fill(kind, 1234, 0, s, n, 8);
    int hbits = 14, hsize = 1 << hbits;
    int32_t *head = malloc(4 * hsize), *link = malloc(4 * n);
    for (int i = 0; i < hsize; i++) head[i] = -1;
    double hist[16] = {0}, cmp_at[16] = {0}, improved = 0, f2 = 0;
    double steps = 0, pass = 0, cmps = 0, cbytes = 0, gsteps = 0, gcmp = 0, gwords = 0, groups = 0, found6 = 0;
    int lsteps[32], lcmp[32], lwords[32];
    for (size_t p = 0; p < n; p++) {
        uint32_t lo = ld32(s + p), hi16 = s[p + 4] | (s[p + 5] << 8);
        uint32_t h = gzm_hash6(lo, hi16, hbits);
        int32_t prev = head[h]; link[p] = (prev >= 0 && p - prev <= 32767) ? prev : -1; head[h] = (int32_t)p;
        int maxlen = n - p < 258 ? (int)(n - p) : 258, best = 2, limit = p < 32768 ? (int)p : 32768, nice = nicep < maxlen ? nicep : maxlen;
        int c = maxlen >= 6 ? link[p] : -1, d = depth, st = 0, nc = 0, nw = 0, cut = 0;
        while (c >= 0 && (int)p - c <= limit && d-- > 0) {
            st++;
            if (maxlen > best && s[c + best] == s[p + best]) {
                int l = common_len(s + c, s + p, maxlen);
                nc++; nw += l / 4 + 1; pass++; cbytes += l; cmp_at[st]++; if (l > best) improved++; if (maxlen > best + 1 && best >= 3 && s[c + best - 1] != s[p + best - 1]) f2++;
                if (l > best) { best = l; if (l >= nice) break; if (good && !cut && best >= good) { d >>= 2; cut = 1; } }
            }
            c = link[c];
        }
        if (best >= 6) found6++;
        steps += st; cmps += nc; hist[st]++;
        int ln = p & 31; lsteps[ln] = st; lcmp[ln] = nc; lwords[ln] = nw;
        if (ln == 31) {
            int ms = 0, mc = 0, mw = 0;
            for (int i = 0; i < 32; i++) { if (lsteps[i] > ms) ms = lsteps[i]; if (lcmp[i] > mc) mc = lcmp[i]; if (lwords[i] > mw) mw = lwords[i]; }
            gsteps += ms; gcmp += mc; gwords += mw; groups++;
        }
    }
    printf("kind %d depth %d nice %d: per position: chain steps %.2f, compares %.2f, compared bytes %.1f, found>=6 %.3f | per group (max over lanes): steps %.2f, compares %.2f, compare words %.1f\n",
           kind, depth, nicep, steps / n, cmps / n, cbytes / n, found6 / n, gsteps / groups, gcmp / groups, gwords / groups);
    printf("  compares that improved: %.3f per position; compares a second filter byte (best-1) would have skipped: %.3f\n", improved / n, f2 / n);
    printf("  steps histogram:"); for (int i = 0; i <= depth; i++) printf(" %d:%.3f", i, hist[i] / n); printf("\n  compares at step:"); for (int i = 1; i <= depth; i++) printf(" %d:%.3f", i, cmp_at[i] / n); printf("\n");
    return 0;
}
