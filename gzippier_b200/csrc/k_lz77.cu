// k_lz77.cu -- LZ77 match finding + greedy/lazy parse + literal/length/distance histograms.
//
// Replaces, for a run of chunks per CTA: deflateSetDictionary (lib/zlib/deflate.c:376-442),
// INSERT_STRING (:182-185), longest_match (:1236-1377), deflate_fast (:1824-1918),
// deflate_slow (:1926-2049) and _tr_tally (deflate.h:326-342).  Not a port: zlib walks one
// position at a time and skips positions inside matches.  Here EVERY position is inserted, in
// exact stream order, into hash chains keyed on 6 bytes (plus single-entry 4-byte and 3-byte
// tables), which makes match(p) a pure function of the data; every position of a 2 KiB tile is
// searched in parallel (one lane per position), and the greedy/lazy parse -- a pure function
// next(p) of the per-position matches -- is evaluated with pointer doubling instead of a walk.
//
// One persistent CTA (1024 threads, ~226 KB shared memory) per SM takes runs of up to 32
// consecutive chunks from an atomic counter and streams through them: the tables at the end of
// a chunk ARE the next chunk's dictionary, so only the first chunk of a run inserts its 32 KiB
// dictionary.  Per tile k the CTA runs three things at once (one barrier per tile):
//   3 warps      insert tile k+1 into the three tables: the serial part only (head exchange)
//   8 warps      parse + emit tile k-1 (tokens, histograms, block / chunk records)
//   all others   search tile k in 32-position groups handed out dynamically; everybody joins
//                when its own role is done, then turns tile k+1's raw heads into links
// Shared memory holds the whole search state:
//   ring   64 KiB  window bytes, stream position & 0xFFFF
//   link   72 KiB  u16 delta to the previous position in the same 6-byte bucket (0xFFFF = none),
//                  indexed by stream position mod LINK_RING; deltas need no sliding
//   head   32 KiB  u16 last position (mod 64 Ki) per 6-byte bucket -- inserter-private
//   head3/4 8+8 KiB u16 last position per 3-byte / 4-byte bucket (single entry, no chain)
//   c3/c4  8+8 KiB per-position delta to the most recent 3-byte / 4-byte twin, double-buffered per tile
//   mlen/mdist 12 KiB  per-position best match of the last two tiles
//   jump    8 KiB  parse scratch (window exit functions)
// Stale/aliased head entries are detected by re-hashing the bytes at the candidate, which makes
// the chains exactly "all earlier positions of the same bucket, most recent first"
// (oracle/engine_model.c is the sequential statement of the same semantics).
#include "gzb_common.cuh"

namespace gzb {

// cycle counters (debugging aid, tools/lzprof.py), summed over CTAs: 7 pipeline iterations, 8-10 time in insert_tile of the
// 6-byte / 3-byte / 4-byte table warp, 11 time in parse_emit_tile (thread 0 of the parser group)
__device__ unsigned long long g_lzprof[16];

namespace {

constexpr uint32_t FULL = 0xFFFFFFFFu;
constexpr uint32_t NOPOS = 0xFFFFFFFFu;
constexpr int GROUPS_PER_TILE = TILE / 32;
constexpr int RING_MIRROR = 288;         // >= 258 + 3 + 19 (longest compare, see match_len_s) + 4, multiple of 16
constexpr int MRING = 2 * TILE;           // per-position match results kept for two tiles

struct Smem {
    uint32_t ring32[RING / 4 + RING_MIRROR / 4];   // + mirror of the first bytes: compare loops never wrap
    uint16_t link[LINK_RING];
    uint16_t head[1 << HBITS];
    uint16_t head3[1 << H3BITS];
    uint16_t head4[1 << H4BITS];
    uint16_t c3[2][TILE];
    uint16_t c4[2][TILE];
    uint16_t mdist[MRING];                // dist-1, 0xFFFF = no match
    uint8_t  mlen[MRING];                 // len-3
    uint32_t lfreq[288];
    uint32_t dfreq[32];
    uint32_t masks[TILE / 32];            // token-start bitmap of the tile being parsed
    uint32_t seg_prefix[TILE / 32 + 1];
    uint16_t jump[2][TILE];               // pointer doubling: 2^r-th successor of every position (TILE = leaves the tile)
    uint32_t grp_ctr[2];                  // dynamic group scheduler of the search phase (alternating per iteration)
    uint32_t gctr[3];                     // flowing pipeline: group scheduler of tile t at [t % 3]
    uint32_t ins_total;                   // ... tiles inserted, summed over the three inserter warps
    uint32_t parsed;                      // ... tiles of the run parsed
    uint32_t b1_done;                     // ... k + 2 once the search barrier of iteration k has completed
    uint32_t chunk_idx;
    uint32_t entry;                       // next token start (stream position)
    uint32_t ntok;                        // tokens emitted so far in this chunk
    uint32_t blk;                         // current block index
    uint32_t blk_tok0;                    // token index where the current block began
    uint32_t blk_in0;                     // chunk-relative input offset where it began
    uint32_t extra_lit;                   // 1 if the final position needs its own literal
    uint32_t carry_pos, carry_len, carry_dist;   // match of the last position of the tile before the one being parsed
                                          // (its slot in mlen/mdist is rewritten by the concurrent search)
};

__device__ __forceinline__ uint32_t ring_ld32(const uint32_t* ring32, uint32_t pos)
{
    uint32_t a = pos & 0xFFFFu, i = a >> 2;
    uint32_t w0 = ring32[i], w1 = ring32[i + 1];               // i + 1 may be the mirror
    return __funnelshift_r(w0, w1, (a & 3u) * 8u);
}
__device__ __forceinline__ uint32_t ring_ld8(const uint32_t* ring32, uint32_t pos)
{
    return reinterpret_cast<const uint8_t*>(ring32)[pos & 0xFFFFu];
}

// The search loop addresses shared memory by 32-bit shared-space address: one IADD per access, no generic-pointer
// arithmetic, and the compiler has nothing to re-derive inside the loop.
__device__ __forceinline__ uint32_t lds_u8(uint32_t a)  { uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) { uint32_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(a), "h"((uint16_t)v) : "memory"); }
template <int OFF>
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(a), "n"(OFF)); return v; }

// Length of the common prefix of the strings at stream positions p and c, at most maxlen (>= 1); four words per trip
// (reads up to 19 bytes past maxlen: RING_MIRROR covers it).
__device__ __forceinline__ int match_len_s(uint32_t ring_s, uint32_t p, uint32_t c, int maxlen)
{
    const uint32_t ap = p & 0xFFFFu, ac = c & 0xFFFFu;
    uint32_t wp = ring_s + (ap & 0xFFFCu), wc = ring_s + (ac & 0xFFFCu);
    const uint32_t sp = (ap & 3u) * 8u, sc = (ac & 3u) * 8u;
    uint32_t p0 = lds_u32<0>(wp), c0 = lds_u32<0>(wc);
    for (int done = 0; done < maxlen; done += 16, wp += 16u, wc += 16u) {
        uint32_t p1, c1, x;
#define GZB_ML_STEP(K)                                                                        \
        p1 = lds_u32<4 * (K) + 4>(wp); c1 = lds_u32<4 * (K) + 4>(wc);                          \
        x = __funnelshift_r(p0, p1, sp) ^ __funnelshift_r(c0, c1, sc);                        \
        if (x) return min(maxlen, done + 4 * (K) + ((__ffs(x) - 1) >> 3));                    \
        p0 = p1; c0 = c1;
        GZB_ML_STEP(0) GZB_ML_STEP(1) GZB_ML_STEP(2) GZB_ML_STEP(3)
#undef GZB_ML_STEP
    }
    return maxlen;
}

// Copy stream bytes [t*TILE, (t+1)*TILE) into the ring, zero beyond the end of the stream.
// nthr threads of the caller cooperate (tid in [0, nthr)), 16 bytes per thread per step.
__device__ __forceinline__ uint4 tile_fetch(const uint8_t* sp, uint32_t n_s, uint32_t pos)
{
    uint4 v = make_uint4(0, 0, 0, 0);
    if (pos + 16u <= n_s) v = __ldg(reinterpret_cast<const uint4*>(sp + pos));
    else if (pos < n_s) {
        uint32_t w[4] = {0, 0, 0, 0};
        for (uint32_t k = 0; pos + k < n_s; k++) w[k >> 2] |= (uint32_t)sp[pos + k] << ((k & 3u) * 8u);
        v = make_uint4(w[0], w[1], w[2], w[3]);
    }
    return v;
}
__device__ __forceinline__ void tile_store(Smem& S, uint32_t pos, uint4 v)
{
    const uint32_t ra = pos & 0xFFFFu;
    *reinterpret_cast<uint4*>(&S.ring32[ra >> 2]) = v;
    if (ra < (uint32_t)RING_MIRROR) *reinterpret_cast<uint4*>(&S.ring32[(RING + ra) >> 2]) = v;
}
__device__ __forceinline__ void load_tile(Smem& S, const uint8_t* sp, uint32_t n_s, uint32_t t, uint32_t tid, uint32_t nthr)
{
    uint32_t base = t * TILE;
    for (uint32_t o = tid * 16u; o < TILE; o += nthr * 16u) tile_store(S, base + o, tile_fetch(sp, n_s, base + o));
}

// The same copy with cp.async: no registers, no waiting -- issued before a warp's own work, awaited after it.
__device__ __forceinline__ void load_tile_async(uint32_t sbase, const uint8_t* sp, uint32_t n_s, uint32_t t, uint32_t lane)
{
    const uint32_t ring_s = sbase + (uint32_t)offsetof(Smem, ring32);
#pragma unroll
    for (uint32_t j = 0; j < (uint32_t)TILE / 512u; j++) {
        const uint32_t pos = t * TILE + (j * 32u + lane) * 16u;
        const uint32_t nb = pos < n_s ? min(16u, n_s - pos) : 0u;            // bytes past the end are zero-filled
        const uint8_t* src = sp + (nb ? pos : 0u);
        const uint32_t ra = pos & 0xFFFFu;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" :: "r"(ring_s + ra), "l"(src), "r"(nb) : "memory");
        if (ra < (uint32_t)RING_MIRROR)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" :: "r"(ring_s + (uint32_t)RING + ra), "l"(src), "r"(nb) : "memory");
    }
}
__device__ __forceinline__ void load_tile_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// ---- insertion ---------------------------------------------------------------------------------------
// Inserting a tile into a hash table is serial per bucket, and one warp doing 64 steps of 32 positions is
// latency-bound -- all the more so next to 29 busy search warps, since a warp issues in order and every
// instruction of the step waits for its turn.  So the step is stripped to what IS serial:
//   hash_tile    (all threads, before)  bucket of every position of the tile, for the three tables
//   insert_tile  (one warp per table)   old = head[h]; head[h] = p; store `old` raw.  Lanes of a step that share
//                                       a bucket are detected by write-then-read-back and ordered with ballots
//                                       (MATCH.ANY would tell the same, at ~500 cycles per step)
//   finish_tile  (all threads, after)   raw head -> delta; sweep_tables ages the entries, so a range test is all it
//                                       takes for a chain to be exactly "all earlier positions of the bucket
//                                       within 32 Ki, most recent first"
// KIND 0: 6-byte buckets -> S.link (chains); KIND 1 / 2: 3-byte / 4-byte buckets -> S.c3 (within TOO_FAR) / S.c4.
template <int KIND>
__device__ __forceinline__ uint32_t bucket_of(uint32_t v, uint32_t v2)
{
    return KIND == 0 ? hash6(v, v2 & 0xFFFFu) : KIND == 1 ? hash3(v) : hash4(v);
}

// N (2 or 4) consecutive positions per thread, o0 a multiple of N: three ring words cover the N + 7 bytes they need, and
// the N buckets of a table go out as one store.
template <int N>
__device__ __forceinline__ void hash_span(Smem& S, uint32_t t, uint32_t o0)
{
    // the bucket of a position is parked where its link / c3 / c4 entry will go: those slots are dead until insertion
    static_assert(N == 2 || N == 4, "hash_span");
    const uint32_t a = (t * TILE + o0) & 0xFFFFu, i = a >> 2, sh = (a & 3u) * 8u;     // sh = 0 (N == 4), 0 or 16 (N == 2)
    const uint32_t w0 = S.ring32[i], w1 = S.ring32[i + 1], w2 = S.ring32[i + 2];      // may be the mirror; zero past the end of the stream
    uint32_t b6[N], b3[N], b4[N];
#pragma unroll
    for (int j = 0; j < N; j++) {
        const uint32_t v = __funnelshift_r(w0, w1, sh + 8u * j), v2 = __funnelshift_r(w1, w2, sh + 8u * j);
        b6[j] = bucket_of<0>(v, v2); b3[j] = bucket_of<1>(v, v2); b4[j] = bucket_of<2>(v, v2);
    }
    const uint32_t slot = (t * TILE) % LINK_RING + o0;                                // a tile's link slots do not wrap
    if (N == 2) {
        *reinterpret_cast<uint32_t*>(&S.link[slot]) = b6[0] | (b6[1] << 16);
        *reinterpret_cast<uint32_t*>(&S.c3[t & 1][o0]) = b3[0] | (b3[1] << 16);
        *reinterpret_cast<uint32_t*>(&S.c4[t & 1][o0]) = b4[0] | (b4[1] << 16);
    } else {
        *reinterpret_cast<uint2*>(&S.link[slot]) = make_uint2(b6[0] | (b6[1] << 16), b6[2] | (b6[3] << 16));
        *reinterpret_cast<uint2*>(&S.c3[t & 1][o0]) = make_uint2(b3[0] | (b3[1] << 16), b3[2] | (b3[3] << 16));
        *reinterpret_cast<uint2*>(&S.c4[t & 1][o0]) = make_uint2(b4[0] | (b4[1] << 16), b4[2] | (b4[3] << 16));
    }
}
__device__ void hash_tile_n(Smem& S, uint32_t t, uint32_t idx, uint32_t nthr)
{
    for (uint32_t q = idx; q < (uint32_t)TILE / 4u; q += nthr) hash_span<4>(S, t, q * 4u);
}
__device__ void hash_tile(Smem& S, uint32_t t, uint32_t tid)
{
    static_assert(TILE == 2 * LZ_THREADS, "hash_tile: one pair of positions per thread");
    hash_span<2>(S, t, tid * 2u);
}

// One step = 32 consecutive positions, lane = position.  The loop is written on shared-space addresses and kept as
// short as it goes: a step costs its instruction count times the issue interval of a warp that shares its scheduler
// with seven busy ones.
// raw head (16-bit position) of position p16 -> what the search reads: see finish_pos
template <int KIND>
__device__ __forceinline__ uint32_t to_delta(uint32_t p16, uint32_t raw)
{
    const uint32_t d = (p16 - raw) & 0xFFFFu;
    if (KIND == 0) return d - 1u < 32767u ? d : 0xFFFFu;
    return d - 1u < (KIND == 1 ? (uint32_t)TOO_FAR : 32767u) ? d : 0u;
}

// DIRECT: store the final delta instead of the raw head (no finish pass: the flowing pipeline needs the links of a tile
// final as soon as its inserters are through).
template <int KIND, bool DIRECT>
__device__ void insert_tile(Smem& S, uint32_t sbase, uint32_t t, uint32_t n_s, uint32_t lane)
{
    (void)S;
    const uint32_t base = t * TILE;                                           // < n_s
    const uint32_t nsteps = min((uint32_t)GROUPS_PER_TILE, (n_s - base + 31u) / 32u);   // nothing past the end is ever searched
    const uint32_t table_s = sbase + (uint32_t)(KIND == 0 ? offsetof(Smem, head) : KIND == 1 ? offsetof(Smem, head3) : offsetof(Smem, head4));
    // where the buckets are parked and the raw heads go; a tile's link slots are contiguous (LINK_RING % TILE == 0)
    static_assert(LINK_RING % TILE == 0, "a tile's link slots must not wrap");
    uint32_t dst = sbase + (KIND == 0 ? (uint32_t)offsetof(Smem, link) + 2u * (base % LINK_RING + lane)
                                      : (uint32_t)(KIND == 1 ? offsetof(Smem, c3) : offsetof(Smem, c4)) + 2u * ((t & 1u) * TILE + lane));
    uint32_t p16 = (base + lane) & 0xFFFFu;                                   // + 32 per step: no wrap inside a tile
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t ta = table_s + 2u * lds_u16(dst);
    for (uint32_t i = 0; i < nsteps; i++, dst += 64u, p16 += 32u) {
        const uint32_t hn = lds_u16(dst + 64u);                              // next step's bucket (garbage after the last step, unused)
        uint32_t raw = lds_u16(ta);
        // Write, read back: a lane that does not find its own position shares its bucket with another lane of the step.
        // The losers then write once more: a bucket shared by exactly TWO lanes -- the usual case -- is settled by these
        // two exchanges (each lane has read the other's position); only buckets with three or more lanes of the step
        // need MATCH.ANY (~450 cycles).
        __syncwarp();
        sts_u16(ta, p16);
        __syncwarp();
        const uint32_t r1 = lds_u16(ta);
        const bool lost = r1 != p16;
        if (__any_sync(FULL, lost)) {                            // uniform
            __syncwarp();
            if (lost) sts_u16(ta, p16);
            __syncwarp();
            const uint32_t r2 = lds_u16(ta);
            if (!__any_sync(FULL, lost && r2 != p16)) {
                const bool paired = lost || r2 != p16;
                const uint32_t partner = lost ? r1 : r2;
                const bool partner_lower = partner < p16;                    // same tile: no wrap
                if (paired && partner_lower) raw = partner;                  // the other lane of the bucket precedes me
                __syncwarp();
                if (paired && partner_lower && !lost) sts_u16(ta, p16);      // the later position owns the bucket
            } else {
                const uint32_t same = __match_any_sync(FULL, ta);
                const uint32_t lower = same & lt_mask;
                if (lower) raw = p16 - (lane - (uint32_t)(31 - __clz(lower)));   // the nearest lower lane of the bucket
                if ((same >> lane) == 1u && lower) sts_u16(ta, p16);             // the last position of the step owns the bucket
            }
        }
        sts_u16(dst, DIRECT ? to_delta<KIND>(p16, raw) : raw);
        ta = table_s + 2u * hn;
        __syncwarp();
    }
}

// Raw head (a 16-bit position) -> delta.  No check against the bytes is needed: sweep_tables keeps every table entry
// either younger than 32 Ki positions -- then it is a position of this bucket, the most recent one -- or at an age in
// [32 Ki, 64 Ki), which the range test rejects.
__device__ __forceinline__ void finish_pos(Smem& S, uint32_t t, uint32_t n_s, uint32_t o)
{
    const uint32_t base = t * TILE, p = base + o;
    if (p >= n_s) return;
    uint32_t slot = base % LINK_RING + o; if (slot >= LINK_RING) slot -= LINK_RING;
    const uint32_t d6 = (p - (uint32_t)S.link[slot]) & 0xFFFFu;
    const uint32_t d3 = (p - (uint32_t)S.c3[t & 1][o]) & 0xFFFFu;
    const uint32_t d4 = (p - (uint32_t)S.c4[t & 1][o]) & 0xFFFFu;
    S.link[slot] = (uint16_t)(d6 - 1u < 32767u ? d6 : 0xFFFFu);
    S.c3[t & 1][o] = (uint16_t)(d3 - 1u < (uint32_t)TOO_FAR ? d3 : 0u);
    S.c4[t & 1][o] = (uint16_t)(d4 - 1u < 32767u ? d4 : 0u);
}
__device__ __forceinline__ void finish_tile(Smem& S, uint32_t t, uint32_t n_s, uint32_t tid)
{
#pragma unroll
    for (int j = 0; j < TILE / LZ_THREADS; j++) finish_pos(S, t, n_s, tid + (uint32_t)j * LZ_THREADS);
}

// Ageing of the 16-bit table entries.  Before tile t is inserted, one eighth of every table is swept: an entry older
// than 32767 positions is replaced by the position exactly 32 Ki back.  Every entry is visited once per 16 Ki positions,
// so no entry ever reaches the age of 64 Ki at which its 16 bits would alias a recent position; the tables start a
// run filled with 0x8000, which is that marker for position 0.
constexpr uint32_t TABLE_FILL = 0x80008000u;
__device__ __forceinline__ uint32_t sweep_pair(uint32_t w, uint32_t p16)
{
    const uint32_t mark = p16 ^ 0x8000u;
    uint32_t lo = w & 0xFFFFu, hi = w >> 16;
    if (((p16 - lo) & 0xFFFFu) > 32767u) lo = mark;
    if (((p16 - hi) & 0xFFFFu) > 32767u) hi = mark;
    return lo | (hi << 16);
}
__device__ __forceinline__ void sweep_tables_n(Smem& S, uint32_t t, uint32_t idx, uint32_t nthr)
{
    const uint32_t p16 = (t * TILE) & 0xFFFFu, slice = t & 7u;
    uint32_t* h6 = reinterpret_cast<uint32_t*>(S.head) + slice * LZ_THREADS;
    for (uint32_t i = idx; i < (uint32_t)LZ_THREADS; i += nthr) h6[i] = sweep_pair(h6[i], p16);
    for (uint32_t i = idx; i < 512u; i += nthr) {
        uint32_t* hx = reinterpret_cast<uint32_t*>(i < 256u ? S.head3 : S.head4) + slice * 256u + (i & 255u);
        *hx = sweep_pair(*hx, p16);
    }
}
__device__ __forceinline__ void sweep_tables(Smem& S, uint32_t t, uint32_t tid)
{
    const uint32_t p16 = (t * TILE) & 0xFFFFu, slice = t & 7u;
    static_assert((1u << HBITS) / 2 / 8 == LZ_THREADS && (1u << H3BITS) / 2 / 8 == 256 && (1u << H4BITS) / 2 / 8 == 256, "sweep slices");
    uint32_t* h6 = reinterpret_cast<uint32_t*>(S.head) + slice * LZ_THREADS + tid;
    *h6 = sweep_pair(*h6, p16);
    if (tid < 512u) {
        uint32_t* hx = reinterpret_cast<uint32_t*>(tid < 256u ? S.head3 : S.head4) + slice * 256u + (tid & 255u);
        *hx = sweep_pair(*hx, p16);
    }
}

// Best match for the 32 positions of one group (one warp).  Semantics: engine_model.c gzm_find_matches.
__device__ void search_group(Smem& S, uint32_t sbase, const LevelParams& lp, uint32_t t, uint32_t g, uint32_t n_s, uint32_t lane)
{
    uint32_t p = t * TILE + g * 32u + lane;
    int maxlen = (p < n_s) ? min(258, (int)(n_s - p)) : 0;
    int best = 2; uint32_t bdist = 0;
    // The chain walk alternates two converged phases:
    //   A: WALK_STEPS lockstep chain steps; a candidate that passes the one-byte filter is only noted
    //      (up to two per lane; a lane with two noted candidates sits out the rest of the phase)
    //   B: the noted candidates are measured, first noted first, by all lanes that hold one
    // Filtering in A uses the best length as of the last B, which is a weaker (still necessary)
    // condition; B re-applies the exact filter.  Candidates are judged in chain order with the
    // "strictly longer wins" rule, so the result equals the one-candidate-at-a-time walk.
    constexpr int WALK_STEPS = 4;
    const uint32_t ring_s = sbase + (uint32_t)offsetof(Smem, ring32), link_s = sbase + (uint32_t)offsetof(Smem, link);
    const uint32_t limit = min((uint32_t)MAXD, p);
    const int nice = min(lp.nice, maxlen);
    constexpr uint32_t END = 0xFFFFu;                           // delta of a chain that is over: d + END > limit
    uint32_t slot2 = (p % LINK_RING) * 2u;                      // byte offset of the link slot the walk stands on
    uint32_t delta = maxlen >= 6 ? lds_u16(link_s + slot2) : END;   // the last five positions of a chunk walk no chain
    uint32_t d = 0;
    int depth = lp.depth;
    uint32_t pb = p + (uint32_t)best;                           // filter byte: the one after the best match so far
    uint32_t scan_end = lds_u8(ring_s + (pb & 0xFFFFu));
    while (__any_sync(FULL, delta != END)) {
        uint32_t pend0 = 0, pend1 = 0;
#pragma unroll
        for (int st = 0; st < WALK_STEPS; st++) {
            const uint32_t nd = d + delta;
            if (pend1 == 0u) {
                if (nd > limit) delta = END;
                else {
                    d = nd;
                    slot2 -= 2u * delta; if ((int32_t)slot2 < 0) slot2 += 2u * LINK_RING;
                    const uint32_t b = lds_u8(ring_s + ((pb - d) & 0xFFFFu));
                    delta = lds_u16(link_s + slot2);
                    if (--depth == 0) delta = END;
                    if (b == scan_end) { if (pend0) pend1 = d; else pend0 = d; }
                }
            }
        }
        // measure: first noted first; the second one only if it still passes the filter
        while (__any_sync(FULL, pend0 != 0u)) {
            if (pend0) {
                const int l = match_len_s(ring_s, p, p - pend0, maxlen);
                if (l > best) {
                    best = l; bdist = pend0; pb = p + (uint32_t)best;
                    if (l >= nice || l >= maxlen) { delta = END; pend1 = 0; }
                    else {
                        scan_end = lds_u8(ring_s + (pb & 0xFFFFu));
                        if (pend1 && lds_u8(ring_s + ((pb - pend1) & 0xFFFFu)) != scan_end) pend1 = 0;
                    }
                }
                pend0 = pend1; pend1 = 0;
            }
        }
    }
    if (maxlen >= 3) {
        if (best < 6 && lp.use_h4) {
            uint32_t d4 = S.c4[t & 1][g * 32u + lane];
            if (d4 != 0u && d4 <= limit && maxlen >= 4) {
                int l = match_len_s(ring_s, p, p - d4, maxlen);
                if (l >= 4 && l > best) { best = l; bdist = d4; }
            }
        }
        if (best < 3 && lp.use_h3) {
            uint32_t d3 = S.c3[t & 1][g * 32u + lane];
            if (d3 != 0u && d3 <= min(limit, (uint32_t)TOO_FAR)) {
                int l = match_len_s(ring_s, p, p - d3, maxlen);
                if (l >= 3) { best = l; bdist = d3; }
            }
        }
        if (best == 3 && bdist > (uint32_t)TOO_FAR) best = 2;
    }
    uint32_t mi = p & (MRING - 1);
    if (best >= 3) { S.mlen[mi] = (uint8_t)(best - 3); S.mdist[mi] = (uint16_t)(bdist - 1u); }
    else { S.mlen[mi] = 0; S.mdist[mi] = 0xFFFFu; }
}

// next(p): the parse rule (engine_model.c gzm_parse).  Returns the step from p and whether p is a literal.
__device__ __forceinline__ uint32_t parse_step(const Smem& S, const LevelParams& lp, uint32_t p, uint32_t n_s, bool& lit, uint32_t& dist)
{
    uint32_t md, L;
    if (p == S.carry_pos) { md = S.carry_dist; L = S.carry_len + 3u; }
    else { const uint32_t mi = p & (MRING - 1); md = S.mdist[mi]; L = (uint32_t)S.mlen[mi] + 3u; }
    lit = (md == 0xFFFFu);
    dist = md + 1u;
    if (!lit && lp.lazy_parse && L < (uint32_t)lp.max_lazy && p + 1u < n_s) {
        const uint32_t mj = (p + 1u) & (MRING - 1);
        if (S.mdist[mj] != 0xFFFFu && (uint32_t)S.mlen[mj] + 3u > L) lit = true;
    }
    return lit ? 1u : L;
}

// ---- parse + emission of one tile, by the parser group (warps 0..PARSE_WARPS-1, named barrier 1) -------
// PARSE_WARPS (8 or 16) is a template parameter of the kernel: the cheap searches of the low levels finish long before an
// 8-warp parser does, the deep ones want the warps for searching.
// Roles by warp: LevelParams::role, filled in by lz77_role_map() below.
enum { ROLE_INS0 = 0, ROLE_INS1 = 1, ROLE_INS2 = 2, ROLE_LOAD = 3, ROLE_SEARCH = 4, ROLE_PARSE0 = 8 };

template <int PG> __device__ __forceinline__ void group_sync_n() { asm volatile("bar.sync 1, %0;" :: "n"(PG) : "memory"); }
__device__ __forceinline__ void flush_block(Smem& S, BlockRec* brec, uint32_t in_end, uint32_t tid, uint32_t nthr)
{
    // histograms of the block that just ended -> global; reset for the next one
    for (uint32_t i = tid; i < 288u; i += nthr) { brec->lfreq[i] = S.lfreq[i]; S.lfreq[i] = 0; }
    if (tid < 32u) { brec->dfreq[tid] = S.dfreq[tid]; S.dfreq[tid] = 0; }
    if (tid == 0) {
        brec->tok_begin = S.blk_tok0; brec->tok_count = S.ntok - S.blk_tok0;
        brec->in_off = S.blk_in0; brec->in_bytes = in_end - S.blk_in0;
    }
}

// Token starts of tile k = the positions reachable from the entry point along next(), a pure function of the
// per-position matches.  Evaluated in three steps by the parser group:
//   1. per 32-position window (one warp, lane = position): where does a parse that enters the window at this
//      position leave it?  Pointer doubling with shuffles, 5 rounds.
//   2. one thread hops from window to window along those exit functions: the entry position of every window.
//   3. per entered window: the positions visited from its entry, again by doubling in registers -> mark bitmap.
// Then every marked position emits its token; index = prefix popcount of the marks.
// D / n_s: stream positions where the tile's chunk starts / ends.  The last tile of a chunk also closes the chunk's
// last block and its record, and resets the per-chunk counters for the next chunk of the run.
template <int PARSE_WARPS>
__device__ void parse_emit_tile(Smem& S, const LevelParams& lp, int k, uint32_t D, uint32_t n_s, uint32_t nblocks,
                                uint32_t* __restrict__ tok, BlockRec* __restrict__ brecs, ChunkRec* __restrict__ crec, uint32_t gtid)
{
    constexpr int PG = PARSE_WARPS * 32;              // threads of the parser group
    constexpr int PPT = (TILE + PG - 1) / PG;         // positions per parser thread
    auto group_sync = [] { group_sync_n<PG>(); };
    const uint32_t lane = gtid & 31u, warp = gtid >> 5;
    const bool last = ((uint32_t)(k + 1) * TILE >= n_s);
    const int base = k * TILE - 1;                        // decidable positions: [base, hi) (lazy needs mlen[p+1])
    const int hi = last ? (int)n_s : base + TILE;
    const int hi_w = min(hi, base + TILE);
    // block boundary: tokens emitted from this tile on belong to the next block
    const uint32_t q0 = (uint32_t)k * TILE - D;           // chunk-relative start of the tile
    if (q0 != 0u && (q0 % BLOCK_SPAN) == 0u && S.blk + 1u < nblocks) {
        flush_block(S, &brecs[S.blk], S.entry - D, gtid, PG);
        group_sync();
        if (gtid == 0) { S.blk++; S.blk_tok0 = S.ntok; S.blk_in0 = S.entry - D; }
        group_sync();
    }
    const long long pt0 = clock64();
    const int entry = (int)S.entry;                       // >= base by construction
    const int lim = hi_w - base;                          // positions [0, lim) of the tile can start a token
    constexpr int NWIN = TILE / 32, WPW = NWIN / PARSE_WARPS;
    uint16_t* wexit = S.jump[0];                          // [o]: where the parse entering at o leaves o's window
    uint16_t* went = S.jump[1];                           // [w]: entry position of window w, TILE = not entered
    // per position o = gtid + i * PG: its match token (0 = literal) | next(o) relative to its window (32 = leaves it) << 24
    uint32_t pk[WPW];
    // 1.
#pragma unroll
    for (int i = 0; i < WPW; i++) {
        const int w = (int)warp + i * PARSE_WARPS, o = w * 32 + (int)lane;
        int nx = TILE + 512;
        uint32_t tokv = 0;
        if (o < lim) {
            bool lit; uint32_t dd;
            const uint32_t step = parse_step(S, lp, (uint32_t)(base + o), n_s, lit, dd);
            nx = o + (int)step;
            if (!lit) tokv = tok_match(step, dd);
        }
        uint32_t j = (nx < lim && nx - w * 32 < 32) ? (uint32_t)(nx - w * 32) : 32u;
        pk[i] = tokv | (j << 24);
        int x = nx;
#pragma unroll
        for (int r = 0; r < 5; r++) {
            const int xs = __shfl_sync(FULL, x, (int)(j & 31u));
            const uint32_t js = __shfl_sync(FULL, j, (int)(j & 31u));
            if (j < 32u) { x = xs; j = js; }
        }
        wexit[o] = (uint16_t)x;
        if (lane == 0) went[w] = (uint16_t)TILE;
    }
    const long long pt1 = clock64();
    group_sync();
    // 2.
    if (gtid == 0) {
        int e = entry - base;
        while (e < lim) { went[e >> 5] = (uint16_t)e; e = wexit[e]; }
        e += base;
        uint32_t extra = 0;
        if (last && e == base + TILE && e < (int)n_s) { extra = 1; e += 1; }   // final byte of a full last tile
        S.entry = (uint32_t)e; S.extra_lit = extra;
    }
    const long long pt2 = clock64();
    group_sync();
    // 3.
#pragma unroll
    for (int i = 0; i < WPW; i++) {
        const int w = (int)warp + i * PARSE_WARPS;
        const uint32_t e = went[w];
        uint32_t m = 0;
        if (e < (uint32_t)TILE) {                           // uniform per warp
            m = 1u << (e & 31u);
            uint32_t j = (pk[i] >> 24) & 63u;
#pragma unroll
            for (int r = 0; r < 5; r++) {
                m |= __reduce_or_sync(FULL, (((m >> lane) & 1u) && j < 32u) ? 1u << j : 0u);
                const uint32_t js = __shfl_sync(FULL, j, (int)(j & 31u));
                if (j < 32u) j = js;
            }
        }
        if (lane == 0) S.masks[w] = m;
    }
    const long long pt3 = clock64();
    group_sync();
    if (warp == 0) {
        const uint32_t c0 = (uint32_t)__popc(S.masks[2 * lane]), c1 = (uint32_t)__popc(S.masks[2 * lane + 1]);
        uint32_t incl = c0 + c1;
        for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(FULL, incl, o); if (lane >= (uint32_t)o) incl += v; }
        S.seg_prefix[2 * lane] = incl - c0 - c1;
        S.seg_prefix[2 * lane + 1] = incl - c1;
        if (lane == 31) S.seg_prefix[TILE / 32] = incl;
    }
    group_sync();
    {
        const uint32_t ntok0 = S.ntok;
#pragma unroll
        for (int r = 0; r < PPT; r++) {
            const uint32_t o = gtid + (uint32_t)r * PG;
            const uint32_t seg = o >> 5, bit = o & 31u;
            const uint32_t mk = S.masks[seg];
            if ((mk >> bit) & 1u) {
                const uint32_t idx = ntok0 + S.seg_prefix[seg] + (uint32_t)__popc(mk & ((1u << bit) - 1u));
                uint32_t tv = pk[r] & 0x80FFFFFFu;
                if (tv == 0u) { tv = ring_ld8(S.ring32, (uint32_t)(base + (int)o)); atomicAdd(&S.lfreq[tv], 1u); }
                else {
                    atomicAdd(&S.lfreq[257u + len_code(((tv >> 16) & 0xFFu) + 3u)], 1u);
                    atomicAdd(&S.dfreq[dist_code((tv & 0xFFFFu) + 1u)], 1u);
                }
                tok[idx] = tv;
            }
        }
        if (gtid == 0 && S.extra_lit) {
            const uint32_t tv = ring_ld8(S.ring32, (uint32_t)(base + TILE));
            atomicAdd(&S.lfreq[tv], 1u);
            tok[ntok0 + S.seg_prefix[TILE / 32]] = tv;
        }
    }
    const long long pt4 = clock64();
    group_sync();
    if (gtid == 0) {
        atomicAdd(&g_lzprof[12], (unsigned long long)(pt1 - pt0)); atomicAdd(&g_lzprof[13], (unsigned long long)(pt2 - pt1));
        atomicAdd(&g_lzprof[14], (unsigned long long)(pt3 - pt2)); atomicAdd(&g_lzprof[15], (unsigned long long)(pt4 - pt3));
        S.ntok += S.seg_prefix[TILE / 32] + S.extra_lit; S.extra_lit = 0;
        // the parse of the next tile starts at this tile's last position, whose slot in mlen/mdist the search that runs
        // beside it rewrites: stash it now
        const uint32_t cp = (uint32_t)(k + 1) * TILE - 1u;
        S.carry_pos = cp; S.carry_len = S.mlen[cp & (MRING - 1)]; S.carry_dist = S.mdist[cp & (MRING - 1)];
    }
    if (last) {
        group_sync();
        flush_block(S, &brecs[S.blk], n_s - D, gtid, PG);
        group_sync();
        if (gtid == 0) {
            crec->n_bytes = n_s - D; crec->n_blocks = nblocks; crec->n_tokens = S.ntok;
            S.ntok = 0; S.blk = 0; S.blk_tok0 = 0; S.blk_in0 = 0;
        }
    }
}

__device__ __forceinline__ void bar_sync_n(uint32_t id, uint32_t n) { asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void bar_arrive_n(uint32_t id, uint32_t n) { asm volatile("bar.arrive %0, %1;" :: "r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void wait_ge(const uint32_t* flag, uint32_t v, uint32_t lane, uint32_t ns)
{
    if (lane == 0) while (*(volatile const uint32_t*)flag < v) __nanosleep(ns);
    __syncwarp();
    __threadfence_block();
}
// the groups of tile k, handed out dynamically
__device__ __forceinline__ void search_tile(Smem& S, uint32_t sbase, const LevelParams& lp, uint32_t k, uint32_t ce, uint32_t* ctr, uint32_t lane)
{
    for (;;) {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(ctr, 1u);
        g = __shfl_sync(FULL, g, 0);
        if (g >= (uint32_t)GROUPS_PER_TILE) break;
        if (k * TILE + g * 32u >= ce) continue;
        search_group(S, sbase, lp, k, g, ce, lane);
    }
}

// One CTA takes a RUN of `run` consecutive chunks and treats them as one stream: the hash tables and chains it has
// at the end of a chunk are the next chunk's dictionary, so only the first chunk of a run pays for inserting its
// 32 KiB dictionary (deflateSetDictionary) -- the parse, the tokens and the blocks stay per chunk.
template <int PARSE_WARPS, bool FLOW>
__global__ void __launch_bounds__(LZ_THREADS, 1)
k_lz77(const uint8_t* __restrict__ in, uint64_t n_bytes, uint32_t halo, uint32_t chunk, uint32_t dict, uint32_t nchunks, uint32_t run,
       uint32_t* __restrict__ tokens, BlockRec* __restrict__ blocks, ChunkRec* __restrict__ chunks, LevelParams lp,
       uint32_t* __restrict__ work_ctr)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem& S = *reinterpret_cast<Smem*>(smem_raw);
    uint32_t sbase;                                  // shared-space address of S, pinned in a register (volatile: not rematerialised)
    asm volatile("{ .reg .u64 t; cvta.to.shared.u64 t, %1; cvt.u32.u64 %0, t; }" : "=r"(sbase) : "l"(smem_raw));
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t bpc = (chunk + BLOCK_SPAN - 1) / BLOCK_SPAN;
    const uint32_t nruns = (nchunks + run - 1) / run;
    const unsigned long long role_w = warp < 8u ? lp.role_b0 : warp < 16u ? lp.role_b1 : warp < 24u ? lp.role_b2 : lp.role_b3;
    const int role = (int)((role_w >> ((warp & 7u) * 8u)) & 0xFFu);              // lz77_role_map

    for (;;) {
        __syncthreads();
        if (tid == 0) S.chunk_idx = atomicAdd(work_ctr, 1u);
        __syncthreads();
        if (S.chunk_idx >= nruns) break;
        const uint32_t c0 = S.chunk_idx * run;                                       // first chunk of the run
        const uint32_t rc = min(run, nchunks - c0);                                  // chunks in the run

        const uint64_t lo = (uint64_t)c0 * chunk;
        const uint32_t nr = (uint32_t)min((uint64_t)rc * chunk, n_bytes - lo);        // bytes of the run
        const uint32_t D = dict ? (uint32_t)min((uint64_t)dict, lo + halo) : 0u;
        const uint8_t* sp = in + lo - D;                 // stream = dictionary || chunks of the run
        const uint32_t n_s = D + nr;
        const uint32_t ntiles = (n_s + TILE - 1) / TILE;
        const uint32_t first_ctile = D / TILE;           // D is a multiple of TILE (0 or 32768)

        unsigned long long pf[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        // ---- per-run reset ----
        for (uint32_t i = tid; i < (1u << HBITS) / 2; i += LZ_THREADS) reinterpret_cast<uint32_t*>(S.head)[i] = TABLE_FILL;
        for (uint32_t i = tid; i < (1u << H3BITS) / 2; i += LZ_THREADS) reinterpret_cast<uint32_t*>(S.head3)[i] = TABLE_FILL;
        for (uint32_t i = tid; i < (1u << H4BITS) / 2; i += LZ_THREADS) reinterpret_cast<uint32_t*>(S.head4)[i] = TABLE_FILL;
        for (uint32_t i = tid; i < 288u; i += LZ_THREADS) S.lfreq[i] = 0;
        if (tid < 32u) S.dfreq[tid] = 0;
        if (tid == 0) { S.entry = D; S.ntok = 0; S.blk = 0; S.blk_tok0 = 0; S.blk_in0 = 0; S.extra_lit = 0; S.grp_ctr[1] = 0; S.gctr[0] = 0; S.gctr[1] = 0; S.gctr[2] = 0; S.ins_total = 0; S.parsed = 0; S.b1_done = 0; S.carry_pos = NOPOS; S.carry_len = 0; S.carry_dist = 0xFFFFu; }
        for (uint32_t t = 0; t < 3; t++) load_tile(S, sp, n_s, t, tid, LZ_THREADS);
        __syncthreads();
        hash_tile(S, 0u, tid);
        __syncthreads();

        // ---- pipeline: iteration k inserts tile k+1, searches tile k and parses + emits tile k-1, all at once ----
        if constexpr (FLOW) {
            // Flowing variant.  The role warps (3 inserters, loader, parsers) run the iterations in step, as below; the pure
            // search warps never wait at a barrier: they only ARRIVE at the search barrier of iteration k (named barrier
            // 2 + (k & 1)) when they are through with tile k and go straight on to tile k+1, whose links are final as soon
            // as its inserters are done (DIRECT).  What they must wait for is told by three counters: ins_total (tile
            // inserted), parsed (the tile whose mlen/mdist slots they are about to rewrite has been parsed) and b1_done (the
            // previous use of the barrier they are about to arrive at is over).  The barrier section -- parking the buckets
            // of tile k+2, ageing the tables -- is the role warps' business (named barrier 4).
            const uint32_t NRT = (uint32_t)__popc(lp.role_mask) * 32u;                          // threads of the role warps
            const uint32_t rtid = (uint32_t)__popc(lp.role_mask & ((1u << warp) - 1u)) * 32u + lane;
            for (int k = -1; k <= (int)ntiles; k++) {
                const bool do_insert = k + 1 < (int)ntiles;
                const bool do_hash = k + 2 < (int)ntiles;
                const bool do_parse = k - 1 >= (int)first_ctile;
                const bool do_search = k >= (int)first_ctile && k < (int)ntiles;
                const uint32_t b1 = 2u + ((uint32_t)k & 1u);
                uint32_t ce = 0;
                if (do_search) { const uint32_t ci = ((uint32_t)k * TILE - D) / chunk; ce = min(D + (ci + 1u) * chunk, n_s); }
                if (role == ROLE_SEARCH) {
                    const long long w0 = clock64();
                    wait_ge(&S.b1_done, (uint32_t)max(k, 0), lane, (uint32_t)lp.spin_ns);
                    const long long w1 = clock64();
                    pf[1] += (unsigned long long)(w1 - w0);
                    if (do_search) {
                        wait_ge(&S.ins_total, 3u * (uint32_t)(k + 1), lane, (uint32_t)lp.spin_ns);
                        const long long w2 = clock64();
                        if (k - 2 >= (int)first_ctile) wait_ge(&S.parsed, (uint32_t)(k - 1 - (int)first_ctile), lane, (uint32_t)lp.spin_ns);
                        const long long ph0 = clock64();
                        pf[2] += (unsigned long long)(w2 - w1); pf[3] += (unsigned long long)(ph0 - w2);
                        search_tile(S, sbase, lp, (uint32_t)k, ce, &S.gctr[(uint32_t)k % 3u], lane);
                        pf[0] += (unsigned long long)(clock64() - ph0);
                    }
                    __threadfence_block();
                    bar_arrive_n(b1, LZ_THREADS);
                } else {
                    if (role <= ROLE_INS2) {
                        const bool loads = role == ROLE_INS0 && lp.ins0_loads;
                        if (loads) load_tile_async(sbase, sp, n_s, (uint32_t)(k + 3), lane);
                        if (do_insert) {
                            long long c0 = clock64();
                            const uint32_t t = (uint32_t)(k + 1);
                            if (role == ROLE_INS0) insert_tile<0, true>(S, sbase, t, n_s, lane);
                            else if (role == ROLE_INS1) insert_tile<1, true>(S, sbase, t, n_s, lane);
                            else insert_tile<2, true>(S, sbase, t, n_s, lane);
                            pf[6] += (unsigned long long)(clock64() - c0);
                            __syncwarp();
                            if (lane == 0) { __threadfence_block(); atomicAdd(&S.ins_total, 1u); }
                        }
                        if (loads) load_tile_wait();
                    } else if (role == ROLE_LOAD) {
                        load_tile(S, sp, n_s, (uint32_t)(k + 3), lane, 32);
                    } else if (do_parse) {
                        long long cl0 = clock64();
                        const uint32_t ci = ((uint32_t)(k - 1) * TILE - D) / chunk;
                        const uint32_t cs = D + ci * chunk, cend = min(cs + chunk, n_s);
                        const uint32_t ncb = cend - cs;
                        parse_emit_tile<PARSE_WARPS>(S, lp, k - 1, cs, cend, (ncb + BLOCK_SPAN - 1) / BLOCK_SPAN, tokens + (size_t)(c0 + ci) * chunk,
                                        blocks + (size_t)(c0 + ci) * bpc, &chunks[c0 + ci], (uint32_t)(role - ROLE_PARSE0) * 32u + lane);
                        pf[6] += (unsigned long long)(clock64() - cl0);
                        if (role == ROLE_PARSE0 && lane == 0) { __threadfence_block(); *(volatile uint32_t*)&S.parsed = (uint32_t)(k - (int)first_ctile); }
                    }
                    // lend a hand with tile k -- at the deep levels, where the search is several times longer than the roles; at
                    // the others a group taken now would only hold up the role warps' next iteration
                    if (do_search && lp.flow > 1) search_tile(S, sbase, lp, (uint32_t)k, ce, &S.gctr[(uint32_t)k % 3u], lane);
                    __threadfence_block();
                    bar_sync_n(b1, LZ_THREADS);
                    if (do_hash) { hash_tile_n(S, (uint32_t)(k + 2), rtid, NRT); sweep_tables_n(S, (uint32_t)(k + 2), rtid, NRT); }
                    if (rtid == 0) { S.gctr[(uint32_t)(k + 2) % 3u] = 0; __threadfence_block(); *(volatile uint32_t*)&S.b1_done = (uint32_t)(k + 2); }
                    bar_sync_n(4u, NRT);
                }
                pf[7]++;
            }
        } else
        for (int k = -1; k <= (int)ntiles; k++) {
            const bool do_insert = k + 1 < (int)ntiles;
            const bool do_hash = k + 2 < (int)ntiles;
            const bool do_parse = k - 1 >= (int)first_ctile;                      // k - 1 < ntiles by the loop bound
            const bool do_search = k >= (int)first_ctile && k < (int)ntiles;
            const uint32_t par = (uint32_t)k & 1u;
            if (tid == 0) S.grp_ctr[par ^ 1u] = 0;                                // next iteration's counter
            // roles: three serial inserter warps (raw heads), one loader, eight parser warps; everybody else -- and
            // everybody who is done -- searches
            if (role <= ROLE_INS2) {
                const bool loads = role == ROLE_INS0 && lp.ins0_loads;
                if (loads) load_tile_async(sbase, sp, n_s, (uint32_t)(k + 3), lane);
                if (do_insert) {
                    long long c0 = clock64();
                    const uint32_t t = (uint32_t)(k + 1);
                    if (role == ROLE_INS0) insert_tile<0, false>(S, sbase, t, n_s, lane);
                    else if (role == ROLE_INS1) insert_tile<1, false>(S, sbase, t, n_s, lane);
                    else insert_tile<2, false>(S, sbase, t, n_s, lane);
                    pf[6] += (unsigned long long)(clock64() - c0);
                }
                if (loads) load_tile_wait();
            } else if (role == ROLE_LOAD) {
                load_tile(S, sp, n_s, (uint32_t)(k + 3), lane, 32);   // zero-fills past the end of the stream
            } else if (do_parse && role >= ROLE_PARSE0) {
                long long cl0 = clock64();
                const uint32_t ci = ((uint32_t)(k - 1) * TILE - D) / chunk;              // chunk of the run that tile k-1 belongs to
                const uint32_t cs = D + ci * chunk, ce = min(cs + chunk, n_s);
                const uint32_t ncb = ce - cs;
                parse_emit_tile<PARSE_WARPS>(S, lp, k - 1, cs, ce, (ncb + BLOCK_SPAN - 1) / BLOCK_SPAN, tokens + (size_t)(c0 + ci) * chunk,
                                blocks + (size_t)(c0 + ci) * bpc, &chunks[c0 + ci], (uint32_t)(role - ROLE_PARSE0) * 32u + lane);
                pf[6] += (unsigned long long)(clock64() - cl0);
            }
            const long long ph0 = clock64();
            if (do_search) {
                const uint32_t ci = ((uint32_t)k * TILE - D) / chunk;
                const uint32_t ce = min(D + (ci + 1u) * chunk, n_s);                       // end of tile k's chunk
                for (;;) {
                    uint32_t g = 0;
                    if (lane == 0) g = atomicAdd(&S.grp_ctr[par], 1u);
                    g = __shfl_sync(FULL, g, 0);
                    if (g >= (uint32_t)GROUPS_PER_TILE) break;
                    if ((uint32_t)k * TILE + g * 32u >= ce) continue;
                    search_group(S, sbase, lp, (uint32_t)k, g, ce, lane);
                }
            }
            const long long ph1 = clock64(), ph2 = ph1;
            const long long ph3 = clock64();
            __syncthreads();
            // between the barriers the tables are at rest: raw heads of tile k+1 -> links, buckets of tile k+2 parked (in
            // slots that the search of tile k was still reading), one slice of the tables aged
            if (do_insert) finish_tile(S, (uint32_t)(k + 1), n_s, tid);
            if (do_hash) { hash_tile(S, (uint32_t)(k + 2), tid); sweep_tables(S, (uint32_t)(k + 2), tid); }
            __syncthreads();
            const long long ph4 = clock64();
            pf[0] += (unsigned long long)(ph1 - ph0); pf[1] += (unsigned long long)(ph2 - ph1);
            pf[2] += (unsigned long long)(ph3 - ph2); pf[3] += (unsigned long long)(ph4 - ph3);
            pf[7]++;
        }

        // ---- run epilogue ----
        if (tid == 0) {
            atomicAdd(&g_lzprof[7], pf[7]);
            for (int i = 0; i < 4; i++) atomicAdd(&g_lzprof[i], pf[i]);
        }
        if (lane == 0 && role == ROLE_INS0) atomicAdd(&g_lzprof[8], pf[6]);
        if (lane == 0 && role == ROLE_INS1) atomicAdd(&g_lzprof[9], pf[6]);
        if (lane == 0 && role == ROLE_INS2) atomicAdd(&g_lzprof[10], pf[6]);
        if (lane == 0 && role == ROLE_PARSE0) atomicAdd(&g_lzprof[11], pf[6]);
    }
}

}  // namespace

cudaError_t lz77_read_prof(unsigned long long* out16, int reset)
{
    cudaError_t e = cudaMemcpyFromSymbol(out16, g_lzprof, sizeof(unsigned long long) * 16);
    if (e == cudaSuccess && reset) { unsigned long long z[16] = {0}; e = cudaMemcpyToSymbol(g_lzprof, z, sizeof(z)); }
    return e;
}

cudaError_t lz77_configure()
{
    cudaError_t e = cudaFuncSetAttribute(k_lz77<8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_lz77<8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_lz77<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
}

// Roles by warp.  A warp's scheduler (SMSP) is warp & 3.  preset = layout + 10 * loader:
//   layout 0: the three serial inserters sit together on SMSP 3 with four of the parser warps (which mostly wait at
//             their group barrier), so that the latency-bound inserters get issue slots that eight search warps per
//             scheduler would otherwise take; the other parsers on SMSP 2 (8) or everywhere (16)
//   layout 1: the 8 parsers two per scheduler: 6 / 6 / 6 / 3 search warps instead of 8 / 8 / 4 / 1
//   layout 2, 3: the inserters' scheduler holds 5 (4) parsers and no (one) search warp: 7 / 7 / 7 / 0 (7 / 7 / 6 / 1)
//   loader 0: a loader warp (warp 15) brings tile k+3 into the ring
//   loader 1: no loader warp (one more search warp, on the scheduler with the most free issue slots): the 6-byte
//             inserter issues the copy as cp.async before it inserts and waits for it afterwards
void lz77_role_map(LevelParams& lp, int preset)
{
    const int pw = lp.parse_warps, layout = preset % 10;
    int np = 0;
    lp.ins0_loads = (preset / 10 || layout == 2) ? 1 : 0;           // layout 2 has no room for a loader warp
    unsigned long long b[4] = {0, 0, 0, 0};
    lp.role_mask = 0;
    for (int w = 0; w < 32; w++) {
        const int sm = w & 3, q = w >> 2;
        int r = ROLE_SEARCH;
        if (sm == 3 && q < 3) r = q;
        else if (sm == 3 && q == 3 && (layout != 2 || !lp.ins0_loads)) r = lp.ins0_loads ? ROLE_SEARCH : ROLE_LOAD;
        else if (pw == 16) { if (q >= 4) r = ROLE_PARSE0; }
        else if (layout == 1) { if (q >= 6) r = ROLE_PARSE0; }
        else if (layout == 2) { if (sm == 3 ? q >= 3 : q >= 7) r = ROLE_PARSE0; }                 // 7 / 7 / 7 / 0
        else if (layout == 3) { if (sm == 3 ? q >= 4 : sm == 2 ? q >= 6 : q >= 7) r = ROLE_PARSE0; } // 7 / 7 / 6 / 1
        else if (sm >= 2 && q >= 4) r = ROLE_PARSE0;
        if (r == ROLE_PARSE0) r += np++;
        b[w >> 3] |= (unsigned long long)r << ((w & 7) * 8);
        if (r != ROLE_SEARCH) lp.role_mask |= 1u << w;
    }
    lp.role_b0 = b[0]; lp.role_b1 = b[1]; lp.role_b2 = b[2]; lp.role_b3 = b[3];
}

void launch_lz77(const uint8_t* in, uint64_t n_bytes, uint32_t halo, uint32_t chunk, uint32_t dict, uint32_t nchunks,
                 uint32_t* tokens, BlockRec* blocks, ChunkRec* chunks, LevelParams lp, uint32_t* work_ctr, int n_sm, cudaStream_t st)
{
    if (nchunks == 0) return;
    // chunks per run: up to 32, whichever keeps the SMs busiest; independent chunks (--rsyncable: no dictionary) are
    // never chained
    // (cost of a choice: waves of runs over the SMs x tiles per run; the 16 dictionary tiles of a run are inserted
    // but neither searched nor parsed, which makes them about as dear as 10 full ones)
    uint32_t run = 1;
    const int max_run = getenv("GZB_MAX_RUN") ? atoi(getenv("GZB_MAX_RUN")) : 32;
    if (dict) {
        uint64_t best_cost = ~0ull;
        for (uint32_t r = 1; r <= (uint32_t)max_run; r++) {
            const uint64_t waves = ((nchunks + r - 1) / r + (uint32_t)n_sm - 1) / (uint32_t)n_sm;
            const uint64_t cost = waves * (r * 64u + 10u);
            if (cost < best_cost) { best_cost = cost; run = r; }
        }
    }
    if (dict && getenv("GZB_FORCE_RUN") && atoi(getenv("GZB_FORCE_RUN")) > 0) run = (uint32_t)atoi(getenv("GZB_FORCE_RUN"));   // tests
    const uint32_t nruns = (nchunks + run - 1) / run;
    const uint32_t grid = nruns < (uint32_t)n_sm ? nruns : (uint32_t)n_sm;
    if (lp.parse_warps == 16) k_lz77<16, false><<<grid, LZ_THREADS, sizeof(Smem), st>>>(in, n_bytes, halo, chunk, dict, nchunks, run, tokens, blocks, chunks, lp, work_ctr);
    else if (lp.flow) k_lz77<8, true><<<grid, LZ_THREADS, sizeof(Smem), st>>>(in, n_bytes, halo, chunk, dict, nchunks, run, tokens, blocks, chunks, lp, work_ctr);
    else k_lz77<8, false><<<grid, LZ_THREADS, sizeof(Smem), st>>>(in, n_bytes, halo, chunk, dict, nchunks, run, tokens, blocks, chunks, lp, work_ctr);
}

}  // namespace gzb
static_assert(sizeof(gzb::Smem) <= 232448, "k_lz77: shared memory of one CTA (227 KB)");
