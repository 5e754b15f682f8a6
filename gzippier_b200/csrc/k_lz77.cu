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
// One persistent CTA (1024 threads, ~223 KB shared memory) per SM takes runs of up to 32
// consecutive chunks from an atomic counter and streams through them: the tables at the end of
// a chunk ARE the next chunk's dictionary, so only the first chunk of a run inserts its 32 KiB
// dictionary.  The stream is cut into 2 KiB tiles, and every tile goes through four stations
// that are joined by shared-memory MBARRIERS (one object per station and tile slot, NSTAGE slots):
//   load    one lane of the 6-byte inserter issues a TMA bulk copy (cp.async.bulk) of tile t+3 into the
//           byte ring; the copy's complete_tx flips bar_load[t+3]
//   insert  3 warps, one per hash table: wait for the bytes of tile t (+ the first bytes of t+1) and for
//           the search of tile t-2 (whose link / twin slots are reused), age one eighth of their
//           table, insert the 2048 positions in stream order, arrive at bar_ins[t]
//   search  every other warp (and the role warps when they are through, if the level says so): wait
//           for bar_ins[t] and for the parse of tile t-2 (whose match slots are reused), take
//           32-position groups from a shared counter, arrive at bar_srch[t] once per group
//   parse   8 or 16 warps: wait for bar_srch[t] (64 arrivals), turn the per-position matches into
//           tokens, histograms, block and chunk records, arrive at bar_parse[t]
// A waiting warp sits in mbarrier.try_wait (suspended by the hardware, no issue slots); nobody
// polls and there is no CTA-wide barrier inside a run.
// Shared memory holds the whole search state:
//   ring   64 KiB  window bytes, stream position & 0xFFFF
//   link   72 KiB  u16 delta to the previous position in the same 6-byte bucket (0xFFFF = none),
//                  indexed by stream position mod LINK_RING; deltas need no sliding
//   head   32 KiB  u16 last position (mod 64 Ki) per 6-byte bucket -- inserter-private
//   head3/4 8+16 KiB u32 last position + 1 per 3-byte / 4-byte bucket (single entry, no chain; kept with atomicMax)
//   c3/c4  8+8 KiB per-position delta to the most recent 3-byte / 4-byte twin, double-buffered per tile
//   mlen/mdist 12 KiB  per-position best match of the last two tiles
//   jump    4 KiB  parse scratch (window exit functions)
// (oracle/engine_model.c is the sequential statement of the same semantics.)
#include "gzb_common.cuh"

namespace gzb {

// cycle counters (debugging aid, tools/lzprof.py; compiled in with -DGZB_LZPROF only), summed over CTAs:
// 0 search (warp 28), 1-3 its waits (counter slot, tile inserted, tile parsed), 7 tiles, 8-10 time in insert_tile of the
// 6-byte / 3-byte / 4-byte table warp, 11 time in parse_emit_tile (thread 0 of the parser group), 12-15 its phases
__device__ unsigned long long g_lzprof[16];
#ifdef GZB_LZPROF
#define LZPROF(x) x
#else
#define LZPROF(x)
#endif

namespace {

constexpr uint32_t FULL = 0xFFFFFFFFu;
constexpr uint32_t NOPOS = 0xFFFFFFFFu;
constexpr int GROUPS_PER_TILE = TILE / 32;
constexpr int RING_MIRROR = 288;         // >= 258 + 3 + 19 (longest compare, see match_len_s) + 4, multiple of 16
constexpr int MRING = 2 * TILE;           // per-position match results kept for two tiles
constexpr int NSTAGE = 4;                 // tile slots of the pipeline's barriers
constexpr int LOAD_AHEAD = 3;             // tile t+3 is copied in while tile t is inserted

struct Smem {
    uint32_t ring32[RING / 4 + RING_MIRROR / 4];   // + mirror of the first bytes: compare loops never wrap
    uint16_t link[LINK_RING];
    uint16_t head[1 << HBITS];
    uint32_t head3[1 << H3BITS];          // stream position + 1 of the bucket's most recent string, 0 = none yet
    uint32_t head4[1 << H4BITS];
    uint16_t c3[2][TILE];
    uint16_t c4[2][TILE];
    uint16_t mdist[MRING];                // dist-1, 0xFFFF = no match
    uint8_t  mlen[MRING];                 // len-3
    uint32_t lfreq[288];
    uint32_t dfreq[32];
    uint32_t masks[TILE / 32];            // token-start bitmap of the tile being parsed
    uint32_t seg_prefix[TILE / 32 + 1];
    uint16_t wexit[TILE];                 // parse: where a parse that enters a 32-position window at this position leaves it
    uint16_t went[TILE / 32];             // parse: entry position of every window (TILE = not entered)
    unsigned long long bar_load[NSTAGE];  // tile's bytes are in the ring           (1 arrival + the copy's bytes)
    unsigned long long bar_ins[NSTAGE];   // tile inserted into the three tables    (3 arrivals)
    unsigned long long bar_srch[NSTAGE];  // all groups of the tile searched         (GROUPS_PER_TILE arrivals)
    unsigned long long bar_parse[NSTAGE]; // tile's matches consumed by the parser   (1 arrival)
    unsigned long long bar_ctr[NSTAGE];   // the slot's group counter is back to 0   (1 arrival)
    uint32_t gctr[NSTAGE];                // group scheduler of the tile in the slot
    uint32_t seg_c0, seg_rc;              // the segment (consecutive chunks that are not stored) the CTA works on; seg_rc == 0: no more work
    uint32_t entry;                       // next token start (stream position)
    uint32_t ntok;                        // tokens emitted so far in this chunk
    uint32_t blk;                         // current block index
    uint32_t blk_tok0;                    // token index where the current block began
    uint32_t blk_in0;                     // chunk-relative input offset where it began
    uint32_t extra_lit;                   // 1 if the final position needs its own literal
    uint32_t carry_pos, carry_len, carry_dist;   // match of the last position of the tile before the one being parsed
                                          // (its slot in mlen/mdist is rewritten by the concurrent search)
};

// ---- mbarrier / TMA bulk copy (PTX ISA 8.x; SASS: SYNCS.*, UBLKCP) ----
__device__ __forceinline__ void mbar_init(uint32_t a, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(a), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t a) { asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" :: "r"(a) : "memory"); }
__device__ __forceinline__ void mbar_arrive_tx(uint32_t a, uint32_t bytes) { asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}" :: "r"(a), "r"(bytes) : "memory"); }
// wait until the phase with this parity is over; the warp is suspended by the hardware between tries
__device__ __forceinline__ void mbar_wait(uint32_t a, uint32_t parity)
{
    // suspend-time hint: longer than a tile takes, so a waiting warp issues one instruction per wait, not a polling loop
    asm volatile("{\n .reg .pred p;\nLAB_WAIT:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n @p bra DONE;\n bra LAB_WAIT;\nDONE:\n}"
                 :: "r"(a), "r"(parity), "r"(1000000u) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst_s, const void* src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(dst_s), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// tile t's slot and the parity of its phase
__device__ __forceinline__ uint32_t slot_of(uint32_t t) { return t % NSTAGE; }
__device__ __forceinline__ uint32_t par_of(uint32_t t) { return (t / NSTAGE) & 1u; }
#define GZB_BAR(sbase, member, t) ((sbase) + (uint32_t)offsetof(Smem, member) + 8u * slot_of(t))

__device__ __forceinline__ uint32_t ring_ld8(const uint32_t* ring32, uint32_t pos)
{
    return reinterpret_cast<const uint8_t*>(ring32)[pos & 0xFFFFu];
}

// The hot loops address shared memory by 32-bit shared-space address: one IADD per access, no generic-pointer
// arithmetic, and the compiler has nothing to re-derive inside the loop.
__device__ __forceinline__ uint32_t lds_u8(uint32_t a)  { uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) { uint32_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(a), "h"((uint16_t)v) : "memory"); }
template <int OFF>
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(a), "n"(OFF)); return v; }

// The 16 bytes at stream position p, kept in registers for the whole search of p: the first compare trip of every
// candidate -- the one that settles most of them -- loads the candidate's side only.
struct Probe { uint32_t q0, q1, q2, q3; };
__device__ __forceinline__ Probe probe_of(uint32_t ring_s, uint32_t p)
{
    const uint32_t ap = p & 0xFFFFu, wp = ring_s + (ap & 0xFFFCu), sp = (ap & 3u) * 8u;
    const uint32_t w0 = lds_u32<0>(wp), w1 = lds_u32<4>(wp), w2 = lds_u32<8>(wp), w3 = lds_u32<12>(wp), w4 = lds_u32<16>(wp);
    Probe r;
    r.q0 = __funnelshift_r(w0, w1, sp); r.q1 = __funnelshift_r(w1, w2, sp); r.q2 = __funnelshift_r(w2, w3, sp); r.q3 = __funnelshift_r(w3, w4, sp);
    return r;
}

// Length of the common prefix of the strings at stream positions p and c, at most maxlen (>= 1); four words per trip
// (reads up to 19 bytes past maxlen: RING_MIRROR covers it).
__device__ __forceinline__ int match_len_s(uint32_t ring_s, uint32_t q0, uint32_t q1, uint32_t q2, uint32_t q3, uint32_t p, uint32_t c, int maxlen)
{
    const uint32_t ac = c & 0xFFFFu;
    uint32_t wc = ring_s + (ac & 0xFFFCu);
    const uint32_t sc = (ac & 3u) * 8u;
    uint32_t c0 = lds_u32<0>(wc), c1, x;
#define GZB_ML_FIRST(K, Q)                                                                    \
    c1 = lds_u32<4 * (K) + 4>(wc);                                                            \
    x = (Q) ^ __funnelshift_r(c0, c1, sc);                                                    \
    if (x) return min(maxlen, 4 * (K) + ((__ffs(x) - 1) >> 3));                               \
    c0 = c1;
    GZB_ML_FIRST(0, q0) GZB_ML_FIRST(1, q1) GZB_ML_FIRST(2, q2) GZB_ML_FIRST(3, q3)
#undef GZB_ML_FIRST
    if (maxlen <= 16) return maxlen;
    const uint32_t ap = p & 0xFFFFu;
    uint32_t wp = ring_s + (ap & 0xFFFCu) + 16u;
    const uint32_t sp = (ap & 3u) * 8u;
    uint32_t p0 = lds_u32<0>(wp);
    wc += 16u;
    for (int done = 16; done < maxlen; done += 16, wp += 16u, wc += 16u) {
        uint32_t p1;
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

// ---- tile copy -------------------------------------------------------------------------------------------
// Stream bytes [t*TILE, (t+1)*TILE) go into the ring.  A full tile is one TMA bulk copy issued by one lane (plus a second
// one for the mirror of the ring's first bytes when the tile sits at the ring's start); the tile that holds the end of
// the stream, and the one after it, are written by the warp with zero fill -- the hash of the last positions reads
// zeros past the end, and a bulk copy must not read past the caller's buffer.
__device__ __forceinline__ uint4 tile_fetch(const uint8_t* sp, uint32_t n_s, uint32_t pos)
{
    uint4 v = make_uint4(0, 0, 0, 0);
    if (pos + 16u <= n_s) v = __ldg(reinterpret_cast<const uint4*>(sp + pos));
    else if (pos < n_s) {
        unsigned long long lo = 0, hi = 0;
#pragma unroll
        for (uint32_t k = 0; k < 8u; k++) {
            if (pos + k < n_s) lo |= (unsigned long long)sp[pos + k] << (k * 8u);
            if (pos + 8u + k < n_s) hi |= (unsigned long long)sp[pos + 8u + k] << (k * 8u);
        }
        v = make_uint4((uint32_t)lo, (uint32_t)(lo >> 32), (uint32_t)hi, (uint32_t)(hi >> 32));
    }
    return v;
}
__device__ __forceinline__ void tile_store(Smem& S, uint32_t pos, uint4 v)
{
    const uint32_t ra = pos & 0xFFFFu;
    *reinterpret_cast<uint4*>(&S.ring32[ra >> 2]) = v;
    if (ra < (uint32_t)RING_MIRROR) *reinterpret_cast<uint4*>(&S.ring32[(RING + ra) >> 2]) = v;
}
__device__ void issue_load(Smem& S, uint32_t sbase, const uint8_t* sp, uint32_t n_s, uint32_t t, uint32_t lane)
{
    const uint32_t base = t * TILE, ra = base & 0xFFFFu, bar = GZB_BAR(sbase, bar_load, t);
    if (base + TILE <= n_s) {
        if (lane == 0) {
            const uint32_t ring_s = sbase + (uint32_t)offsetof(Smem, ring32);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the slot's old bytes were read through the generic proxy
            mbar_arrive_tx(bar, (uint32_t)TILE + (ra == 0u ? (uint32_t)RING_MIRROR : 0u));
            bulk_g2s(ring_s + ra, sp + base, TILE, bar);
            if (ra == 0u) bulk_g2s(ring_s + (uint32_t)RING, sp + base, RING_MIRROR, bar);
        }
    } else {
        for (uint32_t o = lane * 16u; o < TILE; o += 512u) tile_store(S, base + o, tile_fetch(sp, n_s, base + o));
        __syncwarp();
        if (lane == 0) mbar_arrive(bar);
    }
}

// ---- insertion ---------------------------------------------------------------------------------------
// Inserting a tile into a hash table is serial per bucket: one warp per table does 64 steps of 32 positions
// (lane = position), old = head[h]; head[h] = p.  Lanes of a step that share a bucket are detected by
// write-then-read-back and ordered with a second exchange (MATCH.ANY -- ~250 cycles per warp-op and scheduler on
// B200, tools/micro_insert.cu -- only when three or more lanes share a bucket).  The bucket of the NEXT step is hashed
// from the ring inside the step, in the shadow of the exchange's shared-memory round trips.
// Table entries are 16-bit positions; sweep_slice keeps every entry either younger than 32 Ki positions -- then it is
// a position of this bucket, the most recent one -- or at an age in [32 Ki, 64 Ki), which the range test in to_delta
// rejects: a chain is exactly "all earlier positions of the bucket within 32 Ki, most recent first".
// KIND 0: 6-byte buckets -> S.link (chains); KIND 1 / 2: 3-byte / 4-byte buckets -> S.c3 (within TOO_FAR) / S.c4.
template <int KIND>
__device__ __forceinline__ uint32_t bucket_at(uint32_t ring_s, uint32_t p)
{
    const uint32_t a = p & 0xFFFFu, wa = ring_s + (a & 0xFFFCu), sh = (a & 3u) * 8u;
    const uint32_t w0 = lds_u32<0>(wa), w1 = lds_u32<4>(wa);              // may be the mirror; zero past the end of the stream
    const uint32_t v = __funnelshift_r(w0, w1, sh);
    if (KIND == 0) return hash6(v, __funnelshift_r(w1, lds_u32<8>(wa), sh) & 0xFFFFu);
    return KIND == 1 ? hash3(v) : hash4(v);
}

// raw head (16-bit position) of position p16 -> what the search reads
template <int KIND>
__device__ __forceinline__ uint32_t to_delta(uint32_t p16, uint32_t raw)
{
    const uint32_t d = (p16 - raw) & 0xFFFFu;
    if (KIND == 0) return d - 1u < 32767u ? d : 0xFFFFu;
    return d - 1u < (KIND == 1 ? (uint32_t)TOO_FAR : 32767u) ? d : 0u;
}

// Ageing of the 16-bit table entries.  Before tile t is inserted, one eighth of the table is swept: an entry older
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
template <int KIND>
__device__ __forceinline__ void sweep_slice(Smem& S, uint32_t t, uint32_t lane)
{
    static_assert(KIND == 0, "only the 16-bit chain table ages");
    constexpr uint32_t WORDS = (1u << HBITS) / 2u / 8u;   // words per slice
    static_assert(WORDS % 128u == 0, "sweep_slice: four words per lane and trip");
    const uint32_t p16 = (t * TILE) & 0xFFFFu;
    uint4* tab = reinterpret_cast<uint4*>(S.head) + (t & 7u) * (WORDS / 4u);
    for (uint32_t i = lane; i < WORDS / 4u; i += 32u) {
        uint4 v = tab[i];
        v.x = sweep_pair(v.x, p16); v.y = sweep_pair(v.y, p16); v.z = sweep_pair(v.z, p16); v.w = sweep_pair(v.w, p16);
        tab[i] = v;
    }
    __syncwarp();
}

template <int KIND>
__device__ void insert_tile(Smem& S, uint32_t sbase, uint32_t t, uint32_t n_s, uint32_t lane)
{
    sweep_slice<KIND>(S, t, lane);
    const uint32_t base = t * TILE;                                           // < n_s
    const uint32_t nsteps = min((uint32_t)GROUPS_PER_TILE, (n_s - base + 31u) / 32u);   // nothing past the end is ever searched
    const uint32_t ring_s = sbase + (uint32_t)offsetof(Smem, ring32);
    static_assert(KIND == 0, "the exchange step serves the 6-byte chain table; the single-entry tables use insert_tile_max");
    const uint32_t table_s = sbase + (uint32_t)offsetof(Smem, head);
    // where the deltas go; a tile's link slots are contiguous (LINK_RING % TILE == 0)
    static_assert(LINK_RING % TILE == 0, "a tile's link slots must not wrap");
    uint32_t dst = sbase + (KIND == 0 ? (uint32_t)offsetof(Smem, link) + 2u * (base % LINK_RING + lane)
                                      : (uint32_t)(KIND == 1 ? offsetof(Smem, c3) : offsetof(Smem, c4)) + 2u * ((t & 1u) * TILE + lane));
    uint32_t p = base + lane;                                                 // + 32 per step
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t ta = table_s + 2u * bucket_at<KIND>(ring_s, p);
    for (uint32_t i = 0; i < nsteps; i++, dst += 64u, p += 32u) {
        const uint32_t p16 = p & 0xFFFFu;
        uint32_t raw = lds_u16(ta);
        // Write, read back: a lane that does not find its own position shares its bucket with another lane of the step.
        // The losers then write once more: a bucket shared by exactly TWO lanes -- the usual case -- is settled by these
        // two exchanges (each lane has read the other's position); only buckets with three or more lanes of the step
        // need MATCH.ANY.
        __syncwarp();
        sts_u16(ta, p16);
        __syncwarp();
        const uint32_t r1 = lds_u16(ta);
        const uint32_t tn = table_s + 2u * bucket_at<KIND>(ring_s, p + 32u);   // next step's bucket (unused after the last step)
        const bool lost = r1 != p16;
        if (__any_sync(FULL, lost)) {                            // uniform
            __syncwarp();
            if (lost) sts_u16(ta, p16);
            __syncwarp();
            const uint32_t r2 = lds_u16(ta);
            if (!__any_sync(FULL, lost && r2 != p16)) {
                const bool paired = lost || r2 != p16;
                const uint32_t partner = lost ? r1 : r2;
                const bool partner_lower = ((p16 - partner) & 0xFFFFu) < 32u;  // same step: the partner is at most 31 positions away
                if (paired && partner_lower) raw = partner;                  // the other lane of the bucket precedes me
                __syncwarp();
                if (paired && partner_lower && !lost) sts_u16(ta, p16);      // the later position owns the bucket
            } else {
                const uint32_t same = __match_any_sync(FULL, ta);
                const uint32_t lower = same & lt_mask;
                if (lower) raw = (p16 - (lane - (uint32_t)(31 - __clz(lower)))) & 0xFFFFu;   // the nearest lower lane of the bucket
                if ((same >> lane) == 1u && lower) sts_u16(ta, p16);                          // the last position of the step owns the bucket
            }
        }
        sts_u16(dst, to_delta<KIND>(p16, raw));
        ta = tn;
        __syncwarp();
    }
}

// The single-entry tables (KIND 1 / 2: most recent 3-byte / 4-byte twin) need no chain, so they need no exchange either:
// an entry is the stream position + 1 of the bucket's most recent string, and a step is ONE shared-memory atomicMax per
// lane.  Whatever order the hardware serves the lanes in, the table ends the step holding the highest position of every
// bucket, and consecutive steps do not depend on one another through registers, so they pipeline (the exchange of
// insert_tile costs two dependent shared-memory round trips per step, and 57 % / 37 % of the steps of text have lanes that
// share a 3-byte / 4-byte bucket and pay two more).  The value a lane gets back is the table's value before ITS update.
// When the lanes that share a bucket are served in ascending lane order -- what the hardware does, as far as measured --
// that value is exactly what the lane wants: the position of the nearest lower lane of its bucket, or for the lowest one
// the bucket's string from before the step.  Any other order hands some lane a position ABOVE its own (a higher lane of
// its bucket went first), which is what the step checks; only then MATCH.ANY groups the lanes (~400 cycles), a lane's
// twin is the nearest lower lane of its group, and the lowest lane takes the one value of its group that is not a
// position of this step (held by whichever lane was served first).
template <int KIND>
__device__ void insert_tile_max(Smem& S, uint32_t sbase, uint32_t t, uint32_t n_s, uint32_t lane)
{
    static_assert(KIND == 1 || KIND == 2, "single-entry tables only");
    const uint32_t base = t * TILE;
    const uint32_t nsteps = min((uint32_t)GROUPS_PER_TILE, (n_s - base + 31u) / 32u);
    const uint32_t ring_s = sbase + (uint32_t)offsetof(Smem, ring32);
    uint32_t* table = KIND == 1 ? S.head3 : S.head4;
    uint16_t* dst = (KIND == 1 ? S.c3[t & 1u] : S.c4[t & 1u]) + lane;
    constexpr uint32_t REACH = KIND == 1 ? (uint32_t)TOO_FAR : 32767u;
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll 2
    for (uint32_t i = 0; i < nsteps; i++) {
        const uint32_t p1 = base + i * 32u + lane + 1u;                          // this lane's position + 1
        const uint32_t b = bucket_at<KIND>(ring_s, p1 - 1u);
        uint32_t prev = atomicMax(&table[b], p1);
        if (__any_sync(FULL, prev >= p1)) {                                      // served out of order (see above)
            const uint32_t same = __match_any_sync(FULL, b);
            // exactly one lane of every group -- the one served first -- holds the bucket's value from before the step
            const uint32_t holders = __ballot_sync(FULL, prev <= base + i * 32u);
            const uint32_t first = __shfl_sync(FULL, prev, __ffs(same & holders) - 1);
            const uint32_t lower = same & lt_mask;
            prev = lower ? base + i * 32u + (uint32_t)(31 - __clz(lower)) + 1u : first;
        }
        const uint32_t d = p1 - prev;                                            // prev == 0 (none yet): d = p1 > REACH only if p1 > REACH ...
        dst[i * 32u] = (uint16_t)((prev != 0u && d <= REACH) ? d : 0u);          // ... so test prev explicitly
    }
    __syncwarp();
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
    const Probe pr = probe_of(ring_s, p);
    const uint32_t limit = min((uint32_t)MAXD, p);
    const int nice = min(lp.nice, maxlen);
    constexpr uint32_t END = 0xFFFFu;                           // delta of a chain that is over: d + END > limit
    uint32_t slot2 = (p % LINK_RING) * 2u;                      // byte offset of the link slot the walk stands on
    uint32_t delta = maxlen >= 6 ? lds_u16(link_s + slot2) : END;   // the last five positions of a chunk walk no chain
    uint32_t d = 0;
    int depth = lp.depth;
    uint32_t pb = p + (uint32_t)best;                           // filter byte: the one after the best match so far
    uint32_t scan_end = (pr.q0 >> 16) & 0xFFu;                  // = ring[p + 2]
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
                const int l = match_len_s(ring_s, pr.q0, pr.q1, pr.q2, pr.q3, p, p - pend0, maxlen);
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
                int l = match_len_s(ring_s, pr.q0, pr.q1, pr.q2, pr.q3, p, p - d4, maxlen);
                if (l >= 4 && l > best) { best = l; bdist = d4; }
            }
        }
        if (best < 3 && lp.use_h3) {
            uint32_t d3 = S.c3[t & 1][g * 32u + lane];
            if (d3 != 0u && d3 <= min(limit, (uint32_t)TOO_FAR)) {
                int l = match_len_s(ring_s, pr.q0, pr.q1, pr.q2, pr.q3, p, p - d3, maxlen);
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

// ---- parse + emission of one tile, by the parser group (named barrier 1) -------
// PARSE_WARPS (8 or 16) is a template parameter of the kernel: the cheap searches of the low levels finish long before an
// 8-warp parser does, the deep ones want the warps for searching.
// Roles by warp: LevelParams::role_b*, filled in by lz77_role_map() below.
enum { ROLE_INS0 = 0, ROLE_INS1 = 1, ROLE_INS2 = 2, ROLE_SEARCH = 4, ROLE_PARSE0 = 8 };

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
// done_bar: the tile's bar_parse, arrived at as soon as nobody reads the tile's match slots any more.
template <int PARSE_WARPS>
__device__ void parse_emit_tile(Smem& S, const LevelParams& lp, int k, uint32_t D, uint32_t n_s, uint32_t nblocks,
                                uint32_t* __restrict__ tok, BlockRec* __restrict__ brecs, ChunkRec* __restrict__ crec, uint32_t gtid, uint32_t done_bar)
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
    LZPROF(const long long pt0 = clock64();)
    const int entry = (int)S.entry;                       // >= base by construction
    const int lim = hi_w - base;                          // positions [0, lim) of the tile can start a token
    constexpr int NWIN = TILE / 32, WPW = NWIN / PARSE_WARPS;
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
        S.wexit[o] = (uint16_t)x;
        if (lane == 0) S.went[w] = (uint16_t)TILE;
    }
    LZPROF(const long long pt1 = clock64();)
    group_sync();
    // 2.
    if (gtid == 0) {
        // the parse of the next tile starts at this tile's last position, whose slot in mlen/mdist the search two tiles on
        // rewrites: stash it, then let that search go
        const uint32_t cp = (uint32_t)(k + 1) * TILE - 1u;
        const uint32_t cl = S.mlen[cp & (MRING - 1)], cd = S.mdist[cp & (MRING - 1)];
        mbar_arrive(done_bar);
        int e = entry - base;
        while (e < lim) { S.went[e >> 5] = (uint16_t)e; e = S.wexit[e]; }
        e += base;
        uint32_t extra = 0;
        if (last && e == base + TILE && e < (int)n_s) { extra = 1; e += 1; }   // final byte of a full last tile
        S.entry = (uint32_t)e; S.extra_lit = extra;
        S.carry_pos = cp; S.carry_len = cl; S.carry_dist = cd;                 // read again only after the next group_sync
    }
    LZPROF(const long long pt2 = clock64();)
    group_sync();
    // 3.
#pragma unroll
    for (int i = 0; i < WPW; i++) {
        const int w = (int)warp + i * PARSE_WARPS;
        const uint32_t e = S.went[w];
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
    LZPROF(const long long pt3 = clock64();)
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
    LZPROF(const long long pt4 = clock64();)
    group_sync();
    if (gtid == 0) {
        LZPROF(atomicAdd(&g_lzprof[12], (unsigned long long)(pt1 - pt0)); atomicAdd(&g_lzprof[13], (unsigned long long)(pt2 - pt1));
               atomicAdd(&g_lzprof[14], (unsigned long long)(pt3 - pt2)); atomicAdd(&g_lzprof[15], (unsigned long long)(pt4 - pt3));)
        S.ntok += S.seg_prefix[TILE / 32] + S.extra_lit; S.extra_lit = 0;
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
    group_sync();                                       // the next tile's parse reads what thread 0 has just written
}

// The groups of tile t, handed out dynamically to the NP warps that take part in the search of every tile.  Each of
// them claims until the counter runs over (exactly one claim >= GROUPS_PER_TILE per warp and tile); the warp that draws
// the last of those resets the counter and opens the slot for tile t + NSTAGE.
// Dictionary tiles (t < first_ctile) are claimed and signed off without a search, so that every tile passes every station.
__device__ __forceinline__ void search_tile(Smem& S, uint32_t sbase, const LevelParams& lp, uint32_t t, uint32_t first_ctile, uint32_t D,
                                            uint32_t chunk, uint32_t n_s, uint32_t np, uint32_t lane, bool prof)
{
    (void)prof;
    const uint32_t s = slot_of(t);
    LZPROF(const long long w0 = clock64();)
    if (t >= (uint32_t)NSTAGE) mbar_wait(GZB_BAR(sbase, bar_ctr, t), par_of(t - NSTAGE));
    LZPROF(const long long w1 = clock64();)
    mbar_wait(GZB_BAR(sbase, bar_ins, t), par_of(t));
    LZPROF(const long long w2 = clock64();)
    if (t >= 2u) mbar_wait(GZB_BAR(sbase, bar_parse, t - 2u), par_of(t - 2u));
    LZPROF(const long long w3 = clock64();)
    uint32_t ce = 0;
    if (t >= first_ctile) { const uint32_t ci = (t * TILE - D) / chunk; ce = min(D + (ci + 1u) * chunk, n_s); }   // end of the tile's chunk
    const uint32_t done_bar = GZB_BAR(sbase, bar_srch, t);
    for (;;) {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(&S.gctr[s], 1u);
        g = __shfl_sync(FULL, g, 0);
        if (g >= (uint32_t)GROUPS_PER_TILE) {
            if (g == (uint32_t)GROUPS_PER_TILE + np - 1u && lane == 0) { S.gctr[s] = 0; mbar_arrive(GZB_BAR(sbase, bar_ctr, t)); }
            break;
        }
        if (t * TILE + g * 32u < ce) search_group(S, sbase, lp, t, g, ce, lane);
        __syncwarp();
        if (lane == 0) mbar_arrive(done_bar);
    }
    LZPROF(if (prof && lane == 0) { atomicAdd(&g_lzprof[0], (unsigned long long)(clock64() - w3)); atomicAdd(&g_lzprof[1], (unsigned long long)(w1 - w0));
                                     atomicAdd(&g_lzprof[2], (unsigned long long)(w2 - w1)); atomicAdd(&g_lzprof[3], (unsigned long long)(w3 - w2)); atomicAdd(&g_lzprof[7], 1ull); })
}

// One CTA takes a RUN of `run` consecutive chunks and treats them as one stream: the hash tables and chains it has
// at the end of a chunk are the next chunk's dictionary, so only the first chunk of a run pays for inserting its
// 32 KiB dictionary (deflateSetDictionary) -- the parse, the tokens and the blocks stay per chunk.
template <int PARSE_WARPS>
__global__ void __launch_bounds__(LZ_THREADS, 1)
k_lz77(const uint8_t* __restrict__ in, uint64_t n_bytes, uint32_t halo, uint32_t chunk, uint32_t dict, uint32_t nchunks, uint32_t run,
       uint32_t* __restrict__ tokens, BlockRec* __restrict__ blocks, ChunkRec* chunks, LevelParams lp,
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
    // warps that take part in the search of every tile
    const uint32_t np = 32u - (uint32_t)__popc(lp.role_mask) + (lp.help_ins ? 3u : 0u) + (lp.help_parse ? (uint32_t)PARSE_WARPS : 0u);

    bool fresh = true;                                   // (thread 0) the barrier objects have never been initialised
    uint32_t cur_c = 0, cur_end = 0;                     // (thread 0) what is left of the run of chunks last taken from the counter
    for (;;) {
        __syncthreads();                                 // everybody is through with the previous segment
        if (tid == 0) {
            // next SEGMENT: the chunks k_scan marked stored are skipped -- they need neither match finder nor parser -- and
            // what follows one starts a stream of its own (tables reset, its 32 KiB dictionary inserted again)
            uint32_t c0s = 0, rcs = 0;
            for (;;) {
                if (cur_c >= cur_end) {
                    const uint32_t idx = atomicAdd(work_ctr, 1u);
                    if (idx >= nruns) break;
                    cur_c = idx * run; cur_end = min(cur_c + run, nchunks);
                }
                while (cur_c < cur_end && chunks[cur_c].stored) cur_c++;
                if (cur_c < cur_end) {
                    c0s = cur_c;
                    while (cur_c < cur_end && !chunks[cur_c].stored) cur_c++;
                    rcs = cur_c - c0s;
                    break;
                }
            }
            S.seg_c0 = c0s; S.seg_rc = rcs;
            for (uint32_t i = 0; i < 5u * (uint32_t)NSTAGE; i++) {               // the five arrays lie back to back
                const uint32_t a = sbase + (uint32_t)offsetof(Smem, bar_load) + 8u * i;
                static_assert(offsetof(Smem, bar_ctr) == offsetof(Smem, bar_load) + 4 * 8 * NSTAGE, "barrier arrays are contiguous");
                if (!fresh) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" :: "r"(a) : "memory");
                const uint32_t kind = i / (uint32_t)NSTAGE;                       // load, ins, srch, parse, ctr
                mbar_init(a, kind == 1u ? 3u : kind == 2u ? (uint32_t)GROUPS_PER_TILE : 1u);
            }
            fresh = false;
            for (uint32_t i = 0; i < (uint32_t)NSTAGE; i++) S.gctr[i] = 0;
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            // the parse enters the segment's first chunk right behind its dictionary
            const uint64_t lo0 = (uint64_t)c0s * chunk;
            S.entry = (rcs && dict) ? (uint32_t)min((uint64_t)dict, lo0 + halo) : 0u;
            S.ntok = 0; S.blk = 0; S.blk_tok0 = 0; S.blk_in0 = 0; S.extra_lit = 0;
            S.carry_pos = NOPOS; S.carry_len = 0; S.carry_dist = 0xFFFFu;
        }
        // ---- per-run reset ----
        for (uint32_t i = tid; i < (1u << HBITS) / 2; i += LZ_THREADS) reinterpret_cast<uint32_t*>(S.head)[i] = TABLE_FILL;
        for (uint32_t i = tid; i < (1u << H3BITS); i += LZ_THREADS) S.head3[i] = 0;
        for (uint32_t i = tid; i < (1u << H4BITS); i += LZ_THREADS) S.head4[i] = 0;
        for (uint32_t i = tid; i < 288u; i += LZ_THREADS) S.lfreq[i] = 0;
        if (tid < 32u) S.dfreq[tid] = 0;
        __syncthreads();
        if (S.seg_rc == 0u) break;
        const uint32_t c0 = S.seg_c0;                                                // first chunk of the segment
        const uint32_t rc = S.seg_rc;                                                // chunks in it

        const uint64_t lo = (uint64_t)c0 * chunk;
        const uint32_t nr = (uint32_t)min((uint64_t)rc * chunk, n_bytes - lo);        // bytes of the run
        const uint32_t D = dict ? (uint32_t)min((uint64_t)dict, lo + halo) : 0u;
        const uint8_t* sp = in + lo - D;                 // stream = dictionary || chunks of the run
        const uint32_t n_s = D + nr;
        const uint32_t ntiles = (n_s + TILE - 1) / TILE;
        const uint32_t first_ctile = D / TILE;           // D is a multiple of TILE (0 or 32768)

        // ---- the pipeline: every warp walks the tiles of the run in its own role; the stations meet at the mbarriers ----
        if (role <= ROLE_INS2) {
            // tiles 0 .. ntiles are copied in (the one past the end is all zeros): LOAD_AHEAD of them before the first insertion
            if (role == ROLE_INS0) for (uint32_t t = 0; t < (uint32_t)LOAD_AHEAD && t <= ntiles; t++) issue_load(S, sbase, sp, n_s, t, lane);
            for (uint32_t t = 0; t < ntiles; t++) {
                if (role == ROLE_INS0 && t + LOAD_AHEAD <= ntiles) issue_load(S, sbase, sp, n_s, t + LOAD_AHEAD, lane);
                if (t == 0u) mbar_wait(GZB_BAR(sbase, bar_load, 0u), 0u);
                mbar_wait(GZB_BAR(sbase, bar_load, t + 1u), par_of(t + 1u));          // the last positions hash into the next tile
                if (t >= 2u) mbar_wait(GZB_BAR(sbase, bar_srch, t - 2u), par_of(t - 2u));   // its link / twin slots are written now
                LZPROF(const long long i0 = clock64();)
                if (role == ROLE_INS0) insert_tile<0>(S, sbase, t, n_s, lane);
                else if (role == ROLE_INS1) insert_tile_max<1>(S, sbase, t, n_s, lane);
                else insert_tile_max<2>(S, sbase, t, n_s, lane);
                __syncwarp();
                if (lane == 0) mbar_arrive(GZB_BAR(sbase, bar_ins, t));
                LZPROF(if (lane == 0) atomicAdd(&g_lzprof[8 + role], (unsigned long long)(clock64() - i0));)
                if (lp.help_ins && t >= 1u) search_tile(S, sbase, lp, t - 1u, first_ctile, D, chunk, n_s, np, lane, false);
            }
            if (lp.help_ins) search_tile(S, sbase, lp, ntiles - 1u, first_ctile, D, chunk, n_s, np, lane, false);
        } else if (role >= ROLE_PARSE0) {
            const uint32_t gtid = (uint32_t)(role - ROLE_PARSE0) * 32u + lane;
            for (uint32_t t = 0; t < ntiles; t++) {
                if (lp.help_parse) search_tile(S, sbase, lp, t, first_ctile, D, chunk, n_s, np, lane, false);
                mbar_wait(GZB_BAR(sbase, bar_srch, t), par_of(t));
                if (t >= first_ctile) {
                    LZPROF(const long long cl0 = clock64();)
                    const uint32_t ci = (t * TILE - D) / chunk;                      // chunk of the run that tile t belongs to
                    const uint32_t cs = D + ci * chunk, cend = min(cs + chunk, n_s);
                    parse_emit_tile<PARSE_WARPS>(S, lp, (int)t, cs, cend, (cend - cs + BLOCK_SPAN - 1) / BLOCK_SPAN, tokens + (size_t)(c0 + ci) * chunk,
                                                 blocks + (size_t)(c0 + ci) * bpc, &chunks[c0 + ci], gtid, GZB_BAR(sbase, bar_parse, t));
                    LZPROF(if (gtid == 0) atomicAdd(&g_lzprof[11], (unsigned long long)(clock64() - cl0));)
                } else if (gtid == 0) mbar_arrive(GZB_BAR(sbase, bar_parse, t));
            }
        } else {
            for (uint32_t t = 0; t < ntiles; t++) search_tile(S, sbase, lp, t, first_ctile, D, chunk, n_s, np, lane, warp == 28u);
        }
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
    cudaError_t e = cudaFuncSetAttribute(k_lz77<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_lz77<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
}

// Roles by warp.  A warp's scheduler (SMSP) is warp & 3.
//   layout 0: the three serial inserters sit together on SMSP 3 with parser warps (which mostly wait at their group
//             barrier), so that the latency-bound inserters get issue slots that eight search warps per scheduler
//             would otherwise take; the other parsers on SMSP 2 (8) or everywhere (16)
//   layout 1: the 8 parsers two per scheduler: 6 / 6 / 6 / 3 search warps
//   layout 2: the three inserters on three different schedulers, the parsers two per scheduler
//   layout 3, 4: see below (issue priority by warp id)
void lz77_role_map(LevelParams& lp, int layout)
{
    const int pw = lp.parse_warps;
    int np = 0;
    unsigned long long b[4] = {0, 0, 0, 0};
    lp.role_mask = 0;
    for (int w = 0; w < 32; w++) {
        const int sm = w & 3, q = w >> 2;
        int r = ROLE_SEARCH;
        if (layout == 2) {
            if (q == 0 && sm >= 1) r = sm - 1;
            else if (pw == 16 ? q >= 4 : q >= 6) r = ROLE_PARSE0;
        } else if (layout == 3) {
            // a scheduler picks the eligible warp with the HIGHEST id first: the latency-bound inserters get the top ids
            // (29-31, three schedulers), the parsers the ids below them, the throughput-bound search warps the rest
            if (w >= 29) r = w - 29;
            else if (pw == 16 ? (w >= 12 && w < 28) : (w >= 20 && w < 28)) r = ROLE_PARSE0;
        } else if (layout == 4) {
            // as 3, but the parsers below the search warps
            if (w >= 29) r = w - 29;
            else if (pw == 16 ? w < 16 : w < 8) r = ROLE_PARSE0;
        } else {
            if (sm == 3 && q < 3) r = q;
            else if (pw == 16) { if (q >= 4) r = ROLE_PARSE0; }
            else if (layout == 1) { if (q >= 6) r = ROLE_PARSE0; }
            else if (sm >= 2 && q >= 4) r = ROLE_PARSE0;
        }
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
    // chunks per run: up to max_run, whichever keeps the SMs busiest; independent chunks (--rsyncable: no dictionary) are
    // never chained
    // (cost of a choice: waves of runs over the SMs x tiles per run; the 16 dictionary tiles of a run are inserted
    // but neither searched nor parsed, which makes them about as dear as 10 full ones)
    uint32_t run = 1;
    if (dict) {
        uint64_t best_cost = ~0ull;
        for (uint32_t r = 1; r <= (uint32_t)lp.max_run; r++) {
            const uint64_t waves = ((nchunks + r - 1) / r + (uint32_t)n_sm - 1) / (uint32_t)n_sm;
            const uint64_t cost = waves * (r * 64u + 10u);
            if (cost < best_cost) { best_cost = cost; run = r; }
        }
        if (lp.force_run > 0) run = (uint32_t)lp.force_run;      // tests
    }
    const uint32_t nruns = (nchunks + run - 1) / run;
    const uint32_t grid = nruns < (uint32_t)n_sm ? nruns : (uint32_t)n_sm;
    if (lp.parse_warps == 16) k_lz77<16><<<grid, LZ_THREADS, sizeof(Smem), st>>>(in, n_bytes, halo, chunk, dict, nchunks, run, tokens, blocks, chunks, lp, work_ctr);
    else k_lz77<8><<<grid, LZ_THREADS, sizeof(Smem), st>>>(in, n_bytes, halo, chunk, dict, nchunks, run, tokens, blocks, chunks, lp, work_ctr);
}

}  // namespace gzb
static_assert(sizeof(gzb::Smem) <= 232448, "k_lz77: shared memory of one CTA (227 KB)");
