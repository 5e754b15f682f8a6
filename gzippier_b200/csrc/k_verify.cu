// k_verify.cu -- device-side round-trip check of the engine's own output (GZB_F_VERIFY).
//
// the round-trip check of a multi-GiB member on the CPU dominates benchmark wall time; a
// chunk-parallel decoder is possible for OUR output because every chunk's fragment starts on a known byte.  The
// decoding side of the reference (inflateGZIP -> zlib inflate, src/unzip.c:254-370, lib/zlib/inflate.c) is out of
// scope as a product; this kernel is the inflate a verifier needs and no more: it decodes RFC 1951 blocks and
// COMPARES what they say with the input that is still in HBM, without materialising the output:
//   literal        -> must equal in[pos]
//   match (l, d)   -> in[pos .. pos+l) must equal in[pos-d .. pos-d+l), d within the chunk + its dictionary
//                     (everything before pos has already been verified equal to the input, so comparing the
//                      input with itself is comparing the decoder's window with the input)
//   stored block   -> LEN/NLEN consistent, bytes equal
// and checks the framing the reference's decoder relies on: BFINAL only on the last block of the last chunk,
// fragments byte-aligned and exactly frag_bytes long, every chunk decoding to exactly its n_bytes.
//
// One warp per chunk.  The code tables are built by the whole warp, per block, from the code lengths IN THE STREAM (not
// from BlockRec): a canonical prefix code (RFC 1951 3.2.2) is kept as its codes in ascending order, left-justified to
// 15 bits, with the symbol and length beside each; decoding a symbol is then "take the next 15 bits most significant
// first, find the last entry that is not greater" -- a binary search, no bit-serial walk.  Lane 0 owns the bit reader;
// the 32 lanes compare match and stored bytes together.
#include "gzb_common.cuh"

namespace gzb {

namespace {

constexpr int VERIFY_WARPS = 8;
constexpr uint32_t FULLM = 0xFFFFFFFFu;

// A canonical prefix code in decode order: entry i has the i-th smallest code; lj[i] = that code, most significant bit
// first, shifted up to 15 bits; ent[i] = symbol | length << 12.  (Codes sorted by (length, symbol) ascend when left-justified.)
struct Code { uint16_t* lj; uint16_t* ent; int n; };
struct WarpTabs { uint16_t lit_lj[288], lit_ent[288], dst_lj[32], dst_ent[32], aux_lj[32], aux_ent[32]; uint8_t lens[320]; uint8_t cl[32]; };

struct Bits {
    const uint8_t* p; uint64_t nbytes; uint64_t byte; uint64_t buf; int have; int err;
};
__device__ __forceinline__ void refill(Bits& b)
{
    while (b.have <= 56) {
        uint64_t v = 0;
        if (b.byte < b.nbytes) v = b.p[b.byte];
        b.byte++;                                   // counting past the end: detected by consumed() > nbytes
        b.buf |= v << b.have; b.have += 8;
    }
}
__device__ __forceinline__ uint32_t take(Bits& b, int n)   // n <= 16; LSB first (RFC 1951 3.1.1)
{
    if (b.have < n) refill(b);
    const uint32_t v = (uint32_t)(b.buf & ((1ull << n) - 1ull));
    b.buf >>= n; b.have -= n;
    return v;
}
__device__ __forceinline__ uint64_t consumed_bits(const Bits& b) { return b.byte * 8ull - (uint64_t)b.have; }

// Whole warp: the code for lens[0, n) (n <= 288).  Returns the Kraft balance: 0 complete, > 0 incomplete (code space left,
// in units of 2^-15), < 0 over-subscribed (no table is built).  Symbols of one length are ranked with ballots, as k_huffman
// assigns its codes.
__device__ int build_code(Code& c, const uint8_t* lens, int n, uint32_t lane)
{
    uint32_t len[9];
#pragma unroll
    for (int k = 0; k < 9; k++) { const int i = k * 32 + (int)lane; len[k] = i < n ? lens[i] : 0u; }
    const uint32_t lt = (1u << lane) - 1u;
    int used = 0, kraft = 0;
    uint32_t code = 0;                              // first code of the current length, as a number of that many bits
    for (uint32_t l = 1; l <= 15u; l++) {
        code <<= 1;
        uint32_t cnt = 0;
#pragma unroll
        for (int k = 0; k < 9; k++) {
            const uint32_t m = __ballot_sync(FULLM, len[k] == l);
            if (len[k] == l && kraft >= 0) {
                const uint32_t r = cnt + (uint32_t)__popc(m & lt);
                c.lj[used + r] = (uint16_t)((code + r) << (15u - l));
                c.ent[used + r] = (uint16_t)((uint32_t)(k * 32 + (int)lane) | (l << 12));
            }
            cnt += (uint32_t)__popc(m);
        }
        kraft += (int)(cnt << (15u - l));
        if (kraft > (1 << 15)) kraft = -1;           // over-subscribed: stop filling (the entries would overflow the table)
        if (kraft < 0) break;
        used += (int)cnt; code += cnt;
    }
    c.n = used;
    __syncwarp();
    return kraft < 0 ? -1 : (1 << 15) - kraft;
}
// lane 0: next symbol, or -1 if the bits are no code of c
__device__ int decode_sym(Bits& b, const Code& c)
{
    if (b.have < 15) refill(b);
    if (c.n == 0) return -1;
    const uint32_t x = __brev((uint32_t)b.buf) >> 17;          // the next 15 bits, first bit most significant
    int lo = 0, hi = c.n - 1;                                  // the last entry with lj <= x (lj[0] == 0)
    while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (c.lj[mid] <= x) lo = mid; else hi = mid - 1; }
    const uint32_t e = c.ent[lo], l = e >> 12;
    if ((x ^ c.lj[lo]) >> (15u - l)) return -1;                // only in an incomplete code: bits past the last code
    b.buf >>= l; b.have -= (int)l;
    return (int)(e & 0xFFFu);
}

__constant__ uint16_t c_lbase[29] = {3,4,5,6,7,8,9,10,11,13,15,17,19,23,27,31,35,43,51,59,67,83,99,115,131,163,195,227,258};
__constant__ uint8_t  c_lext[29]  = {0,0,0,0,0,0,0,0,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,4,5,5,5,5,0};
__constant__ uint16_t c_dbase[30] = {1,2,3,4,5,7,9,13,17,25,33,49,65,97,129,193,257,385,513,769,1025,1537,2049,3073,4097,6145,8193,12289,16385,24577};
__constant__ uint8_t  c_dext[30]  = {0,0,0,0,1,1,2,2,3,3,4,4,5,5,6,6,7,7,8,8,9,9,10,10,11,11,12,12,13,13};
__constant__ uint8_t  c_order[19] = {16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15};

// all lanes: are a[0..n) and b[0..n) equal?  (a, b: global memory)
__device__ __forceinline__ bool same_bytes(const uint8_t* a, const uint8_t* b, uint32_t n, uint32_t lane)
{
    bool ok = true;
    for (uint32_t i = lane; i < n; i += 32u) ok = ok && (a[i] == b[i]);
    return __all_sync(FULLM, ok);
}

// error codes (first failing chunk is reported): 1 truncated, 2 bad block type, 3 stored LEN/NLEN, 4 bad code lengths,
// 5 bad symbol, 6 distance too far back, 7 output overrun, 8 literal mismatch, 9 match mismatch, 10 stored mismatch,
// 11 BFINAL misplaced, 12 fragment length / alignment, 13 short output
__global__ void __launch_bounds__(VERIFY_WARPS * 32)
k_verify(const uint8_t* __restrict__ in, uint32_t chunk, uint32_t dict, uint32_t halo, const ChunkRec* __restrict__ chunks,
         uint32_t nchunks, const uint8_t* __restrict__ out, int final, uint64_t chunk0, Cursor* __restrict__ cur)
{
    __shared__ WarpTabs tabs[VERIFY_WARPS];
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint32_t c = blockIdx.x * VERIFY_WARPS + w;
    if (c >= nchunks) return;
    WarpTabs& T = tabs[w];
    const ChunkRec cr = chunks[c];
    const uint8_t* base = in + (uint64_t)c * chunk;                      // the chunk's bytes; its dictionary lies before
    const uint32_t n = cr.n_bytes;
    const uint32_t back = dict ? (uint32_t)min((uint64_t)dict, (uint64_t)c * chunk + halo) : 0u;   // bytes a match may reach back before the chunk
    const bool last_chunk = final && c == nchunks - 1u;
    Bits b; b.p = out + cr.out_byte; b.nbytes = cr.frag_bytes; b.byte = 0; b.buf = 0; b.have = 0; b.err = 0;
    uint32_t pos = 0;
    int err = 0;
    bool done = false;                                                  // BFINAL seen
    // every lane follows the same control flow; lane 0 owns the bit reader and broadcasts what it decoded
    for (;;) {
        // a fragment is a run of blocks: up to BFINAL in the last chunk, up to its last byte in the others
        uint32_t hdr = 8u;                                                // 8: no more blocks
        if (lane == 0 && !done && consumed_bits(b) < cr.frag_bytes * 8ull) hdr = take(b, 3);
        hdr = __shfl_sync(FULLM, hdr, 0);
        if (hdr == 8u) break;
        const uint32_t bfinal = hdr & 1u, type = hdr >> 1;
        if (bfinal) { done = true; if (!last_chunk) { err = 11; break; } }
        if (type == 3u) { err = 2; break; }
        if (type == 0u) {
            uint32_t len = 0, e = 0; uint64_t at = 0;
            if (lane == 0) {
                const int drop = b.have & 7; b.buf >>= drop; b.have -= drop;     // to the byte boundary
                len = take(b, 16); const uint32_t nlen = take(b, 16);
                if ((len ^ 0xFFFFu) != nlen) e = 3;
                at = consumed_bits(b) >> 3;
                if (at + len > b.nbytes) e = 1;
            }
            e = __shfl_sync(FULLM, e, 0); len = __shfl_sync(FULLM, len, 0); at = __shfl_sync(FULLM, at, 0);
            if (e) { err = (int)e; break; }
            if (pos + len > n) { err = 7; break; }
            if (!same_bytes(b.p + at, base + pos, len, lane)) { err = 10; break; }
            pos += len;
            if (lane == 0) { b.byte = at + len; b.buf = 0; b.have = 0; }
            continue;
        }
        // code tables of the block
        uint32_t e = 0;
        Code lit = {T.lit_lj, T.lit_ent, 0}, dst = {T.dst_lj, T.dst_ent, 0};
        if (type == 1u) {
            for (uint32_t i = lane; i < 288u; i += 32u) T.lens[i] = (uint8_t)(i < 144u ? 8 : i < 256u ? 9 : i < 280u ? 7 : 8);
            __syncwarp();
            build_code(lit, T.lens, 288, lane);
            __syncwarp();
            T.lens[lane] = 5;
            __syncwarp();
            build_code(dst, T.lens, 30, lane);
        } else {
            int hlit = 0, hdist = 0;
            if (lane == 0) {
                hlit = (int)take(b, 5) + 257; hdist = (int)take(b, 5) + 1;
                const int hclen = (int)take(b, 4) + 4;
                if (hlit > 286 || hdist > 30) e = 4;
                for (int i = 0; i < 19; i++) T.cl[i] = 0;
                for (int i = 0; i < hclen && !e; i++) T.cl[c_order[i]] = (uint8_t)take(b, 3);
            }
            e = __shfl_sync(FULLM, e, 0); hlit = __shfl_sync(FULLM, hlit, 0); hdist = __shfl_sync(FULLM, hdist, 0);
            __syncwarp();
            Code aux = {T.aux_lj, T.aux_ent, 0};
            if (!e && build_code(aux, T.cl, 19, lane) != 0) e = 4;
            if (!e && lane == 0) {
                int idx = 0;
                while (!e && idx < hlit + hdist) {
                    const int sym = decode_sym(b, aux);
                    if (sym < 0) { e = 5; break; }
                    if (sym < 16) T.lens[idx++] = (uint8_t)sym;
                    else {
                        int rep, val = 0;
                        if (sym == 16) { if (!idx) { e = 4; break; } val = T.lens[idx - 1]; rep = 3 + (int)take(b, 2); }
                        else if (sym == 17) rep = 3 + (int)take(b, 3);
                        else rep = 11 + (int)take(b, 7);
                        if (idx + rep > hlit + hdist) { e = 4; break; }
                        while (rep--) T.lens[idx++] = (uint8_t)val;
                    }
                }
                if (!e && T.lens[256] == 0) e = 4;
            }
            e = __shfl_sync(FULLM, e, 0);
            __syncwarp();
            if (!e) {
                // a complete code, or -- the one exception RFC 1951 allows -- a single code of length 1
                int r = build_code(lit, T.lens, hlit, lane);
                if (r < 0 || (r > 0 && !(lit.n == 1 && r == (1 << 14)))) e = 4;
                r = build_code(dst, T.lens + hlit, hdist, lane);
                if (r < 0 || (r > 0 && !(dst.n <= 1 && (dst.n == 0 || r == (1 << 14))))) e = 4;
            }
        }
        if (e) { err = (int)e; break; }
        __syncwarp();
        // symbols
        for (;;) {
            // lane 0 decodes literals on its own until something needs the warp: a match, the end of the block, an error
            uint32_t what = 0, len = 0, dist = 0, npos = pos;            // what: 0 end of block, 1 match, >= 2 error code
            if (lane == 0) {
                for (;;) {
                    const int sym = decode_sym(b, lit);
                    if (sym < 0) { what = 5; break; }
                    if (sym < 256) {
                        if (npos >= n) { what = 7; break; }
                        if (base[npos] != (uint8_t)sym) { what = 8; break; }
                        npos++;
                        continue;
                    }
                    if (sym == 256) { what = 0; break; }
                    const int ls = sym - 257;
                    if (ls >= 29) { what = 5; break; }
                    len = c_lbase[ls] + take(b, c_lext[ls]);
                    const int ds = decode_sym(b, dst);
                    if (ds < 0 || ds >= 30) { what = 5; break; }
                    dist = c_dbase[ds] + take(b, c_dext[ds]);
                    what = 1;
                    break;
                }
                if (consumed_bits(b) > b.nbytes * 8ull) what = 2;         // ran off the fragment: truncated (2 stands for code 1 here)
            }
            what = __shfl_sync(FULLM, what, 0); npos = __shfl_sync(FULLM, npos, 0);
            pos = npos;
            if (what == 0u) break;
            if (what == 2u) { err = 1; break; }
            if (what > 2u) { err = (int)what; break; }
            len = __shfl_sync(FULLM, len, 0); dist = __shfl_sync(FULLM, dist, 0);
            if (dist > pos + back) { err = 6; break; }
            if (pos + len > n) { err = 7; break; }
            if (!same_bytes(base + pos, base + pos - dist, len, lane)) { err = 9; break; }
            pos += len;
        }
        if (err) break;
    }
    if (!err) {
        uint64_t used = 0;
        if (lane == 0) used = consumed_bits(b);
        used = __shfl_sync(FULLM, used, 0);
        if (pos != n) err = 13;
        else if (last_chunk ? (!done || ((used + 7ull) >> 3) != cr.frag_bytes) : (done || used != cr.frag_bytes * 8ull)) err = 12;
    }
    if (err && lane == 0) {
        if (atomicAdd(&cur->verify_bad, 1u) == 0u) { cur->verify_first = chunk0 + c; cur->verify_code = (uint32_t)err; }
    }
}

}  // namespace

void launch_verify(const uint8_t* in, uint32_t chunk, uint32_t dict, uint32_t halo, const ChunkRec* chunks, uint32_t nchunks,
                   const uint8_t* out, int final, uint64_t chunk0, Cursor* cur, cudaStream_t st)
{
    if (nchunks == 0) return;
    k_verify<<<(nchunks + VERIFY_WARPS - 1) / VERIFY_WARPS, VERIFY_WARPS * 32, 0, st>>>(in, chunk, dict, halo, chunks, nchunks, out, final, chunk0, cur);
}

}  // namespace gzb
