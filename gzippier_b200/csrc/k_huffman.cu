// k_huffman.cu -- per deflate block: length-limited Huffman codes, code-length header, block-type choice.
//
// Replaces build_tree / gen_bitlen / gen_codes (lib/zlib/trees.c:615-697, 486-562, 572-605),
// scan_tree / send_tree / build_bl_tree / send_all_trees (:703-858) and the stored/static/dynamic
// decision of _tr_flush_block (:911-1008).  Not a port: zlib builds the tree with a heap and repairs
// over-long codes by moving leaves; here the symbols are sorted by a warp-wide bitonic network,
// code lengths come from the in-place Moffat-Katajainen pass over the sorted frequencies, and the
// 15-bit (7-bit for the code-length alphabet) limit is enforced on the count-per-length vector with
// a Kraft-sum repair.  oracle/engine_model.c (gzm_plan_block) is the sequential statement.
//
// One warp per block.  The serial parts run on lane 0; they are a few thousand dependent steps per
// block and there are tens of thousands of blocks in flight, so the stage is latency-tolerant
// (measured share of the pipeline: profiles/).
#include "gzb_common.cuh"

namespace gzb {

namespace {

constexpr uint32_t FULL = 0xFFFFFFFFu;
__constant__ uint8_t c_bl_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

struct HuffSmem {
    uint32_t key[512];        // sort keys: freq << 9 | symbol
    uint32_t A[288];          // sorted frequencies -> code lengths (in place)
    uint32_t ftmp[288];       // working copy of a histogram (forced symbols live only here)
    uint32_t lfreq[288];
    uint32_t dfreq[32];
    uint32_t bfreq[19];
    uint8_t  llen[288], dlen[32], blen[19];
    uint16_t bcode[19];
    uint8_t  rsym[320], rext[320];
    uint8_t  header[HDR_BYTES];
};

// ---- bitonic sort of 512 keys by one warp ----
__device__ void warp_sort512(uint32_t* key, uint32_t lane)
{
    for (uint32_t k = 2; k <= 512u; k <<= 1) {
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            for (uint32_t t = lane; t < 256u; t += 32u) {
                uint32_t i = ((t & ~(j - 1u)) << 1) | (t & (j - 1u));    // index with bit j clear
                uint32_t l = i | j;
                uint32_t a = key[i], b = key[l];
                bool up = ((i & k) == 0u);
                if ((a > b) == up) { key[i] = b; key[l] = a; }
            }
            __syncwarp();
        }
    }
}

// 32 keys, one per lane, in registers
__device__ __forceinline__ uint32_t warp_sort32(uint32_t key, uint32_t lane)
{
#pragma unroll
    for (uint32_t k = 2; k <= 32u; k <<= 1) {
#pragma unroll
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            const uint32_t other = __shfl_xor_sync(FULL, key, (int)j);
            const bool up = (lane & k) == 0u, lower = (lane & j) == 0u;
            key = (lower == up) ? min(key, other) : max(key, other);
        }
    }
    return key;
}

// In-place minimum-redundancy code lengths (Moffat & Katajainen); A sorted ascending.
__device__ void min_redundancy(uint32_t* A, int n)
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

// freq[0..nsym) -> lens[0..nsym) limited to maxbits; forces >= 2 coded symbols (trees.c:646-651).
// Whole warp calls; returns the highest coded symbol.  SMALL: nsym <= 32 (sorted in registers).
template <bool SMALL>
__device__ int huff_lengths(HuffSmem& S, const uint32_t* freq_in, int nsym, int maxbits, uint8_t* lens, uint32_t lane)
{
    uint32_t* freq = S.ftmp;
    int used = 0;
    for (uint32_t i = lane; i < (uint32_t)nsym; i += 32u) { const uint32_t f = freq_in[i]; freq[i] = f; used += f != 0u; }
    for (int o = 16; o > 0; o >>= 1) used += __shfl_xor_sync(FULL, used, o);
    __syncwarp();
    if (used < 2 && lane == 0)
        for (int i = 0; used < 2 && i < nsym; i++) if (!freq[i]) { freq[i] = 1; used++; }
    __syncwarp();
    int n = 0;
    if (SMALL) {
        const uint32_t key = (lane < (uint32_t)nsym && freq[lane]) ? ((freq[lane] << 9) | lane) : 0xFFFFFFFFu;
        if (lane < (uint32_t)nsym) lens[lane] = 0;
        S.key[lane] = warp_sort32(key, lane);
        n = __popc(__ballot_sync(FULL, key != 0xFFFFFFFFu));
        __syncwarp();
    } else {
        for (uint32_t i = lane; i < 512u; i += 32u)
            S.key[i] = (i < (uint32_t)nsym && freq[i]) ? ((freq[i] << 9) | i) : 0xFFFFFFFFu;
        for (uint32_t i = lane; i < (uint32_t)nsym; i += 32u) lens[i] = 0;
        __syncwarp();
        warp_sort512(S.key, lane);
        for (uint32_t i = lane; i < 512u; i += 32u) n += S.key[i] != 0xFFFFFFFFu;
        for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(FULL, n, o);
    }
    for (uint32_t i = lane; i < (uint32_t)n; i += 32u) S.A[i] = S.key[i] >> 9;
    __syncwarp();
    int maxsym = 0;
    if (lane == 0) {
        min_redundancy(S.A, n);
        int num[33];
        for (int i = 0; i <= 32; i++) num[i] = 0;
        for (int i = 0; i < n; i++) { uint32_t l = S.A[i]; num[l > 32u ? 32 : l]++; }
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
        int j = n;
        for (int i = 1; i <= maxbits; i++) for (int l = num[i]; l > 0; l--) lens[S.key[--j] & 511u] = (uint8_t)i;
    }
    __syncwarp();
    for (uint32_t i = lane; i < (uint32_t)nsym; i += 32u) if (lens[i]) maxsym = (int)i;      // ascending per lane
    for (int o = 16; o > 0; o >>= 1) maxsym = max(maxsym, __shfl_xor_sync(FULL, maxsym, o));
    return maxsym;
}

// Canonical codes, bit-reversed for LSB-first emission (gen_codes, trees.c:572-605), by the whole warp: symbol
// c * 32 + lane sits in slot c of its lane; per code length the symbols that have it are ranked with ballots, and the
// first code of a length follows from the count of the one before.
template <int NCH>
__device__ void huff_codes(const uint8_t* lens, int nsym, int maxbits, uint16_t* codes, uint32_t lane)
{
    uint32_t len[NCH], cod[NCH];
#pragma unroll
    for (int c = 0; c < NCH; c++) { const uint32_t i = (uint32_t)c * 32u + lane; len[c] = i < (uint32_t)nsym ? lens[i] : 0u; cod[c] = 0; }
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t code = 0, prev_cnt = 0;
    for (uint32_t l = 1; l <= (uint32_t)maxbits; l++) {
        code = (code + prev_cnt) << 1;
        uint32_t cnt = 0;
#pragma unroll
        for (int c = 0; c < NCH; c++) {
            const uint32_t m = __ballot_sync(FULL, len[c] == l);
            if (len[c] == l) cod[c] = __brev(code + cnt + (uint32_t)__popc(m & lt)) >> (32u - l);
            cnt += (uint32_t)__popc(m);
        }
        prev_cnt = cnt;
    }
#pragma unroll
    for (int c = 0; c < NCH; c++) { const uint32_t i = (uint32_t)c * 32u + lane; if (i < (uint32_t)nsym) codes[i] = (uint16_t)cod[c]; }
}

// Run-length coding of a code-length sequence with the symbols 16 / 17 / 18 (RFC 1951 3.2.7), the engine's own policy
// (oracle/engine_model.c rle_lengths states it sequentially; zlib's is scan_tree / send_tree, lib/zlib/trees.c:703-793):
// the sequence is cut into MAXIMAL runs of equal lengths and every run is coded on its own --
//   zeros:  as many "18" (138 zeros) as fit, then one "18" / "17" for 3 or more left over, else the zeros themselves;
//   v != 0: v once, then "16" (repeat 3..6) while 3 or more are left, else one or two plain v.
// Whole warp: run starts by ballot, ONE LANE PER RUN computes how many symbols its run takes, a warp scan turns the
// counts into offsets, and every lane writes its run's symbols.  Returns the number of symbols.
__device__ __forceinline__ int rle_run_count(int v, int L)
{
    if (v == 0) { const int q = L / 138, r = L % 138; return q + (r >= 3 ? 1 : r); }
    const int R = L - 1, q = R / 6, r = R % 6;
    return 1 + q + (r >= 3 ? 1 : r);
}
__device__ int rle_lengths(HuffSmem& S, const uint8_t* lens, int n, uint8_t* sym, uint8_t* ext, uint32_t lane)
{
    uint16_t* runpos = reinterpret_cast<uint16_t*>(S.key);        // start position of every run (S.key is free here)
    // 1. run starts: position i starts a run iff i == 0 or lens[i] != lens[i - 1]
    int nruns = 0;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int i = i0 + (int)lane;
        const bool start = i < n && (i == 0 || lens[i] != lens[i - 1]);
        const uint32_t m = __ballot_sync(FULL, start);
        if (start) runpos[nruns + __popc(m & ((1u << lane) - 1u))] = (uint16_t)i;
        nruns += __popc(m);
    }
    if (lane == 0) runpos[nruns] = (uint16_t)n;
    __syncwarp();
    // 2. one lane per run: symbols needed, offsets by a warp scan (runs in rounds of 32), then the symbols
    int total = 0;
    for (int r0 = 0; r0 < nruns; r0 += 32) {
        const int r = r0 + (int)lane;
        int v = 0, L = 0, cnt = 0;
        if (r < nruns) { const int i = runpos[r]; v = lens[i]; L = (int)runpos[r + 1] - i; cnt = rle_run_count(v, L); }
        int incl = cnt;
        for (int o = 1; o < 32; o <<= 1) { const int x = __shfl_up_sync(FULL, incl, o); if ((int)lane >= o) incl += x; }
        int k = total + incl - cnt;
        if (v == 0) {
            while (L >= 138) { sym[k] = 18; ext[k++] = 138 - 11; L -= 138; }
            if (L >= 11) { sym[k] = 18; ext[k++] = (uint8_t)(L - 11); }
            else if (L >= 3) { sym[k] = 17; ext[k++] = (uint8_t)(L - 3); }
            else while (L-- > 0) { sym[k] = 0; ext[k++] = 0; }
        } else {
            sym[k] = (uint8_t)v; ext[k++] = 0;
            int R = L - 1;
            while (R >= 6) { sym[k] = 16; ext[k++] = 6 - 3; R -= 6; }
            if (R >= 3) { sym[k] = 16; ext[k++] = (uint8_t)(R - 3); }
            else while (R-- > 0) { sym[k] = (uint8_t)v; ext[k++] = 0; }
        }
        total += __shfl_sync(FULL, incl, 31);
    }
    __syncwarp();
    return total;
}

// LSB-first bit writer: bits collect in a 64-bit accumulator and leave a byte at a time
struct BitW { uint8_t* out; uint32_t bitpos; unsigned long long acc; uint32_t nacc; };
__device__ __forceinline__ void putbits(BitW& w, uint32_t v, int n)     // n <= 16, v < 2^n
{
    w.acc |= (unsigned long long)v << w.nacc; w.nacc += (uint32_t)n; w.bitpos += (uint32_t)n;
    while (w.nacc >= 8u) { *w.out++ = (uint8_t)w.acc; w.acc >>= 8; w.nacc -= 8u; }
}
__device__ __forceinline__ void flushbits(BitW& w) { if (w.nacc) { *w.out++ = (uint8_t)w.acc; w.acc = 0; w.nacc = 0; } }

__device__ __forceinline__ uint32_t static_llen(uint32_t i) { return i < 144u ? 8u : i < 256u ? 9u : i < 280u ? 7u : 8u; }

__global__ void __launch_bounds__(32)
k_huffman(BlockRec* __restrict__ blocks, const ChunkRec* __restrict__ chunks, uint32_t bpc)
{
    __shared__ HuffSmem S;
    const uint32_t lane = threadIdx.x;
    const uint32_t c = blockIdx.x / bpc, b = blockIdx.x % bpc;
    if (b >= chunks[c].n_blocks || chunks[c].stored) return;       // stored chunk (k_scan): its block records are complete
    BlockRec* B = &blocks[blockIdx.x];

    for (uint32_t i = lane; i < 288u; i += 32u) S.lfreq[i] = (i < (uint32_t)NLSYM) ? B->lfreq[i] : 0u;
    S.dfreq[lane] = (lane < (uint32_t)NDSYM) ? B->dfreq[lane] : 0u;
    __syncwarp();
    if (lane == 0) S.lfreq[256] = 1;                       // EOB (init_block, trees.c:407-420)
    __syncwarp();

    int maxl = huff_lengths<false>(S, S.lfreq, NLSYM, 15, S.llen, lane);
    int maxd = huff_lengths<true>(S, S.dfreq, NDSYM, 15, S.dlen, lane);
    if (maxl < 256) maxl = 256;
    const int hlit = maxl + 1, hdist = maxd + 1;

    // code-length alphabet: run-length code of the two length sequences (whole warp), histogram of its symbols
    const int k1 = rle_lengths(S, S.llen, hlit, S.rsym, S.rext, lane);
    const int nrle = k1 + rle_lengths(S, S.dlen, hdist, S.rsym + k1, S.rext + k1, lane);
    if (lane < 19u) S.bfreq[lane] = 0;
    __syncwarp();
    for (int i = (int)lane; i < nrle; i += 32) atomicAdd(&S.bfreq[S.rsym[i]], 1u);
    __syncwarp();
    huff_lengths<true>(S, S.bfreq, 19, 7, S.blen, lane);
    huff_codes<1>(S.blen, 19, 7, S.bcode, lane);
    __syncwarp();

    // body sizes in bits under the dynamic and the static code (all lanes, then a warp sum)
    unsigned long long dyn_body = 0, sta_body = 0;
    for (uint32_t i = lane; i < (uint32_t)NLSYM; i += 32u) {
        const uint32_t ex = i >= 257u ? len_extra_bits(i - 257u) : 0u;
        dyn_body += (unsigned long long)S.lfreq[i] * (S.llen[i] + ex);
        sta_body += (unsigned long long)S.lfreq[i] * (static_llen(i) + ex);
    }
    if (lane < (uint32_t)NDSYM) {
        const uint32_t ex = dist_extra_bits(lane);
        dyn_body += (unsigned long long)S.dfreq[lane] * (S.dlen[lane] + ex);
        sta_body += (unsigned long long)S.dfreq[lane] * (5u + ex);
    }
    for (int o = 16; o > 0; o >>= 1) { dyn_body += __shfl_xor_sync(FULL, dyn_body, o); sta_body += __shfl_xor_sync(FULL, sta_body, o); }

    if (lane == 0) {
        int hclen = 19;
        while (hclen > 4 && S.blen[c_bl_order[hclen - 1]] == 0) hclen--;
        BitW w = {S.header, 0, 0ull, 0};
        putbits(w, (uint32_t)(hlit - 257), 5); putbits(w, (uint32_t)(hdist - 1), 5); putbits(w, (uint32_t)(hclen - 4), 4);
        for (int i = 0; i < hclen; i++) putbits(w, S.blen[c_bl_order[i]], 3);
        for (int i = 0; i < nrle; i++) {
            uint32_t sy = S.rsym[i];
            putbits(w, S.bcode[sy], S.blen[sy]);
            if (sy == 16u) putbits(w, S.rext[i], 2); else if (sy == 17u) putbits(w, S.rext[i], 3); else if (sy == 18u) putbits(w, S.rext[i], 7);
        }
        flushbits(w);
        const uint32_t header_bits = w.bitpos;
        // sizes in bits (header included, 3 type bits excluded)
        const unsigned long long dyn = header_bits + dyn_body, sta = sta_body;
        // choice (cf. trees.c:945-993): bytes rounded up, 3 type bits included
        const uint32_t in_bytes = B->in_bytes;
        unsigned long long dyn_b = (dyn + 3 + 7) >> 3, sta_b = (sta + 3 + 7) >> 3;
        unsigned long long nsub = in_bytes ? (in_bytes + 65534u) / 65535u : 1u;
        unsigned long long sto_b = in_bytes + 5ull * nsub;
        unsigned long long bestb = dyn_b < sta_b ? dyn_b : sta_b;
        uint32_t type = (sto_b <= bestb) ? 0u : (sta_b <= dyn_b) ? 1u : 2u;
        B->type = type;
        B->header_bits = header_bits;
        B->body_bits = (type == 2u) ? dyn : (type == 1u) ? sta : 0ull;
        S.key[0] = type;
    }
    __syncwarp();
    const uint32_t type = S.key[0];
    __syncwarp();
    if (type == 1u) {
        for (uint32_t i = lane; i < 288u; i += 32u) S.llen[i] = (uint8_t)static_llen(i);
        S.dlen[lane] = (lane < (uint32_t)NDSYM) ? 5 : 0;
        __syncwarp();
    }
    huff_codes<9>(S.llen, type == 1u ? 288 : NLSYM, 15, B->lcode, lane); huff_codes<1>(S.dlen, NDSYM, 15, B->dcode, lane);
    for (uint32_t i = lane; i < 288u; i += 32u) B->llen[i] = (i < (uint32_t)NLSYM || type == 1u) ? S.llen[i] : 0;
    B->dlen[lane] = (lane < (uint32_t)NDSYM) ? S.dlen[lane] : 0;
    for (uint32_t i = lane; i < (uint32_t)HDR_BYTES; i += 32u) B->header[i] = S.header[i];
}

}  // namespace

void launch_huffman(BlockRec* blocks, const ChunkRec* chunks, uint32_t nchunks, uint32_t bpc, cudaStream_t st)
{
    if (nchunks == 0) return;
    k_huffman<<<nchunks * bpc, 32, 0, st>>>(blocks, chunks, bpc);
}

}  // namespace gzb
