// gzb_common.cuh -- constants, device records and small helpers shared by the kernels.
//
// Data layout in HBM (per sub-batch of chunks; see DESIGN.md "Data layout"):
//   in          caller's input, untouched; chunk c = in[c*CHUNK, (c+1)*CHUNK), its dictionary is the
//               32 KiB before it (no copy: the reference memcpy's it, src/parallel.c:767-774)
//   tokens      u32[nchunks][CHUNK]      LZ77 token stream of each chunk (k_lz77 -> k_pack)
//   blocks      BlockRec[nchunks][BPC]   per deflate block: histograms (k_lz77), code tables + header
//                                        (k_huffman), absolute output bit offset (k_layout)
//   chunks      ChunkRec[nchunks]        per chunk: CRC-32, fragment size and byte offset
//   out         caller's output; k_pack writes every block at its final bit position
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace gzb {

constexpr int      MAX_CHUNK   = 131072;          // IN_BUF_SIZE, src/parallel.c:13
constexpr int      DICT        = 32768;           // DICTIONARY_SIZE, src/parallel.c:15
constexpr int      TILE        = 2048;            // positions per pipeline step of k_lz77
constexpr int      RING        = 65536;           // bytes of window kept in shared memory
constexpr int      LINK_RING   = 32768 + 2 * TILE;// chain links kept in shared memory
constexpr int      MAXD        = 32768;           // farthest match (deflate's limit)
constexpr int      TOO_FAR     = 4096;            // lib/zlib/deflate.c:110-112
constexpr int      HBITS       = 14;              // buckets of the 6-byte chain table
constexpr int      H4BITS      = 12;              // buckets of the 4-byte single-entry table
constexpr int      H3BITS      = 11;              // buckets of the 3-byte single-entry table
constexpr int      BLOCK_SPAN  = 65536;           // input bytes per deflate block
constexpr int      MAX_BPC     = MAX_CHUNK / BLOCK_SPAN;   // blocks per chunk
constexpr int      LZ_THREADS  = 1024;
constexpr int      NLSYM       = 286;
constexpr int      NDSYM       = 30;
constexpr int      HDR_BYTES   = 608;             // worst case 14 + 57 + 316*14 bits = 4495 bits

// Search effort per level.  good/lazy/nice are the reference's configuration_table
// (lib/zlib/deflate.c:134-146); depth replaces max_chain (DESIGN.md "Level semantics").
struct LevelParams {
    int depth, nice, max_lazy, lazy_parse, use_h3, use_h4;
    // ---- scheduling of k_lz77: none of these changes a byte of output (tests/ force each of them) ----
    int parse_warps;                         // 8 or 16: size of the parser group
    int help_ins, help_parse;                // the inserter / parser warps take search groups when their own tile is done
    int max_run, force_run;                  // longest run of chained chunks a launch may pick; > 0: run length forced (tests)
    // what each of the 32 warps of a CTA does (lz77_role_map).  role_b0..3: one byte per warp (warps 0-7, 8-15, ...):
    // 0-2 inserter of the 6-byte / 3-byte / 4-byte table, 4 search, 8 + i parser i; role_mask: bit w set = warp w is
    // not a pure search warp.  (Scalars, not arrays: an array indexed by the warp number would move the whole struct to
    // local memory.)
    unsigned long long role_b0, role_b1, role_b2, role_b3;
    uint32_t role_mask;
};

// One deflate block.  Written by k_lz77 (histograms, token range, input range),
// k_huffman (plan) and k_layout (out_bit, bfinal).
struct BlockRec {
    uint32_t lfreq[288];      // literal/length histogram (_tr_tally, lib/zlib/deflate.h:326-342)
    uint32_t dfreq[32];       // distance histogram
    uint16_t lcode[288];      // bit-reversed canonical codes (gen_codes, trees.c:572-605)
    uint16_t dcode[32];
    uint8_t  llen[288];
    uint8_t  dlen[32];
    uint8_t  header[HDR_BYTES];   // HLIT,HDIST,HCLEN + code-length section (send_all_trees, trees.c:834-858)
    uint32_t header_bits;
    uint32_t type;            // 0 stored, 1 static, 2 dynamic (trees.c:962-993)
    uint32_t tok_begin;       // index into the chunk's token array
    uint32_t tok_count;
    uint32_t in_off;          // first input byte covered, relative to the chunk
    uint32_t in_bytes;        // input bytes covered
    uint32_t bfinal;
    uint32_t pad0;
    uint64_t body_bits;       // size of the chosen encoding in bits, 3 type bits excluded (stored: see k_layout)
    uint64_t out_bit;         // absolute bit position of the block in the output buffer
};

struct ChunkRec {
    uint32_t crc;             // CRC-32 of the chunk's bytes
    uint32_t n_bytes;         // input bytes in the chunk
    uint32_t n_blocks;
    uint32_t n_tokens;
    uint64_t frag_bytes;      // size of the byte-aligned raw-deflate fragment
    uint64_t out_byte;        // absolute byte offset of the fragment in the output buffer
    uint32_t tail_kind;       // 0 none, 1 empty stored block appended for byte alignment
    uint32_t stored;          // k_scan: the chunk is incompressible -- no match finding, its blocks are stored blocks
    uint64_t tail_bit;        // where the alignment block goes
};

// Running state carried across sub-batches on the device (no host sync in between).
struct Cursor {
    uint64_t out_bytes;       // bytes produced so far
    uint64_t in_bytes;        // bytes consumed so far
    uint32_t crc;             // CRC-32 of everything consumed so far
    uint32_t n_chunks;
    uint32_t n_blocks;
    int32_t  status;
    uint32_t verify_bad;      // GZB_F_VERIFY: chunks whose fragment does not decode back to their input (k_verify)
    uint32_t verify_code;     // ... what was wrong with the first one found (codes in k_verify.cu)
    uint64_t verify_first;    // ... and its index
};

// ---- token format: bit31 = match; match: (len-3) << 16 | (dist-1); literal: byte ----
__host__ __device__ __forceinline__ uint32_t tok_match(uint32_t len, uint32_t dist) { return 0x80000000u | ((len - 3u) << 16) | (dist - 1u); }

// ---- deflate symbol maps (RFC 1951 3.2.5; cf. _length_code/_dist_code, lib/zlib/trees.h) ----
__host__ __device__ __forceinline__ uint32_t len_code(uint32_t len)     // 3..258 -> 0..28
{
    uint32_t l = len - 3u;
    if (l < 8u) return l;
    if (len == 258u) return 28u;
#ifdef __CUDA_ARCH__
    uint32_t n = 31u - __clz(l);
#else
    uint32_t n = 31u - (uint32_t)__builtin_clz(l);
#endif
    return 4u * (n - 1u) + ((l >> (n - 2u)) & 3u);
}
__host__ __device__ __forceinline__ uint32_t len_extra_bits(uint32_t code) { return (code < 8u || code == 28u) ? 0u : (code - 4u) >> 2; }
__host__ __device__ __forceinline__ uint32_t len_base(uint32_t code)
{
    if (code < 8u) return code + 3u;
    if (code == 28u) return 258u;
    uint32_t e = (code - 4u) >> 2;
    return 3u + ((4u + (code & 3u)) << e);
}
__host__ __device__ __forceinline__ uint32_t dist_code(uint32_t dist)   // 1..32768 -> 0..29
{
    uint32_t d = dist - 1u;
    if (d < 4u) return d;
#ifdef __CUDA_ARCH__
    uint32_t n = 31u - __clz(d);
#else
    uint32_t n = 31u - (uint32_t)__builtin_clz(d);
#endif
    return 2u * n + ((d >> (n - 1u)) & 1u);
}
__host__ __device__ __forceinline__ uint32_t dist_extra_bits(uint32_t code) { return code < 4u ? 0u : (code >> 1) - 1u; }
__host__ __device__ __forceinline__ uint32_t dist_base(uint32_t code)
{
    if (code < 4u) return code + 1u;
    uint32_t e = (code >> 1) - 1u;
    return 1u + ((2u + (code & 1u)) << e);
}

__host__ __device__ __forceinline__ uint32_t hash6(uint32_t lo, uint32_t hi16) { return (lo * 0x9E3779B1u + hi16 * 0x85EBCA6Bu) >> (32 - HBITS); }
__host__ __device__ __forceinline__ uint32_t hash4(uint32_t v) { return (v * 0x9E3779B1u) >> (32 - H4BITS); }
__host__ __device__ __forceinline__ uint32_t hash3(uint32_t v) { return ((v << 8) * 0x9E3779B1u) >> (32 - H3BITS); }

// CRC-32 (reflected, poly 0xEDB88320) helpers shared by k_crc32 and k_layout.
constexpr uint32_t CRC_POLY = 0xEDB88320u;
// a(x) * b(x) mod P in the reflected domain (bit 31 = x^0)
__host__ __device__ __forceinline__ uint32_t gf2_mul(uint32_t a, uint32_t b)
{
    uint32_t p = 0;
#pragma unroll 1
    for (int i = 0; i < 32; i++) {
        if (a & 0x80000000u) p ^= b;
        a <<= 1;
        b = (b >> 1) ^ ((b & 1u) ? CRC_POLY : 0u);
    }
    return p;
}

// multiply crc by x^(8*len) mod P; x2n[k] = x^(2^k) mod P
__host__ __device__ __forceinline__ uint32_t crc_shift(const uint32_t* x2n, uint32_t crc, uint64_t len)
{
    uint64_t bits = len << 3;
    for (int k = 0; bits; k++, bits >>= 1)
        if (bits & 1u) crc = gf2_mul(x2n[k], crc);
    return crc;
}
constexpr int CRC_TAB_WORDS = 1024 + 64;      // slice-by-4 tables followed by x^(2^k), k < 64

// ---- incompressibility test of k_scan (the same integer arithmetic as oracle/engine_model.c gzm_chunk_is_stored) ----
constexpr int      SCAN_MIN        = 4096;    // shorter chunks always take the normal path
constexpr int      SCAN_SAMPLE     = 256;     // one 16-byte word out of every 256 bytes goes into the byte histogram
constexpr uint32_t SCAN_MIN_BITS_N = 127, SCAN_MIN_BITS_D = 128;   // "incompressible": order-0 entropy >= 8 * 127/128 bits per byte ...
constexpr uint32_t SCAN_SLOTS      = 4096;    // ... and fewer than 1 in 64 of the sampled 8-byte grams seen before (table of fingerprints)
constexpr uint32_t SCAN_MAX_GRAMS  = 3072;
// 256 * log2(1 + m / 256), rounded
__device__ __constant__ const uint8_t c_log2_q8[256] = {
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
// a position is sampled iff the low six bits of its first byte are 0x2A: one in 64 on the near-uniform bytes the test runs on
__host__ __device__ __forceinline__ uint32_t scan_select(uint32_t w) { return (w & 0x3Fu) == 0x2Au; }
__host__ __device__ __forceinline__ uint32_t scan_fingerprint(uint32_t w, uint32_t w2) { return ((w * 0x85EBCA6Bu) ^ (w2 * 0xC2B2AE35u)) | 1u; }

// launchers (defined in the k_*.cu files)
// scan: 0 = CRC-32 only; 1 = CRC-32 + incompressibility test (fills the records of stored chunks); dict / halo as for launch_lz77
void launch_scan(const uint8_t* in, uint64_t n_bytes, uint32_t halo, uint32_t chunk, uint32_t dict, uint32_t nchunks, uint32_t bpc, int scan,
                 ChunkRec* chunks, BlockRec* blocks, const uint32_t* d_tab, cudaStream_t st);
void launch_lz77(const uint8_t* in, uint64_t n_bytes, uint32_t halo, uint32_t chunk, uint32_t dict, uint32_t nchunks,
                 uint32_t* tokens, BlockRec* blocks, ChunkRec* chunks, LevelParams lp, uint32_t* work_ctr, int n_sm, cudaStream_t st);
void launch_huffman(BlockRec* blocks, const ChunkRec* chunks, uint32_t nchunks, uint32_t bpc, cudaStream_t st);
void launch_layout(BlockRec* blocks, ChunkRec* chunks, uint32_t nchunks, uint32_t bpc, int final, Cursor* cur, const uint32_t* d_tab, cudaStream_t st);
void launch_verify(const uint8_t* in, uint32_t chunk, uint32_t dict, uint32_t halo, const ChunkRec* chunks, uint32_t nchunks,
                   const uint8_t* out, int final, uint64_t chunk0, Cursor* cur, cudaStream_t st);
void launch_pack(const uint8_t* in, uint32_t chunk, const uint32_t* tokens, const BlockRec* blocks, const ChunkRec* chunks,
                 uint32_t nchunks, uint32_t bpc, uint8_t* out, cudaStream_t st);
cudaError_t lz77_configure();
void lz77_role_map(LevelParams& lp, int layout);   // fills lp.role_b0..3 / role_mask from lp.parse_warps
cudaError_t lz77_read_prof(unsigned long long* out16, int reset);

}  // namespace gzb
