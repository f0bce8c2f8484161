/*
 * xet_oracle.c -- CPU restatement of the Xet xorb data path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the parity oracle for the CUDA library in zest_b200/csrc.  Nothing in the product path
 * (zest_b200/, include/) may link, import or call it; only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py do, and only as the checker or the timed CPU arm.
 *
 * What it restates.  zest (praveer13/zest) does not contain the arithmetic of this path: it calls the
 * third-party Zig package `zig-xet` (module `xet`, /root/reference/build.zig.zon:6-11, pinned by hash
 * xet-0.1.0-d6TQaooXAwAQHBVsrWAQPNFzJO-xiCla5Ca6a_ibqLvc, fetched at build time, NOT vendored) from
 *   - src/xet_bridge.zig:256-257        XorbReader.init(alloc, data).extractChunkRange(start, end)
 *   - src/parallel_download.zig:65-66   the same call, one per reconstruction term
 *   - src/xet_bridge.zig:154, src/server.zig:203   xet.hashing.hashToHex
 *   - src/bench.zig:207-223             std.crypto.hash.Blake3 over 64 KiB (unkeyed)
 * zig-xet implements the public HuggingFace Xet protocol, so the algorithm restated here is the published
 * one: BLAKE3 (7 rounds, 1 KiB leaves, left-full tree) keyed with the Xet
 * DATA / INTERNAL-NODE keys, the variable-branching Merkle fold, the 8-byte chunk header walk, LZ4 *frame*
 * decode, 4-way byte-group regrouping and the Gearhash content-defined chunker.
 *
 * Parity status: the reference's own tests hold no golden vector for this path, so
 * parity is pinned against HuggingFace's production implementation of the same protocol instead: the
 * vectors and real xorbs under tests/golden/ were produced in the build container by hf_xet 1.5.0
 * (xet-core) and PyPI blake3 1.0.8 with tests/golden/make_golden.py, and tests/test_oracle.py checks this
 * file against every one of them.  Parity with zig-xet itself is by protocol equivalence ("unpinned by the
 * reference").
 *
 * Build: make -C oracle   ->  oracle/libxet_oracle.so   (gcc -O3 -march=x86-64-v3 -pthread)
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/zest_gear_table.h"

#define OZ_API __attribute__((visibility("default")))

/* Error codes mirror include/zest_gpu.h (ZG_ERR_*), so tests can compare verdicts one to one. */
enum {
    OZ_OK = 0,
    OZ_ERR_TRUNCATED = -3,       /* header or payload runs past the blob          */
    OZ_ERR_BAD_HEADER = -4,      /* chunk header version != 0 / None with clen != ulen */
    OZ_ERR_SCHEME = -5,          /* compression scheme not 0/1/2                  */
    OZ_ERR_LZ4_FRAME = -6,       /* bad magic / version / header checksum / dict  */
    OZ_ERR_LZ4_BLOCK = -7,       /* malformed sequence, offset 0 or out of window */
    OZ_ERR_SIZE_MISMATCH = -8,   /* decoded size != header uncompressed_len       */
    OZ_ERR_RANGE = -9,           /* chunk range outside the blob                  */
    OZ_ERR_CAPACITY = -10,       /* caller's output too small                     */
    OZ_ERR_NOLZ4 = -20           /* builder only: liblz4.so.1 not loadable        */
};

/* ------------------------------------------------------------------------------------------------ */
/* A.2  BLAKE3                                                                                      */
/* ------------------------------------------------------------------------------------------------ */

enum { B3_CHUNK_START = 1, B3_CHUNK_END = 2, B3_PARENT = 4, B3_ROOT = 8, B3_KEYED = 16 };

static const uint32_t B3_IV[8] = {0x6A09E667u, 0xBB67AE85u, 0x3C6EF372u, 0xA54FF53Au,
                                  0x510E527Fu, 0x9B05688Cu, 0x1F83D9ABu, 0x5BE0CD19u};
static const uint8_t B3_PERM[16] = {2, 6, 3, 10, 7, 0, 4, 13, 1, 11, 12, 5, 9, 14, 15, 8};

/* Xet domain keys, SURVEY.md A.2 / A.3 (decimal bytes as published). */
static const uint8_t XET_DATA_KEY[32] = {102, 151, 245, 119, 91,  149, 80,  222, 49,  53,  203,
                                         172, 165, 151, 24,  28,  157, 228, 33,  16,  155, 235,
                                         43,  88,  180, 208, 176, 75,  147, 173, 242, 41};
static const uint8_t XET_NODE_KEY[32] = {1,   126, 197, 199, 165, 71,  41,  150, 253, 148, 102,
                                         102, 180, 138, 2,   230, 93,  221, 83,  111, 55,  199,
                                         109, 210, 248, 99,  82,  230, 74,  83,  113, 63};

static inline uint32_t rotr32(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
static inline uint32_t ld32(const uint8_t* p) {
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
static inline void st32(uint8_t* p, uint32_t v) {
    p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}

#define B3_G(a, b, c, d, x, y)                      \
    do {                                            \
        v[a] = v[a] + v[b] + (x);                   \
        v[d] = rotr32(v[d] ^ v[a], 16);             \
        v[c] = v[c] + v[d];                         \
        v[b] = rotr32(v[b] ^ v[c], 12);             \
        v[a] = v[a] + v[b] + (y);                   \
        v[d] = rotr32(v[d] ^ v[a], 8);              \
        v[c] = v[c] + v[d];                         \
        v[b] = rotr32(v[b] ^ v[c], 7);              \
    } while (0)

/* One compression; writes the new 8-word chaining value to out (out may alias cv). */
static void b3_compress(const uint32_t cv[8], const uint32_t block[16], uint64_t counter,
                        uint32_t block_len, uint32_t flags, uint32_t out[8]) {
    uint32_t v[16], m[16], t[16];
    for (int i = 0; i < 8; i++) v[i] = cv[i];
    for (int i = 0; i < 4; i++) v[8 + i] = B3_IV[i];
    v[12] = (uint32_t)counter;
    v[13] = (uint32_t)(counter >> 32);
    v[14] = block_len;
    v[15] = flags;
    memcpy(m, block, sizeof m);
    for (int r = 0;; r++) {
        B3_G(0, 4, 8, 12, m[0], m[1]);
        B3_G(1, 5, 9, 13, m[2], m[3]);
        B3_G(2, 6, 10, 14, m[4], m[5]);
        B3_G(3, 7, 11, 15, m[6], m[7]);
        B3_G(0, 5, 10, 15, m[8], m[9]);
        B3_G(1, 6, 11, 12, m[10], m[11]);
        B3_G(2, 7, 8, 13, m[12], m[13]);
        B3_G(3, 4, 9, 14, m[14], m[15]);
        if (r == 6) break;
        for (int i = 0; i < 16; i++) t[i] = m[B3_PERM[i]];
        memcpy(m, t, sizeof m);
    }
    for (int i = 0; i < 8; i++) out[i] = v[i] ^ v[i + 8];
}

/* Chaining value of one <=1024-byte leaf.  root!=0 puts ROOT on the last block (single-leaf inputs). */
static void b3_leaf(const uint32_t key[8], uint32_t base_flags, const uint8_t* p, size_t len,
                    uint64_t leaf_index, int root, uint32_t out[8]) {
    uint32_t cv[8], m[16];
    memcpy(cv, key, sizeof cv);
    size_t nblocks = len == 0 ? 1 : (len + 63) / 64;
    for (size_t b = 0; b < nblocks; b++) {
        size_t off = b * 64, bl = len - off < 64 ? len - off : 64;
        uint8_t buf[64];
        const uint8_t* src = p + off;
        if (bl < 64) {
            memset(buf, 0, 64);
            if (bl) memcpy(buf, p + off, bl);
            src = buf;
        }
        for (int i = 0; i < 16; i++) m[i] = ld32(src + 4 * i);
        uint32_t fl = base_flags;
        if (b == 0) fl |= B3_CHUNK_START;
        if (b == nblocks - 1) fl |= B3_CHUNK_END | (root ? B3_ROOT : 0);
        b3_compress(cv, m, leaf_index, (uint32_t)bl, fl, cv);
    }
    memcpy(out, cv, sizeof cv);
}

/* Eight full 1 KiB leaves at once (one leaf per SIMD lane, GCC vector extensions -> AVX2): the same compression,
 * written over vectors.  This is only here so that the timed CPU arm is a fair stand-in for the reference's
 * std.crypto BLAKE3 (which is SIMD too); results are identical to b3_leaf and the tests compare both paths. */
typedef uint32_t v8u __attribute__((vector_size(32)));
#define V8_ROTR(x, n) (((x) >> (n)) | ((x) << (32 - (n))))
#define V8_G(a, b, c, d, x, y)             \
    do {                                   \
        v[a] = v[a] + v[b] + (x);          \
        v[d] = V8_ROTR(v[d] ^ v[a], 16);   \
        v[c] = v[c] + v[d];                \
        v[b] = V8_ROTR(v[b] ^ v[c], 12);   \
        v[a] = v[a] + v[b] + (y);          \
        v[d] = V8_ROTR(v[d] ^ v[a], 8);    \
        v[c] = v[c] + v[d];                \
        v[b] = V8_ROTR(v[b] ^ v[c], 7);    \
    } while (0)

static void b3_leaves8(const uint32_t key[8], uint32_t base_flags, const uint8_t* p, uint64_t first_leaf,
                       uint32_t out[8][8]) {
    v8u cv[8];
    for (int i = 0; i < 8; i++) cv[i] = (v8u){key[i], key[i], key[i], key[i], key[i], key[i], key[i], key[i]};
    v8u ctr_lo, ctr_hi;
    for (int j = 0; j < 8; j++) { ctr_lo[j] = (uint32_t)(first_leaf + j); ctr_hi[j] = (uint32_t)((first_leaf + j) >> 32); }
    for (int b = 0; b < 16; b++) {
        v8u m[16], t[16], v[16];
        for (int i = 0; i < 16; i++)
            for (int j = 0; j < 8; j++) m[i][j] = ld32(p + 1024 * j + 64 * b + 4 * i);
        uint32_t fl = base_flags | (b == 0 ? B3_CHUNK_START : 0) | (b == 15 ? B3_CHUNK_END : 0);
        for (int i = 0; i < 8; i++) v[i] = cv[i];
        for (int i = 0; i < 4; i++) v[8 + i] = (v8u){B3_IV[i], B3_IV[i], B3_IV[i], B3_IV[i], B3_IV[i], B3_IV[i], B3_IV[i], B3_IV[i]};
        v[12] = ctr_lo; v[13] = ctr_hi;
        v[14] = (v8u){64, 64, 64, 64, 64, 64, 64, 64};
        v[15] = (v8u){fl, fl, fl, fl, fl, fl, fl, fl};
        for (int r = 0;; r++) {
            V8_G(0, 4, 8, 12, m[0], m[1]);
            V8_G(1, 5, 9, 13, m[2], m[3]);
            V8_G(2, 6, 10, 14, m[4], m[5]);
            V8_G(3, 7, 11, 15, m[6], m[7]);
            V8_G(0, 5, 10, 15, m[8], m[9]);
            V8_G(1, 6, 11, 12, m[10], m[11]);
            V8_G(2, 7, 8, 13, m[12], m[13]);
            V8_G(3, 4, 9, 14, m[14], m[15]);
            if (r == 6) break;
            for (int i = 0; i < 16; i++) t[i] = m[B3_PERM[i]];
            for (int i = 0; i < 16; i++) m[i] = t[i];
        }
        for (int i = 0; i < 8; i++) cv[i] = v[i] ^ v[i + 8];
    }
    for (int j = 0; j < 8; j++)
        for (int i = 0; i < 8; i++) out[j][i] = cv[i][j];
}

/* Full BLAKE3 of `len` bytes under `key` words and mode flags (0 = plain hash with IV as key).
 * The tree is reduced level by level, pairing neighbours and promoting an odd tail unchanged; SURVEY.md
 * A.2 records that this equals BLAKE3's left-full tree (checked against PyPI blake3 in tests). */
static void b3_hash(const uint32_t key[8], uint32_t base_flags, const uint8_t* p, size_t len,
                    uint8_t out[32]) {
    size_t nleaves = len == 0 ? 1 : (len + 1023) / 1024;
    uint32_t(*cvs)[8] = malloc(nleaves * sizeof *cvs);
    size_t i0 = 0;
    for (; i0 + 8 <= nleaves && (i0 + 8) * 1024 <= len && nleaves > 8; i0 += 8)  /* eight full leaves per pass */
        b3_leaves8(key, base_flags, p + i0 * 1024, i0, (uint32_t(*)[8])cvs[i0]);
    for (size_t i = i0; i < nleaves; i++) {
        size_t off = i * 1024, l = len - off < 1024 ? len - off : 1024;
        b3_leaf(key, base_flags, p + off, l, i, nleaves == 1, cvs[i]);
    }
    size_t n = nleaves;
    while (n > 1) {
        size_t w = 0;
        for (size_t i = 0; i + 1 < n; i += 2) {
            uint32_t m[16];
            memcpy(m, cvs[i], 32);
            memcpy(m + 8, cvs[i + 1], 32);
            b3_compress(key, m, 0, 64, base_flags | B3_PARENT | (n == 2 ? B3_ROOT : 0), cvs[w++]);
        }
        if (n & 1) memcpy(cvs[w++], cvs[n - 1], 32);
        n = w;
    }
    for (int i = 0; i < 8; i++) st32(out + 4 * i, cvs[0][i]);
    free(cvs);
}

static void key_words(const uint8_t key[32], uint32_t w[8]) {
    for (int i = 0; i < 8; i++) w[i] = ld32(key + 4 * i);
}

/* Xet chunk hash = BLAKE3.keyed(DATA_KEY, bytes)  (A.2). */
OZ_API void oz_chunk_hash(const uint8_t* data, size_t len, uint8_t out[32]) {
    uint32_t k[8];
    key_words(XET_DATA_KEY, k);
    b3_hash(k, B3_KEYED, data, len, out);
}
/* Generic keyed BLAKE3 (used for the node key and the zero-key file hash). */
OZ_API void oz_blake3_keyed(const uint8_t key[32], const uint8_t* data, size_t len, uint8_t out[32]) {
    uint32_t k[8];
    key_words(key, k);
    b3_hash(k, B3_KEYED, data, len, out);
}
/* Plain BLAKE3 -- what src/bench.zig:207-223 times (std.crypto.hash.Blake3.hash(data, &out, .{})). */
OZ_API void oz_blake3(const uint8_t* data, size_t len, uint8_t out[32]) { b3_hash(B3_IV, 0, data, len, out); }

/* Number of compressions the hash of one `len`-byte input costs. */
OZ_API uint64_t oz_blake3_compressions(uint64_t len) {
    uint64_t leaves = len == 0 ? 1 : (len + 1023) / 1024, c = 0;
    for (uint64_t i = 0; i < leaves; i++) {
        uint64_t l = len - i * 1024 < 1024 ? len - i * 1024 : 1024;
        c += l == 0 ? 1 : (l + 63) / 64;
    }
    return c + leaves - 1;
}

/* ------------------------------------------------------------------------------------------------ */
/* A.1  hash <-> hex (xet.hashing.hashToHex: four little-endian u64, each printed %016x)            */
/* ------------------------------------------------------------------------------------------------ */

OZ_API void oz_hash_to_hex(const uint8_t h[32], char out[65]) {
    static const char* d = "0123456789abcdef";
    for (int g = 0; g < 4; g++)
        for (int b = 0; b < 8; b++) {
            uint8_t v = h[g * 8 + (7 - b)];
            out[g * 16 + 2 * b] = d[v >> 4];
            out[g * 16 + 2 * b + 1] = d[v & 15];
        }
    out[64] = 0;
}

/* ------------------------------------------------------------------------------------------------ */
/* A.3  Merkle fold                                                                                 */
/* ------------------------------------------------------------------------------------------------ */

static size_t merkle_next_cut(const uint8_t (*h)[32], size_t n) {
    if (n <= 2) return n;
    size_t end = n < 9 ? n : 9;
    for (size_t i = 2; i < end; i++) {
        uint64_t last = 0;
        for (int b = 7; b >= 0; b--) last = (last << 8) | h[i][24 + b];
        if (last % 4 == 0) return i + 1;
    }
    return end;
}

/* hashes[n][32], lens[n] -> root[32].  n==0 -> zeros, n==1 -> hashes[0]. */
OZ_API void oz_merkle_root(const uint8_t* hashes, const uint64_t* lens, size_t n, uint8_t root[32]) {
    if (n == 0) { memset(root, 0, 32); return; }
    uint8_t(*h)[32] = malloc(n * 32);
    uint64_t* l = malloc(n * sizeof *l);
    memcpy(h, hashes, n * 32);
    memcpy(l, lens, n * sizeof *l);
    char buf[9 * 96];
    while (n > 1) {
        size_t w = 0, r = 0;
        while (r != n) {
            size_t cut = r + merkle_next_cut((const uint8_t(*)[32])(h + r), n - r);
            size_t pos = 0;
            uint64_t total = 0;
            for (size_t i = r; i < cut; i++) {
                char hex[65];
                oz_hash_to_hex(h[i], hex);
                pos += (size_t)sprintf(buf + pos, "%s : %llu\n", hex, (unsigned long long)l[i]);
                total += l[i];
            }
            uint8_t out[32];
            oz_blake3_keyed(XET_NODE_KEY, (const uint8_t*)buf, pos, out);
            memcpy(h[w], out, 32);
            l[w] = total;
            w++;
            r = cut;
        }
        n = w;
    }
    memcpy(root, h[0], 32);
    free(h);
    free(l);
}

/* File hash = BLAKE3.keyed(32 zero bytes, merkle_root(all chunks)); empty file -> zeros (A.3). */
OZ_API void oz_file_hash(const uint8_t* hashes, const uint64_t* lens, size_t n, uint8_t out[32]) {
    if (n == 0) { memset(out, 0, 32); return; }
    uint8_t root[32], zero[32] = {0};
    oz_merkle_root(hashes, lens, n, root);
    oz_blake3_keyed(zero, root, 32, out);
}

/* Term verification ("range") hash: BLAKE3.keyed(VERIFICATION_KEY, h[s] || h[s+1] || ... || h[e-1]) over the chunk
 * hashes of one reconstruction term.  This is xet-core's mdb_shard chunk_verification
 * (range_hash_from_chunks); the key and the construction were confirmed here against the FileVerificationEntry
 * rows hf_xet 1.5.0 writes into its .mdb shards (tests/golden/make_golden.py: shard_range_hashes). */
static const uint8_t XET_VERIFICATION_KEY[32] = {127, 24,  87,  214, 206, 86,  237, 102, 18,  127, 249,
                                                 19,  231, 165, 195, 243, 164, 205, 38,  213, 181, 219,
                                                 73,  230, 65,  36,  152, 127, 40,  251, 148, 195};
/* bounds[0..n_terms] are chunk indices; term t covers hashes[bounds[t] .. bounds[t+1]). */
OZ_API void oz_range_hashes(const uint8_t* hashes, const uint32_t* bounds, size_t n_terms, uint8_t* out) {
    for (size_t t = 0; t < n_terms; t++)
        oz_blake3_keyed(XET_VERIFICATION_KEY, hashes + 32 * (size_t)bounds[t], 32 * (size_t)(bounds[t + 1] - bounds[t]), out + 32 * t);
}

/* ------------------------------------------------------------------------------------------------ */
/* A.4b  LZ4 frame decode                                                                           */
/* ------------------------------------------------------------------------------------------------ */

static uint32_t xxh32(const uint8_t* p, size_t len, uint32_t seed) {
    const uint32_t P1 = 2654435761u, P2 = 2246822519u, P3 = 3266489917u, P4 = 668265263u, P5 = 374761393u;
    const uint8_t* end = p + len;
    uint32_t h;
    if (len >= 16) {
        uint32_t v1 = seed + P1 + P2, v2 = seed + P2, v3 = seed, v4 = seed - P1;
        do {
            v1 = rotr32(v1 + ld32(p) * P2, 32 - 13) * P1;
            v2 = rotr32(v2 + ld32(p + 4) * P2, 32 - 13) * P1;
            v3 = rotr32(v3 + ld32(p + 8) * P2, 32 - 13) * P1;
            v4 = rotr32(v4 + ld32(p + 12) * P2, 32 - 13) * P1;
            p += 16;
        } while (p + 16 <= end);
        h = rotr32(v1, 31) + rotr32(v2, 25) + rotr32(v3, 20) + rotr32(v4, 14);
    } else
        h = seed + P5;
    h += (uint32_t)len;
    while (p + 4 <= end) { h = rotr32(h + ld32(p) * P3, 32 - 17) * P4; p += 4; }
    while (p < end) { h = rotr32(h + *p * P5, 32 - 11) * P1; p++; }
    h ^= h >> 15; h *= P2; h ^= h >> 13; h *= P3; h ^= h >> 16;
    return h;
}
OZ_API uint32_t oz_xxh32(const uint8_t* p, size_t len, uint32_t seed) { return xxh32(p, len, seed); }

/* The timed CPU arm may hand independent blocks to liblz4's LZ4_decompress_safe (loaded by lz4_load below) instead of
 * the plain decoder in this file: zest's CPU path uses an optimised LZ4 decoder too, so this is the fairer number.
 * Off by default (the oracle proper is the code below); bench.py switches it on and reports which decoder ran.  A
 * block liblz4 rejects is decoded again by lz4_block, so error codes do not depend on the switch. */
typedef int (*lz4_decompress_fn)(const char*, char*, int, int);
static lz4_decompress_fn p_decompress;
static int g_use_liblz4;
static pthread_once_t lz4_once = PTHREAD_ONCE_INIT;
static void lz4_load_once(void);
OZ_API int oz_set_lz4_decoder(int use_liblz4) {
    lz4_load_once();
    g_use_liblz4 = use_liblz4 && p_decompress != NULL;
    return g_use_liblz4;
}

/* One LZ4 block: src[0..slen) -> dst[dpos..), matches may reach back to dst[window_lo..).  Returns the
 * new dpos or a negative error.  Byte-serial match copy (overlap semantics of A.4b). */
static int64_t lz4_block(const uint8_t* src, size_t slen, uint8_t* dst, size_t dpos, size_t dcap,
                         size_t window_lo) {
    size_t ip = 0;
    if (slen == 0) return OZ_ERR_LZ4_BLOCK;
    for (;;) {
        if (ip >= slen) return OZ_ERR_LZ4_BLOCK;
        uint32_t tok = src[ip++];
        size_t lit = tok >> 4;
        if (lit == 15) {
            uint8_t b;
            do {
                if (ip >= slen) return OZ_ERR_LZ4_BLOCK;
                b = src[ip++];
                lit += b;
            } while (b == 255);
        }
        if (lit > slen - ip) return OZ_ERR_LZ4_BLOCK;
        if (lit > dcap - dpos) return OZ_ERR_SIZE_MISMATCH;
        memcpy(dst + dpos, src + ip, lit);
        ip += lit;
        dpos += lit;
        if (ip == slen) return (int64_t)dpos; /* last sequence: literals only */
        if (slen - ip < 2) return OZ_ERR_LZ4_BLOCK;
        size_t off = (size_t)src[ip] | ((size_t)src[ip + 1] << 8);
        ip += 2;
        size_t ml = (tok & 15);
        if (ml == 15) {
            uint8_t b;
            do {
                if (ip >= slen) return OZ_ERR_LZ4_BLOCK;
                b = src[ip++];
                ml += b;
            } while (b == 255);
        }
        ml += 4;
        if (off == 0 || off > dpos - window_lo) return OZ_ERR_LZ4_BLOCK;
        if (ml > dcap - dpos) return OZ_ERR_SIZE_MISMATCH;
        if (off >= ml) memcpy(dst + dpos, dst + dpos - off, ml); /* disjoint: plain copy */
        else if (off >= 8) {                                      /* overlapping, period >= 8: 8-byte steps are exact */
            size_t i = 0;
            for (; i + 8 <= ml; i += 8) memcpy(dst + dpos + i, dst + dpos + i - off, 8);
            for (; i < ml; i++) dst[dpos + i] = dst[dpos + i - off];
        } else
            for (size_t i = 0; i < ml; i++) dst[dpos + i] = dst[dpos + i - off]; /* byte-serial overlap semantics */
        dpos += ml;
    }
}

/* LZ4 frame -> exactly `ulen` bytes in dst.  Accepts multi-block, linked or independent blocks, stored
 * blocks, and frames carrying block/content checksums or a content-size field (checksums are skipped,
 * not verified: the chunk's BLAKE3 hash is the integrity check).  Rejects dictionaries, skippable frames
 * and bad header checksums. */
static int lz4_frame(const uint8_t* src, size_t slen, uint8_t* dst, size_t ulen) {
    if (slen < 7 || ld32(src) != 0x184D2204u) return OZ_ERR_LZ4_FRAME;
    uint8_t flg = src[4], bd = src[5];
    if ((flg >> 6) != 1 || (flg & 0x02) || (bd & 0x8F)) return OZ_ERR_LZ4_FRAME;
    int indep = (flg >> 5) & 1, bsum = (flg >> 4) & 1, csize = (flg >> 3) & 1, csum = (flg >> 2) & 1,
        dict = flg & 1;
    if (dict) return OZ_ERR_LZ4_FRAME;
    unsigned bcode = (bd >> 4) & 7;
    if (bcode < 4) return OZ_ERR_LZ4_FRAME;
    size_t bmax = (size_t)1 << (8 + 2 * bcode);
    size_t ip = 6 + (csize ? 8 : 0);
    if (slen < ip + 1) return OZ_ERR_LZ4_FRAME;
    if (((xxh32(src + 4, ip - 4, 0) >> 8) & 0xFF) != src[ip]) return OZ_ERR_LZ4_FRAME;
    if (csize) {
        uint64_t cs = 0;
        for (int b = 7; b >= 0; b--) cs = (cs << 8) | src[6 + b];
        if (cs != ulen) return OZ_ERR_SIZE_MISMATCH;
    }
    ip++;
    size_t dpos = 0;
    for (;;) {
        if (slen - ip < 4) return OZ_ERR_LZ4_FRAME;
        uint32_t bs = ld32(src + ip);
        ip += 4;
        if (bs == 0) break; /* EndMark */
        int raw = bs >> 31;
        bs &= 0x7FFFFFFFu;
        if (bs > bmax || bs > slen - ip) return OZ_ERR_LZ4_FRAME;
        if (raw) {
            if (bs > ulen - dpos) return OZ_ERR_SIZE_MISMATCH;
            memcpy(dst + dpos, src + ip, bs);
            dpos += bs;
        } else {
            size_t lo = indep ? dpos : 0; /* linked blocks may reach back (offsets are <= 65535 anyway) */
            int64_t r = -1;
            if (g_use_liblz4 && indep && ulen - dpos <= 0x7E000000u) {
                int d = p_decompress((const char*)src + ip, (char*)dst + dpos, (int)bs, (int)(ulen - dpos));
                if (d >= 0) r = (int64_t)dpos + d;
            }
            if (r < 0) r = lz4_block(src + ip, bs, dst, dpos, ulen, lo);
            if (r < 0) return (int)r;
            if ((size_t)r - dpos > bmax) return OZ_ERR_LZ4_FRAME;
            dpos = (size_t)r;
        }
        ip += bs;
        if (bsum) {
            if (slen - ip < 4) return OZ_ERR_LZ4_FRAME;
            ip += 4;
        }
    }
    if (csum) {
        if (slen - ip < 4) return OZ_ERR_LZ4_FRAME;
        ip += 4;
    }
    if (dpos != ulen) return OZ_ERR_SIZE_MISMATCH;
    return OZ_OK;
}

/* ------------------------------------------------------------------------------------------------ */
/* A.4  xorb chunk stream                                                                           */
/* ------------------------------------------------------------------------------------------------ */

/* Walk the [8-byte header][payload] stream.  cap = capacity of the output arrays (comp_off needs
 * cap entries: offset of each chunk's HEADER).  Returns n_chunks or a negative error.  Arrays may be
 * NULL to count only. */
OZ_API int64_t oz_xorb_index(const uint8_t* blob, size_t len, size_t cap, uint64_t* hdr_off,
                             uint32_t* clen, uint32_t* ulen, uint8_t* scheme, uint64_t* total_ulen) {
    size_t pos = 0, n = 0;
    uint64_t tot = 0;
    while (pos < len) {
        if (len - pos < 8) return OZ_ERR_TRUNCATED;
        const uint8_t* h = blob + pos;
        if (h[0] != 0) return OZ_ERR_BAD_HEADER;
        uint32_t c = h[1] | (h[2] << 8) | ((uint32_t)h[3] << 16);
        uint32_t u = h[5] | (h[6] << 8) | ((uint32_t)h[7] << 16);
        uint8_t s = h[4];
        if (s > 2) return OZ_ERR_SCHEME;
        if (s == 0 && c != u) return OZ_ERR_BAD_HEADER;
        if (c > len - pos - 8) return OZ_ERR_TRUNCATED;
        if (n < cap) {
            if (hdr_off) hdr_off[n] = pos;
            if (clen) clen[n] = c;
            if (ulen) ulen[n] = u;
            if (scheme) scheme[n] = s;
        } else if (hdr_off || clen || ulen || scheme)
            return OZ_ERR_CAPACITY;
        tot += u;
        pos += 8 + (size_t)c;
        n++;
    }
    if (total_ulen) *total_ulen = tot;
    return (int64_t)n;
}

/* Decode one chunk payload into out[0..ulen). */
OZ_API int oz_decode_chunk(const uint8_t* payload, uint32_t clen, uint8_t scheme, uint8_t* out,
                           uint32_t ulen) {
    if (scheme == 0) {
        if (clen != ulen) return OZ_ERR_BAD_HEADER;
        memcpy(out, payload, ulen);
        return OZ_OK;
    }
    if (scheme == 1) return lz4_frame(payload, clen, out, ulen);
    if (scheme == 2) {
        uint8_t* tmp = malloc(ulen ? ulen : 1);
        int r = lz4_frame(payload, clen, tmp, ulen);
        if (r == OZ_OK) {
            size_t base = 0;
            for (unsigned k = 0; k < 4; k++) {
                size_t pl = ulen / 4 + (k < ulen % 4);
                for (size_t j = 0; j < pl; j++) out[4 * j + k] = tmp[base + j];
                base += pl;
            }
        }
        free(tmp);
        return r;
    }
    return OZ_ERR_SCHEME;
}

/* XorbReader.extractChunkRange(start, end): concatenated decoded bytes of chunks [start, end).
 * Returns bytes written or a negative error. */
OZ_API int64_t oz_extract_chunk_range(const uint8_t* blob, size_t len, uint32_t start, uint32_t end,
                                      uint8_t* out, size_t out_cap) {
    size_t pos = 0, idx = 0, w = 0;
    if (start > end) return OZ_ERR_RANGE;
    while (idx < end) {
        if (pos >= len) return OZ_ERR_RANGE;
        if (len - pos < 8) return OZ_ERR_TRUNCATED;
        const uint8_t* h = blob + pos;
        if (h[0] != 0) return OZ_ERR_BAD_HEADER;
        uint32_t c = h[1] | (h[2] << 8) | ((uint32_t)h[3] << 16);
        uint32_t u = h[5] | (h[6] << 8) | ((uint32_t)h[7] << 16);
        if (h[4] > 2) return OZ_ERR_SCHEME;
        if (c > len - pos - 8) return OZ_ERR_TRUNCATED;
        if (idx >= start) {
            if (u > out_cap - w) return OZ_ERR_CAPACITY;
            int r = oz_decode_chunk(h + 8, c, h[4], out + w, u);
            if (r) return r;
            w += u;
        }
        pos += 8 + (size_t)c;
        idx++;
    }
    return (int64_t)w;
}

/* ------------------------------------------------------------------------------------------------ */
/* A.5  Gearhash CDC                                                                                */
/* ------------------------------------------------------------------------------------------------ */

static const uint64_t GEAR[256] = ZEST_GEAR_TABLE_INIT;
#define CDC_MIN 8192u
#define CDC_MAX 131072u
#define CDC_MASK 0xFFFF000000000000ULL

/* Chunk END offsets (exclusive, cumulative) of data[0..len).  Returns the count or OZ_ERR_CAPACITY. */
OZ_API int64_t oz_cdc(const uint8_t* data, size_t len, uint64_t* ends, size_t cap) {
    size_t n = 0, start = 0;
    while (start < len) {
        uint64_t h = 0;
        size_t i = start, lim = len - start < CDC_MAX ? len : start + CDC_MAX, cut = lim;
        for (; i < lim; i++) {
            h = (h << 1) + GEAR[data[i]];
            if (i + 1 - start >= CDC_MIN && (h & CDC_MASK) == 0) { cut = i + 1; break; }
        }
        if (n >= cap) return OZ_ERR_CAPACITY;
        ends[n++] = cut;
        start = cut;
    }
    return (int64_t)n;
}

/* ------------------------------------------------------------------------------------------------ */
/* Multi-threaded drivers (the timed CPU baseline; SURVEY.md 8d / BASELINE.md rows C1-C3)           */
/* ------------------------------------------------------------------------------------------------ */

typedef struct {
    /* shared work list: one entry per chunk */
    size_t n;
    const uint8_t* const* src;  /* payload (verify) or data (hash) per chunk   */
    const uint32_t* clen;
    const uint32_t* ulen;
    const uint8_t* scheme;      /* NULL: hash only (src is plain data)         */
    uint8_t* hashes;            /* n x 32                                      */
    int* status;                /* per chunk                                   */
    size_t next;
    pthread_mutex_t mu;
} work_t;

static void* work_main(void* arg) {
    work_t* w = arg;
    uint8_t* tmp = malloc(1 << 24);
    for (;;) {
        pthread_mutex_lock(&w->mu);
        size_t lo = w->next, hi = lo + 16 < w->n ? lo + 16 : w->n;
        w->next = hi;
        pthread_mutex_unlock(&w->mu);
        if (lo >= w->n) break;
        for (size_t i = lo; i < hi; i++) {
            if (w->scheme) {
                int r = oz_decode_chunk(w->src[i], w->clen[i], w->scheme[i], tmp, w->ulen[i]);
                w->status[i] = r;
                if (r == OZ_OK) oz_chunk_hash(tmp, w->ulen[i], w->hashes + 32 * i);
            } else {
                oz_chunk_hash(w->src[i], w->ulen[i], w->hashes + 32 * i);
                w->status[i] = OZ_OK;
            }
        }
    }
    free(tmp);
    return NULL;
}

static void run_work(work_t* w, int threads) {
    if (threads < 1) threads = 1;
    pthread_mutex_init(&w->mu, NULL);
    w->next = 0;
    pthread_t* t = malloc(sizeof(pthread_t) * (size_t)threads);
    for (int i = 1; i < threads; i++) pthread_create(&t[i], NULL, work_main, w);
    work_main(w);
    for (int i = 1; i < threads; i++) pthread_join(t[i], NULL);
    free(t);
    pthread_mutex_destroy(&w->mu);
}

/* Keyed chunk hash of n chunks data[off[i]..off[i+1]) on `threads` threads. */
OZ_API void oz_hash_chunks_mt(const uint8_t* data, const uint64_t* off, size_t n, uint8_t* hashes,
                              int threads) {
    const uint8_t** src = malloc(sizeof(*src) * (n ? n : 1));
    uint32_t* ul = malloc(sizeof(*ul) * (n ? n : 1));
    int* st = malloc(sizeof(*st) * (n ? n : 1));
    for (size_t i = 0; i < n; i++) { src[i] = data + off[i]; ul[i] = (uint32_t)(off[i + 1] - off[i]); }
    work_t w = {.n = n, .src = src, .ulen = ul, .hashes = hashes, .status = st};
    run_work(&w, threads);
    free(src); free(ul); free(st);
}

/* Full verify of a batch of xorbs: walk, decode every chunk, hash it, fold, compare with expected.
 * verdict[x] = 1 match, 0 mismatch, <0 error code.  roots: n_xorbs x 32.  chunk_hashes (optional) must
 * hold the total chunk count x 32.  Returns total uncompressed bytes verified, or a negative error if
 * a blob cannot even be walked (the verdict array says which). */
OZ_API int64_t oz_verify_xorbs_mt(const uint8_t* const* blobs, const size_t* lens, const uint8_t* expected,
                                  size_t n_xorbs, int8_t* verdict, uint8_t* roots, uint8_t* chunk_hashes,
                                  int threads) {
    size_t total = 0;
    size_t* first = malloc(sizeof(size_t) * (n_xorbs + 1));
    for (size_t x = 0; x < n_xorbs; x++) {
        int64_t c = oz_xorb_index(blobs[x], lens[x], 0, NULL, NULL, NULL, NULL, NULL);
        first[x] = total;
        verdict[x] = c < 0 ? (int8_t)c : 0;
        total += c < 0 ? 0 : (size_t)c;
    }
    first[n_xorbs] = total;
    size_t alloc = total ? total : 1;
    const uint8_t** src = malloc(sizeof(*src) * alloc);
    uint64_t* ho = malloc(sizeof(uint64_t) * alloc);
    uint32_t* cl = malloc(4 * alloc);
    uint32_t* ul = malloc(4 * alloc);
    uint8_t* sc = malloc(alloc);
    int* st = malloc(sizeof(int) * alloc);
    uint8_t* hs = chunk_hashes ? chunk_hashes : malloc(32 * alloc);
    for (size_t x = 0; x < n_xorbs; x++) {
        size_t c = first[x + 1] - first[x];
        if (!c) continue;
        oz_xorb_index(blobs[x], lens[x], c, ho + first[x], cl + first[x], ul + first[x], sc + first[x], NULL);
        for (size_t i = 0; i < c; i++) src[first[x] + i] = blobs[x] + ho[first[x] + i] + 8;
    }
    work_t w = {.n = total, .src = src, .clen = cl, .ulen = ul, .scheme = sc, .hashes = hs, .status = st};
    run_work(&w, threads);
    int64_t bytes = 0;
    for (size_t x = 0; x < n_xorbs; x++) {
        if (verdict[x] < 0) continue;
        size_t c = first[x + 1] - first[x];
        int bad = 0;
        uint64_t* l64 = malloc(sizeof(uint64_t) * (c ? c : 1));
        for (size_t i = 0; i < c; i++) {
            if (st[first[x] + i]) { bad = st[first[x] + i]; break; }
            l64[i] = ul[first[x] + i];
            bytes += ul[first[x] + i];
        }
        if (bad) verdict[x] = (int8_t)bad;
        else {
            oz_merkle_root(hs + 32 * first[x], l64, c, roots + 32 * x);
            verdict[x] = expected ? (memcmp(roots + 32 * x, expected + 32 * x, 32) == 0) : 1;
        }
        free(l64);
    }
    if (!chunk_hashes) free(hs);
    free(first); free(src); free(ho); free(cl); free(ul); free(sc); free(st);
    return bytes;
}

/* ------------------------------------------------------------------------------------------------ */
/* Synthetic xorb builder (input generator for tests and bench; not part of the verified path).     */
/* LZ4 block compression comes from the system liblz4.so.1 (dlopen; no headers on the image).       */
/* ------------------------------------------------------------------------------------------------ */

typedef int (*lz4_compress_fn)(const char*, char*, int, int);
typedef int (*lz4_bound_fn)(int);
typedef void* (*lz4_create_fn)(void);
typedef int (*lz4_free_fn)(void*);
typedef int (*lz4_continue_fn)(void*, const char*, char*, int, int, int);
static lz4_compress_fn p_compress;
static lz4_bound_fn p_bound;
static lz4_create_fn p_create;
static lz4_free_fn p_free;
static lz4_continue_fn p_continue;
static void lz4_load(void) {
    void* h = dlopen("liblz4.so.1", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return;
    p_compress = (lz4_compress_fn)dlsym(h, "LZ4_compress_default");
    p_bound = (lz4_bound_fn)dlsym(h, "LZ4_compressBound");
    p_create = (lz4_create_fn)dlsym(h, "LZ4_createStream");
    p_free = (lz4_free_fn)dlsym(h, "LZ4_freeStream");
    p_continue = (lz4_continue_fn)dlsym(h, "LZ4_compress_fast_continue");
    p_decompress = (lz4_decompress_fn)dlsym(h, "LZ4_decompress_safe");
}
static void lz4_load_once(void) { pthread_once(&lz4_once, lz4_load); }

/* Frame `ulen` bytes the way lz4_flex does for xet-core (A.4): FLG 0x60, BD 0x40/0x50, HC, blocks of at most
 * block_max bytes (0 = one block of the frame's maximum), EndMark.  `fflags` exercises the optional parts of the
 * frame format that xet-core does not emit but a decoder must accept (A.4b): 1 content-size field, 2 block
 * checksums, 4 content checksum, 8 linked blocks (matches may reach into earlier blocks).  Returns frame size. */
enum { FF_CSIZE = 1, FF_BSUM = 2, FF_CSUM = 4, FF_LINKED = 8 };
static size_t lz4_frame_encode(const uint8_t* src, size_t ulen, uint8_t* dst, size_t block_max, int fflags) {
    size_t o = 0;
    st32(dst, 0x184D2204u);
    dst[4] = (uint8_t)(0x40 | ((fflags & FF_LINKED) ? 0 : 0x20) | ((fflags & FF_BSUM) ? 0x10 : 0) |
                       ((fflags & FF_CSIZE) ? 0x08 : 0) | ((fflags & FF_CSUM) ? 0x04 : 0));
    size_t fmax = ulen > 65536 ? 262144 : 65536;
    dst[5] = ulen > 65536 ? 0x50 : 0x40;
    o = 6;
    if (fflags & FF_CSIZE) {
        for (int b = 0; b < 8; b++) dst[o + b] = (uint8_t)((uint64_t)ulen >> (8 * b));
        o += 8;
    }
    dst[o] = (uint8_t)(xxh32(dst + 4, o - 4, 0) >> 8);
    o++;
    if (block_max == 0 || block_max > fmax) block_max = fmax;
    void* stream = ((fflags & FF_LINKED) && p_create && p_continue) ? p_create() : NULL;
    for (size_t p = 0; p < ulen;) {
        size_t n = ulen - p < block_max ? ulen - p : block_max;
        int c = stream ? p_continue(stream, (const char*)src + p, (char*)dst + o + 4, (int)n, p_bound((int)n), 1)
                       : p_compress((const char*)src + p, (char*)dst + o + 4, (int)n, p_bound((int)n));
        size_t bl;
        if (c <= 0 || (size_t)c >= n) { /* stored block */
            st32(dst + o, (uint32_t)n | 0x80000000u);
            memcpy(dst + o + 4, src + p, n);
            bl = n;
        } else {
            st32(dst + o, (uint32_t)c);
            bl = (size_t)c;
        }
        o += 4 + bl;
        if (fflags & FF_BSUM) { st32(dst + o, xxh32(dst + o - bl, bl, 0)); o += 4; }
        p += n;
    }
    if (stream) p_free(stream);
    st32(dst + o, 0);
    o += 4;
    if (fflags & FF_CSUM) { st32(dst + o, xxh32(src, ulen, 0)); o += 4; }
    return o;
}

/* Serialize chunks data[ends[i-1]..ends[i]) into one xorb chunk stream.  policy: 0 none, 1 lz4, 2 bg4-lz4;
 * a chunk whose frame is not smaller than the raw bytes is stored with scheme 0 (what xet-core does).
 * block_max, fflags: see lz4_frame_encode (with fflags != 0 the framed form is kept even if it is not smaller).
 * out must hold n*8 + len + n*32 + len/200 bytes.  Returns blob size. */
OZ_API int64_t oz_build_xorb(const uint8_t* data, const uint64_t* ends, size_t n, int policy,
                             size_t block_max, int fflags, uint8_t* out, size_t out_cap) {
    pthread_once(&lz4_once, lz4_load);
    if (policy && (!p_compress || !p_bound)) return OZ_ERR_NOLZ4;
    size_t o = 0, start = 0;
    uint8_t* grp = malloc(1 << 24);
    uint8_t* frm = malloc((1 << 24) + (1 << 20));
    for (size_t i = 0; i < n; i++) {
        size_t ul = ends[i] - start;
        const uint8_t* src = data + start;
        size_t cl = ul;
        int scheme = 0;
        if (policy && ul > 0) {
            const uint8_t* in = src;
            if (policy == 2) {
                size_t base = 0;
                for (unsigned k = 0; k < 4; k++) {
                    size_t pl = ul / 4 + (k < ul % 4);
                    for (size_t j = 0; j < pl; j++) grp[base + j] = src[4 * j + k];
                    base += pl;
                }
                in = grp;
            }
            size_t f = lz4_frame_encode(in, ul, frm, block_max, fflags);
            if (f < ul || fflags) { cl = f; scheme = policy; }
        }
        if (o + 8 + cl > out_cap) { free(grp); free(frm); return OZ_ERR_CAPACITY; }
        uint8_t* h = out + o;
        h[0] = 0;
        h[1] = (uint8_t)cl; h[2] = (uint8_t)(cl >> 8); h[3] = (uint8_t)(cl >> 16);
        h[4] = (uint8_t)scheme;
        h[5] = (uint8_t)ul; h[6] = (uint8_t)(ul >> 8); h[7] = (uint8_t)(ul >> 16);
        memcpy(h + 8, scheme ? frm : src, cl);
        o += 8 + cl;
        start = ends[i];
    }
    free(grp);
    free(frm);
    return (int64_t)o;
}


/* ------------------------------------------------------------------------------------------------ */
/* Fast synthetic bf16 bytes for the large bench workloads (input generator only).                  */
/* ------------------------------------------------------------------------------------------------ */

/* Acklam's rational approximation of the inverse normal CDF (relative error < 1.2e-9). */
static double inv_norm_cdf(double p) {
    static const double a[6] = {-3.969683028665376e+01, 2.209460984245205e+02, -2.759285104469687e+02,
                                1.383577518672690e+02, -3.066479806614716e+01, 2.506628277459239e+00};
    static const double b[5] = {-5.447609879822406e+01, 1.615858368580409e+02, -1.556989798598866e+02,
                                6.680131188771972e+01, -1.328068155288572e+01};
    static const double c[6] = {-7.784894002430293e-03, -3.223964580411365e-01, -2.400758277161838e+00,
                                -2.549732539343734e+00, 4.374664141464968e+00, 2.938163982698783e+00};
    static const double d[4] = {7.784695709041462e-03, 3.224671290700398e-01, 2.445134137142996e+00,
                                3.754408661907416e+00};
    double q, r;
    if (p < 0.02425) {
        q = __builtin_sqrt(-2 * __builtin_log(p));
        return (((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) /
               ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1);
    }
    if (p > 1 - 0.02425) return -inv_norm_cdf(1 - p);
    q = p - 0.5;
    r = q * q;
    return (((((a[0] * r + a[1]) * r + a[2]) * r + a[3]) * r + a[4]) * r + a[5]) * q /
           (((((b[0] * r + b[1]) * r + b[2]) * r + b[3]) * r + b[4]) * r + 1);
}

static inline uint64_t splitmix64(uint64_t* s) {
    uint64_t z = (*s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

/* nbytes of bf16 "weights": float32 sigma * Phi^-1(u), u one of 65,536 equally likely quantiles drawn by a
 * counter-based generator (splitmix64 keyed by seed), truncated to the top 16 bits.  Same distribution class as
 * synth_bf16 in xet_oracle.py (float32 N(0, sigma) >> 16; BG4-LZ4 ratio 0.86) at about 1 ns per
 * value, so that 16 GB - 140 GB sets can be generated in seconds per rank.  Any (seed) stream can be produced by
 * any thread: the big workloads give every xorb its own seed. */
OZ_API void oz_synth_bf16_q16(uint8_t* out, size_t nbytes, uint64_t seed, float sigma) {
    uint16_t* tab = malloc(65536 * sizeof(uint16_t));
    for (int i = 0; i < 65536; i++) {
        float v = (float)(sigma * inv_norm_cdf((i + 0.5) / 65536.0));
        uint32_t bits;
        memcpy(&bits, &v, 4);
        tab[i] = (uint16_t)(bits >> 16);
    }
    uint64_t st = seed * 0xD1342543DE82EF95ULL + 0x2545F4914F6CDD1DULL;
    size_t n = nbytes / 2, i = 0;
    uint16_t* o = (uint16_t*)out; /* callers pass 2-byte aligned buffers (numpy / malloc) */
    for (; i + 4 <= n; i += 4) {
        uint64_t r = splitmix64(&st);
        o[i] = tab[r & 0xffff]; o[i + 1] = tab[(r >> 16) & 0xffff];
        o[i + 2] = tab[(r >> 32) & 0xffff]; o[i + 3] = tab[r >> 48];
    }
    if (i < n || (nbytes & 1)) {
        uint64_t r = splitmix64(&st);
        for (; i < n; i++, r >>= 16) o[i] = tab[r & 0xffff];
        if (nbytes & 1) out[nbytes - 1] = (uint8_t)tab[r & 0xffff];
    }
    free(tab);
}
