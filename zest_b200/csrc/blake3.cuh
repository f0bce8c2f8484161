// BLAKE3 compression function for sm_100a, one compression per thread, everything in registers.
//
// Replaces, for the Xet chunk hash, what zest reaches through zig-xet's `hashing` module
// (call sites /root/reference/src/xet_bridge.zig:256-257, src/parallel_download.zig:65-66) and what
// src/bench.zig:207-223 times with std.crypto.hash.Blake3.
//
// Design notes (B200): the hash is integer-ALU bound (7 rounds x 8 G x 12 instructions + 8 feed-forward
// xors ~= 680 issued instructions per 64-byte block).  The message schedule is applied at compile time
// (all seven rounds unrolled, permuted indices as template constants) so no register moves are spent on
// it; rotations by 16 and 8 are byte permutes (PRMT), by 12 and 7 funnel shifts (SHF); a + b + m folds
// into one IADD3.
#pragma once
#include <cstdint>

namespace zg {

enum : uint32_t { B3_CHUNK_START = 1, B3_CHUNK_END = 2, B3_PARENT = 4, B3_ROOT = 8, B3_KEYED = 16 };

__device__ __constant__ const uint32_t kB3IV[8] = {0x6A09E667u, 0xBB67AE85u, 0x3C6EF372u, 0xA54FF53Au,
                                                   0x510E527Fu, 0x9B05688Cu, 0x1F83D9ABu, 0x5BE0CD19u};

__device__ __forceinline__ uint32_t rotr16(uint32_t x) { return __byte_perm(x, 0, 0x1032); }
__device__ __forceinline__ uint32_t rotr8(uint32_t x) { return __byte_perm(x, 0, 0x0321); }
__device__ __forceinline__ uint32_t rotr12(uint32_t x) { return __funnelshift_r(x, x, 12); }
__device__ __forceinline__ uint32_t rotr7(uint32_t x) { return __funnelshift_r(x, x, 7); }

// The integer ALU pipe (IADD3 / LOP3 / SHF / PRMT, 16 lanes/clk/SMSP) is the bound: ncu shows it ~82 % busy with the
// FMA pipe ~9 % busy when every add is an IADD3.  IMAD runs on the FMA pipe at the same rate, so the adds are written
// as x * one + y with `one` an opaque kernel parameter (ptxas folds a literal 1 back into IADD3).  a + b + m becomes
// (a + m) + b: the first IMAD is off the dependency chain (a is old by then), the second replaces the IADD3.
// Per G: 8 ALU-pipe + 6 FMA-pipe instructions instead of 10 + 2.
#define ZG_B3_ADD(x, y) (FMA_ADDS ? ((x) * one + (y)) : ((x) + (y)))
#define ZG_B3_G(a, b, c, d, mx, my)              \
    do {                                         \
        a = ZG_B3_ADD(ZG_B3_ADD(a, (mx)), b);    \
        d = rotr16(d ^ a);                       \
        c = ZG_B3_ADD(c, d);                     \
        b = rotr12(b ^ c);                       \
        a = ZG_B3_ADD(ZG_B3_ADD(a, (my)), b);    \
        d = rotr8(d ^ a);                        \
        c = ZG_B3_ADD(c, d);                     \
        b = rotr7(b ^ c);                        \
    } while (0)

// One round with message words addressed through the round's schedule s0..s15 (compile-time).
#define ZG_B3_ROUND(s0, s1, s2, s3, s4, s5, s6, s7, s8, s9, s10, s11, s12, s13, s14, s15) \
    ZG_B3_G(v0, v4, v8, v12, m[s0], m[s1]);                                                \
    ZG_B3_G(v1, v5, v9, v13, m[s2], m[s3]);                                                \
    ZG_B3_G(v2, v6, v10, v14, m[s4], m[s5]);                                               \
    ZG_B3_G(v3, v7, v11, v15, m[s6], m[s7]);                                               \
    ZG_B3_G(v0, v5, v10, v15, m[s8], m[s9]);                                               \
    ZG_B3_G(v1, v6, v11, v12, m[s10], m[s11]);                                             \
    ZG_B3_G(v2, v7, v8, v13, m[s12], m[s13]);                                              \
    ZG_B3_G(v3, v4, v9, v14, m[s14], m[s15]);

// cv <- compress(cv, m, counter, block_len, flags); only the 8-word chaining value is produced.
// `one` must hold 1 at run time (see above); FMA_ADDS = false gives the plain-add form.
template <bool FMA_ADDS = true>
__device__ __forceinline__ void b3_compress(uint32_t (&cv)[8], const uint32_t (&m)[16], uint32_t counter_lo,
                                            uint32_t counter_hi, uint32_t block_len, uint32_t flags, uint32_t one = 1) {
    uint32_t v0 = cv[0], v1 = cv[1], v2 = cv[2], v3 = cv[3], v4 = cv[4], v5 = cv[5], v6 = cv[6], v7 = cv[7];
    uint32_t v8 = 0x6A09E667u, v9 = 0xBB67AE85u, v10 = 0x3C6EF372u, v11 = 0xA54FF53Au;
    uint32_t v12 = counter_lo, v13 = counter_hi, v14 = block_len, v15 = flags;
    // schedules = successive applications of the BLAKE3 message permutation (2 6 3 10 7 0 4 13 1 11 12 5 9 14 15 8)
    ZG_B3_ROUND(0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15)
    ZG_B3_ROUND(2, 6, 3, 10, 7, 0, 4, 13, 1, 11, 12, 5, 9, 14, 15, 8)
    ZG_B3_ROUND(3, 4, 10, 12, 13, 2, 7, 14, 6, 5, 9, 0, 11, 15, 8, 1)
    ZG_B3_ROUND(10, 7, 12, 9, 14, 3, 13, 15, 4, 0, 11, 2, 5, 8, 1, 6)
    ZG_B3_ROUND(12, 13, 9, 11, 15, 10, 14, 8, 7, 2, 5, 3, 0, 1, 6, 4)
    ZG_B3_ROUND(9, 14, 11, 5, 8, 12, 15, 1, 13, 3, 0, 10, 2, 6, 4, 7)
    ZG_B3_ROUND(11, 15, 5, 0, 1, 9, 8, 6, 14, 10, 2, 12, 3, 4, 7, 13)
    cv[0] = v0 ^ v8;  cv[1] = v1 ^ v9;  cv[2] = v2 ^ v10; cv[3] = v3 ^ v11;
    cv[4] = v4 ^ v12; cv[5] = v5 ^ v13; cv[6] = v6 ^ v14; cv[7] = v7 ^ v15;
}

// Parent node: cv_out = compress(key, left || right, 0, 64, base | PARENT [| ROOT]).
__device__ __forceinline__ void b3_parent(uint32_t (&out)[8], const uint32_t (&key)[8], const uint32_t (&l)[8],
                                          const uint32_t (&r)[8], uint32_t flags) {
    uint32_t m[16];
#pragma unroll
    for (int i = 0; i < 8; i++) { m[i] = l[i]; m[8 + i] = r[i]; }
#pragma unroll
    for (int i = 0; i < 8; i++) out[i] = key[i];
    b3_compress<false>(out, m, 0, 0, 64, flags);
}

struct B3Key {
    uint32_t w[8];
    uint32_t base_flags;  // B3_KEYED for the Xet hashes, 0 for plain BLAKE3 (w = IV)
    uint32_t one;         // always 1: the opaque multiplier that keeps the adds on the FMA pipe (see ZG_B3_ADD)
};

}  // namespace zg
