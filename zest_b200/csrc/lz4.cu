// Xorb chunk decode on sm_100a: scheme 0 (stored), 1 (LZ4 frame), 2 (4-way byte-grouped LZ4 frame).
//
// Reference behaviour replaced: zig-xet XorbReader.extractChunkRange -- the per-chunk decompress loop that zest
// runs once per reconstruction term (/root/reference/src/xet_bridge.zig:256-257,
// src/parallel_download.zig:65-66).  Format: SURVEY.md A.4 / A.4b; oracle: oracle/xet_oracle.c (lz4_frame,
// lz4_block, oz_decode_chunk), whose checks are mirrored one for one and in the same order so that a malformed
// chunk gets the same error code on both sides.
//
// Mapping.  One warp decodes one chunk at a time (chunks are handed out longest-first through a ticket counter,
// 24 single-warp CTAs per SM).  LZ4's token stream is serial, and on real Xet data the sequences are tiny (text:
// ~5,900 per 64 KiB chunk, ~11 output bytes each; the exponent planes of bf16 tensors: ~14 bytes each), so a lane-serial
// walk leaves 31 lanes idle and costs ~100 warp instructions per sequence.  The walk is therefore done for EVERY byte
// position at once, in rounds over the next 32 x 68 stream bytes of the block:
//   tables  : the bytes are staged in shared memory (cp.async); lane k takes the 68 positions of segment k and works out,
//             for each of them, how long a sequence starting there would be (token, at most one length byte each,
//             literals, offset) -- the HOP table, one byte per position, 0 where no such sequence fits -- and, walking
//             its segment backwards, where the chain of hops from each position leaves the segment -- the EXIT table.
//             The same ~30 instructions per position on every lane, whatever the data; nothing speculative, nothing
//             thrown away, no dependent chain longer than two shared-memory reads.
//   compose : the true chain enters segment 0 at the known position; one exit-table look-up per segment carries it
//             across the round (32 dependent shared-memory reads, uniform across the warp), which gives every lane the
//             position at which the true chain enters its segment.
//   list    : each lane follows the hop table from that entry to the end of its segment (<= 23 hops), counting and
//             noting the token positions; a warp scan orders them, and the positions go to a list in stream order.
//   execute : 32 consecutive true sequences per step, one per lane, re-read from the staged bytes: output positions by
//             a warp scan of the lengths, the offset/size checks of the oracle on those positions, then literals (from
//             the staged window) and matches whose source precedes the step in one pass (all loads before all stores),
//             and the matches that read the step's own output in rounds, each as soon as the matches it reads from are
//             done (dependency masks by binary search over the lanes).  Runs longer than 16 bytes are copied by the
//             whole warp.
//   serial  : frame and block headers, the last sequence of a block, stored blocks, sequences longer than 255 stream
//             bytes (long literal runs -- the mantissa planes of byte-grouped bf16 are one 16 KiB run each -- and
//             matches of more than 273 bytes) and every error go through a plain one-sequence-at-a-time walker that
//             mirrors the oracle check for check; it runs uniformly on all lanes (same instructions, no broadcasts) and
//             feeds the same execute step.  A table round that stops early hands over to it; if rounds keep stopping
//             early (runs of zeros: every sequence is long) the walker's turns get longer.
// Round 1 walked the stream speculatively instead (lane k parsed from the first byte of segment k and chains were
// joined where they met): on bf16 planes 2.8 sequences were parsed for every true one and parse + join were 51 % of
// all instructions (profiles/r2_ncu_decode_bf16_baseline_by_region.txt); the tables cost the same on any data.
// Output goes to global memory; byte-group regrouping is done by the warp when a scheme-2 chunk completes, reading
// the planes back from L2.
//
// Roofline: HBM-bound in principle (compressed bytes in + decoded bytes out).  In practice the kernel is bound by
// instruction issue and by the latency of each chunk's chain; bench.py reports achieved GB/s over the algorithmic
// bytes, profiles/ holds the ncu summaries.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <vector>

#include "kernels.h"

namespace zg {

namespace {

constexpr uint32_t kFull = 0xffffffffu;
constexpr uint32_t kShort = 16;       // literal runs / matches up to this length are copied by their own lane
constexpr uint32_t kMaxPiece = 65535;
#ifndef ZG_NP
#define ZG_NP 1
#endif
#ifndef ZG_TABLE_UNROLL
#define ZG_TABLE_UNROLL 1
#endif
constexpr int kNP = ZG_NP;            // groups of 32 sequences per execute step
constexpr int kTableUnroll = ZG_TABLE_UNROLL;  // words of a segment per trip of the table loop
constexpr uint32_t kSerialCap = 16;   // sequences per round of the serial walker
constexpr uint32_t kSeg = 68;          // stream bytes per lane and round: 17 words, so 32 lanes start in 32 banks
constexpr uint32_t kSpan = 32 * kSeg;  // stream bytes parsed per round
constexpr uint32_t kHopMax = 255;      // longest sequence (token .. last length byte) the table holds; longer ones go serial
constexpr uint32_t kWinBytes = 2432;   // staged stream bytes: the span + the longest hop, rounded up to 16
constexpr uint32_t kSlots = 26;        // list slots per lane (a 68-byte segment holds at most 23 sequences); 13 words: no bank conflicts
constexpr uint32_t kStop = 255;        // exit-table mark: the chain from here ends inside the segment
constexpr uint32_t kNone = 0xffffffffu;
constexpr uint32_t kSpecMin = 32;      // blocks (or block remainders) shorter than this go to the serial walker

enum : int { ST_IDLE = 0, ST_START, ST_BLOCK_HDR, ST_IN_BLOCK, ST_AFTER_BLOCK, ST_FINISH, ST_EXHAUSTED };

struct WarpShared {
    uint4 win[kWinBytes / 16];   // staged stream window (window byte w = stream byte ip - a0 + w, a0 = alignment of ip)
    uint8_t nxt[kSpan];          // per window byte: stream length of the sequence whose token sits there (0: none);
                                 // after the chain walk: the true token positions in stream order (u16 each)
    uint8_t ex[kSpan];           // per window byte: where the chain from there leaves its lane's segment, relative to the
                                 // segment end (kStop: it ends inside); after compose: per-lane position slots (u16)
    uint32_t srec[kSerialCap * 3];  // serial walker records: literal source, lit | ml << 16, offset
};

__device__ __forceinline__ uint32_t rd8(const uint8_t* p) { return __ldg(p); }
__device__ __forceinline__ uint32_t rd32(const uint8_t* p) {
    return rd8(p) | (rd8(p + 1) << 8) | (rd8(p + 2) << 16) | (rd8(p + 3) << 24);
}
__device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

// xxh32 of fewer than 16 bytes (the LZ4 frame descriptor is 2..14 bytes), seed 0.
__device__ uint32_t xxh32_small(const uint8_t* p, uint32_t len) {
    const uint32_t P1 = 2654435761u, P2 = 2246822519u, P3 = 3266489917u, P4 = 668265263u, P5 = 374761393u;
    uint32_t h = P5 + len, i = 0;
    for (; i + 4 <= len; i += 4) h = rotl32(h + rd32(p + i) * P3, 17) * P4;
    for (; i < len; i++) h = rotl32(h + rd8(p + i) * P5, 11) * P1;
    h ^= h >> 15; h *= P2; h ^= h >> 13; h *= P3; h ^= h >> 16;
    return h;
}

// Eight stream bytes at any alignment as one u64 (three aligned word loads, issued together).  Touches the aligned
// words up to 11 bytes past p; `hi` is the end of what may be read (the compressed buffer rounded up to 16 bytes, which
// the ABI asks callers to keep readable): words at or beyond it read as zero, so nothing past `hi` is ever touched.
__device__ __forceinline__ uint64_t load_window(const uint8_t* p, const uint8_t* hi) {
    uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
    const uint32_t* h = reinterpret_cast<const uint32_t*>(hi);
    uint32_t sh = (uint32_t)(a & 3) * 8;
    uint32_t w0 = w < h ? __ldg(w) : 0u, w1 = w + 1 < h ? __ldg(w + 1) : 0u, w2 = w + 2 < h ? __ldg(w + 2) : 0u;
    return (uint64_t)__funnelshift_r(w0, w1, sh) | ((uint64_t)__funnelshift_r(w1, w2, sh) << 32);
}

// Up to 16 bytes from any alignment into four words: all loads are issued before anything is stored, so a lane
// pays one memory latency per copy instead of one per byte.  Only aligned words holding wanted bytes are read.
// Written as predicated PTX with the address formed once: left to itself the compiler re-derives the 64-bit
// address under every byte's predicate (~7 instructions per byte stored).
template <bool READ_ONLY, int K>
__device__ __forceinline__ uint32_t ld_word_if(unsigned long long w, uint32_t span) {  // word K of the run, if it holds wanted bytes
    uint32_t v;
    if (READ_ONLY)
        asm volatile("{ .reg .pred p; setp.gt.u32 p, %1, %2; mov.u32 %0, 0; @p ld.global.nc.u32 %0, [%3+%2]; }"
                     : "=r"(v) : "r"(span), "n"(4 * K), "l"(w) : "memory");
    else
        asm volatile("{ .reg .pred p; setp.gt.u32 p, %1, %2; mov.u32 %0, 0; @p ld.global.u32 %0, [%3+%2]; }"
                     : "=r"(v) : "r"(span), "n"(4 * K), "l"(w) : "memory");
    return v;
}
template <bool READ_ONLY>
__device__ __forceinline__ void load16(const uint8_t* s, uint32_t n, uint32_t (&x)[4]) {
#ifdef ZG_DIAG_FAST_LOADS  // timing experiment only (output is wrong): no global load, the address stands in for the data
    if (!READ_ONLY) {
        const uint32_t v = (uint32_t)reinterpret_cast<uintptr_t>(s) * 2654435761u + n;
        x[0] = v; x[1] = v >> 3; x[2] = v >> 7; x[3] = v >> 11;
        return;
    }
#endif
    const uintptr_t a = reinterpret_cast<uintptr_t>(s);
    const unsigned long long w = __cvta_generic_to_global(reinterpret_cast<const void*>(a & ~(uintptr_t)3));
    const uint32_t sh = (uint32_t)(a & 3) * 8, span = n ? (uint32_t)(a & 3) + n : 0;
    uint32_t v[5];
    v[0] = ld_word_if<READ_ONLY, 0>(w, span); v[1] = ld_word_if<READ_ONLY, 1>(w, span); v[2] = ld_word_if<READ_ONLY, 2>(w, span);
    v[3] = ld_word_if<READ_ONLY, 3>(w, span); v[4] = ld_word_if<READ_ONLY, 4>(w, span);
#pragma unroll
    for (int k = 0; k < 4; k++) x[k] = __funnelshift_r(v[k], v[k + 1], sh);
}
// The same from the staged stream window in shared memory (literal runs of a table round).
template <int K>
__device__ __forceinline__ uint32_t lds_word_if(uint32_t w, uint32_t span) {
    uint32_t v;
    asm volatile("{ .reg .pred p; setp.gt.u32 p, %1, %2; mov.u32 %0, 0; @p ld.shared.u32 %0, [%3+%2]; }"
                 : "=r"(v) : "r"(span), "n"(4 * K), "r"(w) : "memory");
    return v;
}
__device__ __forceinline__ void load16_win(const uint8_t* s, uint32_t n, uint32_t (&x)[4]) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(s);
    const uint32_t w = a & ~3u, sh = (a & 3u) * 8, span = n ? (a & 3u) + n : 0;
    uint32_t v[5];
    v[0] = lds_word_if<0>(w, span); v[1] = lds_word_if<1>(w, span); v[2] = lds_word_if<2>(w, span);
    v[3] = lds_word_if<3>(w, span); v[4] = lds_word_if<4>(w, span);
#pragma unroll
    for (int k = 0; k < 4; k++) x[k] = __funnelshift_r(v[k], v[k + 1], sh);
}
template <int I>
__device__ __forceinline__ void st_byte_if(unsigned long long g, uint32_t n, uint32_t word) {  // byte I of the run, if I < n
    asm volatile("{ .reg .pred p; setp.gt.u32 p, %0, %1; @p st.global.u8 [%2+%1], %3; }"
                 :: "r"(n), "n"(I), "l"(g), "r"(word >> (8 * (I & 3))) : "memory");
}
__device__ __forceinline__ void store16(uint8_t* d, uint32_t n, const uint32_t (&x)[4]) {
#ifdef ZG_DIAG_FAST_STORES  // timing experiment only (output is wrong): one aligned word store instead of n byte stores
    if (n) *reinterpret_cast<uint32_t*>(reinterpret_cast<uintptr_t>(d) & ~(uintptr_t)3) = x[0] ^ x[1] ^ x[2] ^ x[3];
    return;
#endif
    const unsigned long long g = __cvta_generic_to_global(d);
    st_byte_if<0>(g, n, x[0]); st_byte_if<1>(g, n, x[0]); st_byte_if<2>(g, n, x[0]); st_byte_if<3>(g, n, x[0]);
    st_byte_if<4>(g, n, x[1]); st_byte_if<5>(g, n, x[1]); st_byte_if<6>(g, n, x[1]); st_byte_if<7>(g, n, x[1]);
    st_byte_if<8>(g, n, x[2]); st_byte_if<9>(g, n, x[2]); st_byte_if<10>(g, n, x[2]); st_byte_if<11>(g, n, x[2]);
    st_byte_if<12>(g, n, x[3]); st_byte_if<13>(g, n, x[3]); st_byte_if<14>(g, n, x[3]); st_byte_if<15>(g, n, x[3]);
}

// Stored chunks (scheme 0: data that does not compress) are one copy: 16 bytes per lane and trip.  The destination is
// brought to a 16-byte boundary by the caller; the source is read as two aligned 16-byte vectors and shifted by its
// misalignment -- whole words by two levels of selects, the rest by a funnel shift -- so a trip costs three memory
// instructions per 16 bytes instead of twelve, and four trips are in flight per lane (2.2 -> 2.6 TB/s on stored chunks).
// Tried for the long literal runs inside LZ4 blocks as well (the 16 KiB mantissa planes of bf16): 2 % slower there -- the
// call in the middle of the execute step spills registers -- so those keep the word-wise copy.
template <bool SRC_IS_OUTPUT>
__device__ __noinline__ void warp_copy_vec(uint4* dst, const uint8_t* src, uint32_t nvec, uint32_t lane) {
    const uintptr_t sa = reinterpret_cast<uintptr_t>(src);
    const uint4* sv = reinterpret_cast<const uint4*>(sa & ~(uintptr_t)15);
    const uint32_t mis = (uint32_t)(sa & 15), ws = mis >> 2, bs = (mis & 3) * 8;
#pragma unroll 4
    for (uint32_t k = lane; k < nvec; k += 32) {
        const uint4 a = SRC_IS_OUTPUT ? sv[k] : __ldg(sv + k);
        uint4 b = make_uint4(0, 0, 0, 0);
        if (mis) b = SRC_IS_OUTPUT ? sv[k + 1] : __ldg(sv + k + 1);  // uniform: the whole run has one misalignment
        uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        if (ws & 2) {
#pragma unroll
            for (int i = 0; i < 6; i++) w[i] = w[i + 2];
        }
        if (ws & 1) {
#pragma unroll
            for (int i = 0; i < 5; i++) w[i] = w[i + 1];
        }
        dst[k] = make_uint4(__funnelshift_r(w[0], w[1], bs), __funnelshift_r(w[1], w[2], bs), __funnelshift_r(w[2], w[3], bs),
                            __funnelshift_r(w[3], w[4], bs));
    }
}

// Whole-warp copy of non-overlapping ranges, any alignment.  Destination words are written aligned; the source
// is read as aligned words and funnel-shifted.  Reads stay inside the 4-byte words that hold source bytes (the vector
// path for long runs: inside the 16-byte blocks that hold them -- the buffers are readable to the next multiple of 16).
template <bool SRC_IS_OUTPUT>
__device__ __forceinline__ void warp_copy(uint8_t* dst, const uint8_t* src, uint32_t len, uint32_t lane) {
    uint32_t head = (uint32_t)(-(intptr_t)reinterpret_cast<uintptr_t>(dst)) & 3u;
    if (head > len) head = len;
    if (lane < head) dst[lane] = SRC_IS_OUTPUT ? src[lane] : __ldg(src + lane);
    dst += head; src += head; len -= head;
    uint32_t nw = len >> 2;
    uintptr_t sa = reinterpret_cast<uintptr_t>(src);
    uint32_t sh = (uint32_t)(sa & 3) * 8;
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(sa & ~(uintptr_t)3);
    uint32_t* dw = reinterpret_cast<uint32_t*>(dst);
    if (sh == 0) {
#pragma unroll 4
        for (uint32_t k = lane; k < nw; k += 32) dw[k] = SRC_IS_OUTPUT ? sw[k] : __ldg(sw + k);
    } else {
#pragma unroll 4
        for (uint32_t k = lane; k < nw; k += 32) {
            uint32_t lo = SRC_IS_OUTPUT ? sw[k] : __ldg(sw + k);
            uint32_t hi = SRC_IS_OUTPUT ? sw[k + 1] : __ldg(sw + k + 1);
            dw[k] = __funnelshift_r(lo, hi, sh);
        }
    }
    uint32_t tail = len & 3;
    if (lane < tail) dst[4 * nw + lane] = SRC_IS_OUTPUT ? src[4 * nw + lane] : __ldg(src + 4 * nw + lane);
}

// Whole-warp LZ4 match copy with overlap semantics: out[dst+i] = out[dst+i-off].  The valid span doubles each
// round, so a 128 KiB run of one byte takes 17 rounds.
__device__ __forceinline__ void warp_match(uint8_t* dst, uint32_t off, uint32_t len, uint32_t lane) {
    const uint8_t* src = dst - off;
    uint32_t avail = off;
    while (len) {
        uint32_t n = len < avail ? len : avail;
        warp_copy<true>(dst, src, n, lane);
        __syncwarp();
        dst += n; len -= n;
        avail += n;  // n is a whole number of periods except for the final piece
    }
}

// Four bytes at any alignment (two aligned words + funnel shift).
__device__ __forceinline__ uint32_t ld4_unaligned(const uint8_t* p) {
    uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
    uint32_t sh = (uint32_t)(a & 3) * 8;
    uint32_t lo = w[0], hi = sh ? w[1] : 0u;
    return __funnelshift_r(lo, hi, sh);
}

// out[4j+k] = plane_k[j]  (SURVEY.md A.4).  tmp holds the four planes back to back.  Each lane turns 4 bytes of
// every plane into 16 output bytes with a 4x4 byte transpose (8 PRMT) and one 16-byte store when the destination
// allows it (the verify layout always does).
__device__ __forceinline__ void warp_regroup(uint8_t* out, const uint8_t* tmp, uint32_t n, uint32_t lane) {
    const uint32_t q = n >> 2, r = n & 3;
    const uint32_t p1 = q + (0 < r), p2 = p1 + q + (1 < r), p3 = p2 + q + (2 < r);
    const uint32_t q4 = q >> 2;  // groups of four quads = 16 output bytes
    const uint32_t oa = (uint32_t)(reinterpret_cast<uintptr_t>(out) & 15);
#pragma unroll 4
    for (uint32_t g = lane; g < q4; g += 32) {
        const uint32_t j = 4 * g;
        const uint32_t a = ld4_unaligned(tmp + j), b = ld4_unaligned(tmp + p1 + j), c = ld4_unaligned(tmp + p2 + j),
                       d = ld4_unaligned(tmp + p3 + j);
        const uint32_t t0 = __byte_perm(a, b, 0x5140), t1 = __byte_perm(a, b, 0x7362);
        const uint32_t u0 = __byte_perm(c, d, 0x5140), u1 = __byte_perm(c, d, 0x7362);
        const uint32_t o0 = __byte_perm(t0, u0, 0x5410), o1 = __byte_perm(t0, u0, 0x7632);
        const uint32_t o2 = __byte_perm(t1, u1, 0x5410), o3 = __byte_perm(t1, u1, 0x7632);
        uint8_t* o = out + 16 * (size_t)g;
        if (oa == 0) *reinterpret_cast<uint4*>(o) = make_uint4(o0, o1, o2, o3);
        else if ((oa & 3) == 0) {
            uint32_t* ow = reinterpret_cast<uint32_t*>(o);
            ow[0] = o0; ow[1] = o1; ow[2] = o2; ow[3] = o3;
        } else {
            const uint32_t v[4] = {o0, o1, o2, o3};
#pragma unroll
            for (int i = 0; i < 16; i++) o[i] = (uint8_t)(v[i >> 2] >> (8 * (i & 3)));
        }
    }
    // the last 0..3 quads and the 0..3 odd bytes
    for (uint32_t j = 4 * q4 + lane; j < q; j += 32) {
        out[4 * j] = tmp[j]; out[4 * j + 1] = tmp[p1 + j]; out[4 * j + 2] = tmp[p2 + j]; out[4 * j + 3] = tmp[p3 + j];
    }
    if (lane == 0) {
        if (r > 0) out[4 * q] = tmp[q];
        if (r > 1) out[4 * q + 1] = tmp[p1 + q];
        if (r > 2) out[4 * q + 2] = tmp[p2 + q];
    }
}


#ifdef ZG_LZ4_PROFILE  // developer build: execute-step counters, read by tools/stage_bench.py (ZG_PROF=1)
__device__ unsigned long long g_prof[8];  // cycles in pass 1 / long copies / match rounds; steps; rounds; pending per round; long runs
#define ZG_PF(i, v) do { if ((threadIdx.x & 31) == 0) atomicAdd(&g_prof[i], (unsigned long long)(v)); } while (0)
#else
#define ZG_PF(i, v)
#endif

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Stage kWinBytes of the stream starting at the 16-byte-aligned address A (clipped at the padded end of the
// compressed buffer; what is not staged keeps stale bytes, which lie beyond every block end).  Asynchronous: wait +
// __syncwarp before use.
__device__ __forceinline__ void stage_window(WarpShared& ws, uintptr_t A, const uint8_t* comp_hi, uint32_t lane) {
    const uint8_t* g = reinterpret_cast<const uint8_t*>(A);
#pragma unroll
    for (uint32_t i = lane; i < kWinBytes / 16; i += 32)
        if (g + 16 * i + 16 <= comp_hi) cp_async16(&ws.win[i], g + 16 * i);
}

// Hop table + exit table of one lane's segment, window bytes [seg0, seg0 + kSeg), highest position first.
//   nxt[w] = length in the stream of the sequence whose token is at w, had a sequence started there: token, at most one
//            literal-length byte, the literals, the offset, at most one match-length byte.  0 when that is not a
//            sequence the round can take: it would end past the block (which covers the block's last, literals-only
//            sequence), its offset is 0, a length needs more than one extra byte, or it is longer than kHopMax.  Purely
//            positional: the checks that need the output position are made at execute time.
//   ex[w]  = where the chain of hops from w first leaves the segment, relative to the segment end (< kHopMax), or kStop
//            if it runs into a 0 before that.  One look-up per segment then carries the true chain across the round.
// Every position costs the same ~30 instructions on every lane, whatever the data: no speculation, nothing discarded.
__device__ __forceinline__ void build_tables(WarpShared& ws, uint32_t seg0, uint32_t bend_w) {
    const uint8_t* const win = reinterpret_cast<const uint8_t*>(ws.win);
    const uint32_t* const ww = reinterpret_cast<const uint32_t*>(ws.win) + seg0 / 4;
    const uint32_t segend = seg0 + kSeg;
    uint32_t hi = ww[kSeg / 4];
    // One word (four positions) per trip: unrolling all 17 would put ~3,000 instructions in the hot loop of 24 warps
    // that are each somewhere else in the kernel, and the instruction cache does not hold that (measured: "no
    // instruction" became the top stall reason).
#pragma unroll kTableUnroll
    for (int i = kSeg / 4 - 1; i >= 0; i--) {
        const uint32_t cur = ww[i];
        uint32_t packed = 0;
#pragma unroll
        for (int j = 3; j >= 0; j--) {
            const uint32_t w = seg0 + 4 * i + j;
            const uint32_t b0 = (cur >> (8 * j)) & 0xffu;
            const uint32_t b1 = j < 3 ? (cur >> (8 * j + 8)) & 0xffu : hi & 0xffu;
            const uint32_t lit4 = b0 >> 4;
            const uint32_t adv = lit4 == 15u ? b1 + 16u : lit4;  // literal bytes + the length byte, if any
            const uint32_t o = w + 1u + adv;                      // where the offset sits
            const uint8_t* q = win + min(o, w + (kHopMax - 2u));   // clamped: anything further is too long anyway
            const uint32_t off = (uint32_t)q[0] | ((uint32_t)q[1] << 8), e2 = q[2];
            const bool mx = (b0 & 15u) == 15u;
            const uint32_t nip = o + 2u + (mx ? 1u : 0u);
            const uint32_t d = nip - w;
            const bool ok = off != 0u && !(mx && e2 == 255u) && nip <= bend_w && d <= kHopMax;
            const uint32_t dd = ok ? d : 0u;
            packed |= dd << (8 * j);
            // exit of the chain from w: through nip if that is still inside the segment
            const uint32_t inside = ws.ex[min(nip, segend - 1u)];
            const uint32_t x = !ok ? kStop : (nip >= segend ? nip - segend : inside);
            ws.ex[w] = (uint8_t)x;
        }
        reinterpret_cast<uint32_t*>(ws.nxt)[seg0 / 4 + i] = packed;
        hi = cur;
    }
}

// The sequence at window byte w, in full (execute step; w is a true token position, so it is one the table accepted).
__device__ __forceinline__ void seq_at(const uint8_t* win, uint32_t w, uint32_t& lit, uint32_t& ml, uint32_t& off, uint32_t& lit_w) {
    const uint32_t b0 = win[w], b1 = win[w + 1];
    const uint32_t lit4 = b0 >> 4;
    const bool lx = lit4 == 15u, mx = (b0 & 15u) == 15u;
    lit = lit4 + (lx ? b1 : 0u);
    lit_w = w + 1u + (lx ? 1u : 0u);
    const uint8_t* q = win + lit_w + lit;
    off = (uint32_t)q[0] | ((uint32_t)q[1] << 8);
    ml = (b0 & 15u) + 4u + (mx ? (uint32_t)q[2] : 0u);
}

// The matches of up to 32 sequences in stream order, one per lane (ml = 0: none), whose sources may lie in the
// output of those same sequences; everything else they read is final.  md (where the match goes) ascends with the
// lane and the ranges [md, md + ml) are disjoint, so the lanes that write what lane i reads are a contiguous run
// jlo .. jhi-1: jlo = number of ranges that end at or before the source, jhi = number that start before its end
// (two 5-step binary searches by shuffle).  A match is copied as soon as none of those is still pending, so the
// number of rounds is the depth of the dependency chains (1-3), not the number of matches.
__device__ __forceinline__ void match_rounds(uint8_t* lz, uint32_t md, uint32_t ml, uint32_t off, uint32_t lane) {
    uint8_t* mp = lz + md;
    uint32_t pm = __ballot_sync(kFull, ml != 0);
    if (!pm) return;
    const uint32_t s = md - off, e = min(s + ml, md), endv = md + ml;
    uint32_t jlo = 0, jhi = 0;
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1) {
        const uint32_t bj = __shfl_sync(kFull, endv, jlo + step - 1);
        const uint32_t aj = __shfl_sync(kFull, md, jhi + step - 1);
        if (bj <= s) jlo += step;
        if (aj < e) jhi += step;
    }
    // jhi <= lane for ascending md; the final mask keeps the lowest pending lane ready (progress) whatever the input
    const uint32_t dep = (jhi > jlo ? ((1u << jhi) - 1u) & ~((1u << jlo) - 1u) : 0u) & ((1u << lane) - 1u);
    for (;;) {
        ZG_PF(4, 1);
        ZG_PF(5, __popc(pm));
        __syncwarp();
        const bool ready = ml != 0 && (dep & pm) == 0;
        uint32_t lm = __ballot_sync(kFull, ready && ml > kShort);  // long matches: whole warp, one after the other
        while (lm) {
            const int l = __ffs(lm) - 1;
            lm &= lm - 1;
            warp_match(lz + __shfl_sync(kFull, md, l), __shfl_sync(kFull, off, l), __shfl_sync(kFull, ml, l), lane);
        }
        const bool go = ready && ml <= kShort;
        const bool fast = go && off >= ml;
        uint32_t x[4];
        load16<false>(mp - off, fast ? ml : 0u, x);
        store16(mp, fast ? ml : 0u, x);
        if (go && !fast) {  // self-overlapping short match: byte-serial
            const uint8_t* g = mp - off;
            for (uint32_t b = 0; b < ml; b++) mp[b] = g[b];
        }
        if (ready) ml = 0;
        pm = __ballot_sync(kFull, ml != 0);
        if (!pm) break;
    }
}

// One execute step: NP x 32 consecutive sequences of one chunk in stream order, lane l holding sequences l, 32 + l, ...
// (lanes without a sequence pass lit = ml = 0).  dst is the sequence's output position, op0 the position of the
// step's first byte; everything below op0 is final.  Pass 1 copies every literal run and every match whose source
// lies wholly below op0 (most of them) with all loads of the step in flight together; the matches that read the
// step's own output follow in match_rounds.
template <int NP, bool LIT_WIN>
__device__ __forceinline__ void exec_group(const uint8_t* src, const uint8_t* win, uint8_t* lz, uint32_t op0, const uint32_t (&lit_src)[NP],
                                           const uint32_t (&lit_w)[NP], const uint32_t (&lit)[NP], uint32_t (&ml)[NP],
                                           const uint32_t (&off)[NP], const uint32_t (&dst)[NP], uint32_t lane) {
    __syncwarp();
#ifdef ZG_LZ4_PROFILE
    const long long pt0 = clock64();
#endif
    uint32_t x[NP][4], y[NP][4], n[NP], k[NP];
    bool old[NP];
#pragma unroll
    for (int q = 0; q < NP; q++) {
        const uint32_t md = dst[q] + lit[q];  // where the match goes
        n[q] = lit[q] <= kShort ? lit[q] : 0u;
        old[q] = ml[q] != 0 && md - off[q] + ml[q] <= op0;
        k[q] = (old[q] && ml[q] <= kShort) ? ml[q] : 0u;
    }
#pragma unroll
    for (int q = 0; q < NP; q++) {
        if (LIT_WIN) load16_win(win + lit_w[q], n[q], x[q]);  // short literal runs of a table round sit in the staged window
        else load16<true>(src + lit_src[q], n[q], x[q]);
    }
#pragma unroll
    for (int q = 0; q < NP; q++) load16<false>(lz + dst[q] + lit[q] - off[q], k[q], y[q]);
#pragma unroll
    for (int q = 0; q < NP; q++) store16(lz + dst[q], n[q], x[q]);
#pragma unroll
    for (int q = 0; q < NP; q++) store16(lz + dst[q] + lit[q], k[q], y[q]);
#ifdef ZG_LZ4_PROFILE
    const long long pt1 = clock64();
    ZG_PF(0, pt1 - pt0);
    ZG_PF(3, 1);
#endif
#pragma unroll
    for (int q = 0; q < NP; q++) {
        uint32_t lm = __ballot_sync(kFull, lit[q] > kShort);  // long literal runs: whole-warp copies
        ZG_PF(6, __popc(lm));
        while (lm) {
            const int l = __ffs(lm) - 1;
            lm &= lm - 1;
            warp_copy<false>(lz + __shfl_sync(kFull, dst[q], l), src + __shfl_sync(kFull, lit_src[q], l),
                             __shfl_sync(kFull, lit[q], l), lane);
        }
        lm = __ballot_sync(kFull, old[q] && ml[q] > kShort);  // long matches from below the step: the same
        ZG_PF(7, __popc(lm));
        while (lm) {
            const int l = __ffs(lm) - 1;
            lm &= lm - 1;
            const uint32_t m_md = __shfl_sync(kFull, dst[q] + lit[q], l);
            warp_copy<true>(lz + m_md, lz + m_md - __shfl_sync(kFull, off[q], l), __shfl_sync(kFull, ml[q], l), lane);
        }
        if (old[q]) ml[q] = 0;
    }
#ifdef ZG_LZ4_PROFILE
    const long long pt2 = clock64();
    ZG_PF(1, pt2 - pt1);
#endif
#ifndef ZG_DIAG_NO_ROUNDS  // (timing experiment: the matches that read the step's own output are not copied)
#pragma unroll
    for (int q = 0; q < NP; q++) match_rounds(lz, dst[q] + lit[q], ml[q], off[q], lane);
#endif
#ifdef ZG_LZ4_PROFILE
    ZG_PF(2, clock64() - pt2);
#endif
}

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, uint32_t lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(kFull, v, d);
        if ((int)lane >= d) v += t;
    }
    return v;
}

#ifndef ZG_MIN_CTAS
#define ZG_MIN_CTAS 24
#endif
__global__ void __launch_bounds__(32, ZG_MIN_CTAS) lz4_decode_kernel(const uint8_t* __restrict__ comp, const ChunkDesc* __restrict__ chunks,
                                                        uint32_t n_chunks, uint8_t* out, uint8_t* tmp,
                                                        int32_t* __restrict__ status, uint32_t* ticket, const uint8_t* comp_hi, uint32_t keep_planes) {
    __shared__ __align__(16) WarpShared ws;
    const uint32_t lane = threadIdx.x;
    const uint8_t* const win = reinterpret_cast<const uint8_t*>(ws.win);
    uint16_t* const slots = reinterpret_cast<uint16_t*>(ws.ex);   // per-lane token positions (the exit table is dead by then)
    uint16_t* const list = reinterpret_cast<uint16_t*>(ws.nxt);   // true token positions in stream order (the hop table is dead by then)
    static_assert(32 * kSlots * 2 <= kSpan && (kSpan / 3 + 1) * 2 <= kSpan, "slots and list must fit the tables they alias");

    // Chunk state.  Warp-uniform: every lane holds the same values and the serial walker runs identically on all
    // lanes (same cost as running it on one, and nothing has to be broadcast).
    int state = ST_IDLE;
    uint32_t ci = 0, clen = 0, ulen = 0, scheme = 0;
    const uint8_t* src = nullptr;  // chunk payload
    uint8_t* lz = nullptr;         // where the LZ4 output goes (final position, or the plane buffer for scheme 2)
    uint8_t* fin_out = nullptr;    // final position of the chunk
    uint32_t ip = 0, bend = 0, op = 0, win_lo = 0, blk_op = 0, bmax = 0, fflags = 0;  // fflags: 1 indep, 2 bsum, 4 csum
    uint32_t p_lit = 0, p_lsrc = 0, p_mlen = 0, p_off = 0;
    int err = 0;
    uint32_t serial_rounds = 0;  // serial rounds to run before the next table round (the last one stopped early)
    uint32_t backoff = 1;        // ... and how many next time, should that one stop early too

#ifdef ZG_LZ4_PROFILE
    long long t_parse = 0, t_res = 0, t_exec = 0, t_serial = 0, n_spec = 0, n_serial = 0, n_true = 0, n_parsed = 0, tq;
#define ZG_T0() tq = clock64()
#define ZG_T(acc) do { long long now__ = clock64(); acc += now__ - tq; tq = now__; } while (0)
#else
#define ZG_T0()
#define ZG_T(acc)
#endif
    for (;;) {
        ZG_T0();
        if (state == ST_IDLE) {
            uint32_t t = 0;
            if (lane == 0) t = atomicAdd(ticket, 1u);
            t = __shfl_sync(kFull, t, 0);
            if (t >= n_chunks) break;
            const uint32_t k = chunks[t].order;  // longest-first work order (fill_decode_order)
            ci = k < n_chunks ? k : t;
            const ChunkDesc d = chunks[ci];
            clen = d.clen; ulen = d.ulen; scheme = d.scheme;
            src = comp + d.src_off;
            fin_out = out + d.dst_off;
            lz = scheme == 2 ? tmp + d.dst_off : fin_out;
            ip = 0; op = 0; err = 0; p_lit = p_mlen = 0;
            serial_rounds = 0; backoff = 1;
            state = ST_START;
        }

        if (state == ST_IN_BLOCK && (p_lit | p_mlen) == 0 && serial_rounds == 0 && bend - ip >= kSpecMin) {
            // ================= table round: up to kSpan stream bytes of the block =================
            const uintptr_t a_ip = reinterpret_cast<uintptr_t>(src) + ip;
            const uintptr_t wbase = a_ip & ~(uintptr_t)15;
            const uint32_t a0 = (uint32_t)(a_ip - wbase);         // window byte of stream position ip
            const uint32_t bend_w = bend - ip + a0;                // window byte of the block end (may lie beyond the window)
            const uint32_t nseg = min(32u, (min(bend_w, kSpan) + kSeg - 1) / kSeg);
            __syncwarp();  // the previous round's list is read until its last execute step
            stage_window(ws, wbase, comp_hi, lane);
            cp_async_wait_all();
            __syncwarp();

            // ---- tables: every lane its own segment (lanes past the block end have nothing to do)
            const uint32_t seg0 = lane * kSeg, segend = seg0 + kSeg;
            if (lane < nseg) build_tables(ws, seg0, bend_w);
            __syncwarp();
            ZG_T(t_parse);

            // ---- compose: carry the true chain from segment to segment (uniform; one look-up per segment)
            uint32_t e = a0, entry = kNone, stop_seg = 32;
#pragma unroll 1
            for (uint32_t k = 0; k < nseg; k++) {
                const uint32_t kend = (k + 1) * kSeg;
                if (e >= kend) continue;  // a long sequence jumped this segment
                if (lane == k) entry = e;
                const uint32_t x = ws.ex[e];
                if (x == kStop) { stop_seg = k; break; }
                e = kend + x;
            }
            __syncwarp();  // the exit table is dead from here: its space becomes the position slots

            // ---- chain walk: each lane lists the true token positions of its segment, then all in stream order
            uint32_t p = entry, cnt = 0;
#pragma unroll 1
            for (uint32_t it = 0; it < kSeg / 3 + 1; it++) {
                const bool act = p < segend;  // entry == kNone: never
                if (!__any_sync(kFull, act)) break;
                if (act) {
                    const uint32_t d = ws.nxt[p];
                    if (d == 0) p |= 0x80000000u;  // stopped here (only the lane of stop_seg gets here)
                    else { slots[lane * kSlots + cnt] = (uint16_t)p; cnt++; p += d; }
                }
            }
            const uint32_t incl = warp_incl_scan(cnt, lane);
            const uint32_t T = __shfl_sync(kFull, incl, 31);
            const bool stopped = stop_seg < 32;
            // where the stream goes on: in front of the sequence that stopped the chain, or where the chain left the span
            const uint32_t next_w = stopped ? (__shfl_sync(kFull, p, stop_seg & 31) & 0x7fffffffu) : e;
            __syncwarp();  // every lane is done reading the hop table: its space becomes the list
            {
                const uint32_t base = incl - cnt;
                for (uint32_t j = 0; j < cnt; j++) list[base + j] = slots[lane * kSlots + j];
            }
            __syncwarp();
            ZG_T(t_res);
#ifdef ZG_LZ4_PROFILE
            n_parsed += nseg * kSeg;
#endif

            // ---- execute: 32 consecutive true sequences per step
            const uint32_t ip0 = ip;
            uint32_t gop = op;
            bool bad = false;
#pragma unroll 1
            for (uint32_t g0 = 0; g0 < T && !bad; g0 += 32 * kNP) {
                uint32_t prel[kNP], lit[kNP], ml[kNP], off[kNP], dst[kNP], lit_src[kNP], lit_w[kNP];
                const uint32_t step0 = gop;
#pragma unroll
                for (int q = 0; q < kNP; q++) {
                    const uint32_t r = g0 + 32 * q + lane;
                    prel[q] = lit[q] = ml[q] = off[q] = lit_src[q] = lit_w[q] = 0;
                    if (r < T) {
                        const uint32_t w = list[r];
                        seq_at(win, w, lit[q], ml[q], off[q], lit_w[q]);
                        prel[q] = w - a0;
                        lit_src[q] = ip0 + lit_w[q] - a0;
                    }
                }
#pragma unroll
                for (int q = 0; q < kNP; q++) {
                    const uint32_t len = lit[q] + ml[q];
                    const uint32_t sc = warp_incl_scan(len, lane);
                    dst[q] = gop + sc - len;
                    // the oracle's checks that need output positions; a failing sequence goes to the serial walker
                    const bool anomaly = !bad && len != 0 && (off[q] > dst[q] + lit[q] - win_lo || dst[q] + len > ulen);
                    const uint32_t am = __ballot_sync(kFull, anomaly);
                    uint32_t adv = __shfl_sync(kFull, sc, 31);
                    if (bad) { lit[q] = ml[q] = 0; adv = 0; }
                    else if (am) {
                        const int fa = __ffs(am) - 1;
                        bad = true;
                        ip = ip0 + __shfl_sync(kFull, prel[q], fa);  // resume in front of that sequence
                        adv = __shfl_sync(kFull, dst[q], fa) - gop;
                        if ((int)lane >= fa) lit[q] = ml[q] = 0;
                    }
                    gop += adv;
                }
#ifndef ZG_DIAG_NO_EXEC  // (timing experiment: parse, place and check the sequences, copy nothing)
                exec_group<kNP, true>(src, win, lz, step0, lit_src, lit_w, lit, ml, off, dst, lane);
#endif
#ifdef ZG_LZ4_PROFILE
                for (int q = 0; q < kNP; q++) n_true += __popc(__ballot_sync(kFull, (lit[q] | ml[q]) != 0));
#endif
            }
            op = gop;
            if (!bad) ip = ip0 + next_w - a0;
            // A round that stopped early (a sequence the table does not hold, the block's last sequence, an error) hands
            // over to the serial walker; if that keeps happening the walker gets longer turns (long runs, long matches).
            if (bad || stopped) {
                serial_rounds = T >= 64 ? 1 : backoff;
                backoff = T >= 64 ? 1 : min(backoff * 2, 16u);
            } else
                backoff = 1;
            ZG_T(t_exec);
#ifdef ZG_LZ4_PROFILE
            n_spec++;
#endif
            continue;
        }

        // ================= serial round: up to kSerialCap sequences, every check of the oracle in its order =================
        uint32_t cnt = 0;
        const uint32_t op_start = op;
        bool finished = false;  // chunk complete (ok or error) after this round's records
        if (serial_rounds) serial_rounds--;
#define ZG_EMIT(LSRC, LIT, ML, OFF)                                                                  \
    do {                                                                                             \
        if (lane == 0) {                                                                             \
            ws.srec[3 * cnt] = (LSRC); ws.srec[3 * cnt + 1] = (LIT) | ((ML) << 16); ws.srec[3 * cnt + 2] = (OFF); \
        }                                                                                            \
        cnt++;                                                                                       \
    } while (0)
#pragma unroll 1
        for (uint32_t step = 0; step < kSerialCap && !finished && cnt < kSerialCap; step++) {
            if (state == ST_IN_BLOCK && (p_lit | p_mlen) == 0) {
                // One sequence.  The bytes come from an 8-byte window fetched with one round trip (token, short
                // literal run and offset usually all sit inside it).
                if (ip >= bend) { err = E_LZ4_BLOCK; finished = true; continue; }
                const uint32_t ip0 = ip;
                const uint64_t w8 = load_window(src + ip0, comp_hi);
                const uint32_t tok = (uint32_t)w8 & 0xffu;
                ip++;
                uint32_t lit = tok >> 4;
                bool bad = false;
                if (lit == 15) {
                    uint32_t b;
                    do {
                        if (ip >= bend) { bad = true; break; }
                        b = rd8(src + ip++);
                        lit += b;
                    } while (b == 255);
                }
                if (bad || lit > bend - ip) { err = E_LZ4_BLOCK; finished = true; continue; }
                if (lit > ulen - op) { err = E_SIZE_MISMATCH; finished = true; continue; }
                const uint32_t lsrc = ip;
                ip += lit;
                if (ip == bend) {  // last sequence of the block: literals only
                    p_lit = lit; p_lsrc = lsrc;
                    state = ST_AFTER_BLOCK;
                    continue;
                }
                if (bend - ip < 2) { err = E_LZ4_BLOCK; finished = true; continue; }
                const uint32_t rel = ip - ip0;
                uint32_t off;
                if (rel <= 6) off = (uint32_t)(w8 >> (8 * rel)) & 0xffffu;
                else off = rd8(src + ip) | (rd8(src + ip + 1) << 8);
                ip += 2;
                uint32_t ml = tok & 15;
                if (ml == 15) {
                    uint32_t b;
                    do {
                        if (ip >= bend) { bad = true; break; }
                        b = rd8(src + ip++);
                        ml += b;
                    } while (b == 255);
                }
                if (bad) { err = E_LZ4_BLOCK; finished = true; continue; }
                ml += 4;
                const uint32_t vop = op + lit;
                if (off == 0 || off > vop - win_lo) { err = E_LZ4_BLOCK; finished = true; continue; }
                if (ml > ulen - vop) { err = E_SIZE_MISMATCH; finished = true; continue; }
                if (lit <= kMaxPiece && ml <= kMaxPiece) {
                    ZG_EMIT(lsrc, lit, ml, off);
                    op = vop + ml;
                } else {
                    p_lit = lit; p_lsrc = lsrc; p_mlen = ml; p_off = off;  // split into pieces below
                }
                continue;
            }
            if (p_lit | p_mlen) {
                const uint32_t l = p_lit < kMaxPiece ? p_lit : kMaxPiece;
                const uint32_t m = (l == p_lit) ? (p_mlen < kMaxPiece ? p_mlen : kMaxPiece) : 0u;
                ZG_EMIT(p_lsrc, l, m, p_off);
                op += l + m; p_lsrc += l; p_lit -= l; p_mlen -= m;
                continue;
            }
            switch (state) {
                case ST_START: {
                    if (scheme == 0) {
                        if (clen != ulen) { err = E_BAD_HEADER; finished = true; break; }
                        if (ulen >= 64) {  // the whole chunk in one vector copy (nothing else of this round is pending)
                            const uint32_t head = (uint32_t)(-(intptr_t)reinterpret_cast<uintptr_t>(lz)) & 15u;
                            if (lane < head) lz[lane] = __ldg(src + lane);
                            const uint32_t nvec = (ulen - head) >> 4, done = head + 16 * nvec;
                            warp_copy_vec<false>(reinterpret_cast<uint4*>(lz + head), src + head, nvec, lane);
                            if (lane < ulen - done) lz[done + lane] = __ldg(src + done + lane);
                            op = ulen;
                        } else {
                            p_lit = ulen; p_lsrc = 0;
                        }
                        state = ST_FINISH;
                    } else if (scheme <= 2) {
                        if (clen < 7 || rd32(src) != 0x184D2204u) { err = E_LZ4_FRAME; finished = true; break; }
                        uint32_t flg = rd8(src + 4), bd = rd8(src + 5);
                        if ((flg >> 6) != 1 || (flg & 0x02) || (bd & 0x8F) || (flg & 1)) { err = E_LZ4_FRAME; finished = true; break; }
                        uint32_t bcode = (bd >> 4) & 7;
                        if (bcode < 4) { err = E_LZ4_FRAME; finished = true; break; }
                        bmax = 1u << (8 + 2 * bcode);
                        fflags = ((flg >> 5) & 1) | (((flg >> 4) & 1) << 1) | (((flg >> 2) & 1) << 2);
                        uint32_t csize = (flg >> 3) & 1;
                        ip = 6 + (csize ? 8 : 0);
                        if (clen < ip + 1) { err = E_LZ4_FRAME; finished = true; break; }
                        if (((xxh32_small(src + 4, ip - 4) >> 8) & 0xFF) != rd8(src + ip)) { err = E_LZ4_FRAME; finished = true; break; }
                        if (csize && (rd32(src + 6) != ulen || rd32(src + 10) != 0)) { err = E_SIZE_MISMATCH; finished = true; break; }
                        ip++;
                        state = ST_BLOCK_HDR;
                    } else { err = E_SCHEME; finished = true; }
                    break;
                }
                case ST_BLOCK_HDR: {
                    if (clen - ip < 4) { err = E_LZ4_FRAME; finished = true; break; }
                    uint32_t bs = rd32(src + ip);
                    ip += 4;
                    if (bs == 0) {  // EndMark
                        if (fflags & 4) {
                            if (clen - ip < 4) { err = E_LZ4_FRAME; finished = true; break; }
                            ip += 4;
                        }
                        if (op != ulen) err = E_SIZE_MISMATCH;
                        finished = true;
                        break;
                    }
                    uint32_t raw = bs >> 31;
                    bs &= 0x7FFFFFFFu;
                    if (bs > bmax || bs > clen - ip) { err = E_LZ4_FRAME; finished = true; break; }
                    if (raw) {
                        if (bs > ulen - op) { err = E_SIZE_MISMATCH; finished = true; break; }
                        p_lit = bs; p_lsrc = ip;
                        ip += bs;
                        blk_op = op + bs;  // stored blocks are not subject to the decoded-size check
                        state = ST_AFTER_BLOCK;
                    } else {
                        blk_op = op;
                        win_lo = (fflags & 1) ? op : 0u;
                        bend = ip + bs;
                        state = ST_IN_BLOCK;
                        // leave the round here when the block is long enough for the speculative walk
                        if (bs >= kSpecMin) step = kSerialCap;
                    }
                    break;
                }
                case ST_AFTER_BLOCK: {  // pending literals of the block have been emitted: op is the block end
                    if (op - blk_op > bmax && blk_op <= op) { err = E_LZ4_FRAME; finished = true; break; }
                    if (fflags & 2) {
                        if (clen - ip < 4) { err = E_LZ4_FRAME; finished = true; break; }
                        ip += 4;
                    }
                    state = ST_BLOCK_HDR;
                    break;
                }
                case ST_FINISH: finished = true; break;
                default: finished = true; break;
            }
        }
#undef ZG_EMIT
        __syncwarp();
        {
            uint32_t lit_src = 0, lit = 0, ml = 0, off = 0;
            if (lane < cnt) {
                lit_src = ws.srec[3 * lane];
                const uint32_t lens = ws.srec[3 * lane + 1];
                lit = lens & 0xffffu; ml = lens >> 16;
                off = ws.srec[3 * lane + 2];
            }
            if (cnt) {  // uniform
                const uint32_t len = lit + ml;
                const uint32_t sc = warp_incl_scan(len, lane);
                const uint32_t a_src[1] = {lit_src}, a_lit[1] = {lit}, a_off[1] = {off}, a_dst[1] = {op_start + sc - len};
                uint32_t a_ml[1] = {ml};
                exec_group<1, false>(src, win, lz, op_start, a_src, a_src, a_lit, a_ml, a_off, a_dst, lane);
                __syncwarp();  // srec is rewritten next round
            }
        }
        ZG_T(t_serial);
#ifdef ZG_LZ4_PROFILE
        n_serial++;
#endif
        // ---- chunk finished this round: regroup (scheme 2) and publish the status
        if (finished) {
            __syncwarp();
            if (err == 0 && scheme == 2 && !keep_planes) warp_regroup(fin_out, lz, ulen, lane);
#ifdef ZG_DIAG_HASH_LOAD  // timing experiment only: integer work of the size and shape a BLAKE3 of the chunk would be
            {             // (784 thread-instructions per 64 bytes per lane, four independent columns, add / xor / rotate), no loads
                uint32_t va[4], vb[4], vc[4], vd[4];
#pragma unroll
                for (int k = 0; k < 4; k++) { va[k] = lane + ulen + k; vb[k] = va[k] * 3u; vc[k] = vb[k] ^ 0x9e3779b9u; vd[k] = vc[k] + 7u; }
                const uint32_t iters = (uint32_t)(((uint64_t)(ulen / 32u) * 784u) / (64u * 48u));
#pragma unroll 1
                for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        va[k] += vb[k]; vd[k] = __funnelshift_l(vd[k] ^ va[k], vd[k] ^ va[k], 16); vc[k] += vd[k]; vb[k] = __funnelshift_l(vb[k] ^ vc[k], vb[k] ^ vc[k], 12);
                        va[k] += vb[k]; vd[k] = __funnelshift_l(vd[k] ^ va[k], vd[k] ^ va[k], 8);  vc[k] += vd[k]; vb[k] = __funnelshift_l(vb[k] ^ vc[k], vb[k] ^ vc[k], 7);
                    }
                }
                uint32_t x = 0;
#pragma unroll
                for (int k = 0; k < 4; k++) x ^= va[k] ^ vb[k] ^ vc[k] ^ vd[k];
                if (x == 0x12345678u) status[ci] = err;  // keeps the work alive (the value is the one lane 0 writes anyway)
            }
#endif
            if (lane == 0) status[ci] = err;
            state = ST_IDLE;
        }
    }
#ifdef ZG_LZ4_PROFILE
    if (lane == 0) {
        unsigned long long* pr = reinterpret_cast<unsigned long long*>(ticket) + 2;
        atomicAdd(pr + 0, (unsigned long long)t_parse); atomicAdd(pr + 1, (unsigned long long)t_res);
        atomicAdd(pr + 2, (unsigned long long)t_exec); atomicAdd(pr + 3, (unsigned long long)t_serial);
        atomicAdd(pr + 4, (unsigned long long)n_spec); atomicAdd(pr + 5, (unsigned long long)n_serial);
        atomicAdd(pr + 6, (unsigned long long)n_true); atomicAdd(pr + 7, (unsigned long long)n_parsed);
    }
#endif
}

}  // namespace

#ifdef ZG_LZ4_PROFILE
extern "C" __attribute__((visibility("default"))) void zg_lz4_profile_read(unsigned long long* out8, int reset) {
    cudaMemcpyFromSymbol(out8, g_prof, sizeof(g_prof));
    if (reset) { unsigned long long z[8] = {0}; cudaMemcpyToSymbol(g_prof, z, sizeof(z)); }
}
#endif

void fill_decode_order(ChunkDesc* chunks, size_t n) {
    std::vector<uint32_t> idx(n);
    for (size_t i = 0; i < n; i++) idx[i] = (uint32_t)i;
    std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) {
        return (uint64_t)chunks[a].clen + chunks[a].ulen > (uint64_t)chunks[b].clen + chunks[b].ulen;
    });
    for (size_t t = 0; t < n; t++) chunks[t].order = idx[t];
}

size_t decode_scratch_bytes(size_t n_chunks, uint64_t out_bytes) {
    (void)n_chunks;
    return 256 + align_up((size_t)out_bytes + 16, 256);  // ticket counter + plane buffer for scheme-2 chunks
}

cudaError_t launch_decode(const uint8_t* comp, uint64_t comp_bytes, const ChunkDesc* chunks, size_t n, uint8_t* out,
                          uint64_t out_bytes, int32_t* status, void* scratch, size_t scratch_bytes, cudaStream_t st, bool keep_planes) {
    if (n == 0) return cudaSuccess;
    if (scratch_bytes < decode_scratch_bytes(n, out_bytes) || n >= 0xffffffffull) return cudaErrorInvalidValue;
    uint8_t* sb = static_cast<uint8_t*>(scratch);
    uint32_t* ticket = reinterpret_cast<uint32_t*>(sb);
    uint8_t* tmp = sb + 256;
    cudaError_t e = cudaMemsetAsync(ticket, 0, 256, st);
    if (e != cudaSuccess) return e;
    // One warp per CTA, one chunk per warp at a time; the grid is what the machine keeps resident (shared memory
    // allows 24 CTAs per SM) and the ticket counter feeds it, longest chunks first.  ZEST_LZ4_WARPS overrides.
    static int ctas_per_sm = 0, sms = 0;  // same answer from every thread that gets here first
    if (ctas_per_sm == 0) {
        int dev = 0, occ = 0, n_sm = 0;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, lz4_decode_kernel, 32, 0) == cudaSuccess && occ > 0 && n_sm > 0) {
            sms = n_sm;
            ctas_per_sm = occ;
        } else {
            sms = 148;
            ctas_per_sm = 24;
        }
    }
    size_t warps = (size_t)sms * ctas_per_sm;
    if (const char* v = getenv("ZEST_LZ4_WARPS")) { long w = atol(v); if (w > 0) warps = (size_t)w; }
    if (warps > n) warps = n;
    const uint8_t* comp_hi = comp + align_up((size_t)comp_bytes, 16);
    lz4_decode_kernel<<<(unsigned)warps, 32, 0, st>>>(comp, chunks, (uint32_t)n, out, tmp, status, ticket, comp_hi, keep_planes ? 1u : 0u);
    return cudaGetLastError();
}

}  // namespace zg
