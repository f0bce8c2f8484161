// Xorb chunk decode on sm_100a: scheme 0 (stored), 1 (LZ4 frame), 2 (4-way byte-grouped LZ4 frame).
//
// Reference behaviour replaced: zig-xet XorbReader.extractChunkRange -- the per-chunk decompress loop that zest
// runs once per reconstruction term (/root/reference/src/xet_bridge.zig:256-257,
// src/parallel_download.zig:65-66).  Format: SURVEY.md A.4 / A.4b; oracle: oracle/xet_oracle.c (lz4_frame,
// lz4_block, oz_decode_chunk), whose checks are mirrored one for one and in the same order so that a malformed
// chunk gets the same error code on both sides.
//
// Mapping.  LZ4 token parsing is serial inside a chunk, and on real Xet data the sequences are tiny (bf16
// byte-grouped: ~4,100 sequences per 64 KiB chunk, ~8 output bytes each outside the two incompressible planes),
// so the per-sequence instruction count decides throughput.  A warp therefore works on up to 32 chunks at once:
//   parse   : LANE c walks the token stream of chunk slot c on its own (no warp-wide work per token) and appends
//             validated (literal source, destination, literal length, match length, offset) records to a
//             shared-memory list -- all format and bounds checks happen here, on positions only;
//   execute : for each slot, LANE s performs sequence s of the list: short literal runs and short matches are
//             copied by their own lane, long ones (the 16-40 KB incompressible planes, zero runs) by the whole
//             warp with word-wide, funnel-shifted copies; matches wait on earlier matches of the same batch only
//             when their source overlaps them (frontier test), so most batches finish in one or two rounds.
// Chunks are handed out through a global ticket counter, so slots refill as soon as a chunk ends.  Output goes to
// global memory (a 64 MiB xorb stays resident in the 126 MB L2); byte-group regrouping is done by the warp when a
// scheme-2 chunk completes, reading the planes back from L2.
//
// Roofline: HBM-bound in principle (compressed bytes in + decoded bytes out); in practice bounded by issue slots
// and L1 wavefronts per sequence.  bench.py reports achieved GB/s over those algorithmic bytes.
#include <algorithm>
#include <cstdlib>
#include <vector>

#include "kernels.h"

namespace zg {

namespace {

constexpr int kWarps = 1;          // warps per CTA (no block-level cooperation; small CTAs spread over all SMs)
constexpr int kSeqCap = 16;        // records parsed per slot per round (two slots share one 32-lane execute step)
constexpr uint32_t kShort = 16;    // literal runs / matches up to this length are copied by their own lane
constexpr uint32_t kMaxPiece = 65535;

enum : int { ST_IDLE = 0, ST_START, ST_BLOCK_HDR, ST_IN_BLOCK, ST_AFTER_BLOCK, ST_FINISH, ST_EXHAUSTED };

struct Seq {           // 16 bytes
    uint32_t lit_src;  // literal source, relative to the chunk payload
    uint32_t dst;      // destination of the literals, relative to the chunk's LZ4 output
    uint32_t lens;     // literal length | match length << 16
    uint32_t off;      // match offset (1..65535)
};

struct WarpShared {
    Seq rec[kSeqCap * 32];         // record s of slot c lives at [s * 32 + (c ^ s)] (conflict-free both ways)
    unsigned long long src[32];    // per slot: payload base
    unsigned long long lz[32];     // per slot: LZ4 output base
    uint32_t pend[32];             // per slot: records whose match is still to be copied (bit s)
    uint32_t frontier[32];         // per slot: output offset below which everything is final this round
    uint32_t w0[32];               // per slot: output offset where this round's records start
    uint8_t first[32];             // per slot: index of the first pending match
    uint8_t cnt[32];               // per slot: records parsed this round
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

// Eight stream bytes at any alignment as one u64 (three aligned word loads, issued together).  May touch the
// aligned words just past p + 7; the compressed buffer is padded by 16 bytes for that.
__device__ __forceinline__ uint64_t load_window(const uint8_t* p) {
    uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
    uint32_t sh = (uint32_t)(a & 3) * 8;
    uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
    return (uint64_t)__funnelshift_r(w0, w1, sh) | ((uint64_t)__funnelshift_r(w1, w2, sh) << 32);
}

// Up to 16 bytes from any alignment into four words: all loads are issued before anything is stored, so a lane
// pays one memory latency per copy instead of one per byte.  Only aligned words holding wanted bytes are read.
template <bool READ_ONLY>
__device__ __forceinline__ void load16(const uint8_t* s, uint32_t n, uint32_t (&x)[4]) {
    uintptr_t a = reinterpret_cast<uintptr_t>(s);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
    uint32_t sh = (uint32_t)(a & 3) * 8, span = n ? (uint32_t)(a & 3) + n : 0;
    uint32_t v[5];
#pragma unroll
    for (int k = 0; k < 5; k++) {
        v[k] = 0;
        if (span > 4u * k) v[k] = READ_ONLY ? __ldg(w + k) : w[k];
    }
#pragma unroll
    for (int k = 0; k < 4; k++) x[k] = __funnelshift_r(v[k], v[k + 1], sh);
}
__device__ __forceinline__ void store16(uint8_t* d, uint32_t n, const uint32_t (&x)[4]) {
#pragma unroll
    for (int i = 0; i < 16; i++)
        if ((uint32_t)i < n) d[i] = (uint8_t)(x[i >> 2] >> (8 * (i & 3)));
}

// Whole-warp copy of non-overlapping ranges, any alignment.  Destination words are written aligned; the source
// is read as aligned words and funnel-shifted.  Reads stay inside the 4-byte words that hold source bytes.
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
        for (uint32_t k = lane; k < nw; k += 32) dw[k] = SRC_IS_OUTPUT ? sw[k] : __ldg(sw + k);
    } else {
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

// ---- execute steps.  A 32-lane step covers two slots x 16 records (slot = base + lane / 16, record = lane % 16); NP = 2
// issues two such steps back to back so the loads of four slots are in flight together, NP = 1 is used when the
// second pair of slots is empty (small batches run with 1-2 slots per warp).

// Pass 1: every literal run, and every match whose source lies wholly before this round's output.
template <int NP>
__device__ __forceinline__ void exec_pass1(WarpShared& ws, uint32_t it, uint32_t lane) {
    const uint32_t half = lane >> 4, s = lane & 15;
    Seq e[NP];
    uint32_t lit[NP], ml[NP], n[NP], k[NP], c[NP], x[NP][4], y[NP][4];
    const uint8_t* sp[NP];
    uint8_t* dp[NP];
    bool old[NP];
#pragma unroll
    for (int q = 0; q < NP; q++) {
        c[q] = 2 * it + half + 2 * q;
        e[q] = Seq{0, 0, 0, 0};
        if (s < ws.cnt[c[q]]) e[q] = ws.rec[s * 32 + (c[q] ^ s)];
        lit[q] = e[q].lens & 0xffffu; ml[q] = e[q].lens >> 16;
        sp[q] = reinterpret_cast<const uint8_t*>(ws.src[c[q]]) + e[q].lit_src;
        dp[q] = reinterpret_cast<uint8_t*>(ws.lz[c[q]]) + e[q].dst;
        n[q] = lit[q] <= kShort ? lit[q] : 0u;
        // "old" match: short, and its source ends at or before the first byte this round produces for the slot
        old[q] = ml[q] != 0 && ml[q] <= kShort && e[q].dst + lit[q] - e[q].off + ml[q] <= ws.w0[c[q]];
        k[q] = old[q] ? ml[q] : 0u;
    }
#pragma unroll
    for (int q = 0; q < NP; q++) load16<true>(sp[q], n[q], x[q]);
#pragma unroll
    for (int q = 0; q < NP; q++) load16<false>(dp[q] + lit[q] - e[q].off, k[q], y[q]);
#pragma unroll
    for (int q = 0; q < NP; q++) store16(dp[q], n[q], x[q]);
#pragma unroll
    for (int q = 0; q < NP; q++) store16(dp[q] + lit[q], k[q], y[q]);
#pragma unroll
    for (int q = 0; q < NP; q++) {
        if (old[q]) atomicAnd(&ws.pend[c[q]], ~(1u << s));
        uint32_t lm = __ballot_sync(0xffffffffu, lit[q] > kShort);  // long literal runs: whole-warp copies
        while (lm) {
            int l = __ffs(lm) - 1;
            lm &= lm - 1;
            uint8_t* d = reinterpret_cast<uint8_t*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(dp[q]), l));
            const uint8_t* g = reinterpret_cast<const uint8_t*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(sp[q]), l));
            warp_copy<false>(d, g, __shfl_sync(0xffffffffu, lit[q], l), lane);
        }
    }
}

// Pass 2, one round: matches that read this round's own output and whose source now lies below their slot's frontier.
template <int NP>
__device__ __forceinline__ void exec_pass2(WarpShared& ws, uint32_t it, uint32_t lane) {
    const uint32_t half = lane >> 4, s = lane & 15;
    Seq e[NP];
    uint32_t ml[NP], md[NP], c[NP], x[NP][4];
    uint8_t* dp[NP];
    bool pd[NP], fs[NP], go[NP], fast[NP];
#pragma unroll
    for (int q = 0; q < NP; q++) {
        c[q] = 2 * it + half + 2 * q;
        e[q] = Seq{0, 0, 0, 0};
        pd[q] = (ws.pend[c[q]] >> s) & 1;
        if (pd[q]) e[q] = ws.rec[s * 32 + (c[q] ^ s)];
        ml[q] = e[q].lens >> 16;
        md[q] = e[q].dst + (e[q].lens & 0xffffu);
        fs[q] = pd[q] && s == ws.first[c[q]];
        dp[q] = reinterpret_cast<uint8_t*>(ws.lz[c[q]]) + md[q];
        // long matches run when they are first in their slot, with the whole warp
        uint32_t lm = __ballot_sync(0xffffffffu, fs[q] && ml[q] > kShort);
        while (lm) {
            int l = __ffs(lm) - 1;
            lm &= lm - 1;
            uint8_t* d = reinterpret_cast<uint8_t*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(dp[q]), l));
            warp_match(d, __shfl_sync(0xffffffffu, e[q].off, l), __shfl_sync(0xffffffffu, ml[q], l), lane);
        }
        go[q] = pd[q] && ml[q] <= kShort && (fs[q] || md[q] - e[q].off + ml[q] <= ws.frontier[c[q]]);
        fast[q] = go[q] && e[q].off >= ml[q];
    }
#pragma unroll
    for (int q = 0; q < NP; q++) load16<false>(dp[q] - e[q].off, fast[q] ? ml[q] : 0u, x[q]);
#pragma unroll
    for (int q = 0; q < NP; q++) store16(dp[q], fast[q] ? ml[q] : 0u, x[q]);
#pragma unroll
    for (int q = 0; q < NP; q++) {
        if (go[q] && !fast[q]) {  // self-overlapping short match: byte-serial
            const uint8_t* g = dp[q] - e[q].off;
            for (uint32_t b = 0; b < ml[q]; b++) dp[q][b] = g[b];
        }
        if (go[q] || (fs[q] && ml[q] > kShort)) atomicAnd(&ws.pend[c[q]], ~(1u << s));
    }
}

__global__ void __launch_bounds__(kWarps * 32) lz4_decode_kernel(const uint8_t* __restrict__ comp,
                                                                 const ChunkDesc* __restrict__ chunks, uint32_t n_chunks,
                                                                 uint8_t* out, uint8_t* tmp, int32_t* __restrict__ status,
                                                                 uint32_t* ticket, uint32_t slots) {
    __shared__ __align__(16) WarpShared s_ws[kWarps];
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    WarpShared& ws = s_ws[wid];

    // per-slot (per-lane) parser state
    int state = lane < slots ? ST_IDLE : ST_EXHAUSTED;
    uint32_t ci = 0, clen = 0, ulen = 0, scheme = 0;
    const uint8_t* src = nullptr;  // chunk payload
    uint8_t* lz = nullptr;         // where the LZ4 output goes (final position, or the plane buffer for scheme 2)
    uint8_t* fin_out = nullptr;    // final position of the chunk
    uint32_t ip = 0, bend = 0, op = 0, win_lo = 0, blk_op = 0, bmax = 0, fflags = 0;  // fflags: 1 indep, 2 bsum, 4 csum
    uint32_t p_lit = 0, p_lsrc = 0, p_mlen = 0, p_off = 0;
    int err = 0;

#ifdef ZG_LZ4_PROFILE
    long long t_parse = 0, t_p1 = 0, t_p2 = 0, t_fin = 0, n_rounds = 0, n_p2 = 0, tq;
#define ZG_T0() tq = clock64()
#define ZG_T(acc) do { long long now__ = clock64(); acc += now__ - tq; tq = now__; } while (0)
#else
#define ZG_T0()
#define ZG_T(acc)
#endif
    for (;;) {
        ZG_T0();
        // ---- refill idle slots from the ticket counter
        uint32_t idle = __ballot_sync(0xffffffffu, state == ST_IDLE);
        if (idle) {
            uint32_t base = 0;
            int leader = __ffs(idle) - 1;
            if ((int)lane == leader) base = atomicAdd(ticket, (uint32_t)__popc(idle));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (state == ST_IDLE) {
                uint32_t t = base + __popc(idle & ((1u << lane) - 1));
                if (t < n_chunks) {
                    const uint32_t k = chunks[t].order;  // longest-first work order (fill_decode_order)
                    ChunkDesc d = chunks[k < n_chunks ? k : t];
                    ci = k < n_chunks ? k : t; clen = d.clen; ulen = d.ulen; scheme = d.scheme;
                    src = comp + d.src_off;
                    fin_out = out + d.dst_off;
                    lz = scheme == 2 ? tmp + d.dst_off : fin_out;
                    ws.src[lane] = reinterpret_cast<unsigned long long>(src);
                    ws.lz[lane] = reinterpret_cast<unsigned long long>(lz);
                    ip = 0; op = 0; err = 0; p_lit = p_mlen = 0;
                    state = ST_START;
                } else
                    state = ST_EXHAUSTED;
            }
        }
        if (__all_sync(0xffffffffu, state == ST_EXHAUSTED)) break;

        // ---- parse: every active lane appends up to kSeqCap records for its own chunk
        uint32_t cnt = 0, pmask = 0;
        ws.w0[lane] = op;
        bool finished = false;  // chunk complete (ok or error) after this round's records
        // Lock-step matters: lanes execute together only while they follow the same path, and a lane that falls behind
        // would otherwise finish its quota alone.  So the loop runs a fixed number of steps for the whole warp and every
        // active lane makes ONE step per iteration -- normally one sequence on the hot path; block/frame boundaries,
        // length-extension bytes, oversized pieces and errors take the general path of that iteration and rejoin.
        uint32_t wv0 = 0, wv1 = 0, wv2 = 0, wip = 0xffffffffu;  // raw words of the 8-byte window at position wip
#pragma unroll 1
        for (int step = 0; step < kSeqCap; step++) {
            const bool act = !finished && state != ST_IDLE && state != ST_EXHAUSTED && cnt < (uint32_t)kSeqCap;
            if (!__any_sync(0xffffffffu, act)) break;
            if (!act) continue;
            bool handled = false;
            if (state == ST_IN_BLOCK && (p_lit | p_mlen) == 0 && bend - ip >= 48 && ulen - op >= 40) {
                // Hot path: far from the end of the block and of the output.  If there are no length-extension bytes the
                // only checks left of the oracle's list are those on the offset; the rest is implied by the two margins
                // (such a sequence consumes <= 17 input bytes and produces <= 32 output bytes).  Software-pipelined: the
                // serial chain is window -> token -> literal length -> next position -> next window, so the next
                // window's words are requested as soon as the literal length is known, carried RAW into the next step,
                // and the record bookkeeping runs in the shadow of that load.
                const uintptr_t a = reinterpret_cast<uintptr_t>(src + ip);
                if (wip != ip) {
                    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
                    wv0 = __ldg(w); wv1 = __ldg(w + 1); wv2 = __ldg(w + 2);
                }
                const uint32_t sh0 = (uint32_t)(a & 3) * 8;
                const uint32_t lo = __funnelshift_r(wv0, wv1, sh0), hi = __funnelshift_r(wv1, wv2, sh0);
                const uint64_t w64 = ((uint64_t)hi << 32) | lo;
                const uint32_t tok = lo & 0xffu;
                const uint32_t lit = tok >> 4, mln = tok & 15u;
                if (lit != 15u && mln != 15u) {
                    // A second sequence that lies wholly inside the same 8-byte window (literal runs of 0-2 bytes are
                    // the common case on both text and byte-grouped bf16) is taken in the same step: one load round
                    // trip then yields two records.
                    const uint32_t k1 = 3 + lit;
                    uint32_t lit2 = 0, mln2 = 0, off2 = 0;
                    bool two = false;
                    if (k1 <= 5 && cnt + 2 <= (uint32_t)kSeqCap && ulen - op >= 72) {
                        const uint32_t r = (uint32_t)(w64 >> (8 * k1));
                        lit2 = (r >> 4) & 15u; mln2 = r & 15u;
                        if (lit2 + k1 <= 5 && mln2 != 15u) {
                            two = true;
                            off2 = (uint32_t)(w64 >> (8 * (k1 + 1 + lit2))) & 0xffffu;
                        }
                    }
                    const uint32_t nip = ip + k1 + (two ? 3 + lit2 : 0u);
                    const uint32_t* nw = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(src + nip) & ~(uintptr_t)3);
                    const uint32_t n0 = __ldg(nw), n1 = __ldg(nw + 1), n2 = __ldg(nw + 2);
                    uint32_t off;
                    if (lit <= 5) off = (uint32_t)(w64 >> (8 + 8 * lit)) & 0xffffu;
                    else off = rd8(src + ip + 1 + lit) | (rd8(src + ip + 2 + lit) << 8);
                    const uint32_t vop = op + lit, ml = mln + 4;
                    if (off == 0 || off > vop - win_lo) { err = E_LZ4_BLOCK; finished = true; }
                    else {
                        Seq e{ip + 1, op, lit | (ml << 16), off};
                        ws.rec[cnt * 32 + (lane ^ cnt)] = e;
                        pmask |= 1u << cnt;
                        cnt++;
                        op = vop + ml;
                        if (two) {
                            const uint32_t vop2 = op + lit2, ml2 = mln2 + 4;
                            if (off2 == 0 || off2 > vop2 - win_lo) { err = E_LZ4_BLOCK; finished = true; }
                            else {
                                Seq e2{ip + k1 + 1, op, lit2 | (ml2 << 16), off2};
                                ws.rec[cnt * 32 + (lane ^ cnt)] = e2;
                                pmask |= 1u << cnt;
                                cnt++;
                                op = vop2 + ml2;
                            }
                        }
                        if (!finished) { ip = nip; wv0 = n0; wv1 = n1; wv2 = n2; wip = nip; }
                    }
                    handled = true;
                }
            }
            if (handled) continue;
            wip = 0xffffffffu;
            if (state == ST_IN_BLOCK && (p_lit | p_mlen) == 0) {
                // One sequence.  The checks and their order are those of the oracle's lz4_block; the bytes come from
                // an 8-byte window fetched with one round trip (token, short literal run and offset usually all sit
                // inside it).
                if (ip >= bend) { err = E_LZ4_BLOCK; finished = true; continue; }
                const uint32_t ip0 = ip;
                const uint64_t win = load_window(src + ip0);
                const uint32_t tok = (uint32_t)win & 0xffu;
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
                if (rel <= 6) off = (uint32_t)(win >> (8 * rel)) & 0xffffu;
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
                    Seq e{lsrc, op, lit | (ml << 16), off};
                    ws.rec[cnt * 32 + (lane ^ cnt)] = e;
                    pmask |= 1u << cnt;
                    cnt++;
                    op = vop + ml;
                } else {
                    p_lit = lit; p_lsrc = lsrc; p_mlen = ml; p_off = off;  // split into pieces below
                }
                continue;
            }
            if (p_lit | p_mlen) {
                uint32_t l = p_lit < kMaxPiece ? p_lit : kMaxPiece;
                uint32_t m = (l == p_lit) ? (p_mlen < kMaxPiece ? p_mlen : kMaxPiece) : 0u;
                Seq e{p_lsrc, op, l | (m << 16), p_off};
                ws.rec[cnt * 32 + (lane ^ cnt)] = e;
                if (m) pmask |= 1u << cnt;
                cnt++;
                op += l + m; p_lsrc += l; p_lit -= l; p_mlen -= m;
                continue;
            }
            switch (state) {
                case ST_START: {
                    if (scheme == 0) {
                        if (clen != ulen) { err = E_BAD_HEADER; finished = true; break; }
                        p_lit = ulen; p_lsrc = 0;
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
        ws.cnt[lane] = (uint8_t)cnt;
        ws.pend[lane] = pmask;
        __syncwarp();
        ZG_T(t_parse);
        const uint32_t have = __ballot_sync(0xffffffffu, cnt > 0);  // slots with records this round

        // ---- execute, pass 1: every literal run, and every match whose source lies wholly before this round's
        // output (no dependence on anything written this round: ~91 % of matches on text, ~98 % on byte-grouped bf16)
#pragma unroll 1
        for (uint32_t it = 0; it < 16; it += 2) {
            const uint32_t m4 = (have >> (2 * it)) & 0xfu;
            if (m4 == 0) continue;
            if (m4 >> 2) exec_pass1<2>(ws, it, lane);
            else exec_pass1<1>(ws, it, lane);
        }
        __syncwarp();
        ZG_T(t_p1);

        // ---- execute, pass 2: the matches that read this round's own output.  Per slot they are ordered; a match may
        // run as soon as its source lies below the slot's frontier (start of the first match still pending), so every
        // round retires at least one match per slot and usually most of them.
        for (;;) {
            uint32_t p = ws.pend[lane];
            if (p) {
                uint32_t f = (uint32_t)__ffs(p) - 1;
                Seq e = ws.rec[f * 32 + (lane ^ f)];
                ws.first[lane] = (uint8_t)f;
                ws.frontier[lane] = e.dst + (e.lens & 0xffffu);
            }
            const uint32_t anyp = __ballot_sync(0xffffffffu, p != 0);
            if (!anyp) break;
            __syncwarp();
#ifdef ZG_LZ4_PROFILE
            n_p2++;
#endif
#pragma unroll 1
            for (uint32_t it = 0; it < 16; it += 2) {
                const uint32_t m4 = (anyp >> (2 * it)) & 0xfu;
                if (m4 == 0) continue;
                if (m4 >> 2) exec_pass2<2>(ws, it, lane);
                else exec_pass2<1>(ws, it, lane);
            }
            __syncwarp();
        }

        ZG_T(t_p2);
        // ---- chunks that finished this round: regroup (scheme 2) and publish the status
        uint32_t fin = __ballot_sync(0xffffffffu, finished);
        while (fin) {
            int c = __ffs(fin) - 1;
            fin &= fin - 1;
            int c_err = __shfl_sync(0xffffffffu, err, c);
            uint32_t c_scheme = __shfl_sync(0xffffffffu, scheme, c);
            if (c_err == 0 && c_scheme == 2) {
                uint8_t* c_out = reinterpret_cast<uint8_t*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(fin_out), c));
                const uint8_t* c_lz = reinterpret_cast<const uint8_t*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(lz), c));
                warp_regroup(c_out, c_lz, __shfl_sync(0xffffffffu, ulen, c), lane);
            }
        }
        if (finished) {
            status[ci] = err;
            state = ST_IDLE;
        }
        __syncwarp();
        ZG_T(t_fin);
#ifdef ZG_LZ4_PROFILE
        n_rounds++;
#endif
    }
#ifdef ZG_LZ4_PROFILE
    if (lane == 0) {
        unsigned long long* pr = reinterpret_cast<unsigned long long*>(ticket) + 2;
        atomicAdd(pr + 0, (unsigned long long)t_parse); atomicAdd(pr + 1, (unsigned long long)t_p1);
        atomicAdd(pr + 2, (unsigned long long)t_p2); atomicAdd(pr + 3, (unsigned long long)t_fin);
        atomicAdd(pr + 4, (unsigned long long)n_rounds); atomicAdd(pr + 5, (unsigned long long)n_p2);
    }
#endif
}

}  // namespace

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
                          uint64_t out_bytes, int32_t* status, void* scratch, size_t scratch_bytes, cudaStream_t st) {
    (void)comp_bytes;
    if (n == 0) return cudaSuccess;
    if (scratch_bytes < decode_scratch_bytes(n, out_bytes) || n >= 0xffffffffull) return cudaErrorInvalidValue;
    uint8_t* sb = static_cast<uint8_t*>(scratch);
    uint32_t* ticket = reinterpret_cast<uint32_t*>(sb);
    uint8_t* tmp = sb + 256;
    cudaError_t e = cudaMemsetAsync(ticket, 0, 256, st);
    if (e != cudaSuccess) return e;
    // Slots (chunks in flight) per warp: few when the batch is small, so that every SM gets many warps (the token walk
    // and the copy rounds are latency chains; only other warps hide them); up to 32 for large batches, where the
    // serial token walk is then spread over all lanes.  Chunks are handed out longest-first through the ticket
    // counter, and for batches larger than the machine the launch keeps about half as many slots as chunks: a slot
    // that finishes a long chunk picks up a short one, so the tail stays balanced.  Measured on B200 (profiles/): 8.5k text chunks -> 2 slots, 15.7k
    // bf16 chunks -> 4 slots.  ZEST_LZ4_SLOTS / ZEST_LZ4_WARPS / ZEST_LZ4_OVERSUB override (tuning knobs).
    size_t want_warps = 148 * 16, max_warps = 148 * 24;
    size_t oversub = n > max_warps ? 2 : 1;  // a batch that fits the machine at one chunk per slot runs at single-chunk latency
    if (const char* e = getenv("ZEST_LZ4_WARPS")) { long v = atol(e); if (v > 0) want_warps = max_warps = (size_t)v; }
    if (const char* e = getenv("ZEST_LZ4_OVERSUB")) { long v = atol(e); if (v > 0) oversub = (size_t)v; }
    uint32_t slots = (uint32_t)((n + want_warps * oversub - 1) / (want_warps * oversub));
    if (const char* e = getenv("ZEST_LZ4_SLOTS")) { long v = atol(e); if (v > 0) slots = (uint32_t)v; }
    if (slots < 1) slots = 1;
    if (slots > 32) slots = 32;
    size_t warps = (n + slots * oversub - 1) / (slots * oversub);
    if (warps > max_warps) warps = max_warps;
    if (warps < 1) warps = 1;
    unsigned blocks = (unsigned)((warps + kWarps - 1) / kWarps);
    lz4_decode_kernel<<<blocks, kWarps * 32, 0, st>>>(comp, chunks, (uint32_t)n, out, tmp, status, ticket, slots);
    return cudaGetLastError();
}

}  // namespace zg
