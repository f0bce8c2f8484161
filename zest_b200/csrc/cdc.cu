// Gearhash content-defined chunking on sm_100a (seed / dedup path).
//
// Reference behaviour replaced: zig-xet `chunking` (Gearhash CDC, 8 KiB min / 64 KiB target / 128 KiB max --
// /root/reference/src/config.zig:12, DESIGN.md:267-273).  zest has no call site for it today; the
// sink is storage.writeChunk / XorbCache.put (/root/reference/src/storage.zig:102-107, src/swarm.zig:57-148).
// Algorithm: SURVEY.md A.5; oracle: oracle/xet_oracle.c (oz_cdc).
//
// Parallel formulation.  h = (h << 1) + T[b] over u64 forgets a byte after 64 steps, and a cut is only legal 8192
// bytes into a chunk, so "position p satisfies h & MASK == 0" depends on bytes p-63..p alone: candidates are a
// property of the position, not of the chunking.  Pass 1 (all SMs) evaluates the windowed hash at every position:
// a CTA stages a 32 KiB tile (+64-byte halo) in shared memory with coalesced 16-byte loads, each thread rolls
// through its own 256-byte segment after a 64-byte warm-up, and the rare hits (2^-16 per byte) are written as
// a short sorted list per tile.  Pass 2 (one warp) applies the min/max rule to the sparse candidate stream, which
// is the only serial part.  A tile with more hits than its list holds is marked dense and re-scanned by pass 2 only
// where a cut is actually needed, so the result is exact for every input.
//
// Roofline: 1 byte read per input byte (HBM-bound in principle); in practice bounded by the shared-memory table
// look-up (one random 8-byte LDS per byte).
#include "../../include/zest_gear_table.h"
#include "kernels.h"

namespace zg {

namespace {

constexpr uint32_t kMin = 8192, kMax = 131072;
constexpr int kThreads = 128;
constexpr int kSeg = 256;                    // bytes per thread
constexpr int kTileBytes = kThreads * kSeg;  // 32 KiB
constexpr int kRow = kSeg + 16;              // padded row: 16-byte reads of 8 consecutive lanes hit distinct banks
constexpr int kListCap = 14;                 // hits kept per tile (u16 each) next to the count -> 32 bytes per tile

__device__ __constant__ const uint64_t kGear[256] = ZEST_GEAR_TABLE_INIT;

struct TileRec {  // 32 bytes
    uint16_t count;  // hits in the tile; > kListCap means "dense, re-scan"
    uint16_t pad;
    uint16_t pos[kListCap];  // ascending positions relative to the tile start
};

__global__ void __launch_bounds__(kThreads) cdc_scan_kernel(const uint8_t* __restrict__ data, uint64_t len,
                                                            TileRec* __restrict__ tiles) {
    __shared__ uint64_t s_gear[256];
    __shared__ __align__(16) uint8_t s_data[(kThreads + 1) * kRow];  // row 0 = halo, row t+1 = thread t's segment
    __shared__ uint32_t s_count;
    __shared__ uint16_t s_hits[64];
    const uint32_t tid = threadIdx.x;
    const uint64_t tile0 = (uint64_t)blockIdx.x * kTileBytes;
    for (int i = tid; i < 256; i += kThreads) s_gear[i] = kGear[i];
    if (tid == 0) s_count = 0;
    // stage [tile0 - 64, tile0 + 32 KiB): 16 bytes per thread per step, zero outside the buffer
    const uint64_t len16 = (len + 15) & ~15ull;  // the buffer is readable up to the next multiple of 16
    for (int v = tid; v < (kTileBytes + 64) / 16; v += kThreads) {
        int64_t g = (int64_t)tile0 - 64 + 16 * (int64_t)v;
        uint4 x = make_uint4(0, 0, 0, 0);
        if (g >= 0 && (uint64_t)g < len16) x = __ldg(reinterpret_cast<const uint4*>(data + g));
        int rel = 16 * v - 64;  // relative to tile0
        int row = rel < 0 ? 0 : 1 + rel / kSeg;
        int col = rel < 0 ? kSeg + rel : rel % kSeg;
        *reinterpret_cast<uint4*>(&s_data[row * kRow + col]) = x;
    }
    __syncthreads();

    // warm-up: the 64 bytes before this thread's segment (tail of the previous row, or the halo row)
    uint64_t h = 0;
    const uint8_t* prev = &s_data[tid * kRow + kSeg - 64];
    const uint8_t* mine = &s_data[(tid + 1) * kRow];
#pragma unroll
    for (int v = 0; v < 4; v++) {
        uint4 x = *reinterpret_cast<const uint4*>(prev + 16 * v);
        uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
        for (int j = 0; j < 16; j++) h = (h << 1) + s_gear[(w[j >> 2] >> (8 * (j & 3))) & 0xff];
    }
    const uint64_t seg0 = tile0 + (uint64_t)tid * kSeg;
#pragma unroll 2
    for (int v = 0; v < kSeg / 16; v++) {
        uint4 x = *reinterpret_cast<const uint4*>(mine + 16 * v);
        uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
        for (int j = 0; j < 16; j++) {
            h = (h << 1) + s_gear[(w[j >> 2] >> (8 * (j & 3))) & 0xff];
            if ((uint32_t)(h >> 48) == 0) {
                uint64_t p = seg0 + 16 * v + j;
                if (p < len) {
                    uint32_t k = atomicAdd(&s_count, 1u);
                    if (k < 64) s_hits[k] = (uint16_t)(p - tile0);
                }
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t c = s_count;
        TileRec r;
        r.count = (uint16_t)(c > 0xffff ? 0xffff : c);
        r.pad = 0;
        for (int i = 0; i < kListCap; i++) r.pos[i] = 0;
        if (c <= (uint32_t)kListCap) {
            for (uint32_t i = 1; i < c; i++) {  // insertion sort of a handful of hits
                uint16_t x = s_hits[i];
                int j = (int)i - 1;
                while (j >= 0 && s_hits[j] > x) { s_hits[j + 1] = s_hits[j]; j--; }
                s_hits[j + 1] = x;
            }
            for (uint32_t i = 0; i < c; i++) r.pos[i] = s_hits[i];
        }
        tiles[blockIdx.x] = r;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Pass 2: the min/max rule over the sparse candidate stream.  The rule is a chain (each cut depends on the previous
// one), but the chain forgets quickly: a walker started at an arbitrary position agrees with the true one as soon as
// both cut at the same candidate, usually within one or two chunks.  So the tile list is split into segments; every
// segment is walked speculatively (as if a cut fell exactly on its first byte) by its own thread; a second parallel
// pass walks every segment from the state its predecessor's speculation ends in until one of its cuts equals one
// of the segment's own speculative cuts, and adopts the rest of that list; one warp checks that every predecessor did
// join (else it repairs serially from there) and a last kernel gathers the pieces in order.
// ---------------------------------------------------------------------------------------------------------------

struct SegInfo {       // 32 bytes per segment
    uint32_t spec_n;   // speculative cuts found
    uint32_t cand_n;   // candidates seen in the segment (0 => only forced cuts, handled arithmetically)
    uint32_t dense;    // a tile overflowed its hit list: the stitcher walks this segment itself
    uint32_t join;     // stitcher: index in the speculative list from which it is adopted
    uint32_t priv_n;   // stitcher: cuts before the join that differ from the speculation
    uint32_t pad;
    uint64_t base;     // stitcher: position of the segment's first cut in the output
};

struct CutSink {  // where a walker puts its cuts
    uint64_t* buf;
    uint32_t n, cap;
    bool overflow;
    __device__ void put(uint64_t e) {
        if (n < cap) buf[n] = e; else overflow = true;
        n++;
    }
};

__device__ __forceinline__ uint32_t tile_pos(const uint4& a, const uint4& b, int i) {
    const uint32_t ww[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    return (ww[(i + 2) >> 1] >> (16 * ((i + 2) & 1))) & 0xffffu;
}

// Apply the rule to one candidate end offset e for a walker in state s; cuts go to `put`.
template <typename Put>
__device__ __forceinline__ void rule_candidate(uint64_t& s, uint64_t e, Put&& put) {
    while (e - s > kMax) { s += kMax; put(s); }
    if (e - s >= kMin) { s = e; put(s); }
}

__global__ void __launch_bounds__(128) cdc_spec_kernel(const TileRec* __restrict__ tiles, uint64_t n_tiles,
                                                       uint64_t len, uint32_t seg_tiles, uint32_t n_segs,
                                                       uint32_t cap, SegInfo* __restrict__ segs,
                                                       uint64_t* __restrict__ spec) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_segs) return;
    const uint64_t t0 = (uint64_t)k * seg_tiles, t1 = t0 + seg_tiles < n_tiles ? t0 + seg_tiles : n_tiles;
    const uint64_t seg_end = t1 * (uint64_t)kTileBytes < len ? t1 * (uint64_t)kTileBytes : len;
    uint64_t s = t0 * (uint64_t)kTileBytes;
    CutSink sink{spec + (size_t)k * cap, 0, cap, false};
    uint32_t cand = 0, dense = 0;
    for (uint64_t t = t0; t < t1; t++) {
        const uint4* q = reinterpret_cast<const uint4*>(tiles + t);
        const uint4 a = __ldg(q);
        const uint32_t c = a.x & 0xffffu;
        if (c == 0) continue;
        if (c > (uint32_t)kListCap) { dense = 1; break; }
        const uint4 b = __ldg(q + 1);
        cand += c;
        const uint64_t tb = t * (uint64_t)kTileBytes;
        for (uint32_t i = 0; i < c; i++) rule_candidate(s, tb + tile_pos(a, b, (int)i) + 1, [&](uint64_t e) { sink.put(e); });
    }
    if (!dense) while (seg_end - s >= kMax && seg_end > s) { s += kMax; sink.put(s); }
    SegInfo r;
    r.spec_n = dense || sink.overflow ? 0u : sink.n;
    r.cand_n = cand;
    r.dense = dense || sink.overflow;
    r.join = 0; r.priv_n = 0; r.pad = 0; r.base = 0;
    segs[k] = r;
}

// True walk of tiles [t0, t1) from state s, with dense tiles re-scanned; every cut goes to `put`, which returns true
// to stop (the stitcher stops at the join).  Returns true if stopped.
template <typename Put>
__device__ bool walk_tiles(const uint8_t* __restrict__ data, uint64_t len, const TileRec* __restrict__ tiles,
                           uint64_t t0, uint64_t t1, uint64_t& s, Put&& put) {
    bool stop = false;
    for (uint64_t t = t0; t < t1 && !stop; t++) {
        const uint4* q = reinterpret_cast<const uint4*>(tiles + t);
        const uint4 a = __ldg(q);
        const uint32_t c = a.x & 0xffffu;
        if (c == 0) continue;
        const uint64_t tb = t * (uint64_t)kTileBytes;
        if (c <= (uint32_t)kListCap) {
            const uint4 b = __ldg(q + 1);
            for (uint32_t i = 0; i < c && !stop; i++)
                rule_candidate(s, tb + tile_pos(a, b, (int)i) + 1, [&](uint64_t e) { if (!stop) stop = put(e); });
        } else {
            // dense tile: roll the hash again, but only from where a cut may next fall
            const uint64_t tend = tb + kTileBytes < len ? tb + kTileBytes : len;
            uint64_t p = tb;
            while (p < tend && !stop) {
                while (p + 1 - s > kMax && !stop) { s += kMax; stop = put(s); }
                if (stop) break;
                if (p + 1 - s < kMin) p = s + kMin - 1;
                if (p >= tend) break;
                uint64_t h = 0;
                for (uint64_t x = p >= 63 ? p - 63 : 0; x <= p; x++) h = (h << 1) + kGear[data[x]];
                for (;;) {
                    if ((h >> 48) == 0) {
                        rule_candidate(s, p + 1, [&](uint64_t e) { if (!stop) stop = put(e); });
                        p++;
                        break;
                    }
                    p++;
                    if (p >= tend) break;
                    if (p + 1 - s > kMax) break;
                    h = (h << 1) + kGear[data[p]];
                }
            }
        }
    }
    return stop;
}

// The true walk of segment k from entry state s: cuts that differ from the segment's speculation go to its private
// list until one coincides with a speculative cut (the join); from there the speculation is adopted.  Fills
// g.join / g.priv_n / g.pad (1 = joined) and returns the state after the segment.
__device__ uint64_t stitch_segment(const uint8_t* __restrict__ data, uint64_t len, const TileRec* __restrict__ tiles,
                                   uint64_t n_tiles, uint32_t seg_tiles, uint32_t cap, uint32_t k, uint64_t s, SegInfo& g,
                                   const uint64_t* __restrict__ spec, uint64_t* __restrict__ priv, bool& overflow) {
    const uint64_t t0 = (uint64_t)k * seg_tiles, t1 = t0 + seg_tiles < n_tiles ? t0 + seg_tiles : n_tiles;
    const uint64_t seg_end = t1 * (uint64_t)kTileBytes < len ? t1 * (uint64_t)kTileBytes : len;
    const uint64_t* sp = spec + (size_t)k * cap;
    CutSink pv{priv + (size_t)k * cap, 0, cap, false};
    uint32_t j = 0;
    bool joined = false;
    auto put = [&](uint64_t e) -> bool {
        if (!g.dense) {
            while (j < g.spec_n && sp[j] < e) j++;
            if (j < g.spec_n && sp[j] == e) { joined = true; return true; }
        }
        pv.put(e);
        return false;
    };
    if (s == t0 * (uint64_t)kTileBytes && !g.dense) {
        joined = true;  // the speculation started from the true state (always so for segment 0)
        j = 0;
    } else {
        if (g.cand_n != 0 || g.dense) walk_tiles(data, len, tiles, t0, t1, s, put);
        if (!joined) while (seg_end - s >= kMax && seg_end > s && !joined) { s += kMax; if (put(s)) break; }
    }
    g.priv_n = pv.n;
    g.join = joined ? j : g.spec_n;
    g.pad = joined ? 1u : 0u;
    if (pv.overflow) overflow = true;
    if (joined && g.spec_n > g.join) s = sp[g.spec_n - 1];
    return s;
}

// Every segment's true walk at once, each entered with the state its predecessor's speculation ends in.  That guess is
// right whenever the predecessor's own true walk joins its speculation -- which is the normal case -- and then the
// exit state of a segment does not depend on how it was entered.
__global__ void __launch_bounds__(128) cdc_join_kernel(const uint8_t* __restrict__ data, uint64_t len,
                                                       const TileRec* __restrict__ tiles, uint64_t n_tiles,
                                                       uint32_t seg_tiles, uint32_t n_segs, uint32_t cap,
                                                       SegInfo* __restrict__ segs, const uint64_t* __restrict__ spec,
                                                       uint64_t* __restrict__ priv) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_segs) return;
    SegInfo g = segs[k];
    uint64_t s = 0;
    if (k > 0) {
        const SegInfo p = segs[k - 1];  // only spec_n / dense are read, which cdc_spec_kernel wrote
        s = (p.dense || p.spec_n == 0) ? (uint64_t)(k - 1) * seg_tiles * (uint64_t)kTileBytes
                                       : spec[(size_t)(k - 1) * cap + p.spec_n - 1];
    }
    bool overflow = false;
    const uint64_t exit_s = stitch_segment(data, len, tiles, n_tiles, seg_tiles, cap, k, s, g, spec, priv, overflow);
    g.base = exit_s;  // carried to the finishing kernel, which turns it into the output position
    if (overflow) g.pad |= 2u;
    // join / priv_n / pad / base only: spec_n, cand_n and dense are read by the neighbour's thread
    segs[k].join = g.join; segs[k].priv_n = g.priv_n; segs[k].pad = g.pad; segs[k].base = g.base;
}

// One warp: validate the chain of guesses (segment k's guess holds iff segment k-1 joined), repair serially from
// the first segment that did not join (pathological inputs only), and turn counts into output positions.
__global__ void __launch_bounds__(32) cdc_finish_kernel(const uint8_t* __restrict__ data, uint64_t len,
                                                        const TileRec* __restrict__ tiles, uint64_t n_tiles,
                                                        uint32_t seg_tiles, uint32_t n_segs, uint32_t cap,
                                                        SegInfo* __restrict__ segs, const uint64_t* __restrict__ spec,
                                                        uint64_t* __restrict__ priv, uint64_t* __restrict__ n_out,
                                                        uint64_t out_cap) {
    const uint32_t lane = threadIdx.x;
    // first segment whose walk did not join (its successor's guess is then wrong)
    uint32_t first_bad = n_segs;
    for (uint32_t k0 = 0; k0 < n_segs && first_bad == n_segs; k0 += 32) {
        const uint32_t k = k0 + lane;
        const bool bad = k < n_segs && (segs[k].pad & 1u) == 0;
        const uint32_t m = __ballot_sync(0xffffffffu, bad);
        if (m) first_bad = k0 + (uint32_t)__ffs(m) - 1;
    }
    bool overflow = false;
    uint64_t s = 0;
    const uint32_t n_ok = first_bad < n_segs ? first_bad + 1 : n_segs;  // segments [0, n_ok) were entered correctly
    if (lane == 0) {
        if (n_ok > 0) s = segs[n_ok - 1].base;  // exit state of the last correctly entered segment
        for (uint32_t k = n_ok; k < n_segs; k++) {  // serial repair
            SegInfo g = segs[k];
            s = stitch_segment(data, len, tiles, n_tiles, seg_tiles, cap, k, s, g, spec, priv, overflow);
            segs[k] = g;
        }
    }
    __syncwarp();
    // output positions: exclusive prefix sum of per-segment counts (32 segments per pass)
    uint64_t total = 0;
    for (uint32_t k0 = 0; k0 < n_segs; k0 += 32) {
        const uint32_t k = k0 + lane;
        uint64_t c = 0;
        if (k < n_segs) { c = (uint64_t)segs[k].priv_n + (segs[k].spec_n - segs[k].join); if (segs[k].pad & 2u) overflow = true; }
        uint64_t x = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint64_t y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= (uint32_t)d) x += y;
        }
        if (k < n_segs) segs[k].base = total + x - c;
        total += __shfl_sync(0xffffffffu, x, 31);
    }
    const uint32_t any_over = __ballot_sync(0xffffffffu, overflow);
    if (lane == 0) {
        // whatever is left after the last cut is the final (possibly short) chunk
        uint64_t extra = 0;
        uint64_t* tail = priv + (size_t)n_segs * cap;  // room for one more entry
        if (len > s) { tail[0] = len; extra = 1; }
        n_out[0] = total + extra;
        n_out[1] = (any_over || total + extra > out_cap) ? 1 : 0;
        n_out[2] = total;  // where the tail entry goes
    }
}

__global__ void __launch_bounds__(128) cdc_gather_kernel(const SegInfo* __restrict__ segs, uint32_t n_segs, uint32_t cap,
                                                         const uint64_t* __restrict__ spec,
                                                         const uint64_t* __restrict__ priv,
                                                         const uint64_t* __restrict__ n_out, uint64_t* __restrict__ ends,
                                                         uint64_t out_cap) {
    const uint32_t k = blockIdx.x;
    if (k == n_segs) {  // the tail entry
        if (threadIdx.x == 0 && n_out[0] > n_out[2] && n_out[2] < out_cap) ends[n_out[2]] = priv[(size_t)n_segs * cap];
        return;
    }
    const SegInfo g = segs[k];
    const uint64_t* pv = priv + (size_t)k * cap;
    const uint64_t* sp = spec + (size_t)k * cap;
    const uint32_t take = g.spec_n - g.join;
    for (uint32_t i = threadIdx.x; i < g.priv_n + take; i += blockDim.x) {
        const uint64_t o = g.base + i;
        if (o < out_cap) ends[o] = i < g.priv_n ? pv[i] : sp[g.join + (i - g.priv_n)];
    }
}

struct CdcLayout { size_t tiles, segs, spec, priv, total; uint32_t seg_tiles, n_segs, cap; uint64_t n_tiles; };
CdcLayout cdc_layout(uint64_t len) {
    CdcLayout L;
    L.n_tiles = (len + kTileBytes - 1) / kTileBytes;
    uint64_t st = (L.n_tiles + 1023) / 1024;  // <= 1024 segments (the stitcher is serial over them), >= 2 MiB each
    if (st < 64) st = 64;
    L.seg_tiles = (uint32_t)st;
    L.n_segs = (uint32_t)((L.n_tiles + st - 1) / st);
    L.cap = L.seg_tiles * (kTileBytes / kMin) + 4;
    size_t o = 0;
    L.tiles = o; o = align_up(o + (size_t)(L.n_tiles + 1) * sizeof(TileRec), 256);
    L.segs = o; o = align_up(o + (size_t)(L.n_segs + 1) * sizeof(SegInfo), 256);
    L.spec = o; o = align_up(o + ((size_t)L.n_segs * L.cap + 1) * 8, 256);
    L.priv = o; o = align_up(o + ((size_t)L.n_segs * L.cap + 2) * 8, 256);
    L.total = o;
    return L;
}

}  // namespace

size_t cdc_scratch_bytes(uint64_t len) { return cdc_layout(len).total; }

cudaError_t launch_cdc(const uint8_t* data, uint64_t len, uint64_t* ends, uint64_t cap, uint64_t* n_out,
                       void* scratch, size_t scratch_bytes, cudaStream_t st) {
    CdcLayout L = cdc_layout(len);
    if (scratch_bytes < L.total || L.n_tiles > 0x7fffffffull) return cudaErrorInvalidValue;
    uint8_t* sb = static_cast<uint8_t*>(scratch);
    TileRec* tiles = reinterpret_cast<TileRec*>(sb + L.tiles);
    SegInfo* segs = reinterpret_cast<SegInfo*>(sb + L.segs);
    uint64_t* spec = reinterpret_cast<uint64_t*>(sb + L.spec);
    uint64_t* priv = reinterpret_cast<uint64_t*>(sb + L.priv);
    if (L.n_tiles) {
        cdc_scan_kernel<<<(unsigned)L.n_tiles, kThreads, 0, st>>>(data, len, tiles);
        cdc_spec_kernel<<<(L.n_segs + 127) / 128, 128, 0, st>>>(tiles, L.n_tiles, len, L.seg_tiles, L.n_segs, L.cap, segs, spec);
    }
    if (L.n_segs) cdc_join_kernel<<<(L.n_segs + 127) / 128, 128, 0, st>>>(data, len, tiles, L.n_tiles, L.seg_tiles, L.n_segs, L.cap, segs, spec, priv);
    cdc_finish_kernel<<<1, 32, 0, st>>>(data, len, tiles, L.n_tiles, L.seg_tiles, L.n_segs, L.cap, segs, spec, priv, n_out, cap);
    cdc_gather_kernel<<<L.n_segs + 1, 128, 0, st>>>(segs, L.n_segs, L.cap, spec, priv, n_out, ends, cap);
    return cudaGetLastError();
}

}  // namespace zg
