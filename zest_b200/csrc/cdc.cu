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

// Serial part: min/max rule over the sparse candidate stream.  One warp; all lanes hold the same walker state,
// lanes load 32 tile records at a time.
struct Walker {
    uint64_t s;      // current chunk start
    uint64_t n;      // cuts emitted
    uint64_t cap;
    uint64_t* ends;
    bool overflow;
    __device__ void emit(uint64_t e, uint32_t lane) {
        if (n < cap) { if (lane == 0) ends[n] = e; }
        else overflow = true;
        n++;
        s = e;
    }
    __device__ void candidate(uint64_t e, uint32_t lane) {  // e = end offset (position + 1)
        while (e - s > kMax) emit(s + kMax, lane);
        if (e - s >= kMin) emit(e, lane);
    }
};

__device__ __forceinline__ void load_tile_rec(const TileRec* __restrict__ tiles, uint64_t t, uint64_t n_tiles, uint4& a, uint4& b) {
    a = make_uint4(0, 0, 0, 0);
    b = a;
    if (t < n_tiles) {
        const uint4* q = reinterpret_cast<const uint4*>(tiles + t);
        a = __ldg(q);
        b = __ldg(q + 1);
    }
}

__global__ void __launch_bounds__(32) cdc_walk_kernel(const uint8_t* __restrict__ data, uint64_t len,
                                                      const TileRec* __restrict__ tiles, uint64_t n_tiles,
                                                      uint64_t* __restrict__ ends, uint64_t cap,
                                                      uint64_t* __restrict__ n_out) {
    __shared__ uint64_t s_cand[32 * kListCap];
    const uint32_t lane = threadIdx.x;
    Walker w{0, 0, cap, ends, false};
    uint4 na, nb;
    load_tile_rec(tiles, lane, n_tiles, na, nb);
    for (uint64_t t0 = 0; t0 < n_tiles; t0 += 32) {
        const uint4 a = na, b = nb;
        load_tile_rec(tiles, t0 + 32 + lane, n_tiles, na, nb);  // next batch's records are in flight while this one is walked
        const uint32_t cnt = a.x & 0xffffu;
        const uint32_t ww[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        const uint64_t tbase = (t0 + lane) * (uint64_t)kTileBytes;
        const uint32_t dense = __ballot_sync(0xffffffffu, cnt > (uint32_t)kListCap);
        if (!dense) {
            // compact the batch's candidates (ascending by construction) into shared memory, then apply the rule
            uint32_t x = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
                if (lane >= (uint32_t)d) x += y;
            }
            const uint32_t total = __shfl_sync(0xffffffffu, x, 31);
            uint32_t o = x - cnt;
#pragma unroll
            for (int i = 0; i < kListCap; i++)
                if ((uint32_t)i < cnt) s_cand[o + i] = tbase + ((ww[(i + 2) >> 1] >> (16 * ((i + 2) & 1))) & 0xffffu) + 1;
            __syncwarp();
            for (uint32_t j = 0; j < total; j++) w.candidate(s_cand[j], lane);
            __syncwarp();
            continue;
        }
        // a dense tile in the batch: take the tiles one by one
        uint32_t nz = __ballot_sync(0xffffffffu, cnt != 0);
        while (nz) {
            int l = __ffs(nz) - 1;
            nz &= nz - 1;
            uint32_t c = __shfl_sync(0xffffffffu, cnt, l);
            uint64_t tb = (t0 + l) * (uint64_t)kTileBytes;
            if (c <= (uint32_t)kListCap) {
#pragma unroll
                for (int i = 0; i < kListCap; i++) {
                    uint32_t p = __shfl_sync(0xffffffffu, (ww[(i + 2) >> 1] >> (16 * ((i + 2) & 1))) & 0xffffu, l);
                    if ((uint32_t)i < c) w.candidate(tb + p + 1, lane);
                }
            } else {
                // dense tile: roll the hash again, but only from where a cut may next fall
                uint64_t tend = tb + kTileBytes < len ? tb + kTileBytes : len;
                uint64_t p = tb;
                while (p < tend) {
                    while (p + 1 - w.s > kMax) w.emit(w.s + kMax, lane);
                    if (p + 1 - w.s < kMin) p = w.s + kMin - 1;
                    if (p >= tend) break;
                    uint64_t h = 0;
                    for (uint64_t q = p >= 63 ? p - 63 : 0; q <= p; q++) h = (h << 1) + kGear[data[q]];
                    for (;;) {
                        if ((h >> 48) == 0) { w.candidate(p + 1, lane); p++; break; }
                        p++;
                        if (p >= tend) break;
                        if (p + 1 - w.s > kMax) break;
                        h = (h << 1) + kGear[data[p]];
                    }
                }
            }
        }
    }
    while (len - w.s > kMax) w.emit(w.s + kMax, lane);
    if (len > w.s) w.emit(len, lane);
    if (lane == 0) { n_out[0] = w.n; n_out[1] = w.overflow ? 1 : 0; }
}

}  // namespace

size_t cdc_scratch_bytes(uint64_t len) {
    uint64_t n_tiles = (len + kTileBytes - 1) / kTileBytes;
    return align_up((size_t)(n_tiles + 1) * sizeof(TileRec), 256);
}

cudaError_t launch_cdc(const uint8_t* data, uint64_t len, uint64_t* ends, uint64_t cap, uint64_t* n_out,
                       void* scratch, size_t scratch_bytes, cudaStream_t st) {
    uint64_t n_tiles = (len + kTileBytes - 1) / kTileBytes;
    if (scratch_bytes < cdc_scratch_bytes(len) || n_tiles > 0x7fffffffull) return cudaErrorInvalidValue;
    TileRec* tiles = static_cast<TileRec*>(scratch);
    if (n_tiles) cdc_scan_kernel<<<(unsigned)n_tiles, kThreads, 0, st>>>(data, len, tiles);
    cdc_walk_kernel<<<1, 32, 0, st>>>(data, len, tiles, n_tiles, ends, cap, n_out);
    return cudaGetLastError();
}

}  // namespace zg
