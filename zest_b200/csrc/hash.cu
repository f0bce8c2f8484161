// Xet chunk hashing on sm_100a: BLAKE3 (keyed or plain) of many variable-length chunks in one pass.
//
// Reference behaviour replaced: zig-xet `hashing` (chunk hash = BLAKE3.keyed(DATA_KEY, bytes)), reached from
// /root/reference/src/xet_bridge.zig:256-257 and src/parallel_download.zig:65-66; plain mode reproduces
// src/bench.zig:207-223.  Algorithm: SURVEY.md A.2; oracle: oracle/xet_oracle.c (b3_hash).
//
// Mapping.  BLAKE3 is a chain of 16 compressions inside every 1 KiB leaf and a binary tree above the leaves, so
// the parallel unit is the leaf.  One THREAD owns an aligned group of G leaves (G = 1, 2, 4 or 8 KiB of
// contiguous input), chains the 16 x G block compressions with state and message words in registers, and folds
// its own complete sub-tree with a small chaining-value stack -- so 32 lanes stay busy through the first log2(G)
// tree levels.  Group nodes (32 B per G KiB, the only scratch traffic) go to a second, tiny kernel that finishes
// each chunk's tree with one thread per chunk (lanes stay dense across chunks) or one warp for very large
// chunks.  Lanes are packed densely over chunks of any size by a prefix sum over groups-per-chunk.
//
// Roofline: integer-ALU bound.  A block costs 7 x 8 x 12 + 8 = 680 ALU-pipe instructions (IADD3/LOP3/SHF/PRMT);
// the input is read exactly once (1 B/B of HBM traffic, >= 4x below the HBM roof at ALU speed).
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "../../include/zest_gpu.h"
#include "kernels.h"

namespace zg {

namespace {

constexpr int kLeaf = 1024;
constexpr int kTile = 1024;  // chunks per scan tile

__device__ __forceinline__ uint32_t groups_of(uint64_t len, int group_bytes) {
    return len == 0 ? 1u : (uint32_t)((len + group_bytes - 1) / group_bytes);
}
__device__ __forceinline__ uint64_t chunk_len(const uint64_t* off, const uint32_t* len, size_t i) {
    return len ? (uint64_t)len[i] : off[i + 1] - off[i];
}

// ---- job set-up: P[i] = first group slot of chunk i (exclusive prefix sum), map[slot] = chunk -------------
__global__ void __launch_bounds__(kTile) hash_count_kernel(const uint64_t* __restrict__ off,
                                                           const uint32_t* __restrict__ len, size_t n,
                                                           int group_bytes, uint32_t* __restrict__ P,
                                                           uint32_t* __restrict__ tile_sum) {
    __shared__ uint32_t warp_tot[32];
    size_t i = (size_t)blockIdx.x * kTile + threadIdx.x;
    uint32_t m = i < n ? groups_of(chunk_len(off, len, i), group_bytes) : 0u;
    uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5, x = m;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
    }
    if (lane == 31) warp_tot[w] = x;
    __syncthreads();
    if (w == 0) {
        uint32_t t = warp_tot[lane], s = t;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, s, d);
            if (lane >= d) s += y;
        }
        warp_tot[lane] = s - t;
        if (lane == 31) tile_sum[blockIdx.x] = s;
    }
    __syncthreads();
    if (i < n) P[i] = warp_tot[w] + x - m;
}

__global__ void __launch_bounds__(1024) hash_tiles_kernel(uint32_t* __restrict__ tile_sum, uint32_t n_tiles,
                                                          uint32_t slots_cap, uint32_t* __restrict__ total_out) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry_s;
    uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_tiles; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint32_t carry = carry_s;  // stable here: the previous iteration ended with a barrier
        uint32_t m = i < n_tiles ? tile_sum[i] : 0u, x = m;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) warp_tot[w] = x;
        __syncthreads();
        if (w == 0) {
            uint32_t t = warp_tot[lane], s = t;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                uint32_t y = __shfl_up_sync(0xffffffffu, s, d);
                if (lane >= d) s += y;
            }
            warp_tot[lane] = s - t;
            if (lane == 31) carry_s = carry + s;
        }
        __syncthreads();
        if (i < n_tiles) tile_sum[i] = carry + warp_tot[w] + x - m;
        __syncthreads();
    }
    // The leaf kernel runs min(total, slots_cap) threads: slots beyond the scratch the caller sized (lengths that
    // add up to more than data_bytes) are never touched; the chunks that own them come out as zero hashes.
    if (threadIdx.x == 0) {
        total_out[0] = carry_s < slots_cap ? carry_s : slots_cap;
        total_out[1] = carry_s;  // the last chunk's end, for the tree kernel
    }
}

__global__ void __launch_bounds__(256) hash_map_kernel(const uint64_t* __restrict__ off,
                                                       const uint32_t* __restrict__ len, size_t n, int group_bytes,
                                                       uint32_t* __restrict__ P,
                                                       const uint32_t* __restrict__ tile_sum,
                                                       uint32_t slots_cap, uint32_t* __restrict__ map,
                                                       uint4* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    uint32_t base = P[i] + tile_sum[i / kTile];
    P[i] = base;
    uint32_t m = groups_of(chunk_len(off, len, i), group_bytes);
    if (base > slots_cap || m > slots_cap - base) {  // does not fit the scratch: no slots, zero hash (see hash_tiles_kernel)
        out[2 * i] = make_uint4(0, 0, 0, 0);
        out[2 * i + 1] = make_uint4(0, 0, 0, 0);
        for (uint32_t t = base; t < slots_cap; t++) map[t] = 0xffffffffu;  // only the first such chunk starts below the cap
        return;
    }
    for (uint32_t k = 0; k < m; k++) map[base + k] = (uint32_t)i;
}

// ---- message loads ----------------------------------------------------------------------------------------
__device__ __forceinline__ void load_block_aligned(const uint8_t* p, uint32_t (&m)[16]) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
    m[0] = a.x; m[1] = a.y; m[2] = a.z; m[3] = a.w;
    m[4] = b.x; m[5] = b.y; m[6] = b.z; m[7] = b.w;
    m[8] = c.x; m[9] = c.y; m[10] = c.z; m[11] = c.w;
    m[12] = d.x; m[13] = d.y; m[14] = d.z; m[15] = d.w;
}

// Any alignment, any length 0..64; bytes past `blen` read as zero.  Only the 16-byte granules that hold bytes
// of the block are touched, so nothing outside the (16-byte padded) buffer is read.
__device__ __noinline__ void load_block_general(const uint8_t* p, uint32_t blen, uint32_t (&m)[16]) {
    uintptr_t a = reinterpret_cast<uintptr_t>(p);
    uint32_t s = (uint32_t)(a & 15);
    const uint4* q = reinterpret_cast<const uint4*>(a - s);
    uint32_t w[21];
    uint32_t span = s + blen;  // bytes of the 80-byte window that matter
#pragma unroll
    for (int g = 0; g < 5; g++) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (blen > 0 && (uint32_t)(16 * g) < span) v = __ldg(q + g);
        w[4 * g] = v.x; w[4 * g + 1] = v.y; w[4 * g + 2] = v.z; w[4 * g + 3] = v.w;
    }
    w[20] = 0;
    if (s & 8) {
#pragma unroll
        for (int i = 0; i < 19; i++) w[i] = w[i + 2];
    }
    if (s & 4) {
#pragma unroll
        for (int i = 0; i < 18; i++) w[i] = w[i + 1];
    }
    uint32_t bs = (s & 3) * 8;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        uint32_t x = __funnelshift_r(w[i], w[i + 1], bs);
        int valid = (int)blen - 4 * i;
        uint32_t mask = valid >= 4 ? 0xffffffffu : (valid <= 0 ? 0u : ((1u << (8 * valid)) - 1u));
        m[i] = x & mask;
    }
}

// One 64-byte block of a chunk that is still BYTE-GROUPED (scheme 2 as the decoder leaves it when only the hash is wanted:
// plane k = the chunk's bytes at index = k mod 4, the four planes back to back).  Output byte 4j + k of the
// chunk is plane_k[j], so block bytes boff .. boff+63 are 16 consecutive bytes of each plane, interleaved: four unaligned
// 16-byte reads and a 4 x 4 byte transpose per group of four words -- ~80 instructions on top of a 784-instruction
// compression, against a separate regroup pass over the whole chunk in the decoder (17 % of its time on bf16).
// pl = the chunk's planes (16-byte aligned), n = the chunk's length, blen = bytes of the block that exist (zero beyond).
__device__ __forceinline__ void load_block_planar(const uint8_t* pl, uint32_t n, uint32_t boff, uint32_t blen, uint32_t (&m)[16]) {
    const uint32_t q = n >> 2, r = n & 3u;
    const uint32_t pstart[4] = {0u, q + (0u < r), 2u * q + (0u < r) + (1u < r), 3u * q + (0u < r) + (1u < r) + (2u < r)};
    const uint32_t j0 = boff >> 2;
    uint32_t w[4][4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        // 16 bytes of plane k from any alignment: the two aligned 16-byte vectors that hold them (the second one is the
        // next block's first: an L1 hit), shifted by whole words with two levels of selects and by bytes with a funnel
        // shift.  Reads stay inside the chunk's 16-byte padded storage plus the 16 bytes of slack behind the buffer.
        const uintptr_t a = reinterpret_cast<uintptr_t>(pl) + pstart[k] + j0;
        const uint4* vp = reinterpret_cast<const uint4*>(a & ~(uintptr_t)15);
        const uint32_t mis = (uint32_t)(a & 15), ws = mis >> 2, bs = (mis & 3u) * 8;
        const bool any = blen > (uint32_t)k;  // plane k has bytes in this block
        uint4 lo = make_uint4(0, 0, 0, 0), hi = make_uint4(0, 0, 0, 0);
        if (any) { lo = vp[0]; if (mis) hi = vp[1]; }
        uint32_t v[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
        if (ws & 2) {
#pragma unroll
            for (int i = 0; i < 6; i++) v[i] = v[i + 2];
        }
        if (ws & 1) {
#pragma unroll
            for (int i = 0; i < 5; i++) v[i] = v[i + 1];
        }
#pragma unroll
        for (int i = 0; i < 4; i++) w[k][i] = __funnelshift_r(v[i], v[i + 1], bs);
    }
#pragma unroll
    for (int g = 0; g < 4; g++) {
        const uint32_t a = w[0][g], b = w[1][g], c = w[2][g], d = w[3][g];
        const uint32_t t0 = __byte_perm(a, b, 0x5140), t1 = __byte_perm(a, b, 0x7362);
        const uint32_t u0 = __byte_perm(c, d, 0x5140), u1 = __byte_perm(c, d, 0x7362);
        m[4 * g] = __byte_perm(t0, u0, 0x5410); m[4 * g + 1] = __byte_perm(t0, u0, 0x7632);
        m[4 * g + 2] = __byte_perm(t1, u1, 0x5410); m[4 * g + 3] = __byte_perm(t1, u1, 0x7632);
    }
    if (blen < 64u) {  // the chunk ends inside this block: what lies beyond (the next plane's bytes, padding) reads as zero
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const int valid = (int)blen - 4 * i;
            m[i] &= valid >= 4 ? 0xffffffffu : (valid <= 0 ? 0u : ((1u << (8 * valid)) - 1u));
        }
    }
}

// ---- leaf kernel: one thread = one aligned group of G leaves ------------------------------------------------
template <int G, bool FMA_ADDS, bool PLANAR>
__global__ void __launch_bounds__(128) hash_leaf_kernel(const uint8_t* __restrict__ data,
                                                        const uint64_t* __restrict__ off,
                                                        const uint32_t* __restrict__ len, const B3Key key,
                                                        const uint32_t* __restrict__ P,
                                                        const uint32_t* __restrict__ map,
                                                        const uint32_t* __restrict__ total_slots,
                                                        uint4* __restrict__ nodes, uint4* __restrict__ out,
                                                        const uint8_t* __restrict__ planes,
                                                        const ChunkDesc* __restrict__ descs) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= *total_slots) return;
    uint32_t ci = map[t];
    if (ci == 0xffffffffu) return;  // slot of a chunk that did not fit the scratch (hash_map_kernel)
    uint32_t k = t - P[ci];
    uint64_t clen = chunk_len(off, len, ci);
    const bool planar = PLANAR && descs[ci].scheme == 2;  // still byte-grouped: read the planes, regroup on the fly
    const uint8_t* base = (planar ? planes : data) + off[ci];
    uint32_t n_leaves = clen == 0 ? 1u : (uint32_t)((clen + kLeaf - 1) / kLeaf);
    uint32_t first_leaf = k * G;
    uint32_t g = n_leaves - first_leaf < (uint32_t)G ? n_leaves - first_leaf : (uint32_t)G;
    bool single_group = n_leaves <= (uint32_t)G;  // this thread produces the chunk's final hash

    constexpr int kDepth = G == 1 ? 1 : (G == 2 ? 2 : (G == 4 ? 3 : 4));
    uint32_t stack[kDepth][8];
    int sp = 0;
    uint32_t cv[8];

#pragma unroll 1
    for (uint32_t l = 0; l < g; l++) {
        uint32_t leaf = first_leaf + l;
        uint64_t lo = (uint64_t)leaf * kLeaf;
        uint32_t llen = clen - lo < (uint64_t)kLeaf ? (uint32_t)(clen - lo) : (uint32_t)kLeaf;
        uint32_t nblk = llen == 0 ? 1u : (llen + 63) / 64;
        const uint8_t* p = base + lo;
        bool fast_align = (reinterpret_cast<uintptr_t>(p) & 15) == 0;
        uint32_t root_leaf = (n_leaves == 1) ? B3_ROOT : 0u;
#pragma unroll
        for (int i = 0; i < 8; i++) cv[i] = key.w[i];
#pragma unroll 1
        for (uint32_t b = 0; b < nblk; b++) {
            uint32_t blen = llen - 64 * b < 64 ? llen - 64 * b : 64;
            uint32_t m[16];
            if (planar) load_block_planar(base, (uint32_t)clen, (uint32_t)lo + 64 * b, blen, m);
            else if (fast_align && blen == 64) load_block_aligned(p + 64 * b, m);
            else load_block_general(p + 64 * b, blen, m);
            uint32_t fl = key.base_flags | (b == 0 ? B3_CHUNK_START : 0u) |
                          (b == nblk - 1 ? (B3_CHUNK_END | root_leaf) : 0u);
            b3_compress<FMA_ADDS>(cv, m, leaf, 0, blen, fl, key.one);
        }
        if (G == 1) break;
        // push, then merge completed sub-trees: after leaf l (count l+1), one merge per trailing zero bit; after
        // the last leaf of the group, merge everything that is left (right to left).
        bool last = (l == g - 1);
        uint32_t cnt = l + 1;
        while (sp > 0 && (last || (cnt & 1) == 0)) {
            uint32_t left[8];
#pragma unroll
            for (int i = 0; i < 8; i++) left[i] = stack[sp - 1][i];
            sp--;
            uint32_t fl = key.base_flags | B3_PARENT | ((last && sp == 0 && single_group) ? B3_ROOT : 0u);
            uint32_t r[8];
#pragma unroll
            for (int i = 0; i < 8; i++) r[i] = cv[i];
            b3_parent(cv, key.w, left, r, fl);
            cnt >>= 1;
        }
        if (!last) {
#pragma unroll
            for (int i = 0; i < 8; i++) stack[sp][i] = cv[i];
            sp++;
        }
    }
    uint4* dst = single_group ? (out + 2 * (size_t)ci) : (nodes + 2 * (size_t)t);
    dst[0] = make_uint4(cv[0], cv[1], cv[2], cv[3]);
    dst[1] = make_uint4(cv[4], cv[5], cv[6], cv[7]);
}

// ---- tree kernel: finish every chunk that has more than one group node -------------------------------------
__device__ __forceinline__ void load_cv(const uint4* p, uint32_t (&cv)[8]) {
    uint4 a = p[0], b = p[1];
    cv[0] = a.x; cv[1] = a.y; cv[2] = a.z; cv[3] = a.w;
    cv[4] = b.x; cv[5] = b.y; cv[6] = b.z; cv[7] = b.w;
}
__device__ __forceinline__ void store_cv(uint4* p, const uint32_t (&cv)[8]) {
    p[0] = make_uint4(cv[0], cv[1], cv[2], cv[3]);
    p[1] = make_uint4(cv[4], cv[5], cv[6], cv[7]);
}

// A team of T lanes (T = 1..32, power of two) folds one chunk's group nodes level by level, in place in the node
// array: pairs are combined left to right, an odd tail is promoted unchanged (== BLAKE3's left-full tree,
// SURVEY.md A.2), ROOT goes on the last parent.  Teams are packed side by side in a warp, so lanes stay busy across
// chunks; any node count works (a 16 MiB chunk just loops).
__global__ void __launch_bounds__(128) hash_tree_kernel(const B3Key key, const uint32_t* __restrict__ P, size_t n,
                                                        uint32_t team, uint32_t slots_cap, uint4* __restrict__ nodes,
                                                        uint4* __restrict__ out) {
    const uint32_t lane = threadIdx.x & 31;
    const size_t gt = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t ci = gt / team;
    const uint32_t lt = (uint32_t)(gt % team);
    const uint32_t tmask = (team == 32 ? 0xffffffffu : ((1u << team) - 1u)) << (lane & ~(team - 1));
    uint32_t base = 0, cnt = 0;
    if (ci < n) {
        base = P[ci];
        const uint32_t end = ci + 1 < n ? P[ci + 1] : *(P + n + 1);  // P[n + 1]: unclamped total
        cnt = (end <= slots_cap && end >= base) ? end - base : 0u;    // a chunk that did not fit has no nodes
    }
    uint4* nd = nodes + 2 * (size_t)base;
    while (cnt > 1) {
        const uint32_t pairs = cnt >> 1;
        const bool root = (cnt == 2);
        for (uint32_t p0 = 0; p0 < pairs; p0 += team) {
            const uint32_t p = p0 + lt;
            uint32_t l[8], r[8], o[8];
            if (p < pairs) { load_cv(nd + 2 * (size_t)(2 * p), l); load_cv(nd + 2 * (size_t)(2 * p + 1), r); }
            __syncwarp(tmask);
            if (p < pairs) {
                b3_parent(o, key.w, l, r, key.base_flags | B3_PARENT | (root ? B3_ROOT : 0u));
                store_cv(root ? out + 2 * ci : nd + 2 * (size_t)p, o);
            }
            __syncwarp(tmask);
        }
        if (cnt & 1) {
            if (lt == 0) { uint32_t t[8]; load_cv(nd + 2 * (size_t)(cnt - 1), t); store_cv(nd + 2 * (size_t)pairs, t); }
            __syncwarp(tmask);
        }
        cnt = pairs + (cnt & 1);
    }
}

struct HashLayout {
    size_t p_off, tiles_off, map_off, nodes_off, total;
    size_t n_tiles, slots_cap;
};
HashLayout hash_layout(uint64_t data_bytes, size_t n, int group) {
    HashLayout L;
    L.n_tiles = (n + kTile - 1) / kTile;
    L.slots_cap = (size_t)(data_bytes / ((uint64_t)kLeaf * group)) + n + 1;
    size_t o = 0;
    L.p_off = o; o = align_up(o + (n + 2) * 4, 256);
    L.tiles_off = o; o = align_up(o + (L.n_tiles + 1) * 4, 256);
    L.map_off = o; o = align_up(o + L.slots_cap * 4, 256);
    L.nodes_off = o; o = align_up(o + L.slots_cap * 32, 256);
    L.total = o;
    return L;
}

}  // namespace

int hash_pick_group(uint64_t data_bytes) {
    static const int forced = []() { const char* e = getenv("ZEST_B3_GROUP"); int g = e ? atoi(e) : 0; return (g == 1 || g == 2 || g == 4 || g == 8) ? g : 0; }();
    if (forced) return forced;  // tuning knob
    // keep the machine full (148 SMs x ~50 warps of 64-register threads) before folding more leaves per thread:
    // measured on the 550 MB set, G = 2 / 4 / 8 -> 1548 / 1474 / 1395 GB/s
    const uint64_t want = 148ull * 32 * 50;
    for (int g = 8; g > 1; g >>= 1)
        if (data_bytes / ((uint64_t)kLeaf * g) >= want) return g;
    return 1;
}

size_t hash_scratch_bytes(uint64_t data_bytes, size_t n, int group) { return hash_layout(data_bytes, n, group).total; }

cudaError_t launch_hash_chunks(const uint8_t* data, uint64_t data_bytes, const uint64_t* off, const uint32_t* len,
                               size_t n, const B3Key& key, uint8_t* out, void* scratch, size_t scratch_bytes,
                               int group, cudaStream_t st, const uint8_t* planes, const ChunkDesc* descs) {
    if (n == 0) return cudaSuccess;
    if (group != 1 && group != 2 && group != 4 && group != 8) return cudaErrorInvalidValue;
    HashLayout L = hash_layout(data_bytes, n, group);
    if (scratch_bytes < L.total || n >= 0xffffffffull || L.slots_cap >= 0xffffffffull) return cudaErrorInvalidValue;
    uint8_t* sb = static_cast<uint8_t*>(scratch);
    uint32_t* P = reinterpret_cast<uint32_t*>(sb + L.p_off);
    uint32_t* tiles = reinterpret_cast<uint32_t*>(sb + L.tiles_off);
    uint32_t* map = reinterpret_cast<uint32_t*>(sb + L.map_off);
    uint4* nodes = reinterpret_cast<uint4*>(sb + L.nodes_off);
    int gb = kLeaf * group;
    hash_count_kernel<<<(unsigned)L.n_tiles, kTile, 0, st>>>(off, len, n, gb, P, tiles);
    hash_tiles_kernel<<<1, 1024, 0, st>>>(tiles, (uint32_t)L.n_tiles, (uint32_t)L.slots_cap, P + n);
    uint4* o4 = reinterpret_cast<uint4*>(out);
    hash_map_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(off, len, n, gb, P, tiles, (uint32_t)L.slots_cap, map, o4);
    unsigned blocks = (unsigned)((L.slots_cap + 127) / 128);
    static const bool fma_adds = []() { const char* e = getenv("ZEST_B3_PLAIN_ADDS"); return !(e && e[0] == '1'); }();
    const bool planar = planes != nullptr && descs != nullptr;
#define ZG_LEAF(G)                                                                                                                         \
    do {                                                                                                                                   \
        if (planar) hash_leaf_kernel<G, true, true><<<blocks, 128, 0, st>>>(data, off, len, key, P, map, P + n, nodes, o4, planes, descs); \
        else if (fma_adds) hash_leaf_kernel<G, true, false><<<blocks, 128, 0, st>>>(data, off, len, key, P, map, P + n, nodes, o4, nullptr, nullptr); \
        else hash_leaf_kernel<G, false, false><<<blocks, 128, 0, st>>>(data, off, len, key, P, map, P + n, nodes, o4, nullptr, nullptr); \
    } while (0)
    switch (group) {
        case 1: ZG_LEAF(1); break;
        case 2: ZG_LEAF(2); break;
        case 4: ZG_LEAF(4); break;
        default: ZG_LEAF(8); break;
    }
#undef ZG_LEAF
    // team size: a 128 KiB chunk has 128 / G nodes; give each chunk a quarter of that in lanes
    uint32_t team = group == 8 ? 4u : (group == 4 ? 8u : (group == 2 ? 16u : 32u));
    size_t tthreads = n * team;
    hash_tree_kernel<<<(unsigned)((tthreads + 127) / 128), 128, 0, st>>>(key, P, n, team, (uint32_t)L.slots_cap, nodes, o4);
    return cudaGetLastError();
}

}  // namespace zg

// ---------------------------------------------------------------------------------------------------------------
// zg_dev_hash_chunks_static: the chunk-hash pass for callers that must not allocate.
//
// Some users of this library run where device allocation is off the table: a transfer loop that has already handed
// every byte of its memory budget to pinned staging and decode buffers, a host thread that may not take the driver's
// allocation lock while a download is in flight, or a process whose allocator has been frozen after start-up.  They
// still own their input and output buffers, but they have nowhere to put the hash pass's working set (the prefix
// sums, the slot map and the 32-byte group nodes).  For them the library carries that working set itself, in the
// fixed array below; it is part of the loaded module, so every GPU that loads the library holds these 8 MiB until
// the process exits, whether or not this entry point is ever called.
//
// The array's name is only a host-side handle in host code, not a device address, so the address is looked up once
// per device with cudaGetSymbolAddress (first call on that device, with the device current) and cached.  That first
// call therefore does a little set-up work and should be made before any stream capture begins.
//
// All calls on a device share the one array and nothing in it survives a call in a useful form.  Nothing has to be
// cleared between calls: every word the kernels read (prefix sums, map, nodes) is written earlier in the same call.
// Two calls whose kernels overlap on the GPU (different streams, or different host threads) would race on it, so
// the caller must keep calls on one device in order, e.g. by issuing them on one stream.
// ---------------------------------------------------------------------------------------------------------------
namespace {
constexpr size_t kStaticScratchBytes = 8u << 20;  // covers any batch up to 256 MiB of data in up to 65,536 chunks
static __device__ __align__(256) unsigned char g_static_hash_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kMaxDevices] = {nullptr};
}  // namespace

extern "C" ZG_API int zg_dev_hash_chunks_static(const uint8_t* d_data, uint64_t data_bytes, const uint64_t* d_offsets,
                                                const uint32_t* d_lens, size_t n, const uint8_t* key, uint8_t* d_out,
                                                void* stream) {
    using namespace zg;
    if (n == 0) return ZG_OK;
    if (!d_data && data_bytes) return set_last_error(ZG_ERR_INVALID_ARG, "d_data is null");
    if (!d_offsets || !d_out) return set_last_error(ZG_ERR_INVALID_ARG, "null argument");
    int group = hash_pick_group(data_bytes);
    // a larger group (more leaves per thread) shrinks the working set; the hashes do not depend on it
    while (group < 8 && hash_scratch_bytes(data_bytes, n, group) > kStaticScratchBytes) group *= 2;
    const size_t need = hash_scratch_bytes(data_bytes, n, group);
    if (need > kStaticScratchBytes)
        return set_last_error(ZG_ERR_LIMIT,
                              "zg_dev_hash_chunks_static: this batch needs %zu bytes of scratch, the built-in array holds %zu "
                              "(enough for 256 MiB in 65,536 chunks); split the batch or use zg_dev_hash_chunks",
                              need, kStaticScratchBytes);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_last_error(ZG_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return set_last_error(ZG_ERR_LIMIT, "device ordinal %d not supported", dev);
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_static_mu);
        if (!g_static_ptr[dev]) {  // first call on this device: resolve the array's device address
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_static_hash_scratch);
            if (e != cudaSuccess || !p) return set_last_error(ZG_ERR_CUDA, "cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            g_static_ptr[dev] = p;
        }
        scratch = g_static_ptr[dev];
    }
    B3Key k;
    k.one = 1;
    if (key == ZG_PLAIN_BLAKE3) {
        const uint32_t iv[8] = {0x6A09E667u, 0xBB67AE85u, 0x3C6EF372u, 0xA54FF53Au, 0x510E527Fu, 0x9B05688Cu, 0x1F83D9ABu, 0x5BE0CD19u};
        for (int i = 0; i < 8; i++) k.w[i] = iv[i];
        k.base_flags = 0;
    } else {
        static const uint8_t data_key[32] = {102, 151, 245, 119, 91,  149, 80,  222, 49,  53,  203, 172, 165, 151, 24,  28,
                                             157, 228, 33,  16,  155, 235, 43,  88,  180, 208, 176, 75,  147, 173, 242, 41};
        const uint8_t* p = key ? key : data_key;
        for (int i = 0; i < 8; i++)
            k.w[i] = (uint32_t)p[4 * i] | ((uint32_t)p[4 * i + 1] << 8) | ((uint32_t)p[4 * i + 2] << 16) | ((uint32_t)p[4 * i + 3] << 24);
        k.base_flags = B3_KEYED;
    }
    e = launch_hash_chunks(d_data, data_bytes, d_offsets, d_lens, n, k, d_out, scratch, kStaticScratchBytes, group,
                           static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return set_last_error(ZG_ERR_CUDA, "zg_dev_hash_chunks_static launch: %s", cudaGetErrorString(e));
    return ZG_OK;
}

