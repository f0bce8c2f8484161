// Xorb Merkle fold and verdict on sm_100a.
//
// Reference behaviour replaced: zig-xet `hashing` aggregated node hash -- "branching factor 4, domain-separation
// keys" (/root/reference/CONTRIBUTING.md:144-146) -- whose result is the xorb content address that zest uses as
// cache key and term.hash (/root/reference/src/xet_bridge.zig:154, src/swarm.zig:57-148).  zest itself never
// compares it; this kernel makes the promised check real.  Algorithm: SURVEY.md A.3; oracle:
// oracle/xet_oracle.c (oz_merkle_root).
//
// Shape of the work: a few hundred nodes per level, 4-5 levels, cuts that depend on the previous cut.  One CTA per
// xorb: all threads classify nodes (hash.u64[3] % 4 == 0), one thread chases the cut chain through shared memory,
// then every group is formatted ("{hex} : {len}\n" per child) and hashed by its own thread with the INTERNAL-NODE
// key.  Latency-bound and ~0.4 % of the hashing work; reported in microseconds per xorb, not against a roofline.
#include "kernels.h"

namespace zg {

namespace {

constexpr int kThreads = 256;
constexpr int kTileNodes = 2048;  // group starts handled per chain pass
constexpr int kMaxGroup = 9;

__device__ __constant__ const uint8_t kNodeKey[32] = {1,   126, 197, 199, 165, 71,  41,  150, 253, 148, 102,
                                                      102, 180, 138, 2,   230, 93,  221, 83,  111, 55,  199,
                                                      109, 210, 248, 99,  82,  230, 74,  83,  113, 63};

struct NodeView {  // one level's nodes: 32-byte hashes + lengths (level 0 reads the caller's arrays directly)
    const uint8_t* h;
    const uint32_t* l32;
    const uint64_t* l64;
    __device__ uint64_t len(size_t i) const { return l64 ? l64[i] : (uint64_t)l32[i]; }
};

__device__ __forceinline__ char hexd(uint32_t v) { return (char)(v < 10 ? '0' + v : 'a' + (v - 10)); }

// Hash one group of `cnt` children (1..9) -> 32-byte node hash, returns total length.
__device__ uint64_t hash_group(const NodeView& in, size_t first, int cnt, const uint32_t (&key)[8], uint8_t* out_h) {
    uint32_t buf32[(kMaxGroup * 88 + 63) / 64 * 16];  // <= 9 x (64 + 3 + 20 + 1) bytes, rounded up to whole blocks
    uint8_t* buf = reinterpret_cast<uint8_t*>(buf32);
    uint32_t pos = 0;
    uint64_t total = 0;
    for (int c = 0; c < cnt; c++) {
        const uint8_t* h = in.h + 32 * (first + c);
        for (int g = 0; g < 4; g++)
            for (int b = 7; b >= 0; b--) {  // each 8-byte group is a little-endian u64 printed %016x (A.1)
                uint32_t v = h[8 * g + b];
                buf[pos++] = hexd(v >> 4);
                buf[pos++] = hexd(v & 15);
            }
        buf[pos++] = ' '; buf[pos++] = ':'; buf[pos++] = ' ';
        uint64_t L = in.len(first + c);
        total += L;
        char dig[20];
        int nd = 0;
        do { dig[nd++] = (char)('0' + (int)(L % 10)); L /= 10; } while (L);
        while (nd) buf[pos++] = dig[--nd];
        buf[pos++] = '\n';
    }
    uint32_t nblk = pos == 0 ? 1 : (pos + 63) / 64;
    for (uint32_t i = pos; i < nblk * 64; i++) buf[i] = 0;
    uint32_t cv[8];
#pragma unroll
    for (int i = 0; i < 8; i++) cv[i] = key[i];
#pragma unroll 1
    for (uint32_t b = 0; b < nblk; b++) {
        uint32_t m[16];
#pragma unroll
        for (int i = 0; i < 16; i++) m[i] = buf32[16 * b + i];
        uint32_t blen = pos - 64 * b < 64 ? pos - 64 * b : 64;
        uint32_t fl = B3_KEYED | (b == 0 ? B3_CHUNK_START : 0u) | (b == nblk - 1 ? (B3_CHUNK_END | B3_ROOT) : 0u);
        b3_compress<false>(cv, m, 0, 0, blen, fl);
    }
    uint32_t* o = reinterpret_cast<uint32_t*>(out_h);
#pragma unroll
    for (int i = 0; i < 8; i++) o[i] = cv[i];
    return total;
}

__global__ void __launch_bounds__(kThreads) merkle_kernel(const uint8_t* __restrict__ hashes,
                                                          const uint32_t* __restrict__ lens32,
                                                          const uint64_t* __restrict__ lens64,
                                                          const uint32_t* __restrict__ first, size_t n_jobs,
                                                          const int32_t* __restrict__ chunk_status,
                                                          const uint8_t* __restrict__ expected,
                                                          uint8_t* __restrict__ roots, int8_t* __restrict__ verdict,
                                                          uint8_t* __restrict__ nodeA_h, uint64_t* __restrict__ nodeA_l,
                                                          uint8_t* __restrict__ nodeB_h, uint64_t* __restrict__ nodeB_l) {
    __shared__ uint8_t flag[kTileNodes + 16];
    __shared__ uint16_t gstart[kTileNodes / 2 + 8];
    __shared__ uint8_t gsize[kTileNodes / 2 + 8];
    __shared__ uint32_t s_ngroups, s_next;
    __shared__ int s_err;
    size_t x = blockIdx.x;
    if (x >= n_jobs) return;
    size_t lo = first[x], hi = first[x + 1], n = hi - lo;
    uint32_t tid = threadIdx.x;
    uint8_t* root = roots + 32 * x;

    // any chunk that failed to decode decides the verdict (first failing chunk, as the oracle reports it)
    if (tid == 0) s_err = 0;
    __syncthreads();
    if (chunk_status) {
        for (size_t i = lo + tid; i < hi; i += kThreads)
            if (chunk_status[i] != 0) atomicMin(&s_err, (int)(i - lo) - 0x40000000);  // order by index
        __syncthreads();
        if (s_err != 0) {
            if (tid < 32) root[tid] = 0;
            if (tid == 0) verdict[x] = (int8_t)chunk_status[lo + (size_t)(s_err + 0x40000000)];
            return;
        }
    }

    uint32_t key[8];
#pragma unroll
    for (int i = 0; i < 8; i++)
        key[i] = kNodeKey[4 * i] | (kNodeKey[4 * i + 1] << 8) | (kNodeKey[4 * i + 2] << 16) | ((uint32_t)kNodeKey[4 * i + 3] << 24);

    NodeView in{hashes + 32 * lo, lens32 ? lens32 + lo : nullptr, lens64 ? lens64 + lo : nullptr};
    size_t region = lo + 2 * x;  // each job owns n + 2 node slots in A and in B
    uint8_t* outH = nodeA_h + 32 * region;
    uint64_t* outL = nodeA_l + region;
    uint8_t* altH = nodeB_h + 32 * region;
    uint64_t* altL = nodeB_l + region;

    while (n > 1) {
        size_t r = 0, w = 0;
        while (r < n) {
            size_t tile_n = n - r < (size_t)kTileNodes ? n - r : (size_t)kTileNodes;
            size_t look = n - r < (size_t)kTileNodes + 16 ? n - r : (size_t)kTileNodes + 16;
            for (size_t j = tid; j < look; j += kThreads) flag[j] = (in.h[32 * (r + j) + 24] & 3) == 0;
            __syncthreads();
            if (tid == 0) {
                uint32_t g = 0;
                size_t pos = 0;
                while (pos < tile_n) {
                    size_t rem = n - (r + pos), size;
                    if (rem <= 2) size = rem;
                    else {
                        size_t end = rem < (size_t)kMaxGroup ? rem : (size_t)kMaxGroup;
                        size = end;
                        for (size_t i = 2; i < end; i++)
                            if (flag[pos + i]) { size = i + 1; break; }
                    }
                    gstart[g] = (uint16_t)pos;
                    gsize[g] = (uint8_t)size;
                    g++;
                    pos += size;
                }
                s_ngroups = g;
                s_next = (uint32_t)pos;
            }
            __syncthreads();
            uint32_t ng = s_ngroups;
            for (uint32_t g = tid; g < ng; g += kThreads)
                outL[w + g] = hash_group(in, r + gstart[g], gsize[g], key, outH + 32 * (w + g));
            w += ng;
            r += s_next;
            __syncthreads();
        }
        __threadfence_block();
        in = NodeView{outH, nullptr, outL};
        uint8_t* th = outH; outH = altH; altH = th;
        uint64_t* tl = outL; outL = altL; altL = tl;
        n = w;
        __syncthreads();
    }

    // n == 1: the root is the single remaining node (a lone chunk hash is returned unchanged); n == 0: zeros
    __shared__ int s_match;
    if (tid == 0) s_match = 1;
    __syncthreads();
    if (tid < 32) {
        uint8_t v = n == 1 ? in.h[tid] : 0;
        root[tid] = v;
        if (expected && expected[32 * x + tid] != v) s_match = 0;
    }
    __syncthreads();
    if (tid == 0) verdict[x] = (int8_t)s_match;
}

}  // namespace

size_t merkle_scratch_bytes(size_t n_total, size_t n_jobs) {
    size_t slots = n_total + 2 * n_jobs + 2;
    return 2 * (align_up(slots * 32, 256) + align_up(slots * 8, 256));
}

cudaError_t launch_merkle(const uint8_t* hashes, const uint32_t* lens32, const uint64_t* lens64,
                          const uint32_t* first, size_t n_jobs, size_t n_total, const int32_t* chunk_status,
                          const uint8_t* expected, uint8_t* roots, int8_t* verdict, void* scratch,
                          size_t scratch_bytes, cudaStream_t st) {
    if (n_jobs == 0) return cudaSuccess;
    if (scratch_bytes < merkle_scratch_bytes(n_total, n_jobs)) return cudaErrorInvalidValue;
    size_t slots = n_total + 2 * n_jobs + 2;
    uint8_t* sb = static_cast<uint8_t*>(scratch);
    uint8_t* aH = sb;
    uint64_t* aL = reinterpret_cast<uint64_t*>(sb + align_up(slots * 32, 256));
    uint8_t* bH = sb + align_up(slots * 32, 256) + align_up(slots * 8, 256);
    uint64_t* bL = reinterpret_cast<uint64_t*>(bH + align_up(slots * 32, 256));
    merkle_kernel<<<(unsigned)n_jobs, kThreads, 0, st>>>(hashes, lens32, lens64, first, n_jobs, chunk_status, expected,
                                                        roots, verdict, aH, aL, bH, bL);
    return cudaGetLastError();
}

}  // namespace zg
