// Device-resident dedup index: Xet chunk hash (32 bytes) -> 64-bit location, in HBM.
//
// SURVEY.md 8(f) item 4.  What it stands in for on the seed/dedup path: zest answers "do we already hold this
// chunk / xorb" with one file-system probe per hash (storage.hasChunk, /root/reference/src/storage.zig:137; the
// XorbRegistry scan, src/storage.zig:145-228) and xet-core keeps a global_dedup_lookup table keyed by chunk hash.
// After zg_cdc_hash the chunk hashes of a shard are already in HBM; this table answers the same question for a
// whole batch of them in one launch.  Semantics (deterministic, checked against a dict-based oracle in tests):
//   insert(hash, loc): the table keeps, per hash, the SMALLEST location ever inserted for it;
//   was_new[i] = 1 iff the hash was absent before this batch and occurrence i carries the location that was kept;
//   lookup(hash) -> location, or ZG_DEDUP_MISSING.
// With locations that grow over time (xorb number << 32 | chunk index) "smallest" is "first seen".
//
// Layout.  Open addressing, linear probing, capacity a power of two, at most half full.  Three parallel arrays so
// that a probe that misses touches 8 bytes per slot: tags (u64: hash bytes 8..15 with the top bit set, 0 = empty),
// keys (32 B), locs (u64, atomicMin) and born (u32: the batch that created the slot).  A slot is claimed by CAS on
// its tag; the claimer then writes the key.  Inside one insert launch other threads match on the tag alone (the
// key may not be written yet); a second launch (dedup_resolve_kernel), ordered after the first on the stream,
// re-probes every inserted hash against the full 32-byte keys, computes was_new and counts tag-only matches whose
// key differs -- two different chunk hashes agreeing in bytes 8..15 AND probing into each other, a 2^-64 event
// that is reported as an error instead of being silently merged.  Lookups always compare the full key.
//
// Roofline: HBM random access.  A hit reads the tag sector, the key sector(s) and the location: ~3 x 32-byte
// sectors per probe; bench: tools/dedup_bench.py.
#include <cuda_runtime.h>

#include <cstdint>

#include "kernels.h"

namespace zg {

namespace {

constexpr unsigned long long kEmpty = 0ull;
constexpr unsigned long long kNoLoc = ~0ull;

__device__ __forceinline__ unsigned long long ld64(const uint8_t* p) {  // hashes are 8-byte aligned (32-byte records)
    return *reinterpret_cast<const unsigned long long*>(p);
}

// One thread per hash.  status[0] |= 1 when the probe ran through the whole table (full).
__global__ void dedup_insert_kernel(unsigned long long* tags, uint8_t* keys, unsigned long long* locs, uint32_t* born,
                                    uint64_t mask, const uint8_t* hashes, const uint64_t* locations, uint64_t n,
                                    uint32_t epoch, unsigned long long* counters) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t* h = hashes + 32 * i;
    const unsigned long long tag = ld64(h + 8) | (1ull << 63);
    uint64_t slot = ld64(h) & mask;
    for (uint64_t probe = 0; probe <= mask; probe++, slot = (slot + 1) & mask) {
        unsigned long long t = tags[slot];
        if (t == kEmpty) {
            t = atomicCAS(&tags[slot], kEmpty, tag);
            if (t == kEmpty) {  // ours: write the key, note the batch
                const unsigned long long* hs = reinterpret_cast<const unsigned long long*>(h);
                unsigned long long* kd = reinterpret_cast<unsigned long long*>(keys + 32 * slot);
                kd[0] = hs[0]; kd[1] = hs[1]; kd[2] = hs[2]; kd[3] = hs[3];
                born[slot] = epoch;
                atomicAdd(&counters[0], 1ull);  // entries
                t = tag;
            }
        }
        if (t == tag) {
            atomicMin(&locs[slot], (unsigned long long)locations[i]);
            return;
        }
    }
    atomicOr(&counters[1], 1ull);  // table full
}

// After the insert launch: full-key check, was_new.
__global__ void dedup_resolve_kernel(const unsigned long long* tags, const uint8_t* keys, const unsigned long long* locs,
                                     const uint32_t* born, uint64_t mask, const uint8_t* hashes, const uint64_t* locations,
                                     uint64_t n, uint32_t epoch, uint8_t* was_new, unsigned long long* counters) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t* h = hashes + 32 * i;
    const unsigned long long* hs = reinterpret_cast<const unsigned long long*>(h);
    const unsigned long long tag = hs[1] | (1ull << 63);
    uint64_t slot = hs[0] & mask;
    for (uint64_t probe = 0; probe <= mask; probe++, slot = (slot + 1) & mask) {
        const unsigned long long t = tags[slot];
        if (t == kEmpty) break;
        if (t == tag) {
            const unsigned long long* kd = reinterpret_cast<const unsigned long long*>(keys + 32 * slot);
            if (kd[0] != hs[0] || kd[1] != hs[1] || kd[2] != hs[2] || kd[3] != hs[3]) atomicAdd(&counters[2], 1ull);  // tag collision
            if (was_new) was_new[i] = (born[slot] == epoch && locs[slot] == (unsigned long long)locations[i]) ? 1 : 0;
            return;
        }
    }
    if (was_new) was_new[i] = 0;  // not inserted (table full)
}

__global__ void dedup_lookup_kernel(const unsigned long long* tags, const uint8_t* keys, const unsigned long long* locs,
                                    uint64_t mask, const uint8_t* hashes, uint64_t n, uint64_t* out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long* hs = reinterpret_cast<const unsigned long long*>(hashes + 32 * i);
    const unsigned long long h0 = hs[0], h1 = hs[1], h2 = hs[2], h3 = hs[3];
    const unsigned long long tag = h1 | (1ull << 63);
    uint64_t slot = h0 & mask;
    unsigned long long r = kNoLoc;
    for (uint64_t probe = 0; probe <= mask; probe++, slot = (slot + 1) & mask) {
        const unsigned long long t = tags[slot];
        if (t == kEmpty) break;
        if (t == tag) {
            const unsigned long long* kd = reinterpret_cast<const unsigned long long*>(keys + 32 * slot);
            if (kd[0] == h0 && kd[1] == h1 && kd[2] == h2 && kd[3] == h3) { r = locs[slot]; break; }
        }
    }
    out[i] = r;
}

}  // namespace

size_t dedup_table_bytes(uint64_t slots) { return (size_t)slots * (8 + 32 + 8 + 4) + 256; }

// Carve the table out of one allocation: tags | locs | keys | born | counters (entries, full flag, collisions).
DedupView dedup_view(void* base, uint64_t slots) {
    DedupView v;
    uint8_t* p = static_cast<uint8_t*>(base);
    v.slots = slots;
    v.tags = reinterpret_cast<unsigned long long*>(p); p += 8 * slots;
    v.locs = reinterpret_cast<unsigned long long*>(p); p += 8 * slots;
    v.keys = p; p += 32 * slots;
    v.born = reinterpret_cast<uint32_t*>(p); p += 4 * slots;
    v.counters = reinterpret_cast<unsigned long long*>(p);  // slots >= 64: every array is a multiple of 256 bytes
    return v;
}

cudaError_t dedup_clear(const DedupView& v, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(v.tags, 0, 8 * v.slots, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(v.locs, 0xff, 8 * v.slots, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(v.counters, 0, 32, st);
    return e;
}

cudaError_t launch_dedup_insert(const DedupView& v, const uint8_t* hashes, const uint64_t* locations, uint64_t n,
                                uint32_t epoch, uint8_t* was_new, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    dedup_insert_kernel<<<blocks, 256, 0, st>>>(v.tags, v.keys, v.locs, v.born, v.slots - 1, hashes, locations, n, epoch, v.counters);
    dedup_resolve_kernel<<<blocks, 256, 0, st>>>(v.tags, v.keys, v.locs, v.born, v.slots - 1, hashes, locations, n, epoch, was_new,
                                                 v.counters);
    return cudaGetLastError();
}

cudaError_t launch_dedup_lookup(const DedupView& v, const uint8_t* hashes, uint64_t n, uint64_t* out, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    dedup_lookup_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(v.tags, v.keys, v.locs, v.slots - 1, hashes, n, out);
    return cudaGetLastError();
}

}  // namespace zg
