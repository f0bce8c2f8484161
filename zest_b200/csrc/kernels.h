// Internal launch interface between the C-ABI layer (api.cu) and the kernel translation units.
// Everything here is asynchronous on the given stream and allocates nothing: scratch comes from the caller.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

#include "blake3.cuh"

namespace zg {

// Error codes (mirror include/zest_gpu.h; kept numerically identical to the oracle's OZ_ERR_*).
enum : int {
    E_OK = 0, E_INVALID = -1, E_CUDA = -2, E_TRUNCATED = -3, E_BAD_HEADER = -4, E_SCHEME = -5, E_LZ4_FRAME = -6,
    E_LZ4_BLOCK = -7, E_SIZE_MISMATCH = -8, E_RANGE = -9, E_CAPACITY = -10, E_NO_DEVICE = -11, E_OOM = -12,
    E_LIMIT = -13
};

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Record a failure for zg_last_error() on the calling thread and return `code` (defined in api.cu).
int set_last_error(int code, const char* fmt, ...);

// ---------------------------------------------------------------------------------------------- hashing
// Leaves per thread in the leaf kernel (1, 2, 4 or 8), chosen from the batch size: small batches need every
// leaf on its own thread to fill 148 SMs, large ones amortise the tree by folding 8 KiB per thread.
int hash_pick_group(uint64_t data_bytes);
size_t hash_scratch_bytes(uint64_t data_bytes, size_t n_chunks, int group);

// Hash n chunks.  Chunk i = data[off[i] .. off[i] + len_i), len_i = len ? len[i] : off[i+1] - off[i].
// data must be 16-byte aligned and readable up to the next multiple of 16 past the last chunk.
// out: n x 32 bytes.  scratch: hash_scratch_bytes(data_bytes, n, group) bytes, 256-byte aligned.
struct ChunkDesc;
// planes + descs (both or neither): chunk i with descs[i].scheme == 2 is read from planes + off[i], still byte-grouped as
// the decoder left it (launch_decode with keep_planes), and regrouped on the fly; every other chunk from data + off[i].
cudaError_t launch_hash_chunks(const uint8_t* data, uint64_t data_bytes, const uint64_t* off, const uint32_t* len,
                               size_t n, const B3Key& key, uint8_t* out, void* scratch, size_t scratch_bytes,
                               int group, cudaStream_t st, const uint8_t* planes = nullptr, const ChunkDesc* descs = nullptr);

// ---------------------------------------------------------------------------------------------- merkle
// One fold job per xorb: chunks [first[x], first[x+1]) of hashes/lens -> roots[x]; verdict from status/expected.
size_t merkle_scratch_bytes(size_t n_chunks_total, size_t n_jobs);
cudaError_t launch_merkle(const uint8_t* hashes, const uint32_t* lens32, const uint64_t* lens64,
                          const uint32_t* first, size_t n_jobs, size_t n_chunks_total, const int32_t* chunk_status,
                          const uint8_t* expected, uint8_t* roots, int8_t* verdict, void* scratch,
                          size_t scratch_bytes, cudaStream_t st);

// ---------------------------------------------------------------------------------------------- decode
struct ChunkDesc {      // one per chunk, built by the host walk (zg_xorb_index) or the device index kernel
    uint64_t src_off;   // payload offset in the compressed buffer
    uint64_t dst_off;   // where the decoded bytes go in the output buffer
    uint32_t clen;
    uint32_t ulen;
    uint32_t scheme;    // 0 none, 1 lz4 frame, 2 bg4 + lz4 frame
    uint32_t order;     // work order: the t-th chunk handed out is chunks[t].order (a permutation; longest first)
};
// Fill chunks[t].order so that work is handed out longest-first (keeps the last warps from finishing alone).
void fill_decode_order(ChunkDesc* chunks, size_t n);
size_t decode_scratch_bytes(size_t n_chunks, uint64_t out_bytes);
// keep_planes: scheme-2 chunks are left byte-grouped in the plane buffer (scratch + 256, at their dst_off) instead of being
// regrouped into out -- for a caller that only hashes them (launch_hash_chunks with planes).
cudaError_t launch_decode(const uint8_t* comp, uint64_t comp_bytes, const ChunkDesc* chunks, size_t n, uint8_t* out,
                          uint64_t out_bytes, int32_t* status, void* scratch, size_t scratch_bytes, cudaStream_t st,
                          bool keep_planes = false);

// ---------------------------------------------------------------------------------------------- cdc
size_t cdc_scratch_bytes(uint64_t len);
// ends: device array of cap u64; n_out: device u64 {count, status}.
cudaError_t launch_cdc(const uint8_t* data, uint64_t len, uint64_t* ends, uint64_t cap, uint64_t* n_out,
                       void* scratch, size_t scratch_bytes, cudaStream_t st);

// ---------------------------------------------------------------------------------------------- dedup index
struct DedupView {  // the table's arrays inside one device allocation of dedup_table_bytes(slots)
    uint64_t slots;                 // power of two, >= 64
    unsigned long long* tags;       // hash bytes 8..15 with the top bit set; 0 = empty
    unsigned long long* locs;       // smallest location inserted for the slot's hash
    uint8_t* keys;                  // 32 bytes per slot
    uint32_t* born;                 // insert batch that created the slot
    unsigned long long* counters;   // [0] entries, [1] table-full flag, [2] tag matches whose key differed
};
size_t dedup_table_bytes(uint64_t slots);
DedupView dedup_view(void* base, uint64_t slots);
cudaError_t dedup_clear(const DedupView& v, cudaStream_t st);
// hashes: n x 32 bytes, locations: n u64, was_new: n bytes or null -- all device pointers.
cudaError_t launch_dedup_insert(const DedupView& v, const uint8_t* hashes, const uint64_t* locations, uint64_t n,
                                uint32_t epoch, uint8_t* was_new, cudaStream_t st);
cudaError_t launch_dedup_lookup(const DedupView& v, const uint8_t* hashes, uint64_t n, uint64_t* out, cudaStream_t st);

}  // namespace zg
