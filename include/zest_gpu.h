/*
 * zest_gpu.h -- C ABI of libzest_gpu.so: B200 (sm_100a) implementation of zest's xorb verify / reconstruct
 * hot path.  Plain pointers and sizes only; no C++ or torch types cross this boundary.
 *
 * What it stands in for.  praveer13/zest reaches this arithmetic through the third-party Zig package zig-xet
 * (module `xet`, build.zig.zon:6-11).  Each entry point below names the reference interface it replaces
 * (paths relative to the zest repository).  A Zig `extern` block binding these symbols is shown in
 * INTEGRATION.md.
 *
 * Conventions
 *   - every function returns ZG_OK (0) or a negative ZG_ERR_*; nothing aborts or throws.  A Zig caller maps a
 *     non-zero return to an error and takes its existing fallback (parallel -> sequential -> zig-xet,
 *     src/main.zig:233-256), exactly as it does for a Zig error union today;
 *   - output buffers are caller-owned, sizes are queried first (no cross-allocator frees; today the callee
 *     returns an allocator-owned slice that the caller frees, src/xet_bridge.zig:254-258);
 *   - a context may be used from many threads at once (zest fans out up to 16*8 = 128 tasks per batch,
 *     src/parallel_download.zig:117,174-182): calls queue for one of the context's lanes (stream + staging);
 *   - there is no CPU fallback: without a usable CUDA device zg_init fails with ZG_ERR_NO_DEVICE.
 *   - hashes are 32 raw bytes; their text form is zg_hash_to_hex (four little-endian u64 printed %016x).
 */
#ifndef ZEST_GPU_H
#define ZEST_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define ZG_API __attribute__((visibility("default")))
#else
#define ZG_API
#endif

#define ZG_OK 0
#define ZG_ERR_INVALID_ARG (-1)
#define ZG_ERR_CUDA (-2)          /* a CUDA call failed; see zg_last_error()                                 */
#define ZG_ERR_TRUNCATED (-3)     /* chunk header or payload runs past the end of the blob                    */
#define ZG_ERR_BAD_HEADER (-4)    /* chunk header version != 0, or scheme 0 with clen != ulen                 */
#define ZG_ERR_SCHEME (-5)        /* compression scheme other than 0 / 1 / 2 (zig-xet's scheme 3 is rejected) */
#define ZG_ERR_LZ4_FRAME (-6)     /* bad magic / version / header checksum / dictionary / block size          */
#define ZG_ERR_LZ4_BLOCK (-7)     /* malformed sequence, offset 0 or outside the window                       */
#define ZG_ERR_SIZE_MISMATCH (-8) /* decoded size differs from the header's uncompressed_len                  */
#define ZG_ERR_RANGE (-9)         /* chunk range outside the blob                                             */
#define ZG_ERR_CAPACITY (-10)     /* caller's output buffer is too small                                      */
#define ZG_ERR_NO_DEVICE (-11)    /* no CUDA device / driver: this library has no CPU path                    */
#define ZG_ERR_OOM (-12)          /* device or pinned allocation failed                                       */
#define ZG_ERR_LIMIT (-13)        /* input exceeds a documented limit of the entry point                      */

typedef struct zg_ctx zg_ctx;

typedef struct zg_opts {
    uint32_t struct_size; /* sizeof(zg_opts), for forward compatibility                         */
    uint32_t n_lanes;     /* concurrent calls served at once (streams); 0 = default (8)        */
    uint64_t reserved[3];
} zg_opts;

/* ---- life cycle ------------------------------------------------------------------------------------------ */

/* Create a context on CUDA device `device` (one context per GPU; one process per GPU under torchrun).
 * New: zest has no device state; the nearest analogue is XetBridge.init (src/xet_bridge.zig:54-70). */
ZG_API int zg_init(int device, const zg_opts* opts, zg_ctx** out);
ZG_API void zg_shutdown(zg_ctx* ctx);
ZG_API const char* zg_strerror(int code);
/* Message of the last failure on the calling thread (empty string if none). */
ZG_API const char* zg_last_error(void);

/* Pinned host memory the caller can fill directly (network / disk reads land here, then cudaMemcpyAsync).
 * Replaces the allocator-owned `[]u8` blob of XorbFetchResult.data (src/xet_bridge.zig:45-52). */
ZG_API int zg_pinned_alloc(zg_ctx* ctx, size_t bytes, void** out);
ZG_API int zg_pinned_free(zg_ctx* ctx, void* p);

/* ---- host-only helpers (no GPU work) ----------------------------------------------------------------------- */

/* xet.hashing.hashToHex (src/xet_bridge.zig:154, src/server.zig:201-203; pitfall CONTRIBUTING.md:134-136). */
ZG_API void zg_hash_to_hex(const uint8_t hash[32], char out[65]);
ZG_API int zg_hex_to_hash(const char* hex, uint8_t out[32]);

/* Walk the [8-byte header][payload] chunk stream of a xorb blob (what XorbReader.init + iteration does,
 * src/xet_bridge.zig:256).  Arrays may be NULL to count only; otherwise they hold `cap` entries.
 * hdr_off[i] = byte offset of chunk i's header.  Returns ZG_ERR_CAPACITY if the blob has more than cap chunks. */
ZG_API int zg_xorb_index(const uint8_t* blob, size_t len, uint32_t cap, uint32_t* n_chunks, uint64_t* hdr_off,
                  uint32_t* clen, uint32_t* ulen, uint8_t* scheme, uint64_t* total_ulen);

/* The "XETBLOB" footer xet-core appends to a stored xorb: xorb hash, every chunk hash, cumulative
 * compressed / uncompressed boundaries, section offsets, and the trailing length word.  With the chunk hashes and root
 * that zg_verify_xorbs / zg_cdc_hash return, this completes a XorbCache-format blob on the seed path
 * (XorbCache.put, src/swarm.zig:57-148).  out receives zg_xorb_footer_bytes(n) bytes to append
 * after the chunk stream. */
ZG_API size_t zg_xorb_footer_bytes(size_t n_chunks);
ZG_API int zg_xorb_footer(const uint8_t xorb_hash[32], const uint8_t* chunk_hashes, const uint32_t* clen,
                          const uint32_t* ulen, size_t n, uint8_t* out, size_t out_cap, size_t* out_len);

/* Size of extractChunkRange(start, end)'s result, so the caller can allocate. */
ZG_API int zg_chunk_range_size(const uint8_t* blob, size_t len, uint32_t start, uint32_t end, uint64_t* out_len);

/* ---- the hot path, host buffers in / host buffers out ------------------------------------------------------ */

/* XorbReader.init(alloc, blob).extractChunkRange(start, end) (src/xet_bridge.zig:256-257,
 * src/parallel_download.zig:65-66): decoded bytes of chunks [start, end), concatenated, into out[0..*out_len).
 * start/end are relative to the blob actually fetched (XorbFetchResult.local_start/local_end,
 * src/xet_bridge.zig:45-52). */
ZG_API int zg_extract_chunk_range(zg_ctx* ctx, const uint8_t* blob, size_t len, uint32_t start, uint32_t end, uint8_t* out,
                           size_t out_cap, uint64_t* out_len);

/* Verify n whole xorbs: decode every chunk, BLAKE3-keyed chunk hashes, Merkle fold, compare with the content
 * address.  This is the integrity check zest's README promises (README.md:232,255) and defers to zig-xet
 * (src/bt_peer.zig:166-169); it belongs right after XetBridge.fetchXorbForTerm returns a full xorb
 * (src/xet_bridge.zig:149-218) and before XorbCache.put (src/swarm.zig:57-148).
 *   expected      n x 32 bytes (the xorb hashes = cache keys = term.hash), or NULL to only compute roots
 *   verdict[i]    1 = root matches, 0 = mismatch, < 0 = ZG_ERR_* of the first chunk that failed to decode
 *   roots         n x 32 bytes, computed Merkle roots (zeros for a xorb that failed)
 *   chunk_hashes  optional, chunk_hash_cap x 32 bytes, all xorbs' chunk hashes back to back
 *   bytes_verified  optional, total uncompressed bytes of the xorbs that decoded */
ZG_API int zg_verify_xorbs(zg_ctx* ctx, const uint8_t* const* blobs, const size_t* lens, const uint8_t* expected, size_t n,
                    int8_t* verdict, uint8_t* roots, uint8_t* chunk_hashes, size_t chunk_hash_cap,
                    uint64_t* bytes_verified);

/* Xet chunk hashes (zig-xet hashing; BLAKE3 keyed with the Xet data key) of chunks
 * data[offsets[i] .. offsets[i+1]), i < n.  key = NULL selects the Xet data key; key = 32 bytes selects another
 * BLAKE3 key; ZG_PLAIN_BLAKE3 selects unkeyed BLAKE3, which is what `zest bench --synthetic` times
 * (std.crypto.hash.Blake3, src/bench.zig:207-223). */
#define ZG_PLAIN_BLAKE3 ((const uint8_t*)(uintptr_t)1)
ZG_API int zg_hash_chunks(zg_ctx* ctx, const uint8_t* data, const uint64_t* offsets, size_t n, const uint8_t* key,
                   uint8_t* out_hashes);

/* Verification ("range") hashes of reconstruction terms: term t covers the chunks bounds[t] .. bounds[t+1]-1 of the
 * chunk-hash list and its hash is BLAKE3.keyed(VERIFICATION_KEY, those 32-byte hashes back to back) -- the value a
 * Xet shard stores beside every term (xet-core mdb_shard FileVerificationEntry).  With it the bytes a term
 * contributed in XetBridge.reconstructToFile (src/xet_bridge.zig:231-264: one extractChunkRange per term) can be
 * checked against the file's metadata: extract, zg_hash_chunks, zg_range_hashes, compare.  zest itself does not
 * verify terms today.  out holds n_terms x 32 bytes. */
ZG_API int zg_range_hashes(zg_ctx* ctx, const uint8_t* chunk_hashes, const uint32_t* bounds /* n_terms + 1 chunk indices */,
                    size_t n_terms, uint8_t* out);

/* Merkle root of (hash, length) pairs -- zig-xet hashing's aggregated node hash (CONTRIBUTING.md:144-146).
 * file_hash != 0 additionally applies the zero-key file-hash step (the value hf_xet.hash_files reports). */
ZG_API int zg_merkle_root(zg_ctx* ctx, const uint8_t* hashes, const uint64_t* lens, size_t n, int file_hash,
                   uint8_t out[32]);

/* Gearhash content-defined chunking (zig-xet chunking; parameters src/config.zig:12): chunk END offsets
 * (exclusive, ascending) of data[0..len).  *n_ends receives the count even when it exceeds cap
 * (then ZG_ERR_CAPACITY is returned and only cap entries are written). */
ZG_API int zg_cdc(zg_ctx* ctx, const uint8_t* data, size_t len, uint64_t* ends, size_t cap, size_t* n_ends);

/* Seed path in one call: CDC + chunk hashes (+ optional Xet file hash) -- what feeds storage.writeChunk
 * (src/storage.zig:102-107).  hashes holds cap x 32 bytes. */
ZG_API int zg_cdc_hash(zg_ctx* ctx, const uint8_t* data, size_t len, uint64_t* ends, size_t cap, size_t* n_ends,
                uint8_t* hashes, uint8_t* file_hash /* 32 bytes or NULL */);

/* ---- the same work on device-resident buffers (asynchronous on `stream`, a cudaStream_t) -------------------- */
/* For callers that already hold the data in HBM (bench.py's device-resident figure, pipelines that overlap
 * their own copies).  All pointers are device pointers unless noted; buffers must be 16-byte aligned and
 * readable up to the next multiple of 16.  Scratch is caller-provided (size from the *_scratch_bytes query),
 * 256-byte aligned; nothing is allocated or synchronised inside these calls. */

typedef struct zg_chunk_desc { /* one per chunk; fill from zg_xorb_index (src_off = hdr_off + 8 + blob base) */
    uint64_t src_off;          /* payload offset inside the compressed buffer                       */
    uint64_t dst_off;          /* where the decoded chunk goes inside the output buffer             */
    uint32_t clen, ulen, scheme;
    uint32_t order;            /* work order, filled by zg_chunk_order: the t-th chunk decoded is descs[t].order */
} zg_chunk_desc;

/* Fill descs[t].order (host memory) with a longest-first permutation of 0..n-1.  Must be called before the
 * descriptors are copied to the device: the decoder hands chunks out in this order so that its last warps do not
 * finish alone.  No counterpart in zest (scheduling detail of the GPU decoder). */
ZG_API void zg_chunk_order(zg_chunk_desc* descs, size_t n);

/* Chunk hashes on the device.  The scratch is sized from data_bytes and n, while the chunk lengths are read on the device:
 * the chunks must not overlap and their lengths must add up to at most data_bytes (true for anything laid out by
 * zg_xorb_index / zg_dev_cdc).  Lengths that add up to more are not allowed to write past the scratch: the kernels clamp
 * to the capacity they were sized for and the chunks that did not fit come out as 32 zero bytes (which fails any
 * verification) -- the same holds for zg_dev_hash_chunks_static and for the ulen fields given to zg_dev_verify_xorbs. */
ZG_API size_t zg_dev_hash_scratch_bytes(uint64_t data_bytes, size_t n);
ZG_API int zg_dev_hash_chunks(zg_ctx* ctx, const uint8_t* d_data, uint64_t data_bytes, const uint64_t* d_offsets,
                       const uint32_t* d_lens /* NULL: lens from offsets[i+1]-offsets[i] */, size_t n,
                       const uint8_t* key /* host pointer, as zg_hash_chunks */, uint8_t* d_out, void* d_scratch,
                       size_t scratch_bytes, void* stream);

/* The same hash pass for callers that cannot allocate device memory: a transfer loop whose memory budget is already
 * spent on staging and decode buffers, a thread that must not take the driver's allocation lock mid-download, a
 * process whose allocator is frozen after start-up.  Inputs and outputs are the caller's device buffers, as for
 * zg_dev_hash_chunks; the working set lives in an 8 MiB array built into the library (held by every GPU that loads
 * it), enough for any batch of up to 256 MiB of data in up to 65,536 chunks -- a batch whose working set
 * (zg_dev_hash_scratch_bytes) exceeds the array is refused with ZG_ERR_LIMIT before anything is launched.  Works on the CUDA device that is current on the calling thread; the first call on a device looks up
 * the array's address and should happen before any stream capture.  Calls on one device share the array: keep them
 * in order (one stream), overlapping calls race.  Calls no cudaMalloc / cudaFree and no allocating library routine.
 * This entry point has no counterpart in the reference. */
ZG_API int zg_dev_hash_chunks_static(const uint8_t* d_data, uint64_t data_bytes, const uint64_t* d_offsets,
                                     const uint32_t* d_lens, size_t n, const uint8_t* key, uint8_t* d_out,
                                     void* stream);

ZG_API size_t zg_dev_decode_scratch_bytes(size_t n_chunks, uint64_t out_bytes);
ZG_API int zg_dev_decode_chunks(zg_ctx* ctx, const uint8_t* d_comp, uint64_t comp_bytes, const zg_chunk_desc* d_chunks,
                         size_t n_chunks, uint8_t* d_out, uint64_t out_bytes, int32_t* d_status, void* d_scratch,
                         size_t scratch_bytes, void* stream);

/* The two halves of zg_dev_verify_xorbs for a caller that only wants hashes, not bytes (and for measuring them apart).
 * zg_dev_decode_for_hash is zg_dev_decode_chunks except that byte-grouped chunks (scheme 2 -- what safetensors xorbs use,
 * CLAUDE.md:22) are NOT regrouped into d_out: their four byte planes stay in the decode scratch, and
 * zg_dev_hash_decoded hashes them from there, interleaving the planes as it reads (the chunk hash is over the regrouped
 * bytes either way).  That saves a read and a write of every decoded byte: 17 % of the decode time on bf16.  d_out holds
 * the bytes of scheme-0/1 chunks only; dst_off must be a multiple of 16 (zg_xorb_index-style layouts are).  Pass the
 * same d_chunks, out_bytes and decode scratch to both.  No counterpart in zest: XorbReader.extractChunkRange always
 * materialises the bytes (src/xet_bridge.zig:256-257); verification has no use for them. */
ZG_API int zg_dev_decode_for_hash(zg_ctx* ctx, const uint8_t* d_comp, uint64_t comp_bytes, const zg_chunk_desc* d_chunks,
                           size_t n_chunks, uint8_t* d_out, uint64_t out_bytes, int32_t* d_status, void* d_scratch,
                           size_t scratch_bytes, void* stream);
ZG_API size_t zg_dev_hash_decoded_scratch_bytes(uint64_t out_bytes, size_t n);
ZG_API int zg_dev_hash_decoded(zg_ctx* ctx, const uint8_t* d_out, uint64_t out_bytes, const zg_chunk_desc* d_chunks, size_t n_chunks,
                        const void* d_decode_scratch, uint8_t* d_hashes, void* d_scratch, size_t scratch_bytes, void* stream);

ZG_API size_t zg_dev_merkle_scratch_bytes(size_t n_chunks, size_t n_xorbs);
/* d_first: n_xorbs + 1 chunk indices delimiting each xorb's chunks in d_hashes / d_lens. */
ZG_API int zg_dev_merkle(zg_ctx* ctx, const uint8_t* d_hashes, const uint32_t* d_lens, const uint32_t* d_first,
                  size_t n_xorbs, size_t n_chunks, const int32_t* d_status, const uint8_t* d_expected,
                  uint8_t* d_roots, int8_t* d_verdict, void* d_scratch, size_t scratch_bytes, void* stream);

/* decode -> hash -> fold -> compare in one call (zg_dev_decode_for_hash, zg_dev_hash_decoded and zg_dev_merkle on one
 * stream, sharing scratch; the decoded bytes are not an output of this call). */
ZG_API size_t zg_dev_verify_scratch_bytes(size_t n_chunks, size_t n_xorbs, uint64_t out_bytes);
ZG_API int zg_dev_verify_xorbs(zg_ctx* ctx, const uint8_t* d_comp, uint64_t comp_bytes, const zg_chunk_desc* d_chunks,
                        size_t n_chunks, const uint32_t* d_first, size_t n_xorbs, const uint8_t* d_expected,
                        uint64_t out_bytes, uint8_t* d_roots, int8_t* d_verdict, uint8_t* d_chunk_hashes,
                        void* d_scratch, size_t scratch_bytes, void* stream);

ZG_API size_t zg_dev_cdc_scratch_bytes(uint64_t len);
/* d_n_out: four u64 {number of cuts, overflow flag, 2 x internal}. */
ZG_API int zg_dev_cdc(zg_ctx* ctx, const uint8_t* d_data, uint64_t len, uint64_t* d_ends, uint64_t cap, uint64_t* d_n_out,
               void* d_scratch, size_t scratch_bytes, void* stream);

/* ---- several GPUs behind one call (SURVEY.md 8b "batch, shards across GPUs", 8e) ------------------------------------ */
/* zest is one process: XetBridge.reconstructToFile walks a file's terms (src/xet_bridge.zig:231-264) and the parallel
 * downloader fans them out to one task group (src/parallel_download.zig:91-134, 174-182).  A group gives that one
 * process every GPU of the box: one context and one worker thread per device.  zg_group_verify_xorbs splits the batch
 * over the devices by zg_partition_xorbs (greedy longest-first by compressed bytes; xorbs are independent, nothing
 * moves between GPUs, no collective), every device verifies its part through its own context (own lanes, streams and
 * PCIe link), and the results come back in the caller's order -- same arguments and same meaning as zg_verify_xorbs.
 * devices = NULL: devices 0 .. n_devices-1; n_devices <= 0: every visible device.  Pinned buffers from zg_pinned_alloc
 * of any of the group's contexts are pinned for all of them.  One group call runs at a time; a context obtained with
 * zg_group_ctx may be used directly (extract, hash, CDC) while no group call is in flight or from other threads. */
typedef struct zg_group zg_group;
ZG_API int zg_group_init(const int* devices, int n_devices, const zg_opts* opts, zg_group** out);
ZG_API void zg_group_shutdown(zg_group* g);
ZG_API int zg_group_size(const zg_group* g);
ZG_API zg_ctx* zg_group_ctx(zg_group* g, int i);
/* part[i] = which of n_parts owns xorb i: longest (by lens[i], the compressed bytes) first, each to the part with the
 * least bytes so far; ties go to the lower index / lower part, so every caller computes the same split. */
ZG_API int zg_partition_xorbs(const size_t* lens, size_t n, int n_parts, int32_t* part);
ZG_API int zg_group_verify_xorbs(zg_group* g, const uint8_t* const* blobs, const size_t* lens, const uint8_t* expected, size_t n,
                          int8_t* verdict, uint8_t* roots, uint8_t* chunk_hashes, size_t chunk_hash_cap,
                          uint64_t* bytes_verified);

/* ---- dedup index --------------------------------------------------------------------------- */
/* A table in HBM from Xet chunk hash to a caller-defined 64-bit location (for instance xorb number << 32 | chunk
 * index).  On the seed/dedup path it answers for a whole batch of hashes what zest asks one file-system probe at a
 * time -- storage.hasChunk (src/storage.zig:137) and the XorbRegistry scan (src/storage.zig:145-228) -- and what
 * xet-core keeps in its global dedup table; the hashes come from zg_cdc_hash and can stay on the device
 * (zg_dev_dedup_*).  Per hash the table keeps the smallest location ever inserted for it ("first seen" when
 * locations grow); was_new[i] = 1 iff the hash was absent before this call and occurrence i carries the location
 * kept.  max_entries fixes the size (the table is kept at most half full; 52 bytes per slot): inserting more returns
 * ZG_ERR_CAPACITY.  Calls on one table must not overlap (the host-buffer calls serialise themselves; device calls
 * are ordered by the caller's stream); errors of device calls (table full, two different hashes agreeing in their
 * 64 tag bits) surface at the next zg_dedup_count / host-buffer insert.  ZG_DEDUP_MISSING is not a valid location. */
typedef struct zg_dedup zg_dedup;
#define ZG_DEDUP_MISSING (~(uint64_t)0)
ZG_API int zg_dedup_create(zg_ctx* ctx, size_t max_entries, zg_dedup** out);
ZG_API void zg_dedup_destroy(zg_dedup* t);
ZG_API int zg_dedup_insert(zg_dedup* t, const uint8_t* hashes /* n x 32 */, const uint64_t* locations, size_t n,
                    uint8_t* was_new /* n bytes or NULL */);
ZG_API int zg_dedup_lookup(zg_dedup* t, const uint8_t* hashes, size_t n, uint64_t* locations /* ZG_DEDUP_MISSING: absent */);
ZG_API int zg_dedup_count(zg_dedup* t, uint64_t* entries);
ZG_API int zg_dev_dedup_insert(zg_dedup* t, const uint8_t* d_hashes, const uint64_t* d_locations, size_t n, uint8_t* d_was_new,
                        void* stream);
ZG_API int zg_dev_dedup_lookup(zg_dedup* t, const uint8_t* d_hashes, size_t n, uint64_t* d_locations, void* stream);

/* ---- measurement aid -------------------------------------------------------------------------------------------- */
/* Measured peak of the SM's integer pipes in Gop/s (thread-level 32-bit operations): the ALU pipe (LOP3 / SHF / IADD3),
 * the FMA pipe (IMAD) and both together.  bench.py reports the BLAKE3 kernel against these (SURVEY.md 8d asks for a
 * calibrated integer roofline).  d_sink: any 4 writable device bytes.  Synchronises on `stream`.  No counterpart in
 * zest. */
ZG_API int zg_dev_int_pipe_peak(void* d_sink, double* alu_gops, double* fma_gops, double* both_gops, void* stream);

/* ---- statistics -------------------------------------------------------------------------------------------- */
typedef struct zg_stats {
    uint64_t calls, kernels_launched, h2d_bytes, d2h_bytes, bytes_decoded, chunks_hashed;
    /* GPU time of the stages of zg_verify_xorbs, summed over its sub-batches (CUDA events on each lane's stream;
     * sub-batches on different lanes overlap, so the sum can exceed the wall time of a call): copies in, LZ4/BG4 decode,
     * chunk hashing, Merkle fold + compare, results out. */
    uint64_t ns_h2d, ns_decode, ns_hash, ns_merkle, ns_d2h;
} zg_stats;
ZG_API int zg_get_stats(zg_ctx* ctx, zg_stats* out);

#ifdef __cplusplus
}
#endif
#endif /* ZEST_GPU_H */
