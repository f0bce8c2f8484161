// C ABI of libzest_gpu.so (declared in include/zest_gpu.h): context, lanes (stream + grow-only device and pinned
// buffers), the host walk of the xorb chunk stream, and the host-buffer entry points that stand where zest calls
// zig-xet today (/root/reference/src/xet_bridge.zig:256-257, src/parallel_download.zig:65-66).
// There is no CPU path in here: every data-plane call ends in the kernels of hash.cu / lz4.cu / merkle.cu / cdc.cu.
#include <atomic>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <new>
#include <vector>

#include "../../include/zest_gpu.h"
#include "kernels.h"

using namespace zg;

static_assert(sizeof(zg_chunk_desc) == sizeof(ChunkDesc), "ABI struct must match the kernel's descriptor");

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}
}  // namespace
namespace zg {
int set_last_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}
}  // namespace zg
namespace {
int cuda_fail(cudaError_t e, const char* what) {
    int code = (e == cudaErrorMemoryAllocation) ? ZG_ERR_OOM : ZG_ERR_CUDA;
    return fail(code, "%s: %s", what, cudaGetErrorString(e));
}
#define CU(call)                                       \
    do {                                               \
        cudaError_t e__ = (call);                      \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
    } while (0)

const uint8_t kDataKey[32] = {102, 151, 245, 119, 91,  149, 80,  222, 49,  53,  203, 172, 165, 151, 24,  28,
                              157, 228, 33,  16,  155, 235, 43,  88,  180, 208, 176, 75,  147, 173, 242, 41};

B3Key make_key(const uint8_t* key) {
    B3Key k;
    if (key == ZG_PLAIN_BLAKE3) {
        const uint32_t iv[8] = {0x6A09E667u, 0xBB67AE85u, 0x3C6EF372u, 0xA54FF53Au,
                                0x510E527Fu, 0x9B05688Cu, 0x1F83D9ABu, 0x5BE0CD19u};
        memcpy(k.w, iv, 32);
        k.base_flags = 0;
        k.one = 1;
        return k;
    }
    const uint8_t* p = key ? key : kDataKey;
    for (int i = 0; i < 8; i++)
        k.w[i] = (uint32_t)p[4 * i] | ((uint32_t)p[4 * i + 1] << 8) | ((uint32_t)p[4 * i + 2] << 16) | ((uint32_t)p[4 * i + 3] << 24);
    k.base_flags = B3_KEYED;
    k.one = 1;
    return k;
}

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes, cudaStream_t st) {
        if (bytes <= cap) return ZG_OK;
        if (p) { CU(cudaStreamSynchronize(st)); CU(cudaFree(p)); p = nullptr; cap = 0; }
        size_t want = align_up(bytes + bytes / 8 + 256, 1 << 20);
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { p = nullptr; (void)cudaGetLastError(); return fail(ZG_ERR_OOM, "cudaMalloc(%zu): %s", want, cudaGetErrorString(e)); }
        cap = want;
        return ZG_OK;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes, cudaStream_t st) {
        if (bytes <= cap) return ZG_OK;
        if (p) { CU(cudaStreamSynchronize(st)); CU(cudaFreeHost(p)); p = nullptr; cap = 0; }
        size_t want = align_up(bytes + bytes / 8 + 256, 1 << 16);
        cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocDefault);
        if (e != cudaSuccess) { p = nullptr; (void)cudaGetLastError(); return fail(ZG_ERR_OOM, "cudaHostAlloc(%zu): %s", want, cudaGetErrorString(e)); }
        cap = want;
        return ZG_OK;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

struct Lane {
    cudaStream_t st = nullptr;
    DevBuf comp, out, meta, res, scratch;  // compressed in, decoded out, descriptors/offsets, results, kernel scratch
    PinBuf hmeta, hres;                    // pinned staging for descriptors and results
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // verify stage marks (created on first use)
    bool busy = false;
};

}  // namespace

struct zg_ctx {
    int device = 0;
    std::vector<Lane> lanes;
    std::mutex mu;
    std::condition_variable cv;
    std::atomic<uint64_t> calls{0}, kernels{0}, h2d{0}, d2h{0}, decoded{0}, hashed{0};
    std::atomic<uint64_t> ns_h2d{0}, ns_decode{0}, ns_hash{0}, ns_merkle{0}, ns_d2h{0};  // GPU time per verify stage (events)
};

namespace {

struct LaneGuard {
    zg_ctx* c;
    Lane* l;
    LaneGuard(zg_ctx* ctx, bool wait = true) : c(ctx), l(nullptr) {
        std::unique_lock<std::mutex> lk(c->mu);
        for (;;) {
            for (auto& x : c->lanes)
                if (!x.busy) { x.busy = true; l = &x; return; }
            if (!wait) return;
            c->cv.wait(lk);
        }
    }
    ~LaneGuard() {
        if (!l) return;
        { std::lock_guard<std::mutex> lk(c->mu); l->busy = false; }
        c->cv.notify_one();
    }
};

// ---- host walk of the chunk stream ----
struct HostIndex {
    std::vector<uint64_t> hdr_off;
    std::vector<uint32_t> clen, ulen;
    std::vector<uint8_t> scheme;
    uint64_t total_ulen = 0;
};

int walk(const uint8_t* blob, size_t len, uint32_t stop_after, HostIndex* ix, uint32_t* n_out) {
    size_t pos = 0;
    uint32_t n = 0;
    uint64_t tot = 0;
    // A stored xorb file (xet-core's local store, zg_xorb_footer / seal_xorb) carries a "XETBLOB" footer after the chunk
    // stream whose last 4 bytes give its length: the stream ends where that footer begins.
    if (len >= 100) {
        const uint8_t* t = blob + len - 4;
        const uint64_t info = (uint64_t)t[0] | ((uint64_t)t[1] << 8) | ((uint64_t)t[2] << 16) | ((uint64_t)t[3] << 24);
        if (info + 4 <= len && info >= 92 && memcmp(blob + len - 4 - info, "XETBLOB", 7) == 0) len -= (size_t)info + 4;
    }
    while (pos < len && n < stop_after) {
        if (len - pos < 8) return fail(ZG_ERR_TRUNCATED, "chunk %u: header runs past the blob", n);
        const uint8_t* h = blob + pos;
        if (h[0] != 0) return fail(ZG_ERR_BAD_HEADER, "chunk %u: header version %u", n, h[0]);
        uint32_t c = h[1] | (h[2] << 8) | ((uint32_t)h[3] << 16);
        uint32_t u = h[5] | (h[6] << 8) | ((uint32_t)h[7] << 16);
        if (h[4] > 2) return fail(ZG_ERR_SCHEME, "chunk %u: compression scheme %u is not supported", n, h[4]);
        if (h[4] == 0 && c != u) return fail(ZG_ERR_BAD_HEADER, "chunk %u: stored chunk with clen != ulen", n);
        if (c > len - pos - 8) return fail(ZG_ERR_TRUNCATED, "chunk %u: payload runs past the blob", n);
        if (ix) { ix->hdr_off.push_back(pos); ix->clen.push_back(c); ix->ulen.push_back(u); ix->scheme.push_back(h[4]); }
        tot += u;
        pos += 8 + (size_t)c;
        n++;
    }
    if (ix) ix->total_ulen = tot;
    *n_out = n;
    return ZG_OK;
}

// Every entry point that takes a context works on that context's device and puts the calling thread's current device
// back when it returns (a thread with torch on cuda:0 may hold a context on device 1; a zg_group worker and its caller
// sit on different devices).
struct DeviceScope {
    int prev = -1, rc = ZG_OK;
    explicit DeviceScope(zg_ctx* ctx) {
        if (!ctx) { rc = fail(ZG_ERR_INVALID_ARG, "null context"); return; }
        int cur = -1;
        if (cudaGetDevice(&cur) != cudaSuccess) { (void)cudaGetLastError(); cur = -1; }
        if (cur == ctx->device) return;
        cudaError_t e = cudaSetDevice(ctx->device);
        if (e != cudaSuccess) { rc = cuda_fail(e, "cudaSetDevice"); return; }
        prev = cur;
    }
    ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceScope(const DeviceScope&) = delete;
    DeviceScope& operator=(const DeviceScope&) = delete;
};

__global__ void split_desc_kernel(const ChunkDesc* __restrict__ d, size_t n, uint64_t* __restrict__ off, uint32_t* __restrict__ len) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { off[i] = d[i].dst_off; len[i] = d[i].ulen; }
}
}  // namespace

extern "C" {

int zg_init(int device, const zg_opts* opts, zg_ctx** out) {
    if (!out) return fail(ZG_ERR_INVALID_ARG, "out is null");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        (void)cudaGetLastError();
        return fail(ZG_ERR_NO_DEVICE, "no CUDA device (%s); libzest_gpu has no CPU path", e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
    }
    if (device < 0 || device >= count) return fail(ZG_ERR_INVALID_ARG, "device %d out of range (0..%d)", device, count - 1);
    // work on `device`, then put the calling thread's device back (whatever way this function is left)
    struct Restore {
        int prev = -1;
        ~Restore() { if (prev >= 0) cudaSetDevice(prev); }
    } restore;
    if (cudaGetDevice(&restore.prev) != cudaSuccess) { (void)cudaGetLastError(); restore.prev = -1; }
    if (restore.prev == device) restore.prev = -1;
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(ZG_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    zg_ctx* c = new (std::nothrow) zg_ctx();
    if (!c) return fail(ZG_ERR_OOM, "out of host memory");
    c->device = device;
    uint32_t nl = (opts && opts->struct_size >= 8 && opts->n_lanes) ? opts->n_lanes : 8;
    if (nl > 64) nl = 64;
    c->lanes.resize(nl);
    for (auto& l : c->lanes) {
        e = cudaStreamCreateWithFlags(&l.st, cudaStreamNonBlocking);
        if (e != cudaSuccess) { zg_shutdown(c); return cuda_fail(e, "cudaStreamCreate"); }
    }
    *out = c;
    return ZG_OK;
}

void zg_shutdown(zg_ctx* c) {
    if (!c) return;
    DeviceScope dev_scope(c);
    for (auto& l : c->lanes) {
        if (l.st) cudaStreamSynchronize(l.st);
        l.comp.release(); l.out.release(); l.meta.release(); l.res.release(); l.scratch.release();
        l.hmeta.release(); l.hres.release();
        for (auto& e : l.ev) if (e) cudaEventDestroy(e);
        if (l.st) cudaStreamDestroy(l.st);
    }
    delete c;
}

const char* zg_strerror(int code) {
    switch (code) {
        case ZG_OK: return "ok";
        case ZG_ERR_INVALID_ARG: return "invalid argument";
        case ZG_ERR_CUDA: return "CUDA error";
        case ZG_ERR_TRUNCATED: return "xorb truncated";
        case ZG_ERR_BAD_HEADER: return "bad chunk header";
        case ZG_ERR_SCHEME: return "unsupported compression scheme";
        case ZG_ERR_LZ4_FRAME: return "bad LZ4 frame";
        case ZG_ERR_LZ4_BLOCK: return "bad LZ4 block";
        case ZG_ERR_SIZE_MISMATCH: return "decoded size mismatch";
        case ZG_ERR_RANGE: return "chunk range out of bounds";
        case ZG_ERR_CAPACITY: return "output buffer too small";
        case ZG_ERR_NO_DEVICE: return "no usable CUDA device";
        case ZG_ERR_OOM: return "out of memory";
        case ZG_ERR_LIMIT: return "input exceeds a limit of this entry point";
        default: return "unknown error";
    }
}

const char* zg_last_error(void) { return g_err; }

int zg_pinned_alloc(zg_ctx* ctx, size_t bytes, void** out) {
    if (!out) return fail(ZG_ERR_INVALID_ARG, "out is null");
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable);  // pinned for every device (zg_group)
    if (e != cudaSuccess) { (void)cudaGetLastError(); *out = nullptr; return fail(ZG_ERR_OOM, "cudaHostAlloc(%zu): %s", bytes, cudaGetErrorString(e)); }
    return ZG_OK;
}
int zg_pinned_free(zg_ctx* ctx, void* p) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (p) CU(cudaFreeHost(p));
    return ZG_OK;
}

void zg_hash_to_hex(const uint8_t h[32], char out[65]) {
    static const char* d = "0123456789abcdef";
    for (int g = 0; g < 4; g++)
        for (int b = 0; b < 8; b++) {
            uint8_t v = h[g * 8 + (7 - b)];
            out[g * 16 + 2 * b] = d[v >> 4];
            out[g * 16 + 2 * b + 1] = d[v & 15];
        }
    out[64] = 0;
}
int zg_hex_to_hash(const char* hex, uint8_t out[32]) {
    if (!hex || strlen(hex) != 64) return fail(ZG_ERR_INVALID_ARG, "hash hex must be 64 characters");
    for (int g = 0; g < 4; g++)
        for (int b = 0; b < 8; b++) {
            int v = 0;
            for (int k = 0; k < 2; k++) {
                char ch = hex[g * 16 + 2 * b + k];
                int d = (ch >= '0' && ch <= '9') ? ch - '0' : (ch >= 'a' && ch <= 'f') ? ch - 'a' + 10 : (ch >= 'A' && ch <= 'F') ? ch - 'A' + 10 : -1;
                if (d < 0) return fail(ZG_ERR_INVALID_ARG, "hash hex has a non-hex character");
                v = v * 16 + d;
            }
            out[g * 8 + (7 - b)] = (uint8_t)v;
        }
    return ZG_OK;
}

int zg_xorb_index(const uint8_t* blob, size_t len, uint32_t cap, uint32_t* n_chunks, uint64_t* hdr_off, uint32_t* clen,
                  uint32_t* ulen, uint8_t* scheme, uint64_t* total_ulen) {
    if (!blob && len) return fail(ZG_ERR_INVALID_ARG, "blob is null");
    HostIndex ix;
    uint32_t n = 0;
    int r = walk(blob, len, 0xffffffffu, &ix, &n);
    if (r) return r;
    if (n_chunks) *n_chunks = n;
    if (total_ulen) *total_ulen = ix.total_ulen;
    if (hdr_off || clen || ulen || scheme) {
        if (n > cap) return fail(ZG_ERR_CAPACITY, "blob has %u chunks, arrays hold %u", n, cap);
        if (hdr_off) memcpy(hdr_off, ix.hdr_off.data(), n * sizeof(uint64_t));
        if (clen) memcpy(clen, ix.clen.data(), n * sizeof(uint32_t));
        if (ulen) memcpy(ulen, ix.ulen.data(), n * sizeof(uint32_t));
        if (scheme) memcpy(scheme, ix.scheme.data(), n);
    }
    return ZG_OK;
}

int zg_chunk_range_size(const uint8_t* blob, size_t len, uint32_t start, uint32_t end, uint64_t* out_len) {
    if (!out_len || (!blob && len)) return fail(ZG_ERR_INVALID_ARG, "null argument");
    if (start > end) return fail(ZG_ERR_RANGE, "chunk range %u..%u is reversed", start, end);
    HostIndex ix;
    uint32_t n = 0;
    int r = walk(blob, len, end, &ix, &n);
    if (r) return r;
    if (n < end) return fail(ZG_ERR_RANGE, "chunk range %u..%u outside a blob of %u chunks", start, end, n);
    uint64_t t = 0;
    for (uint32_t i = start; i < end; i++) t += ix.ulen[i];
    *out_len = t;
    return ZG_OK;
}

// -------------------------------------------------------------------------------------------------------------
// host-buffer entry points
// -------------------------------------------------------------------------------------------------------------

int zg_extract_chunk_range(zg_ctx* ctx, const uint8_t* blob, size_t len, uint32_t start, uint32_t end, uint8_t* out,
                           size_t out_cap, uint64_t* out_len) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if ((!blob && len) || !out_len) return fail(ZG_ERR_INVALID_ARG, "null argument");
    if (start > end) return fail(ZG_ERR_RANGE, "chunk range %u..%u is reversed", start, end);
    HostIndex ix;
    uint32_t n = 0;
    r = walk(blob, len, end, &ix, &n);
    if (r) return r;
    if (n < end) return fail(ZG_ERR_RANGE, "chunk range %u..%u outside a blob of %u chunks", start, end, n);
    uint32_t cnt = end - start;
    uint64_t total = 0;
    for (uint32_t i = start; i < end; i++) total += ix.ulen[i];
    *out_len = total;
    if (total > out_cap) return fail(ZG_ERR_CAPACITY, "range decodes to %llu bytes, buffer holds %zu", (unsigned long long)total, out_cap);
    if (cnt == 0 || total == 0) return ZG_OK;
    if (!out) return fail(ZG_ERR_INVALID_ARG, "out is null");
    ctx->calls++;
    LaneGuard g(ctx);
    Lane& L = *g.l;
    uint64_t src0 = ix.hdr_off[start];
    uint64_t src1 = ix.hdr_off[end - 1] + 8 + ix.clen[end - 1];
    uint64_t cbytes = src1 - src0;
    if ((r = L.comp.ensure(cbytes + 16, L.st))) return r;
    if ((r = L.out.ensure(total + 16, L.st))) return r;
    size_t sbytes = decode_scratch_bytes(cnt, total);
    if ((r = L.scratch.ensure(sbytes, L.st))) return r;
    size_t mbytes = align_up(cnt * sizeof(ChunkDesc), 256) + align_up(cnt * 4, 256);
    if ((r = L.meta.ensure(mbytes, L.st))) return r;
    if ((r = L.hmeta.ensure(mbytes, L.st))) return r;
    ChunkDesc* hd = static_cast<ChunkDesc*>(L.hmeta.p);
    uint64_t dpos = 0;
    for (uint32_t i = 0; i < cnt; i++) {
        uint32_t k = start + i;
        hd[i] = ChunkDesc{ix.hdr_off[k] + 8 - src0, dpos, ix.clen[k], ix.ulen[k], ix.scheme[k], 0};
        dpos += ix.ulen[k];
    }
    fill_decode_order(hd, cnt);
    ChunkDesc* dd = static_cast<ChunkDesc*>(L.meta.p);
    int32_t* dstatus = reinterpret_cast<int32_t*>(static_cast<uint8_t*>(L.meta.p) + align_up(cnt * sizeof(ChunkDesc), 256));
    CU(cudaMemcpyAsync(L.comp.p, blob + src0, cbytes, cudaMemcpyHostToDevice, L.st));
    CU(cudaMemcpyAsync(dd, hd, cnt * sizeof(ChunkDesc), cudaMemcpyHostToDevice, L.st));
    CU(launch_decode(static_cast<const uint8_t*>(L.comp.p), cbytes, dd, cnt, static_cast<uint8_t*>(L.out.p), total, dstatus,
                     L.scratch.p, L.scratch.cap, L.st));
    ctx->kernels += 1;
    if ((r = L.hres.ensure(cnt * 4, L.st))) return r;
    CU(cudaMemcpyAsync(L.hres.p, dstatus, cnt * 4, cudaMemcpyDeviceToHost, L.st));
    CU(cudaMemcpyAsync(out, L.out.p, total, cudaMemcpyDeviceToHost, L.st));
    CU(cudaStreamSynchronize(L.st));
    ctx->h2d += cbytes + cnt * sizeof(ChunkDesc);
    ctx->d2h += total + cnt * 4;
    ctx->decoded += total;
    const int32_t* hs = static_cast<const int32_t*>(L.hres.p);
    for (uint32_t i = 0; i < cnt; i++)
        if (hs[i]) return fail(hs[i], "chunk %u failed to decode: %s", start + i, zg_strerror(hs[i]));
    return ZG_OK;
}

namespace {

// One sub-batch of xorbs in flight on a lane.
struct VerifySlot {
    double t_plan = 0, t_enq = 0;  // ZEST_VERIFY_TRACE: host-side times of planning and enqueueing
    Lane* L = nullptr;
    bool launched = false;
    size_t x0 = 0, nx = 0;
    std::vector<int> walk_err;        // per xorb: ZG_OK or the error of the host walk
    std::vector<uint64_t> ulen_total;  // per xorb
};

uint64_t sub_batch_bytes() {  // ZEST_VERIFY_SUBBATCH_MB: tuning knob
    static const uint64_t v = []() { const char* e = getenv("ZEST_VERIFY_SUBBATCH_MB"); long m = e ? atol(e) : 0; return (uint64_t)(m > 0 ? m : 128) << 20; }();
    return v;
}
int max_verify_lanes() {  // ZEST_VERIFY_LANES: tuning knob
    static const int v = []() { const char* e = getenv("ZEST_VERIFY_LANES"); long m = e ? atol(e) : 0; return (int)(m > 0 && m < 8 ? m : 4); }();
    return v;
}

}  // namespace

int zg_verify_xorbs(zg_ctx* ctx, const uint8_t* const* blobs, const size_t* lens, const uint8_t* expected, size_t n,
                    int8_t* verdict, uint8_t* roots, uint8_t* chunk_hashes, size_t chunk_hash_cap,
                    uint64_t* bytes_verified) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (n && (!blobs || !lens || !verdict || !roots)) return fail(ZG_ERR_INVALID_ARG, "null argument");
    if (bytes_verified) *bytes_verified = 0;
    if (n == 0) return ZG_OK;
    ctx->calls++;
    // Up to four lanes (ZEST_VERIFY_LANES): the copies of one sub-batch overlap the kernels of the others and the host walk of the next.
    // Several sub-batches in flight also matter for the decoder itself: one LZ4 chunk is a serial chain of a few
    // milliseconds, so a single ~128 MB sub-batch cannot fill the GPU; three or four of them on separate streams do.
    constexpr int kMaxSlots = 8;
    std::unique_ptr<LaneGuard> guards[kMaxSlots];
    VerifySlot slots[kMaxSlots];
    int n_slots = 0;
    for (int i = 0; i < kMaxSlots && i < max_verify_lanes(); i++) {
        guards[i].reset(new LaneGuard(ctx, /*wait=*/i == 0));  // the first lane is waited for, the others only if free
        if (!guards[i]->l) { guards[i].reset(); break; }
        slots[n_slots++].L = guards[i]->l;
    }
    size_t hashes_written = 0;
    uint64_t verified = 0;
    // ZEST_VERIFY_TRACE=1 (developer aid): per sub-batch, when its copies, kernels and result download started and ended
    static const bool trace = getenv("ZEST_VERIFY_TRACE") != nullptr;
    cudaEvent_t ev0 = nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    auto now_ms = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count(); };
    if (trace) { cudaEventCreate(&ev0); cudaEventRecord(ev0, slots[0].L->st); }

    auto collect = [&](VerifySlot& S) -> int {  // wait for the slot's sub-batch and hand its results to the caller
        if (!S.launched) return ZG_OK;
        S.launched = false;
        CU(cudaStreamSynchronize(S.L->st));
        {   // stage marks of this sub-batch (lane events): GPU time per stage for zg_get_stats, and the trace
            cudaEvent_t* ev = S.L->ev;
            float dt[5] = {0, 0, 0, 0, 0};
            bool okt = ev[5] != nullptr;
            for (int k = 0; k < 5 && okt; k++) okt = cudaEventElapsedTime(&dt[k], ev[k], ev[k + 1]) == cudaSuccess;
            if (okt) {
                ctx->ns_h2d += (uint64_t)(dt[0] * 1e6f); ctx->ns_decode += (uint64_t)(dt[1] * 1e6f); ctx->ns_hash += (uint64_t)(dt[2] * 1e6f);
                ctx->ns_merkle += (uint64_t)(dt[3] * 1e6f); ctx->ns_d2h += (uint64_t)(dt[4] * 1e6f);
            } else
                (void)cudaGetLastError();
            if (trace && okt) {
                float a = 0;
                cudaEventElapsedTime(&a, ev0, ev[0]);
                fprintf(stderr, "[zest trace] xorbs %zu..%zu: planned %.2f enqueued %.2f | copy %.2f-%.2f decode -%.2f hash -%.2f fold -%.2f results %.2f | collected %.2f ms\n",
                        S.x0, S.x0 + S.nx, S.t_plan, S.t_enq, a, a + dt[0], a + dt[0] + dt[1], a + dt[0] + dt[1] + dt[2],
                        a + dt[0] + dt[1] + dt[2] + dt[3], a + dt[0] + dt[1] + dt[2] + dt[3] + dt[4], now_ms());
            }
        }
        const uint8_t* hr = static_cast<const uint8_t*>(S.L->hres.p);
        memcpy(roots + 32 * S.x0, hr, 32 * S.nx);
        memcpy(verdict + S.x0, hr + align_up(32 * S.nx, 256), S.nx);
        for (size_t i = 0; i < S.nx; i++) {
            if (S.walk_err[i]) {  // never reached the GPU as chunks: the verdict is the walk error
                verdict[S.x0 + i] = (int8_t)S.walk_err[i];
                memset(roots + 32 * (S.x0 + i), 0, 32);
            } else if (verdict[S.x0 + i] >= 0)
                verified += S.ulen_total[i];
        }
        return ZG_OK;
    };

    int rc = ZG_OK;
    size_t x = 0;
    int si = 0;
    std::vector<HostIndex> ixs;
    while (x < n) {
        // ---- plan the next sub-batch on the host (the previous ones are running)
        ixs.clear();
        std::vector<int> walk_err;
        std::vector<uint64_t> utot;
        size_t x0 = x, nc = 0;
        uint64_t ub = 0, cb = 0;
        while (x < n && (x == x0 || ub < sub_batch_bytes())) {
            ixs.emplace_back();
            uint32_t c = 0;
            int wr = (!blobs[x] && lens[x]) ? ZG_ERR_INVALID_ARG : walk(blobs[x], lens[x], 0xffffffffu, &ixs.back(), &c);
            if (wr != ZG_OK) { ixs.back() = HostIndex(); c = 0; }
            else { ub += ixs.back().total_ulen + 16ull * c; cb += align_up(lens[x], 16); nc += c; }
            walk_err.push_back(wr);
            utot.push_back(ixs.back().total_ulen);
            x++;
        }
        size_t nx = x - x0;
        VerifySlot& S = slots[si];
        const double t_planned = now_ms();
        if ((rc = collect(S))) break;
        Lane& L = *S.L;
        S.t_plan = t_planned;
        for (auto& e : L.ev) if (!e) cudaEventCreate(&e);
        cudaEventRecord(L.ev[0], L.st);
        S.x0 = x0; S.nx = nx; S.walk_err = walk_err; S.ulen_total = utot;
        if (chunk_hashes && hashes_written + nc > chunk_hash_cap) {
            rc = fail(ZG_ERR_CAPACITY, "chunk_hashes holds %zu hashes, the batch has more", chunk_hash_cap);
            break;
        }
        // ---- layout: one upload block [desc | first | off | len | expected], one result block [roots | verdict | hashes]
        int group = hash_pick_group(ub);
        size_t sz_dec = align_up(decode_scratch_bytes(nc, ub), 256), sz_hash = align_up(hash_scratch_bytes(ub, nc, group), 256),
               sz_mer = merkle_scratch_bytes(nc, nx);
        size_t o_first = align_up(nc * sizeof(ChunkDesc), 256), o_off = o_first + align_up((nx + 1) * 4, 256),
               o_len = o_off + align_up(nc * 8 + 8, 256), o_exp = o_len + align_up(nc * 4 + 4, 256),
               up_bytes = o_exp + align_up(nx * 32, 256), o_status = up_bytes, meta_total = o_status + align_up(nc * 4 + 4, 256);
        size_t r_verdict = align_up(32 * nx, 256), r_hashes = r_verdict + align_up(nx, 256), res_total = r_hashes + align_up(32 * nc + 32, 256);
        if ((rc = L.comp.ensure(cb + 16, L.st)) || (rc = L.out.ensure(ub + 16, L.st)) || (rc = L.meta.ensure(meta_total, L.st)) ||
            (rc = L.hmeta.ensure(up_bytes, L.st)) || (rc = L.res.ensure(res_total, L.st)) || (rc = L.hres.ensure(r_hashes, L.st)) ||
            (rc = L.scratch.ensure(sz_dec + sz_hash + sz_mer, L.st)))
            break;
        uint8_t* hm = static_cast<uint8_t*>(L.hmeta.p);
        ChunkDesc* hd = reinterpret_cast<ChunkDesc*>(hm);
        uint32_t* hfirst = reinterpret_cast<uint32_t*>(hm + o_first);
        uint64_t* hoff = reinterpret_cast<uint64_t*>(hm + o_off);
        uint32_t* hlen = reinterpret_cast<uint32_t*>(hm + o_len);
        uint8_t* hexp = hm + o_exp;
        uint64_t cpos = 0, dpos = 0;
        size_t k = 0;
        cudaError_t e = cudaSuccess;
        for (size_t i = 0; i < nx && e == cudaSuccess; i++) {
            hfirst[i] = (uint32_t)k;
            const HostIndex& ix = ixs[i];
            if (walk_err[i] == ZG_OK && lens[x0 + i]) {
                for (size_t j = 0; j < ix.clen.size(); j++, k++) {
                    hd[k] = ChunkDesc{cpos + ix.hdr_off[j] + 8, dpos, ix.clen[j], ix.ulen[j], ix.scheme[j], 0};
                    hoff[k] = dpos;
                    hlen[k] = ix.ulen[j];
                    dpos += align_up(ix.ulen[j], 16);
                }
                e = cudaMemcpyAsync(static_cast<uint8_t*>(L.comp.p) + cpos, blobs[x0 + i], lens[x0 + i], cudaMemcpyHostToDevice, L.st);
                ctx->h2d += lens[x0 + i];
                cpos += align_up(lens[x0 + i], 16);
            }
            if (expected) memcpy(hexp + 32 * i, expected + 32 * (x0 + i), 32);
        }
        hfirst[nx] = (uint32_t)k;
        fill_decode_order(hd, k);
        uint8_t* dm = static_cast<uint8_t*>(L.meta.p);
        uint8_t* dres = static_cast<uint8_t*>(L.res.p);
        uint8_t* scr = static_cast<uint8_t*>(L.scratch.p);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dm, hm, up_bytes, cudaMemcpyHostToDevice, L.st);
        cudaEventRecord(L.ev[1], L.st);
        if (e == cudaSuccess)
            e = launch_decode(static_cast<const uint8_t*>(L.comp.p), cpos, reinterpret_cast<ChunkDesc*>(dm), nc, static_cast<uint8_t*>(L.out.p),
                              ub, reinterpret_cast<int32_t*>(dm + o_status), scr, sz_dec, L.st, /*keep_planes=*/true);
        cudaEventRecord(L.ev[2], L.st);
        if (e == cudaSuccess)
            e = launch_hash_chunks(static_cast<const uint8_t*>(L.out.p), ub, reinterpret_cast<uint64_t*>(dm + o_off),
                                   reinterpret_cast<uint32_t*>(dm + o_len), nc, make_key(nullptr), dres + r_hashes, scr + sz_dec, sz_hash,
                                   group, L.st, /*planes=*/scr + 256, reinterpret_cast<const ChunkDesc*>(dm));
        cudaEventRecord(L.ev[3], L.st);
        if (e == cudaSuccess)
            e = launch_merkle(dres + r_hashes, reinterpret_cast<uint32_t*>(dm + o_len), nullptr, reinterpret_cast<uint32_t*>(dm + o_first), nx,
                              nc, reinterpret_cast<int32_t*>(dm + o_status), expected ? dm + o_exp : nullptr, dres,
                              reinterpret_cast<int8_t*>(dres + r_verdict), scr + sz_dec + sz_hash, sz_mer, L.st);
        cudaEventRecord(L.ev[4], L.st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(L.hres.p, dres, r_hashes, cudaMemcpyDeviceToHost, L.st);
        if (e == cudaSuccess && chunk_hashes && nc)
            e = cudaMemcpyAsync(chunk_hashes + 32 * hashes_written, dres + r_hashes, 32 * nc, cudaMemcpyDeviceToHost, L.st);
        cudaEventRecord(L.ev[5], L.st);
        if (trace) S.t_enq = now_ms();
        S.launched = true;  // even on a launch error the stream may hold work: collect() drains it
        if (e != cudaSuccess) { rc = cuda_fail(e, "verify pipeline"); break; }
        hashes_written += nc;
        ctx->kernels += 7;
        ctx->h2d += up_bytes;
        ctx->d2h += r_hashes + (chunk_hashes ? 32 * nc : 0);
        ctx->decoded += dpos;
        ctx->hashed += nc;
        si = (si + 1) % n_slots;
    }
    for (int s = 0; s < n_slots; s++) {
        int cr = collect(slots[(si + s) % n_slots]);
        if (cr && rc == ZG_OK) rc = cr;
    }
    if (trace) {
        cudaEventDestroy(ev0);
        fprintf(stderr, "[zest trace] call done %.2f ms\n", now_ms());
    }
    if (bytes_verified) *bytes_verified = verified;
    return rc;
}

int zg_hash_chunks(zg_ctx* ctx, const uint8_t* data, const uint64_t* offsets, size_t n, const uint8_t* key, uint8_t* out) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (n == 0) return ZG_OK;
    if (!offsets || !out) return fail(ZG_ERR_INVALID_ARG, "null argument");
    for (size_t i = 0; i < n; i++)
        if (offsets[i + 1] < offsets[i]) return fail(ZG_ERR_INVALID_ARG, "offsets must be non-decreasing (at %zu)", i);
    uint64_t lo = offsets[0], hi = offsets[n], bytes = hi - lo;
    if (bytes && !data) return fail(ZG_ERR_INVALID_ARG, "data is null");
    ctx->calls++;
    LaneGuard g(ctx);
    Lane& L = *g.l;
    int group = hash_pick_group(bytes);
    size_t sz = hash_scratch_bytes(bytes, n, group);
    size_t obytes = (n + 1) * 8;
    if ((r = L.out.ensure(bytes + 16, L.st)) || (r = L.meta.ensure(obytes, L.st)) || (r = L.hmeta.ensure(obytes, L.st)) ||
        (r = L.res.ensure(32 * n, L.st)) || (r = L.scratch.ensure(sz, L.st)))
        return r;
    uint64_t* ho = static_cast<uint64_t*>(L.hmeta.p);
    for (size_t i = 0; i <= n; i++) ho[i] = offsets[i] - lo;
    if (bytes) CU(cudaMemcpyAsync(L.out.p, data + lo, bytes, cudaMemcpyHostToDevice, L.st));
    CU(cudaMemcpyAsync(L.meta.p, ho, obytes, cudaMemcpyHostToDevice, L.st));
    CU(launch_hash_chunks(static_cast<const uint8_t*>(L.out.p), bytes, static_cast<const uint64_t*>(L.meta.p), nullptr, n, make_key(key),
                          static_cast<uint8_t*>(L.res.p), L.scratch.p, L.scratch.cap, group, L.st));
    CU(cudaMemcpyAsync(out, L.res.p, 32 * n, cudaMemcpyDeviceToHost, L.st));
    CU(cudaStreamSynchronize(L.st));
    ctx->kernels += 5;
    ctx->h2d += bytes + obytes;
    ctx->d2h += 32 * n;
    ctx->hashed += n;
    return ZG_OK;
}

// xet-core's VERIFICATION_KEY (mdb_shard chunk_verification); pinned against hf_xet shards in tests/golden/vectors.json "terms"
static const uint8_t kVerificationKey[32] = {127, 24,  87,  214, 206, 86,  237, 102, 18,  127, 249, 19,  231, 165, 195, 243,
                                             164, 205, 38,  213, 181, 219, 73,  230, 65,  36,  152, 127, 40,  251, 148, 195};

int zg_range_hashes(zg_ctx* ctx, const uint8_t* chunk_hashes, const uint32_t* bounds, size_t n_terms, uint8_t* out) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (n_terms == 0) return ZG_OK;
    if (!chunk_hashes || !bounds || !out) return fail(ZG_ERR_INVALID_ARG, "null argument");
    std::vector<uint64_t> off(n_terms + 1);
    for (size_t t = 0; t <= n_terms; t++) {
        if (t && bounds[t] < bounds[t - 1]) return fail(ZG_ERR_INVALID_ARG, "bounds must be non-decreasing (at %zu)", t);
        off[t] = 32ull * bounds[t];
    }
    // one keyed BLAKE3 per term over its run of 32-byte chunk hashes: the chunk-hash kernel with another key
    return zg_hash_chunks(ctx, chunk_hashes, off.data(), n_terms, kVerificationKey, out);
}

int zg_merkle_root(zg_ctx* ctx, const uint8_t* hashes, const uint64_t* lens, size_t n, int file_hash, uint8_t out[32]) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (!out || (n && (!hashes || !lens))) return fail(ZG_ERR_INVALID_ARG, "null argument");
    if (n == 0) { memset(out, 0, 32); return ZG_OK; }
    if (n >= 0xffffffffull) return fail(ZG_ERR_LIMIT, "too many nodes");
    ctx->calls++;
    LaneGuard g(ctx);
    Lane& L = *g.l;
    size_t o_len = align_up(32 * n, 256), o_first = o_len + align_up(8 * n, 256), o_off = o_first + 256, total = o_off + 256;
    size_t sz_m = merkle_scratch_bytes(n, 1), sz_h = hash_scratch_bytes(32, 1, 1);
    if ((r = L.meta.ensure(total, L.st)) || (r = L.hmeta.ensure(total, L.st)) || (r = L.res.ensure(512, L.st)) ||
        (r = L.scratch.ensure(sz_m + sz_h, L.st)))
        return r;
    uint8_t* hm = static_cast<uint8_t*>(L.hmeta.p);
    memcpy(hm, hashes, 32 * n);
    memcpy(hm + o_len, lens, 8 * n);
    uint32_t* f = reinterpret_cast<uint32_t*>(hm + o_first);
    f[0] = 0; f[1] = (uint32_t)n;
    uint64_t* off = reinterpret_cast<uint64_t*>(hm + o_off);
    off[0] = 0; off[1] = 32;
    uint8_t* dm = static_cast<uint8_t*>(L.meta.p);
    uint8_t* dres = static_cast<uint8_t*>(L.res.p);  // [root 32 | pad | verdict @64 | file hash @128]
    CU(cudaMemcpyAsync(dm, hm, total, cudaMemcpyHostToDevice, L.st));
    CU(launch_merkle(dm, nullptr, reinterpret_cast<uint64_t*>(dm + o_len), reinterpret_cast<uint32_t*>(dm + o_first), 1, n, nullptr, nullptr,
                     dres, reinterpret_cast<int8_t*>(dres + 64), L.scratch.p, sz_m, L.st));
    ctx->kernels += 1;
    if (file_hash) {
        static const uint8_t zero_key[32] = {0};
        CU(launch_hash_chunks(dres, 32, reinterpret_cast<uint64_t*>(dm + o_off), nullptr, 1, make_key(zero_key), dres + 128,
                              static_cast<uint8_t*>(L.scratch.p) + sz_m, sz_h, 1, L.st));
        ctx->kernels += 5;
    }
    CU(cudaMemcpyAsync(out, dres + (file_hash ? 128 : 0), 32, cudaMemcpyDeviceToHost, L.st));
    CU(cudaStreamSynchronize(L.st));
    ctx->h2d += total;
    ctx->d2h += 32;
    return ZG_OK;
}

namespace {

// CDC of a host buffer already copied to L.out; leaves ends in L.meta (device) and returns the count.
int cdc_on_lane(zg_ctx* ctx, Lane& L, size_t len, uint64_t* ends, size_t cap, size_t* n_ends, size_t dev_cap) {
    size_t sz = cdc_scratch_bytes(len);
    int r;
    if ((r = L.scratch.ensure(sz, L.st)) || (r = L.meta.ensure((dev_cap + 2) * 8 + 256, L.st)) || (r = L.hres.ensure(256, L.st))) return r;
    uint64_t* d_n = static_cast<uint64_t*>(L.meta.p);
    uint64_t* d_ends = d_n + 32;  // 256 bytes in
    CU(launch_cdc(static_cast<const uint8_t*>(L.out.p), len, d_ends + 1, dev_cap, d_n, L.scratch.p, L.scratch.cap, L.st));
    ctx->kernels += 2;
    CU(cudaMemcpyAsync(L.hres.p, d_n, 16, cudaMemcpyDeviceToHost, L.st));
    CU(cudaStreamSynchronize(L.st));
    uint64_t cnt = static_cast<uint64_t*>(L.hres.p)[0];
    *n_ends = (size_t)cnt;
    if (cnt > dev_cap) return fail(ZG_ERR_CAPACITY, "internal boundary buffer too small (%llu > %zu)", (unsigned long long)cnt, dev_cap);
    size_t ncopy = cnt < cap ? (size_t)cnt : cap;
    if (ncopy && ends) CU(cudaMemcpyAsync(ends, d_ends + 1, ncopy * 8, cudaMemcpyDeviceToHost, L.st));
    ctx->d2h += ncopy * 8 + 16;
    return ZG_OK;
}

}  // namespace

int zg_cdc(zg_ctx* ctx, const uint8_t* data, size_t len, uint64_t* ends, size_t cap, size_t* n_ends) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (!n_ends || (len && !data)) return fail(ZG_ERR_INVALID_ARG, "null argument");
    *n_ends = 0;
    if (len == 0) return ZG_OK;
    ctx->calls++;
    LaneGuard g(ctx);
    Lane& L = *g.l;
    if ((r = L.out.ensure(len + 16, L.st))) return r;
    CU(cudaMemcpyAsync(L.out.p, data, len, cudaMemcpyHostToDevice, L.st));
    ctx->h2d += len;
    size_t dev_cap = len / 8192 + 2;
    if ((r = cdc_on_lane(ctx, L, len, ends, cap, n_ends, dev_cap))) return r;
    CU(cudaStreamSynchronize(L.st));
    if (*n_ends > cap) return fail(ZG_ERR_CAPACITY, "%zu boundaries, buffer holds %zu", *n_ends, cap);
    return ZG_OK;
}

int zg_cdc_hash(zg_ctx* ctx, const uint8_t* data, size_t len, uint64_t* ends, size_t cap, size_t* n_ends, uint8_t* hashes,
                uint8_t* file_hash) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (!n_ends || (len && !data)) return fail(ZG_ERR_INVALID_ARG, "null argument");
    *n_ends = 0;
    if (len == 0) { if (file_hash) memset(file_hash, 0, 32); return ZG_OK; }
    ctx->calls++;
    LaneGuard g(ctx);
    Lane& L = *g.l;
    if ((r = L.out.ensure(len + 16, L.st))) return r;
    CU(cudaMemcpyAsync(L.out.p, data, len, cudaMemcpyHostToDevice, L.st));
    ctx->h2d += len;
    size_t dev_cap = len / 8192 + 2;
    if ((r = cdc_on_lane(ctx, L, len, ends, cap, n_ends, dev_cap))) return r;
    size_t n = *n_ends;
    if (n > cap) { cudaStreamSynchronize(L.st); return fail(ZG_ERR_CAPACITY, "%zu boundaries, buffers hold %zu", n, cap); }
    // offsets[0..n] = {0, ends...}: the walk kernel wrote ends at d_ends + 1, so d_ends is the offsets array
    uint64_t* d_off = static_cast<uint64_t*>(L.meta.p) + 32;
    CU(cudaMemsetAsync(d_off, 0, 8, L.st));
    int group = hash_pick_group(len);
    size_t sz_h = hash_scratch_bytes(len, n, group), sz_m = merkle_scratch_bytes(n, 1), sz_f = hash_scratch_bytes(32, 1, 1);
    size_t lens_off = align_up(32 * n, 256);
    if ((r = L.scratch.ensure(sz_h + sz_m + sz_f, L.st)) || (r = L.res.ensure(lens_off + align_up(8 * n, 256) + 1024, L.st))) return r;
    uint8_t* dres = static_cast<uint8_t*>(L.res.p);
    CU(launch_hash_chunks(static_cast<const uint8_t*>(L.out.p), len, d_off, nullptr, n, make_key(nullptr), dres, L.scratch.p, sz_h, group, L.st));
    ctx->kernels += 5;
    ctx->hashed += n;
    if (hashes) { CU(cudaMemcpyAsync(hashes, dres, 32 * n, cudaMemcpyDeviceToHost, L.st)); ctx->d2h += 32 * n; }
    if (file_hash) {
        // lens64[i] = off[i+1] - off[i]: computed on the host from the boundaries already copied back
        CU(cudaStreamSynchronize(L.st));
        if ((r = L.hmeta.ensure(8 * n + 512, L.st))) return r;
        uint64_t* hl = static_cast<uint64_t*>(L.hmeta.p);
        std::vector<uint64_t> tmp;
        const uint64_t* e = ends;
        if (!e) { tmp.resize(n); CU(cudaMemcpy(tmp.data(), d_off + 1, 8 * n, cudaMemcpyDeviceToHost)); e = tmp.data(); }
        for (size_t i = 0; i < n; i++) hl[i] = e[i] - (i ? e[i - 1] : 0);
        uint32_t* hf = reinterpret_cast<uint32_t*>(hl + n);
        hf[0] = 0; hf[1] = (uint32_t)n;
        uint64_t* ho = reinterpret_cast<uint64_t*>(hf + 4);
        ho[0] = 0; ho[1] = 32;
        uint8_t* dl = dres + lens_off;
        uint8_t* dtail = dl + align_up(8 * n, 256);  // [first(16) | off(16) | root @256 | verdict @320 | file hash @384]
        CU(cudaMemcpyAsync(dl, hl, 8 * n, cudaMemcpyHostToDevice, L.st));
        CU(cudaMemcpyAsync(dtail, hf, 32, cudaMemcpyHostToDevice, L.st));
        uint8_t* scr = static_cast<uint8_t*>(L.scratch.p);
        CU(launch_merkle(dres, nullptr, reinterpret_cast<uint64_t*>(dl), reinterpret_cast<uint32_t*>(dtail), 1, n, nullptr, nullptr,
                         dtail + 256, reinterpret_cast<int8_t*>(dtail + 320), scr + sz_h, sz_m, L.st));
        static const uint8_t zero_key[32] = {0};
        CU(launch_hash_chunks(dtail + 256, 32, reinterpret_cast<uint64_t*>(dtail + 16), nullptr, 1, make_key(zero_key), dtail + 384,
                              scr + sz_h + sz_m, sz_f, 1, L.st));
        CU(cudaMemcpyAsync(file_hash, dtail + 384, 32, cudaMemcpyDeviceToHost, L.st));
        ctx->kernels += 6;
    }
    CU(cudaStreamSynchronize(L.st));
    return ZG_OK;
}

// -------------------------------------------------------------------------------------------------------------
// device-resident entry points
// -------------------------------------------------------------------------------------------------------------

size_t zg_dev_hash_scratch_bytes(uint64_t data_bytes, size_t n) { return hash_scratch_bytes(data_bytes, n, hash_pick_group(data_bytes)); }

int zg_dev_hash_chunks(zg_ctx* ctx, const uint8_t* d_data, uint64_t data_bytes, const uint64_t* d_offsets, const uint32_t* d_lens,
                       size_t n, const uint8_t* key, uint8_t* d_out, void* d_scratch, size_t scratch_bytes, void* stream) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    CU(launch_hash_chunks(d_data, data_bytes, d_offsets, d_lens, n, make_key(key), d_out, d_scratch, scratch_bytes,
                          hash_pick_group(data_bytes), static_cast<cudaStream_t>(stream)));
    ctx->kernels += 5;
    ctx->hashed += n;
    return ZG_OK;
}

size_t zg_dev_decode_scratch_bytes(size_t n_chunks, uint64_t out_bytes) { return decode_scratch_bytes(n_chunks, out_bytes); }

int zg_dev_decode_chunks(zg_ctx* ctx, const uint8_t* d_comp, uint64_t comp_bytes, const zg_chunk_desc* d_chunks, size_t n_chunks,
                         uint8_t* d_out, uint64_t out_bytes, int32_t* d_status, void* d_scratch, size_t scratch_bytes, void* stream) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    CU(launch_decode(d_comp, comp_bytes, reinterpret_cast<const ChunkDesc*>(d_chunks), n_chunks, d_out, out_bytes, d_status, d_scratch,
                     scratch_bytes, static_cast<cudaStream_t>(stream)));
    ctx->kernels += 1;
    ctx->decoded += out_bytes;
    return ZG_OK;
}

int zg_dev_decode_for_hash(zg_ctx* ctx, const uint8_t* d_comp, uint64_t comp_bytes, const zg_chunk_desc* d_chunks, size_t n_chunks,
                           uint8_t* d_out, uint64_t out_bytes, int32_t* d_status, void* d_scratch, size_t scratch_bytes, void* stream) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    CU(launch_decode(d_comp, comp_bytes, reinterpret_cast<const ChunkDesc*>(d_chunks), n_chunks, d_out, out_bytes, d_status, d_scratch,
                     scratch_bytes, static_cast<cudaStream_t>(stream), /*keep_planes=*/true));
    ctx->kernels += 1;
    ctx->decoded += out_bytes;
    return ZG_OK;
}

namespace {
// Leaves per thread for the hash that reads byte-grouped chunks as planes: four strided streams per thread instead of one,
// so at most 4 KiB per thread (measured at 8 GB: G = 8 / 4 / 2 -> 5.40 / 5.18 / 5.19 ms).
int planar_group(uint64_t out_bytes) {
    const int g = hash_pick_group(out_bytes);
    return g > 4 ? 4 : g;
}
}  // namespace

size_t zg_dev_hash_decoded_scratch_bytes(uint64_t out_bytes, size_t n) {
    return align_up(n * 8 + 8, 256) + align_up(n * 4 + 4, 256) + hash_scratch_bytes(out_bytes, n, planar_group(out_bytes));
}

int zg_dev_hash_decoded(zg_ctx* ctx, const uint8_t* d_out, uint64_t out_bytes, const zg_chunk_desc* d_chunks, size_t n_chunks,
                        const void* d_decode_scratch, uint8_t* d_hashes, void* d_scratch, size_t scratch_bytes, void* stream) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (n_chunks == 0) return ZG_OK;
    if (!d_chunks || !d_decode_scratch || !d_hashes || !d_scratch) return fail(ZG_ERR_INVALID_ARG, "null argument");
    if (scratch_bytes < zg_dev_hash_decoded_scratch_bytes(out_bytes, n_chunks)) return fail(ZG_ERR_CAPACITY, "scratch too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint8_t* s = static_cast<uint8_t*>(d_scratch);
    const ChunkDesc* dc = reinterpret_cast<const ChunkDesc*>(d_chunks);
    uint64_t* off = reinterpret_cast<uint64_t*>(s);
    uint32_t* len = reinterpret_cast<uint32_t*>(s + align_up(n_chunks * 8 + 8, 256));
    const size_t o_hash = align_up(n_chunks * 8 + 8, 256) + align_up(n_chunks * 4 + 4, 256);
    split_desc_kernel<<<(unsigned)((n_chunks + 255) / 256), 256, 0, st>>>(dc, n_chunks, off, len);
    CU(launch_hash_chunks(d_out, out_bytes, off, len, n_chunks, make_key(nullptr), d_hashes, s + o_hash, scratch_bytes - o_hash,
                          planar_group(out_bytes), st, static_cast<const uint8_t*>(d_decode_scratch) + 256, dc));
    ctx->kernels += 6;
    ctx->hashed += n_chunks;
    return ZG_OK;
}

size_t zg_dev_merkle_scratch_bytes(size_t n_chunks, size_t n_xorbs) { return merkle_scratch_bytes(n_chunks, n_xorbs); }

int zg_dev_merkle(zg_ctx* ctx, const uint8_t* d_hashes, const uint32_t* d_lens, const uint32_t* d_first, size_t n_xorbs, size_t n_chunks,
                  const int32_t* d_status, const uint8_t* d_expected, uint8_t* d_roots, int8_t* d_verdict, void* d_scratch,
                  size_t scratch_bytes, void* stream) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    CU(launch_merkle(d_hashes, d_lens, nullptr, d_first, n_xorbs, n_chunks, d_status, d_expected, d_roots, d_verdict, d_scratch,
                     scratch_bytes, static_cast<cudaStream_t>(stream)));
    ctx->kernels += 1;
    return ZG_OK;
}

namespace {
struct VerifyLayout { size_t dec, hash, mer, off, len, status, total; int group; };
VerifyLayout verify_layout(size_t nc, size_t nx, uint64_t out_bytes) {
    VerifyLayout v;
    v.group = planar_group(out_bytes);
    size_t o = 0;
    v.dec = o; o += align_up(decode_scratch_bytes(nc, out_bytes), 256);
    v.hash = o; o += align_up(hash_scratch_bytes(out_bytes, nc, v.group), 256);
    v.mer = o; o += align_up(merkle_scratch_bytes(nc, nx), 256);
    v.off = o; o += align_up(nc * 8 + 8, 256);
    v.len = o; o += align_up(nc * 4 + 4, 256);
    v.status = o; o += align_up(nc * 4 + 4, 256);
    v.total = o + align_up((size_t)out_bytes + 16, 256);  // decoded bytes live at the end of the scratch
    return v;
}
}  // namespace

size_t zg_dev_verify_scratch_bytes(size_t n_chunks, size_t n_xorbs, uint64_t out_bytes) { return verify_layout(n_chunks, n_xorbs, out_bytes).total; }

int zg_dev_verify_xorbs(zg_ctx* ctx, const uint8_t* d_comp, uint64_t comp_bytes, const zg_chunk_desc* d_chunks, size_t n_chunks,
                        const uint32_t* d_first, size_t n_xorbs, const uint8_t* d_expected, uint64_t out_bytes, uint8_t* d_roots,
                        int8_t* d_verdict, uint8_t* d_chunk_hashes, void* d_scratch, size_t scratch_bytes, void* stream) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (n_xorbs == 0) return ZG_OK;
    if (!d_chunk_hashes) return fail(ZG_ERR_INVALID_ARG, "d_chunk_hashes is required (n_chunks x 32 bytes)");
    VerifyLayout v = verify_layout(n_chunks, n_xorbs, out_bytes);
    if (scratch_bytes < v.total) return fail(ZG_ERR_CAPACITY, "scratch holds %zu bytes, need %zu", scratch_bytes, v.total);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint8_t* s = static_cast<uint8_t*>(d_scratch);
    const ChunkDesc* dc = reinterpret_cast<const ChunkDesc*>(d_chunks);
    uint64_t* off = reinterpret_cast<uint64_t*>(s + v.off);
    uint32_t* len = reinterpret_cast<uint32_t*>(s + v.len);
    int32_t* status = reinterpret_cast<int32_t*>(s + v.status);
    uint8_t* decoded = s + (v.total - align_up((size_t)out_bytes + 16, 256));
    if (n_chunks) split_desc_kernel<<<(unsigned)((n_chunks + 255) / 256), 256, 0, st>>>(dc, n_chunks, off, len);
    // verify only hashes the decoded bytes: byte-grouped chunks stay as planes and the hash regroups them as it reads
    CU(launch_decode(d_comp, comp_bytes, dc, n_chunks, decoded, out_bytes, status, s + v.dec, v.hash - v.dec, st, /*keep_planes=*/true));
    CU(launch_hash_chunks(decoded, out_bytes, off, len, n_chunks, make_key(nullptr), d_chunk_hashes, s + v.hash, v.mer - v.hash, v.group, st,
                          /*planes=*/s + v.dec + 256, dc));
    CU(launch_merkle(d_chunk_hashes, len, nullptr, d_first, n_xorbs, n_chunks, status, d_expected, d_roots, d_verdict, s + v.mer,
                     v.off - v.mer, st));
    ctx->kernels += 8;
    ctx->decoded += out_bytes;
    ctx->hashed += n_chunks;
    return ZG_OK;
}

size_t zg_dev_cdc_scratch_bytes(uint64_t len) { return cdc_scratch_bytes(len); }

int zg_dev_cdc(zg_ctx* ctx, const uint8_t* d_data, uint64_t len, uint64_t* d_ends, uint64_t cap, uint64_t* d_n_out, void* d_scratch,
               size_t scratch_bytes, void* stream) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    CU(launch_cdc(d_data, len, d_ends, cap, d_n_out, d_scratch, scratch_bytes, static_cast<cudaStream_t>(stream)));
    ctx->kernels += 2;
    return ZG_OK;
}

size_t zg_xorb_footer_bytes(size_t n_chunks) { return 96 + 40 * n_chunks; }

int zg_xorb_footer(const uint8_t xorb_hash[32], const uint8_t* chunk_hashes, const uint32_t* clen, const uint32_t* ulen,
                   size_t n, uint8_t* out, size_t out_cap, size_t* out_len) {
    if (!xorb_hash || !out_len || (n && (!chunk_hashes || !clen || !ulen))) return fail(ZG_ERR_INVALID_ARG, "null argument");
    if (n > 0x7fffffffull) return fail(ZG_ERR_LIMIT, "too many chunks");
    const size_t total = zg_xorb_footer_bytes(n);
    *out_len = total;
    if (total > out_cap || !out) return fail(ZG_ERR_CAPACITY, "footer needs %zu bytes, buffer holds %zu", total, out_cap);
    const uint32_t info_len = (uint32_t)(total - 4);
    auto put32 = [](uint8_t* p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24); };
    uint8_t* p = out;
    memcpy(p, "XETBLOB", 7); p[7] = 1; p += 8;
    memcpy(p, xorb_hash, 32); p += 32;
    const uint32_t hashes_at = (uint32_t)(p - out);
    memcpy(p, "XBLBHSH", 7); p[7] = 0; p += 8;
    put32(p, (uint32_t)n); p += 4;
    if (n) memcpy(p, chunk_hashes, 32 * n);
    p += 32 * n;
    const uint32_t bnd_at = (uint32_t)(p - out);
    memcpy(p, "XBLBBND", 7); p[7] = 1; p += 8;
    put32(p, (uint32_t)n); p += 4;
    uint64_t c = 0, u = 0;
    for (size_t i = 0; i < n; i++) { c += 8ull + clen[i]; if (c > 0xffffffffull) return fail(ZG_ERR_LIMIT, "xorb larger than 4 GiB"); put32(p + 4 * i, (uint32_t)c); }
    p += 4 * n;
    for (size_t i = 0; i < n; i++) { u += ulen[i]; if (u > 0xffffffffull) return fail(ZG_ERR_LIMIT, "xorb larger than 4 GiB"); put32(p + 4 * i, (uint32_t)u); }
    p += 4 * n;
    put32(p, (uint32_t)n); p += 4;
    put32(p, info_len - hashes_at); p += 4;
    put32(p, info_len - bnd_at); p += 4;
    memset(p, 0, 16); p += 16;
    put32(p, info_len); p += 4;
    return ZG_OK;
}

void zg_chunk_order(zg_chunk_desc* descs, size_t n) {
    if (descs && n) fill_decode_order(reinterpret_cast<ChunkDesc*>(descs), n);
}

// ---- dedup index (dedup.cu) ----------------------------------------------------------------------------------
struct zg_dedup {
    zg_ctx* ctx = nullptr;
    void* mem = nullptr;
    DedupView v{};
    uint64_t max_entries = 0;
    std::mutex mu;               // host-buffer inserts and lookups on one table are serialised
    std::atomic<uint32_t> epoch{0};
};

static int dedup_check_counters(zg_dedup* t, const unsigned long long c[3]) {
    if (c[2]) return fail(ZG_ERR_LIMIT, "dedup index: %llu chunk hashes agree in 64 tag bits with a different stored hash", c[2]);
    if (c[1] || c[0] > t->max_entries) return fail(ZG_ERR_CAPACITY, "dedup index holds %llu entries, created for %llu", c[0], (unsigned long long)t->max_entries);
    return ZG_OK;
}

int zg_dedup_create(zg_ctx* ctx, size_t max_entries, zg_dedup** out) {
    DeviceScope dev_scope(ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (!out || max_entries == 0) return fail(ZG_ERR_INVALID_ARG, "null argument or zero capacity");
    if (max_entries > (1ull << 36)) return fail(ZG_ERR_LIMIT, "dedup index too large");
    uint64_t slots = 64;
    while (slots < 2 * (uint64_t)max_entries) slots <<= 1;  // at most half full
    std::unique_ptr<zg_dedup> t(new zg_dedup);
    t->ctx = ctx;
    t->max_entries = max_entries;
    cudaError_t e = cudaMalloc(&t->mem, dedup_table_bytes(slots));
    if (e != cudaSuccess) { (void)cudaGetLastError(); return fail(ZG_ERR_OOM, "cudaMalloc(%zu): %s", dedup_table_bytes(slots), cudaGetErrorString(e)); }
    t->v = dedup_view(t->mem, slots);
    LaneGuard g(ctx);
    e = dedup_clear(t->v, g.l->st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(g.l->st);
    if (e != cudaSuccess) { cudaFree(t->mem); return cuda_fail(e, "dedup clear"); }
    *out = t.release();
    return ZG_OK;
}

void zg_dedup_destroy(zg_dedup* t) {
    if (!t) return;
    DeviceScope dev_scope(t->ctx);
    if (t->mem) cudaFree(t->mem);
    delete t;
}

int zg_dedup_insert(zg_dedup* t, const uint8_t* hashes, const uint64_t* locations, size_t n, uint8_t* was_new) {
    if (!t) return fail(ZG_ERR_INVALID_ARG, "null table");
    DeviceScope dev_scope(t->ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (n == 0) return ZG_OK;
    if (!hashes || !locations) return fail(ZG_ERR_INVALID_ARG, "null argument");
    for (size_t i = 0; i < n; i++)
        if (locations[i] == ZG_DEDUP_MISSING) return fail(ZG_ERR_INVALID_ARG, "location %zu is the reserved value", i);
    std::lock_guard<std::mutex> lk(t->mu);
    LaneGuard g(t->ctx);
    Lane& L = *g.l;
    const size_t o_loc = align_up(32 * n, 256), o_new = o_loc + align_up(8 * n, 256), total = o_new + align_up(n, 256);
    if ((r = L.meta.ensure(total, L.st)) || (r = L.hres.ensure(align_up(n, 256) + 64, L.st))) return r;
    uint8_t* d = static_cast<uint8_t*>(L.meta.p);
    CU(cudaMemcpyAsync(d, hashes, 32 * n, cudaMemcpyHostToDevice, L.st));
    CU(cudaMemcpyAsync(d + o_loc, locations, 8 * n, cudaMemcpyHostToDevice, L.st));
    CU(launch_dedup_insert(t->v, d, reinterpret_cast<const uint64_t*>(d + o_loc), n, ++t->epoch, d + o_new, L.st));
    uint8_t* hr = static_cast<uint8_t*>(L.hres.p);
    CU(cudaMemcpyAsync(hr, t->v.counters, 24, cudaMemcpyDeviceToHost, L.st));
    CU(cudaMemcpyAsync(hr + 64, d + o_new, n, cudaMemcpyDeviceToHost, L.st));
    CU(cudaStreamSynchronize(L.st));
    if (was_new) memcpy(was_new, hr + 64, n);
    t->ctx->kernels += 2;
    t->ctx->h2d += 40 * n;
    t->ctx->d2h += n + 24;
    unsigned long long c[3];
    memcpy(c, hr, 24);
    return dedup_check_counters(t, c);
}

int zg_dedup_lookup(zg_dedup* t, const uint8_t* hashes, size_t n, uint64_t* locations) {
    if (!t) return fail(ZG_ERR_INVALID_ARG, "null table");
    DeviceScope dev_scope(t->ctx);
    int r = dev_scope.rc;
    if (r) return r;
    if (n == 0) return ZG_OK;
    if (!hashes || !locations) return fail(ZG_ERR_INVALID_ARG, "null argument");
    std::lock_guard<std::mutex> lk(t->mu);
    LaneGuard g(t->ctx);
    Lane& L = *g.l;
    const size_t o_out = align_up(32 * n, 256);
    if ((r = L.meta.ensure(o_out + 8 * n, L.st))) return r;
    uint8_t* d = static_cast<uint8_t*>(L.meta.p);
    CU(cudaMemcpyAsync(d, hashes, 32 * n, cudaMemcpyHostToDevice, L.st));
    CU(launch_dedup_lookup(t->v, d, n, reinterpret_cast<uint64_t*>(d + o_out), L.st));
    CU(cudaMemcpyAsync(locations, d + o_out, 8 * n, cudaMemcpyDeviceToHost, L.st));
    CU(cudaStreamSynchronize(L.st));
    t->ctx->kernels += 1;
    t->ctx->h2d += 32 * n;
    t->ctx->d2h += 8 * n;
    return ZG_OK;
}

int zg_dedup_count(zg_dedup* t, uint64_t* entries) {
    if (!t || !entries) return fail(ZG_ERR_INVALID_ARG, "null argument");
    DeviceScope dev_scope(t->ctx);
    if (dev_scope.rc) return dev_scope.rc;
    std::lock_guard<std::mutex> lk(t->mu);
    unsigned long long c[3];
    // The inserts whose errors surface here may sit on any stream (the library's lanes and torch's streams are
    // non-blocking ones, which a copy on the legacy default stream does not wait for): drain the whole device first.
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(c, t->v.counters, 24, cudaMemcpyDeviceToHost));
    *entries = c[0];
    return dedup_check_counters(t, c);
}

int zg_dev_dedup_insert(zg_dedup* t, const uint8_t* d_hashes, const uint64_t* d_locations, size_t n, uint8_t* d_was_new, void* stream) {
    if (!t) return fail(ZG_ERR_INVALID_ARG, "null table");
    DeviceScope dev_scope(t->ctx);
    if (dev_scope.rc) return dev_scope.rc;
    if (n == 0) return ZG_OK;
    if (!d_hashes || !d_locations) return fail(ZG_ERR_INVALID_ARG, "null argument");
    CU(launch_dedup_insert(t->v, d_hashes, d_locations, n, ++t->epoch, d_was_new, static_cast<cudaStream_t>(stream)));
    t->ctx->kernels += 2;
    return ZG_OK;
}

int zg_dev_dedup_lookup(zg_dedup* t, const uint8_t* d_hashes, size_t n, uint64_t* d_locations, void* stream) {
    if (!t) return fail(ZG_ERR_INVALID_ARG, "null table");
    DeviceScope dev_scope(t->ctx);
    if (dev_scope.rc) return dev_scope.rc;
    if (n == 0) return ZG_OK;
    if (!d_hashes || !d_locations) return fail(ZG_ERR_INVALID_ARG, "null argument");
    CU(launch_dedup_lookup(t->v, d_hashes, n, d_locations, static_cast<cudaStream_t>(stream)));
    t->ctx->kernels += 1;
    return ZG_OK;
}

int zg_get_stats(zg_ctx* ctx, zg_stats* out) {
    if (!ctx || !out) return fail(ZG_ERR_INVALID_ARG, "null argument");
    out->calls = ctx->calls; out->kernels_launched = ctx->kernels; out->h2d_bytes = ctx->h2d; out->d2h_bytes = ctx->d2h;
    out->bytes_decoded = ctx->decoded; out->chunks_hashed = ctx->hashed;
    out->ns_h2d = ctx->ns_h2d; out->ns_decode = ctx->ns_decode; out->ns_hash = ctx->ns_hash; out->ns_merkle = ctx->ns_merkle;
    out->ns_d2h = ctx->ns_d2h;
    return ZG_OK;
}

}  // extern "C"
