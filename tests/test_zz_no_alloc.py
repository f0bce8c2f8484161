"""zg_dev_hash_chunks_static: the chunk-hash pass for callers that cannot allocate device memory.

Who needs it: a transfer loop that has already committed its whole memory budget to pinned staging and decode
buffers, a thread that must not take the driver's allocation lock while a download is in flight, or a process whose
allocator is frozen after start-up.  Such a caller still owns its input and output buffers but has no place for the
hash pass's working set, so the library keeps that in a fixed 8 MiB device array of its own.

The test drives the entry point through the C ABI (ctypes) and checks it bit for bit (integer work: the tolerance of
every other test here is exact equality) against the existing entry point, zg_dev_hash_chunks, which takes
caller-provided scratch, and against the CPU oracle (keyed BLAKE3 per chunk), so the two CUDA entry points cannot be
wrong together.  Two calls in a row with different batches show that what the first call leaves in the shared array
does not leak into the second.  "Allocates nothing" is checked where an allocation by this library would show: the
library links the CUDA runtime statically and would allocate with cudaMalloc, which torch's allocator counters never
see, so the test compares the driver's free-memory figure (cudaMemGetInfo) around the call.  The file is named so that pytest collects it last: it passes on its
own in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch(rng, total, n_small):
    """A ragged batch: leaf/block edge lengths, empty chunks, then random lengths; offsets are unaligned."""
    lens = [0, 1, 63, 64, 65, 1023, 1024, 1025, 4096, 8191, 8192, 131072, 3, 0, 70000][:n_small]
    while sum(lens) < total - 140_000:
        lens.append(int(rng.integers(0, 131073)))
    lens.append(total - sum(lens))
    data = rng.integers(0, 256, total, dtype=np.uint8)
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    return data, off


def test_static_scratch_hash_matches_regular_entry_point():
    import torch
    import zest_b200 as z
    from zest_b200 import _native
    from zest_b200._native import check

    L = _native.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)            # the entry point works on the calling thread's current device
    ctx = z.Context(0)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rng = np.random.default_rng(2024)
    from oracle import xet_oracle as oz
    torch.cuda.synchronize()
    # first call: the address look-up happens here (outside any stream capture); second call reuses the array
    for total, n_small, key in [(40_000_000, 15, None), (3_000_001, 9, b"\x07" * 32)]:
        data, off = _batch(rng, total, n_small)
        n = off.size - 1
        d_data = torch.from_numpy(np.concatenate([data, np.zeros(16, np.uint8)])).to(dev)
        d_off = torch.from_numpy(off.view(np.int64)).to(dev)
        out_static = torch.zeros(32 * n, dtype=torch.uint8, device=dev)
        out_regular = torch.zeros(32 * n, dtype=torch.uint8, device=dev)
        kbuf = np.frombuffer(key, np.uint8).copy() if key else None
        kp = C.c_void_p(kbuf.ctypes.data) if key else None
        # everything the caller needs is allocated by now; the call itself must not allocate.  The first call on a
        # device may make the driver load the kernels' module, so the free-memory comparison is made around a
        # repeat of the call (same inputs, same outputs).
        check(L.zg_dev_hash_chunks_static(d_data.data_ptr(), total, d_off.data_ptr(), None, n, kp,
                                          out_static.data_ptr(), stream))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        check(L.zg_dev_hash_chunks_static(d_data.data_ptr(), total, d_off.data_ptr(), None, n, kp,
                                          out_static.data_ptr(), stream))
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(dev)[0] == free_before
        scratch = torch.empty(L.zg_dev_hash_scratch_bytes(total, n), dtype=torch.uint8, device=dev)
        check(L.zg_dev_hash_chunks(ctx.handle, d_data.data_ptr(), total, d_off.data_ptr(), None, n, kp,
                                   out_regular.data_ptr(), scratch.data_ptr(), scratch.numel(), stream))
        torch.cuda.synchronize()
        a, b = out_static.cpu().numpy(), out_regular.cpu().numpy()
        assert np.array_equal(a, b)
        assert a.any()
        # ... and against the CPU oracle
        want = np.stack([np.frombuffer(oz.blake3_keyed(key, data[int(off[i]):int(off[i + 1])]) if key
                                       else oz.chunk_hash(data[int(off[i]):int(off[i + 1])]), np.uint8)
                         for i in range(n)])
        assert np.array_equal(a.reshape(-1, 32), want)
    # a batch whose working set does not fit the built-in array is refused with a clear error before anything runs
    n_big = 400_000                                  # 400k (empty) chunks need ~14 MiB of prefix sums, map and nodes
    d_off = torch.zeros(n_big + 1, dtype=torch.int64, device=dev)
    out = torch.zeros(32 * n_big, dtype=torch.uint8, device=dev)
    rc = L.zg_dev_hash_chunks_static(d_data.data_ptr(), 0, d_off.data_ptr(), None, n_big, None, out.data_ptr(), stream)
    assert rc == -13 and b"scratch" in L.zg_last_error()
    torch.cuda.synchronize()
    ctx.close()
