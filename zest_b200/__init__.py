"""zest_b200 -- B200 (sm_100a) implementation of zest's xorb verify / reconstruct hot path.

Host-side mirror of the reference interfaces for this path only:
  xorb.XorbReader.extract_chunk_range   <- xet.xorb.XorbReader.extractChunkRange (src/xet_bridge.zig:256-257)
  hashing.hash_to_hex / hash_chunks / merkle_root / file_hash / range_hashes   <- xet.hashing (src/xet_bridge.zig:154)
  verify.verify_xorbs                   <- the integrity check zest defers to zig-xet (src/bt_peer.zig:166-169)
  group.Group.verify_xorbs              <- the same call over every GPU of the box from one process (src/xet_bridge.zig:231-264)
  chunking.cdc / cdc_hash               <- xet.chunking (src/config.zig:12)
  dedup.DedupIndex                      <- batch form of storage.hasChunk / XorbRegistry (src/storage.zig:137-228)
Everything computes in libzest_gpu.so (hand-written CUDA, C ABI in include/zest_gpu.h).  No CPU fallback.
"""
from ._native import ZestGpuError, LIB_PATH, build  # noqa: F401
from .context import Context, default_context  # noqa: F401
from .hashing import (PLAIN_BLAKE3, blake3, chunk_hash, file_hash, hash_chunks, hash_to_hex, hex_to_hash,  # noqa: F401
                      merkle_root, range_hashes)
from .xorb import XorbIndex, XorbReader, seal_xorb, xorb_footer, xorb_index  # noqa: F401
from .verify import VerifyResult, partition_xorbs, verify_xorbs  # noqa: F401
from .group import Group, partition_native  # noqa: F401
from .chunking import cdc, cdc_hash  # noqa: F401
from .dedup import DedupIndex, shard_of  # noqa: F401

__all__ = ["Context", "default_context", "ZestGpuError", "XorbReader", "XorbIndex", "xorb_index", "xorb_footer", "seal_xorb", "verify_xorbs",
           "VerifyResult", "partition_xorbs", "Group", "partition_native", "hash_chunks", "chunk_hash", "blake3", "merkle_root", "file_hash", "range_hashes",
           "hash_to_hex", "hex_to_hash", "cdc", "cdc_hash", "DedupIndex", "shard_of", "PLAIN_BLAKE3", "build", "LIB_PATH"]
