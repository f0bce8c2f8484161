/* A plain-C caller of libzest_gpu.so, standing where zest's Zig `extern` block would stand: includes
 * the header as C99, links the library, and drives the entry points that need no GPU.  With a GPU it also runs one
 * verify through zg_init / zg_verify_xorbs and one through a two-context zg_group.  Exit code 0 = every check passed;
 * prints "no-device" when zg_init reports ZG_ERR_NO_DEVICE (the CPU-only build box), "gpu-ok" after the GPU part. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "zest_gpu.h"

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "abi_check: %s failed (line %d): %s\n", #c, __LINE__, zg_last_error()); return 1; } } while (0)

int main(int argc, char** argv) {
    /* hash text form: four little-endian u64 printed %016x */
    uint8_t h[32], back[32];
    char hex[65];
    int i;
    for (i = 0; i < 32; i++) h[i] = (uint8_t)i;
    zg_hash_to_hex(h, hex);
    CHECK(strcmp(hex, "0706050403020100" "0f0e0d0c0b0a0908" "1716151413121110" "1f1e1d1c1b1a1918") == 0);
    CHECK(zg_hex_to_hash(hex, back) == ZG_OK && memcmp(h, back, 32) == 0);
    CHECK(zg_hex_to_hash("xyz", back) == ZG_ERR_INVALID_ARG);

    /* chunk-stream walk: two stored chunks (scheme 0) of 3 and 5 bytes */
    {
        const uint8_t blob[] = {0, 3, 0, 0, 0, 3, 0, 0, 'a', 'b', 'c', 0, 5, 0, 0, 0, 5, 0, 0, 'h', 'e', 'l', 'l', 'o'};
        uint32_t n = 0, clen[2], ulen[2];
        uint64_t off[2], total = 0;
        uint8_t scheme[2];
        CHECK(zg_xorb_index(blob, sizeof blob, 2, &n, off, clen, ulen, scheme, &total) == ZG_OK);
        CHECK(n == 2 && total == 8 && off[1] == 11 && clen[1] == 5 && scheme[0] == 0);
        CHECK(zg_xorb_index(blob, sizeof blob - 1, 2, &n, off, clen, ulen, scheme, &total) == ZG_ERR_TRUNCATED);
        CHECK(zg_chunk_range_size(blob, sizeof blob, 1, 2, &total) == ZG_OK && total == 5);
        CHECK(zg_chunk_range_size(blob, sizeof blob, 1, 3, &total) == ZG_ERR_RANGE);
    }
    /* longest-first partition, ties to the lower index / lower part */
    {
        const size_t lens[5] = {10, 30, 20, 30, 5};
        int32_t part[5];
        CHECK(zg_partition_xorbs(lens, 5, 2, part) == ZG_OK);
        CHECK(part[1] == 0 && part[3] == 1 && part[2] == 0 && part[0] == 1 && part[4] == 1);
    }
    /* footer size */
    CHECK(zg_xorb_footer_bytes(0) == 96 && zg_xorb_footer_bytes(10) == 496);
    CHECK(strcmp(zg_strerror(ZG_ERR_LZ4_BLOCK), "bad LZ4 block") == 0);

    {
        zg_ctx* ctx = NULL;
        int r = zg_init(0, NULL, &ctx);
        if (r == ZG_ERR_NO_DEVICE) {
            CHECK(ctx == NULL && strstr(zg_last_error(), "no CPU path") != NULL);
            printf("no-device\n");
            return 0;
        }
        CHECK(r == ZG_OK && ctx != NULL);
        {
            /* one xorb of two stored chunks: verify without an expected hash gives the root; with it, verdict 1;
             * with a wrong one, verdict 0; and the same through a group of two contexts on this device */
            const uint8_t blob[] = {0, 3, 0, 0, 0, 3, 0, 0, 'a', 'b', 'c', 0, 5, 0, 0, 0, 5, 0, 0, 'h', 'e', 'l', 'l', 'o'};
            const uint8_t* blobs[2];
            size_t lens[2];
            int8_t verdict[2] = {9, 9};
            uint8_t roots[64], expected[64], out[8];
            uint64_t nb = 0, got = 0;
            zg_group* g = NULL;
            int devs[2] = {0, 0};
            zg_stats st;
            blobs[0] = blobs[1] = blob;
            lens[0] = lens[1] = sizeof blob;
            CHECK(zg_verify_xorbs(ctx, blobs, lens, NULL, 1, verdict, roots, NULL, 0, &nb) == ZG_OK && verdict[0] == 1 && nb == 8);
            memcpy(expected, roots, 32);
            memcpy(expected + 32, roots, 32);
            expected[32] ^= 1;
            CHECK(zg_verify_xorbs(ctx, blobs, lens, expected, 2, verdict, roots, NULL, 0, &nb) == ZG_OK);
            CHECK(verdict[0] == 1 && verdict[1] == 0 && nb == 16);
            CHECK(zg_extract_chunk_range(ctx, blob, sizeof blob, 0, 2, out, sizeof out, &got) == ZG_OK && got == 8 && memcmp(out, "abchello", 8) == 0);
            CHECK(zg_group_init(devs, 2, NULL, &g) == ZG_OK && zg_group_size(g) == 2 && zg_group_ctx(g, 1) != NULL);
            verdict[0] = verdict[1] = 9;
            CHECK(zg_group_verify_xorbs(g, blobs, lens, expected, 2, verdict, roots, NULL, 0, &nb) == ZG_OK);
            CHECK(verdict[0] == 1 && verdict[1] == 0 && nb == 16 && memcmp(roots, expected, 32) == 0);
            zg_group_shutdown(g);
            CHECK(zg_get_stats(ctx, &st) == ZG_OK && st.calls >= 3 && st.kernels_launched > 0);
        }
        zg_shutdown(ctx);
        printf("gpu-ok\n");
    }
    (void)argc; (void)argv;
    return 0;
}
