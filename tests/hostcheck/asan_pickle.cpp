// TEST ONLY - never part of the product.  The pickle decoder of the native cache reader (csrc/cache_reader.h) run
// over length-prefixed blobs from a file, each in a heap buffer of exactly its size, under AddressSanitizer and
// UBSan (tests/test_host_cpp_sanitized.py builds and feeds it): a cache.db is a file someone else wrote.
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include "cache_reader.h"
int main(int argc, char** argv) {
    FILE* f = fopen(argv[1], "rb");
    if (!f) return 2;
    sgcache::Table T;
    T.text_offsets.assign(1, 0);
    uint32_t len;
    long n = 0;
    uint8_t key[32] = {0};
    while (fread(&len, 4, 1, f) == 1) {
        // exact-size heap buffer so that ASan sees any read past the blob
        uint8_t* buf = (uint8_t*)malloc(len ? len : 1);
        if (len && fread(buf, 1, len, f) != len) break;
        sgcache::Decoded d = sgcache::decode_pickle(buf, len);
        sgcache::append_entry(&T, key, n, d, false);
        free(buf);
        ++n;
    }
    printf("decoded %ld blobs, m=%lld, fallbacks=%lld\n", n, (long long)T.m, (long long)T.fallbacks);
    return 0;
}
