// Memory-safety fuzz of the LZ4 frame codec (reth_b200/csrc/lz4f.h): built with
// -fsanitize=address,undefined by tests/test_lz4f_cpu.py.  Round trips of compressible and
// incompressible inputs through exact-size heap buffers, then mutated frames (byte writes, bit
// flips, truncation) into output buffers of random capacity: every outcome must be a byte count
// within the capacity or a negative status, never an out-of-bounds access.
#include "../../reth_b200/csrc/lz4f.h"
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

int main(int argc, char **argv)
{
    const int iters = argc > 1 ? atoi(argv[1]) : 1000;
    std::mt19937_64 rng(1);
    long decoded = 0, total = 0;
    for (int it = 0; it < iters; ++it) {
        const size_t n = rng() % 200000;
        std::vector<uint8_t> plain(n);
        const int mode = rng() % 3;
        for (size_t i = 0; i < n; ++i)
            plain[i] = mode == 0 ? (uint8_t)rng() : mode == 1 ? (uint8_t)((i / 7) % 5) : (uint8_t)((rng() % 4) * 60);
        std::vector<uint8_t> comp(lz4f::compress_bound((int64_t)n));
        const int64_t c = lz4f::compress(plain.data(), (int64_t)n, comp.data(), (int64_t)comp.size());
        if (c < 0) { printf("compress failed\n"); return 1; }
        std::vector<uint8_t> in(comp.begin(), comp.begin() + c), out(n);
        const int64_t d = lz4f::decompress(in.data(), (int64_t)in.size(), out.data(), (int64_t)out.size());
        if (d != (int64_t)n || memcmp(out.data(), plain.data(), n)) { printf("round trip failed at %d\n", it); return 1; }
        for (int m = 0; m < 8; ++m) {
            std::vector<uint8_t> bad(in);
            const int k = 1 + rng() % 4;
            for (int j = 0; j < k && !bad.empty(); ++j) {
                const int op = rng() % 3;
                if (op == 0) bad[rng() % bad.size()] = (uint8_t)rng();
                else if (op == 1) bad.resize(rng() % bad.size() + 1);
                else bad[rng() % bad.size()] ^= (uint8_t)(1u << (rng() % 8));
            }
            const size_t cap = rng() % (n + 100);
            std::vector<uint8_t> o2(cap);
            const int64_t r = lz4f::decompress(bad.data(), (int64_t)bad.size(), o2.data(), (int64_t)cap);
            ++total;
            decoded += r >= 0;
            if (r > (int64_t)cap) { printf("result beyond the capacity\n"); return 1; }
        }
    }
    printf("ok %ld mutated inputs, %ld decoded\n", total, decoded);
    return 0;
}
