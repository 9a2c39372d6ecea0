// Host check: ssr::std_sort must produce exactly the permutation of libstdc++ std::sort with the reference's
// comparator (value only) on 16-byte {float,size_t} elements, for tie-heavy / presorted / adversarial inputs.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
#include "../../lego-loam_comment_b200/csrc/std_sort_replay.cuh"

struct smoothness_t { float value; size_t ind; };
struct by_value { bool operator()(smoothness_t const& l, smoothness_t const& r) { return l.value < r.value; } };

int main(int argc, char** argv) {
    int cases = argc > 1 ? atoi(argv[1]) : 20000;
    std::mt19937 rng(777);
    long mismatches = 0, heap_cases = 0;
    for (int c = 0; c < cases; ++c) {
        int n = 1 + rng() % 420;
        if (c % 50 == 0) n = 1 + rng() % 2100;
        std::vector<float> v(n);
        int mode = c % 8;
        for (int i = 0; i < n; ++i) {
            switch (mode) {
                case 0: v[i] = (float)(rng() % 5); break;                       // very tie heavy
                case 1: v[i] = (float)(rng() % 50) * 0.01f; break;
                case 2: v[i] = (float)i; break;                                  // presorted
                case 3: v[i] = (float)(n - i); break;                            // reversed
                case 4: v[i] = (float)((i * 7919) % 13); break;
                case 5: { float d = (float)((int)(rng() % 2001) - 1000) * 1e-3f; v[i] = d * d; } break;   // curvature-like squares
                case 6: v[i] = (i % 2) ? (float)i : (float)(n - i); break;      // organ pipe-ish
                default: v[i] = (float)rng() / 4294967296.0f; break;
            }
        }
        if (mode == 6 && n > 32) {   // median-of-3 killer style pattern to drive the heap fallback
            int k = n / 2;
            for (int i = 0; i < k; ++i) { v[i] = (i % 2 == 0) ? (float)(i + 1) : (float)(k + i + (k % 2 == 0 ? 0 : 1)); }
            for (int i = k; i < n; ++i) v[i] = (float)((i - k + 1) * 2);
        }
        std::vector<smoothness_t> a(n);
        std::vector<ssr::Item> b(n);
        for (int i = 0; i < n; ++i) { a[i].value = v[i]; a[i].ind = i; b[i].value = v[i]; b[i].ind = i; }
        std::sort(a.begin(), a.end(), by_value());
        ssr::std_sort(b.data(), n);
        for (int i = 0; i < n; ++i) if ((int)a[i].ind != b[i].ind) { ++mismatches; break; }
    }
    printf("{\"cases\": %d, \"mismatches\": %ld}\n", cases, mismatches);
    return mismatches ? 1 : 0;
}
