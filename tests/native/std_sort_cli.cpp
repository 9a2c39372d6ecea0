// Reads segments from stdin (binary: int nseg, int offsets[nseg+1], float values[total]) and writes the
// permutation libstdc++ std::sort produces with the reference's value-only comparator (utility.h:139-148).
// Used by GPU tests to check the device replay against the real std::sort of this image's GCC.
#include <algorithm>
#include <cstdio>
#include <vector>
struct smoothness_t { float value; size_t ind; };
struct by_value { bool operator()(smoothness_t const& l, smoothness_t const& r) { return l.value < r.value; } };
int main() {
    int nseg; if (fread(&nseg, 4, 1, stdin) != 1) return 1;
    std::vector<int> off(nseg + 1); if (fread(off.data(), 4, nseg + 1, stdin) != (size_t)nseg + 1) return 1;
    std::vector<float> v(off[nseg]); if (fread(v.data(), 4, v.size(), stdin) != v.size()) return 1;
    std::vector<int> out(v.size());
    for (int s = 0; s < nseg; ++s) {
        std::vector<smoothness_t> a(off[s + 1] - off[s]);
        for (size_t i = 0; i < a.size(); ++i) { a[i].value = v[off[s] + i]; a[i].ind = i; }
        std::sort(a.begin(), a.end(), by_value());
        for (size_t i = 0; i < a.size(); ++i) out[off[s] + i] = (int)a[i].ind;
    }
    fwrite(out.data(), 4, out.size(), stdout);
    return 0;
}
