// TEST INFRASTRUCTURE (oracle/): plain C++ restatement of the reference's one native function,
// map_events (utils/modules_to_build/transformations/src/transformation.cpp:35-85, exported as
// transformation.map by src/module.cpp:18-40), with std containers only.  Checker for the CUDA
// path and the single-thread CPU baseline of bench.py; never linked into the product library.
//
//   :41-42  src and dst must have the same length             -> return 1 (ValueError upstream)
//   :50-54  hash map  src pixel -> destinations, in LUT order;  the KEY is an int: the uint64 LUT
//           entry is narrowed on insertion
//   :57-63  pass 1: count the output rows; the pixel of an event is int(y) * W + int(x), the float
//           coordinates truncated toward zero (:62); 32-bit loop counters
//   :71-83  pass 2: for every event in input order and every destination in insertion order one
//           row: (x, y) = (dst % W, dst / W) as floats, columns 2.. copied verbatim
// Events on pixels without an entry are dropped; a pixel with k entries emits k rows.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <unordered_map>
#include <vector>

extern "C" int port_map_events(const float* events, long n, long e, int H, int W, const uint64_t* src,
                               long n_src, const uint64_t* dst, long n_dst, float** out, long* n_out) {
    (void)H;
    if (n_src != n_dst) return 1;
    std::unordered_map<int, std::vector<int>> lut;
    for (unsigned k = 0; k < (unsigned long)n_src; ++k) lut[(int)src[k]].push_back((int)dst[k]);
    auto pixel = [&](const float* ev) { return (int)ev[1] * W + (int)ev[0]; };
    size_t rows = 0;
    for (unsigned i = 0; i < (unsigned long)n; ++i) {
        auto it = lut.find(pixel(events + (size_t)i * e));
        if (it != lut.end()) rows += it->second.size();
    }
    const size_t cells = rows * (size_t)e;
    float* o = static_cast<float*>(std::calloc(cells > 0 ? cells : 1, sizeof(float)));
    size_t j = 0;
    for (unsigned i = 0; i < (unsigned long)n; ++i) {
        const float* ev = events + (size_t)i * e;
        auto it = lut.find(pixel(ev));
        if (it == lut.end()) continue;
        for (int d : it->second) {
            float* r = o + (j++) * (size_t)e;
            r[0] = (float)(d % W);
            r[1] = (float)(d / W);
            std::memcpy(r + 2, ev + 2, (size_t)(e - 2) * sizeof(float));
        }
    }
    *out = o;
    *n_out = (long)rows;
    return 0;
}

extern "C" void port_map_events_free(float* p) { std::free(p); }
