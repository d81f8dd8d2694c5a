// TEST INFRASTRUCTURE (oracle/): C entry point around the reference's map_events
// (utils/modules_to_build/transformations/src/transformation.cpp:35-85, compiled from the
// reference tree by oracle/Makefile) so tests can call it through ctypes -- what
// src/module.cpp:18-40 does through pybind11.  Returns 0, or 1 for std::invalid_argument (the
// pybind binding turns that into ValueError).
#include <transformation.h>
#include <cstdlib>
#include <stdexcept>

extern "C" int ref_map_events(float* events, long n, long e, int H, int W, uint64_t* src, long n_src,
                              uint64_t* dst, long n_dst, float** out, long* n_out) {
    try {
        Eigen::Ref<Event_t> ev(events, n, e);
        Eigen::Ref<VectorXull> s(src, n_src, 1), d(dst, n_dst, 1);
        Event_t r = map_events(ev, std::make_tuple(H, W), s, d);
        *n_out = r.rows();
        const size_t bytes = (size_t)r.rows() * (size_t)r.cols() * sizeof(float);
        *out = static_cast<float*>(std::malloc(bytes ? bytes : 1));
        std::memcpy(*out, r.data(), bytes);
        return 0;
    } catch (const std::invalid_argument&) {
        return 1;
    }
}

extern "C" void ref_map_events_free(float* p) { std::free(p); }
