// The exact bin-edge table of the device function energy_index() (particle-simulation_b200/csrc/pic_device.cuh), checked
// against the reference's energyToIndex formula evaluated in double (src/cross_section.cu:32-35) for every non-negative
// fp32: the index is "the last edge not above E" (what the device's estimate-and-probe resolves to; the estimate itself
// comes from the SFU's lg2.approx and is checked against a bisection of the same table ON THE DEVICE, all 2^31 values:
// picb200_dev_energy_index_sweep).  Also checked here: an estimate computed with a correctly rounded log2 lands within
// 0.02 of the exact s, i.e. the fp32 evaluation of the scaling leaves 0.48 of the 0.5 margin to the SFU.
// The edge table is read from stdin (10001 raw floats, as produced by picb200_host_energy_edges).
//   usage: check_energy_index <stride> < edges.bin      (stride 1 = every positive fp32)
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

static const int kRows = 10000;

static inline int ref_index(float e) {
    double v = std::trunc((std::log10((double)e) + 6) * kRows / 22);
    int idx;
    if (std::isnan(v)) idx = 0;
    else if (v <= -2147483648.0) idx = INT32_MIN;
    else if (v >= 2147483647.0) idx = INT32_MAX;
    else idx = (int)v;
    return idx < 0 ? 0 : (idx >= kRows ? kRows - 1 : idx);
}

static inline int rn_int(float s) {   // cvt.rni.s32.f32 with saturation, NaN -> 0
    if (s != s) return 0;
    if (s >= 2147483648.0f) return INT32_MAX;
    if (s <= -2147483648.0f) return INT32_MIN;
    return (int)std::nearbyintf(s);
}

static inline int dev_index(float e, const float* edge) {
    if (e != e) return 0;
    int lo = 0, hi = kRows;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (e >= edge[mid]) lo = mid; else hi = mid;
    }
    return lo;
}
// the device's scaling of log2(E) in fp32, with the host's (correctly rounded to < 1 ulp) log2f standing in for the SFU
static inline bool estimate_ok(float e) {
    if (!(e > 0.f) || std::isinf(e)) return true;
    const float s = fmaf(log2f(e), 136.83181621090054f, 2727.2727272727275f);
    const double exact = (std::log10((double)e) + 6) * kRows / 22;
    return std::fabs((double)s - exact) < 0.02;
}

int main(int argc, char** argv) {
    const uint32_t stride = argc > 1 ? (uint32_t)strtoul(argv[1], nullptr, 10) : 4099;
    std::vector<float> edge(kRows + 1);
    if (fread(edge.data(), 4, kRows + 1, stdin) != (size_t)kRows + 1) { fprintf(stderr, "short edge table\n"); return 2; }
    long long bad = 0, checked = 0;
    // every positive float including +0, denormals, inf and NaNs (exponent 255)
#pragma omp parallel for reduction(+ : bad, checked) schedule(static)
    for (long long b = 0; b <= 0x7fffffffLL; b += stride) {
        uint32_t u = (uint32_t)b;
        float e;
        memcpy(&e, &u, 4);
        ++checked;
        if (dev_index(e, edge.data()) != ref_index(e) || !estimate_ok(e)) {
            if (bad < 5) fprintf(stderr, "mismatch e=%a dev=%d ref=%d estimate_ok=%d\n", e, dev_index(e, edge.data()), ref_index(e), (int)estimate_ok(e));
            ++bad;
        }
    }
    // +-3 ulp around every edge
    for (int j = 1; j < kRows; ++j) {
        uint32_t c;
        memcpy(&c, &edge[j], 4);
        for (int d = -3; d <= 3; ++d) {
            uint32_t u = c + (uint32_t)d;
            float e;
            memcpy(&e, &u, 4);
            ++checked;
            if (dev_index(e, edge.data()) != ref_index(e)) { ++bad; fprintf(stderr, "edge mismatch j=%d d=%d\n", j, d); }
        }
    }
    printf("checked %lld mismatches %lld\n", checked, bad);
    return bad ? 1 : 0;
}
