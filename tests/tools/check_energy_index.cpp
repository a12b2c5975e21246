// CPU emulation of the device function energy_index() (particle-simulation_b200/csrc/pic_device.cuh)
// with the same IEEE fp32 operations, checked against the reference's energyToIndex formula
// evaluated in double (src/cross_section.cu:32-35).  The edge table is read from stdin
// (10001 raw floats, as produced by picb200_host_energy_edges).
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
    uint32_t b;
    memcpy(&b, &e, 4);
    const int ex = (int)(b >> 23) - 127;
    uint32_t mb = (b & 0x007fffffu) | 0x3f800000u;
    float m;
    memcpy(&m, &mb, 4);
    float p = fmaf(m, 0.155445858836174f, -1.0392581224441528f);
    p = fmaf(p, m, 3.0294783115386963f);
    p = fmaf(p, m, -2.1449406147003174f);
    const float lg = (float)ex + p;
    const float s = fmaf(lg, 136.83181621090054f, 2727.2727272727275f);
    int j = rn_int(s);
    j = j < 0 ? 0 : (j > kRows ? kRows : j);
    int idx = j - ((e < edge[j]) ? 1 : 0);
    idx = idx > kRows - 1 ? kRows - 1 : idx;
    return (e != e) ? 0 : idx;
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
        if (dev_index(e, edge.data()) != ref_index(e)) {
            if (bad < 5) fprintf(stderr, "mismatch e=%a dev=%d ref=%d\n", e, dev_index(e, edge.data()), ref_index(e));
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
