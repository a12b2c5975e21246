// CPU emulation of the device function cell_of_fast() (particle-simulation_b200/csrc/pic_device.cuh)
// with the same IEEE fp32 operations, checked against the reference's cell formula evaluated in double,
// int(double(pos) / 1e-2) (src/grid_operations.cu:21-23), for EVERY fp32 value in [0, 5.12] and a band
// beyond both ends.  The boundary table is read from stdin (513 raw floats, as produced by
// picb200_host_cell_bounds).
//   usage: check_cell_index < bounds.bin
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

static const int kGrid = 512;

static inline int ref_cell(float x) {                      // cell_of(): the reference's formula, clamped to the grid
    int k = (int)((double)x / 1e-2);
    return k < 0 ? 0 : (k > kGrid - 1 ? kGrid - 1 : k);
}

#ifdef ONE_PROBE
static inline int dev_cell(float x, const float* bnd) {    // cell_of_fast() (the default single-probe form), operation for operation
    int k = (int)(x * 99.999969482421875f);                // kCellScaleLow = 100 - 4 ulp: the estimate is never above the index
    k = k < 0 ? 0 : (k > kGrid - 1 ? kGrid - 1 : k);
    k += (x >= bnd[k + 1]) ? 1 : 0;
    return k;
}
#else
static inline int dev_cell(float x, const float* bnd) {    // cell_of_fast() under -DPIC_CELL_TWO_PROBES, operation for operation
    int k = (int)(x * 100.0f);                             // __float2int_rz(__fmul_rn(p, 100.0f)); |x| < 8 here, no saturation
    k = k < 0 ? 0 : (k > kGrid - 1 ? kGrid - 1 : k);
    k -= (x < bnd[k]) ? 1 : 0;
    k += (x >= bnd[k + 1]) ? 1 : 0;
    return k;
}
#endif

int main() {
    std::vector<float> bnd(kGrid + 1);
    if (fread(bnd.data(), 4, kGrid + 1, stdin) != (size_t)kGrid + 1) { fprintf(stderr, "short boundary table\n"); return 2; }
    float hi = 8.0f;                                       // the domain is [0, 5.12) (src/particle_move.cu:37-42); go well past it
    uint32_t bhi;
    memcpy(&bhi, &hi, 4);
    long long bad = 0, checked = 0;
    for (int sign = 0; sign < 2; ++sign) {
#pragma omp parallel for reduction(+ : bad, checked) schedule(static)
        for (long long b = 0; b <= (long long)bhi; ++b) {  // every fp32 with |x| <= 8, denormals and +-0 included
            uint32_t u = (uint32_t)b | (sign ? 0x80000000u : 0u);
            float x;
            memcpy(&x, &u, 4);
            ++checked;
            const int r = ref_cell(x), d = dev_cell(x, bnd.data());
            if (d != r) {
                if (bad < 10) fprintf(stderr, "mismatch x=%.9g dev=%d ref=%d\n", x, d, r);
                ++bad;
            }
        }
    }
    printf("checked %lld mismatches %lld\n", checked, bad);
    return bad ? 1 : 0;
}
