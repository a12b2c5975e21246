// =====================================================================================
// pic_oracle.cpp -- CPU ORACLE for the PIC-MCC electron hot path.   TEST INFRASTRUCTURE.
//
// This file is NOT part of the product.  It is a plain C++17 restatement (no CUDA) of the
// algorithm of MagnusMouritzen/particle-simulation for the one path this repository
// accelerates, used only as (a) the parity checker in tests/, (b) the checker inside
// __graft_entry__.smoke(), and (c) the timed CPU baseline of bench.py.  Nothing under
// particle-simulation_b200/ may include, link or call it.
//
// Every function cites the reference lines it restates (paths relative to the reference
// checkout).  Two things are deliberately NOT the reference's, as BASELINE.json's
// north_star prescribes, and are restated identically (but independently) in the engine:
//   * RNG: cuRAND XORWOW bound to the *slot* (src/random.cu:3-12) is replaced by the
//     counter-based Philox4x32-10 keyed on (seed, particle id, global mobility step).
//     Philox is pinned here by the Random123 known-answer vectors (tests/test_oracle.py).
//   * state precision: fp32 positions/velocities (`Real=float`).  `Real=double` keeps the
//     reference's double/float mix (src/electron.h:12-18) for sensitivity checks.
//
// PARITY PINNING: the reference ships no golden vectors and no CPU implementation of this
// path (every physics function is __device__), and this container has no GPU, so the
// oracle is pinned by (i) Philox KATs, (ii) the reference's own energyToIndex formula
// evaluated literally in double, (iii) the closed-form collision-free fixture
//  and (iv) statistical agreement with the reference CUDA binary measured
// on the B200 box.  Bit-level parity of the RNG stream with the reference
// is "unpinned" by construction (different generator, as north_star requires).
// =====================================================================================
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace orc {

// ---- constants: src/electron.h:8-10, src/cell.h:5-7, src/cell.cu:3-5, src/cross_section.h:9
static const int    kDead        = -2;                    // DEAD
static const double kElectronQ   = -1.602176487e-19;      // electron_charge
static const double kElectronM   = 9.1093837015e-31;      // electron_mass
static const double kCellSize    = 1e-2;                  // cell_size
static const double kEps0        = 8.8541878176E-12;      // epsilon0
static const double kPi          = 3.1415926536;          // pi (the reference's 11-digit value)
static const int    kGrid        = 512;                   // Grid_Size.{x,y,z}
static const int    kTableRows   = 10000;                 // N_STEPS
static const double kMobilityDt  = 1e-12;                 // src/pic.cu:371
static const uint64_t kSeedDefault = 39587;               // src/random.cu:4

// src/cell.cu:5 -- evaluated in the reference's order.
static inline double electric_force_constant() {
    return (kElectronQ * kElectronQ) / (4 * kPi * kEps0 * kCellSize * kCellSize * kElectronM);
}
// src/cell.cu:4 -- Sim_Size is a float3: Grid_Size * cell_size rounded to float.
static inline float sim_size_f() { return (float)(kGrid * kCellSize); }

// ---- Philox4x32-10 (Salmon et al., SC'11; same constants as curand_philox4x32_x.h:88-91)
struct Philox4 { uint32_t w[4]; };
static inline Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                    uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * c0;
        uint64_t p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    return Philox4{{c0, c1, c2, c3}};
}

// Stream tags (4th counter word).  One Philox block serves four consecutive global
// mobility steps: block = g >> 2, word = g & 3.
enum : uint32_t { kTagCollide = 0, kTagInit = 1 };

// curand_uniform's mapping (0,1]: x*2^-32 + 2^-33 (curand_kernel.h `_curand_uniform`),
// as used by src/random.cu:7-12.  Written with an explicit fma so CPU and GPU agree bitwise.
static inline float uniform01(uint32_t w) { return std::fmaf((float)w, 0x1p-32f, 0x1p-33f); }

// Child identity: splitmix64 finaliser of (parent id, global step); bit 63 marks "spawned".
static inline uint64_t child_id(uint64_t parent, uint64_t g) {
    uint64_t z = parent + g * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= (z >> 31);
    return z | 0x8000000000000000ull;
}

// src/cross_section.cu:32-35 -- literally the reference's expression, in double.
// The double->int conversion is made explicit with CUDA's saturating semantics
// (NaN -> 0, -inf -> INT_MIN, +inf -> INT_MAX) because that is what the device code does.
static inline int energy_to_index(double energy) {
    double v = std::trunc((std::log10(energy) + 6) * kTableRows / 22);
    int idx;
    if (std::isnan(v)) idx = 0;
    else if (v <= -2147483648.0) idx = INT32_MIN;
    else if (v >= 2147483647.0) idx = INT32_MAX;
    else idx = (int)v;
    return (idx < 0) ? 0 : ((idx >= kTableRows) ? kTableRows - 1 : idx);
}

struct CSRow { float split_chance, remove_chance; };   // src/cross_section.h:11-16

// src/cross_section.cu:3-28 -- `ifstream >> float` pairs; stops at the first short read.
// Returns rows read.  (The reference carries on with uninitialised memory when the file
// is missing; the oracle zero-fills instead so behaviour is defined.)
static int load_cross_sections(const char* path, std::vector<CSRow>& rows) {
    rows.assign(kTableRows, CSRow{0.f, 0.f});
    std::ifstream in(path);
    if (!in) return -1;
    int i = 0;
    for (; i < kTableRows; ++i) {
        if (!(in >> rows[i].split_chance >> rows[i].remove_chance)) break;
    }
    return i;
}

template <class Real> struct RealTraits;
template <> struct RealTraits<float> {
    static inline float fma(float a, float b, float c) { return std::fmaf(a, b, c); }
};
template <> struct RealTraits<double> {
    static inline double fma(double a, double b, double c) { return std::fma(a, b, c); }
};

// One electron: src/electron.h:12-18 (+ a stable 64-bit id the reference does not have).
template <class Real> struct Electron {
    Real x, y, z, vx, vy, vz;
    float ax, ay, az;
    int ts;
    uint64_t id;
};

template <class Real> class Sim {
public:
    typedef Electron<Real> E;
    std::vector<E> el;              // slots [0, n)
    int capacity = 0;
    int pt = 1;                     // poisson_timestep (mobility steps per Poisson step)
    uint64_t seed = kSeedDefault;
    std::vector<CSRow> table;
    int64_t n_virtual = 0;          // the device counter `n` (may overshoot capacity)
    int64_t total_added = 0, total_removed = 0, pushes = 0;
    int poisson_index = 0;
    int last_n_after_mobility = 0;
    int threads = 1;

    int32_t* charge = nullptr;      // dense 512^3 counts, calloc'd (pages touched lazily)
    std::vector<uint32_t> touched;
    const double efc = electric_force_constant();
    const Real dt = (Real)kMobilityDt;
    const Real hdt = (Real)(kMobilityDt / 2);
    const Real sim = (Real)sim_size_f();

    Sim(int cap, int pt_) : capacity(cap), pt(pt_) { table.assign(kTableRows, CSRow{0.f, 0.f}); }
    ~Sim() { free(charge); }
    Sim(const Sim&) = delete;
    Sim& operator=(const Sim&) = delete;

    // ---- src/particle_move.cu:7-19 (setup_particles): position ~ U(lo,hi] per axis,
    // timestamp -1.  Velocity/acceleration are uninitialised in the reference; 0 here.
    // The three draws are words 0,1,2 of Philox(counter=(id, block 0, kTagInit)).
    void init_uniform(int init_n, uint64_t id_base) {
        el.clear();
        el.reserve(init_n);
        const float lo = (float)((kGrid / 2 - 30) * kCellSize);
        const float hi = (float)((kGrid / 2 + 32) * kCellSize);
        const float range = hi - lo;
        for (int i = 0; i < init_n; ++i) {
            uint64_t id = id_base + (uint64_t)i;
            Philox4 r = philox4x32_10((uint32_t)id, (uint32_t)(id >> 32), 0, kTagInit,
                                      (uint32_t)seed, (uint32_t)(seed >> 32));
            E e{};
            e.x = (Real)std::fmaf(uniform01(r.w[0]), range, lo);
            e.y = (Real)std::fmaf(uniform01(r.w[1]), range, lo);
            e.z = (Real)std::fmaf(uniform01(r.w[2]), range, lo);
            e.ts = -1;
            e.id = id;
            el.push_back(e);
        }
        n_virtual = init_n;
    }

    // ---- src/particle_move.cu:22-39 (leapfrog): kick-drift-kick, electron sign.
    inline void leapfrog(E& e) const {
        typedef RealTraits<Real> T;
        e.vx = T::fma(-(Real)e.ax, hdt, e.vx);
        e.vy = T::fma(-(Real)e.ay, hdt, e.vy);
        e.vz = T::fma(-(Real)e.az, hdt, e.vz);
        e.x = T::fma(e.vx, dt, e.x);
        e.y = T::fma(e.vy, dt, e.y);
        e.z = T::fma(e.vz, dt, e.z);
        e.vx = T::fma(-(Real)e.ax, hdt, e.vx);
        e.vy = T::fma(-(Real)e.ay, hdt, e.vy);
        e.vz = T::fma(-(Real)e.az, hdt, e.vz);
    }

    // ---- src/particle_move.cu:41-52 (checkOutOfBounds)
    inline bool out_of_bounds(E& e) const {
        if (e.x < 0 || e.x >= sim || e.y < 0 || e.y >= sim || e.z < 0 || e.z >= sim) {
            e.ts = kDead;
            return true;
        }
        return false;
    }

    // ---- src/particle_move.cu:55-80 (collider) with the Philox stream.
    // g = global mobility step (poisson_index*pt + t), t = step inside this Poisson step.
    inline bool collide(E& e, E& child, int t, uint64_t g) const {
        typedef RealTraits<Real> T;
        Philox4 r = philox4x32_10((uint32_t)e.id, (uint32_t)(e.id >> 32), (uint32_t)(g >> 2),
                                  kTagCollide, (uint32_t)seed, (uint32_t)(seed >> 32));
        float rnd = uniform01(r.w[g & 3]) * 100.0f;              // randFloat(state, 0, 100)
        Real energy = T::fma(e.vz, e.vz, T::fma(e.vy, e.vy, e.vx * e.vx));
        int idx = energy_to_index((double)energy);
        float split = table[idx].split_chance, remove = table[idx].remove_chance;
        if (rnd < split) {
            child = e;                      // copy BEFORE the velocity flip
            child.ts = t;
            child.id = child_id(e.id, g);
            e.vx = -e.vx; e.vy = -e.vy; e.vz = -e.vz;
            return true;
        } else if (rnd < remove + split) {
            e.ts = kDead;
        }
        return false;
    }

    // ---- src/particle_move.cu:83-87 (updateParticle) = one "push".
    inline bool push(E& e, E& child, int t, uint64_t g) const {
        leapfrog(e);
        if (out_of_bounds(e)) return false;
        return collide(e, child, t, g);
    }

    // ---- src/grid_operations.cu:15-26, 59-72: int(pos / cell_size), double divide, trunc.
    static inline int cell_of(Real p) { return (int)((double)p / kCellSize); }
    static inline size_t cell_index(int ix, int iy, int iz) {
        return ((size_t)iz * kGrid + (size_t)iy) * kGrid + (size_t)ix;
    }

    // ---- src/grid_operations.cu:6-12 (resetGrid) + :15-26 (particlesToGrid)
    void deposit() {
        if (!charge) charge = (int32_t*)calloc((size_t)kGrid * kGrid * kGrid, sizeof(int32_t));
        for (uint32_t c : touched) charge[c] = 0;
        touched.clear();
        const int n = (int)el.size();
        for (int i = 0; i < n; ++i) {
            const E& e = el[i];
            size_t c = cell_index(cell_of(e.x), cell_of(e.y), cell_of(e.z));
            if (charge[c]++ == 0) touched.push_back((uint32_t)c);
        }
    }
    inline int charge_at(int ix, int iy, int iz) const {
        if (ix < 0 || iy < 0 || iz < 0 || ix >= kGrid || iy >= kGrid || iz >= kGrid) return 0;
        return charge[cell_index(ix, iy, iz)];
    }
    // ---- src/grid_operations.cu:29-56 (updateGrid) evaluated for one cell.
    inline void cell_acceleration(int ix, int iy, int iz, float& ax, float& ay, float& az) const {
        double xa = 0, ya = 0, za = 0;
        if (ix != 0) xa -= charge_at(ix - 1, iy, iz);
        if (ix != kGrid - 1) xa += charge_at(ix + 1, iy, iz);
        xa *= efc;
        if (iy != 0) ya -= charge_at(ix, iy - 1, iz);
        if (iy != kGrid - 1) ya += charge_at(ix, iy + 1, iz);
        ya *= efc;
        if (iz != 0) za -= charge_at(ix, iy, iz - 1);
        if (iz != kGrid - 1) za += charge_at(ix, iy, iz + 1);
        za *= efc;
        ax = (float)xa; ay = (float)ya; az = (float)za;
    }
    // ---- src/grid_operations.cu:59-72 (gridToParticles)
    void gather() {
        const int n = (int)el.size();
#pragma omp parallel for schedule(static) num_threads(threads)
        for (int i = 0; i < n; ++i) {
            E& e = el[i];
            cell_acceleration(cell_of(e.x), cell_of(e.y), cell_of(e.z), e.ax, e.ay, e.az);
        }
    }

    // ---- src/pic.cu:214-240 (simulateMany): all remaining steps of one electron.
    // Children are returned to the caller, which applies the capacity rule (:219-229).
    inline void simulate_many(E& e, int t_hi, std::vector<E>& kids, int64_t& npush) const {
        int start_t = std::max(1, e.ts + 1);
        E child;
        for (int t = start_t; t <= t_hi; ++t) {
            uint64_t g = (uint64_t)poisson_index * (uint64_t)pt + (uint64_t)t;
            ++npush;
            if (push(e, child, t, g)) kids.push_back(child);
            else if (e.ts == kDead) break;
        }
    }
    inline void append_children(const std::vector<E>& kids) {
        for (const E& k : kids) {
            if (n_virtual < capacity) {           // `if (*n < capacity)` src/pic.cu:222
                ++n_virtual;
                el.push_back(k);
            }
        }
    }

    // Mobility phase, generation by generation.  Generation 0 = electrons present at the
    // start of the Poisson step; generation k+1 = children of generation k.  This is the
    // set semantics shared by the reference's dynamic / cpuSync / dynamicOld schedulers
    // (src/pic.cu:70-211, :243-248 + host loop :514-523, :291-316).
    void mobility_generations() {
        size_t lo = 0, hi = el.size();
        while (lo < hi) {
            std::vector<std::vector<E>> kids((size_t)threads);
            std::vector<int64_t> np((size_t)threads, 0);
#pragma omp parallel num_threads(threads)
            {
#ifdef _OPENMP
                int tid = omp_get_thread_num(), nt = omp_get_num_threads();
#else
                int tid = 0, nt = 1;
#endif
                size_t span = hi - lo, a = lo + span * (size_t)tid / (size_t)nt,
                       b = lo + span * (size_t)(tid + 1) / (size_t)nt;
                for (size_t i = a; i < b; ++i) simulate_many(el[i], pt, kids[(size_t)tid], np[(size_t)tid]);
            }
            for (int t = 0; t < threads; ++t) { append_children(kids[(size_t)t]); pushes += np[(size_t)t]; }
            lo = hi; hi = el.size();
        }
    }

    // Mobility phase, one step at a time -- the reference's naive scheduler
    // (src/pic.cu:251-288 + host loop :524-532).
    void mobility_stepwise() {
        std::vector<E> kids;
        for (int t = 1; t <= pt; ++t) {
            uint64_t g = (uint64_t)poisson_index * (uint64_t)pt + (uint64_t)t;
            size_t start_n = el.size();
            kids.clear();
            E child;
            for (size_t i = 0; i < start_n; ++i) {
                E& e = el[i];
                if (e.ts == kDead) continue;
                ++pushes;
                if (push(e, child, t, g)) kids.push_back(child);
            }
            append_children(kids);
        }
    }

    // ---- src/pic.cu:320-357 (remove_dead_particles): drop DEAD, survivors get ts = -1.
    // (Order-preserving here; the reference's order is an atomic race.)
    void compact() {
        size_t w = 0;
        for (size_t i = 0; i < el.size(); ++i) {
            if (el[i].ts != kDead) { el[w] = el[i]; el[w].ts = -1; ++w; }
        }
        el.resize(w);
    }

    // ---- src/pic.cu:487-560: one iteration of the host Poisson loop.
    // stepwise != 0 selects the naive scheduler's order of evaluation.
    void poisson_step(int stepwise) {
        int64_t n_before = (int64_t)el.size();
        n_virtual = n_before;
        deposit();                 // resetGrid + particlesToGrid
        gather();                  // updateGrid + gridToParticles
        if (stepwise) mobility_stepwise(); else mobility_generations();
        total_added += n_virtual - n_before;          // src/pic.cu:539-540
        last_n_after_mobility = (int)n_virtual;
        compact();                                    // src/pic.cu:547-551
        total_removed += n_virtual - (int64_t)el.size();   // :552
        n_virtual = (int64_t)el.size();
        ++poisson_index;
    }
};

}  // namespace orc

// =====================================================================================
// C ABI for the tests (ctypes).  pos/vel cross the boundary as double for both precisions.
// =====================================================================================
struct OrcHandle {
    int fp64;
    orc::Sim<float>* f;
    orc::Sim<double>* d;
};
#define ORC_DISPATCH(h, expr) do { if ((h)->fp64) { auto& S = *(h)->d; expr; } else { auto& S = *(h)->f; expr; } } while (0)

extern "C" {

void orc_philox4x32_10(const uint32_t* ctr, const uint32_t* key, uint32_t* out) {
    orc::Philox4 r = orc::philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]);
    memcpy(out, r.w, 16);
}
float orc_uniform01(uint32_t w) { return orc::uniform01(w); }
uint64_t orc_child_id(uint64_t parent, uint64_t g) { return orc::child_id(parent, g); }
int orc_energy_to_index(double e) { return orc::energy_to_index(e); }
double orc_electric_force_constant() { return orc::electric_force_constant(); }
int orc_load_table(const char* path, float* out_rows2) {
    std::vector<orc::CSRow> rows;
    int r = orc::load_cross_sections(path, rows);
    memcpy(out_rows2, rows.data(), sizeof(orc::CSRow) * orc::kTableRows);
    return r;
}

OrcHandle* orc_create(int capacity, int poisson_timestep, int fp64, uint64_t seed, int threads) {
    OrcHandle* h = new OrcHandle{fp64, nullptr, nullptr};
    if (fp64) h->d = new orc::Sim<double>(capacity, poisson_timestep);
    else h->f = new orc::Sim<float>(capacity, poisson_timestep);
    ORC_DISPATCH(h, S.seed = seed; S.threads = threads < 1 ? 1 : threads);
    return h;
}
void orc_destroy(OrcHandle* h) { if (!h) return; delete h->f; delete h->d; delete h; }
void orc_set_table(OrcHandle* h, const float* rows2) {
    ORC_DISPATCH(h, memcpy(S.table.data(), rows2, sizeof(orc::CSRow) * orc::kTableRows));
}
int orc_set_table_file(OrcHandle* h, const char* path) {
    int r = 0;
    ORC_DISPATCH(h, r = orc::load_cross_sections(path, S.table));
    return r;
}
void orc_set_poisson_index(OrcHandle* h, int idx) { ORC_DISPATCH(h, S.poisson_index = idx); }
void orc_init_uniform(OrcHandle* h, int init_n, uint64_t id_base) { ORC_DISPATCH(h, S.init_uniform(init_n, id_base)); }

void orc_upload(OrcHandle* h, int n, const double* x, const double* y, const double* z,
                const double* vx, const double* vy, const double* vz,
                const float* ax, const float* ay, const float* az, const int* ts, const uint64_t* id) {
    ORC_DISPATCH(h, {
        typedef typename std::remove_reference<decltype(S)>::type::E E;
        typedef decltype(E().x) R;
        S.el.resize((size_t)n);
        for (int i = 0; i < n; ++i) {
            E& e = S.el[(size_t)i];
            e.x = (R)x[i]; e.y = (R)y[i]; e.z = (R)z[i];
            e.vx = (R)vx[i]; e.vy = (R)vy[i]; e.vz = (R)vz[i];
            e.ax = ax[i]; e.ay = ay[i]; e.az = az[i];
            e.ts = ts[i]; e.id = id[i];
        }
        S.n_virtual = n;
    });
}
int orc_n(OrcHandle* h) { int n = 0; ORC_DISPATCH(h, n = (int)S.el.size()); return n; }
void orc_download(OrcHandle* h, double* x, double* y, double* z, double* vx, double* vy, double* vz,
                  float* ax, float* ay, float* az, int* ts, uint64_t* id) {
    ORC_DISPATCH(h, {
        for (size_t i = 0; i < S.el.size(); ++i) {
            const auto& e = S.el[i];
            x[i] = e.x; y[i] = e.y; z[i] = e.z; vx[i] = e.vx; vy[i] = e.vy; vz[i] = e.vz;
            ax[i] = e.ax; ay[i] = e.ay; az[i] = e.az; ts[i] = e.ts; id[i] = e.id;
        }
    });
}
// counters: [0]=n (device counter incl. overshoot) [1]=n after last mobility phase
// [2]=total added [3]=total removed [4]=pushes [5]=poisson index
void orc_counters(OrcHandle* h, int64_t* out) {
    ORC_DISPATCH(h, {
        out[0] = S.n_virtual; out[1] = S.last_n_after_mobility; out[2] = S.total_added;
        out[3] = S.total_removed; out[4] = S.pushes; out[5] = S.poisson_index;
    });
}
void orc_poisson_step(OrcHandle* h, int stepwise) { ORC_DISPATCH(h, S.poisson_step(stepwise)); }
// individual phases, for kernel-level parity tests
void orc_op_deposit(OrcHandle* h) { ORC_DISPATCH(h, S.deposit()); }
void orc_op_gather(OrcHandle* h) { ORC_DISPATCH(h, S.gather()); }
void orc_op_mobility(OrcHandle* h, int stepwise) {
    ORC_DISPATCH(h, { S.n_virtual = (int64_t)S.el.size(); if (stepwise) S.mobility_stepwise(); else S.mobility_generations(); });
}
void orc_op_compact(OrcHandle* h) { ORC_DISPATCH(h, { S.compact(); S.n_virtual = (int64_t)S.el.size(); }); }
int orc_charge_at(OrcHandle* h, int ix, int iy, int iz) {
    int c = 0;
    ORC_DISPATCH(h, c = !S.charge ? 0 : S.charge_at(ix, iy, iz));
    return c;
}
// dense copy of the charge counts in the box [lo, hi] (inclusive), x fastest.
void orc_charge_box(OrcHandle* h, const int* lo, const int* hi, int32_t* out) {
    ORC_DISPATCH(h, {
        size_t k = 0;
        for (int z = lo[2]; z <= hi[2]; ++z)
            for (int y = lo[1]; y <= hi[1]; ++y)
                for (int x = lo[0]; x <= hi[0]; ++x) out[k++] = !S.charge ? 0 : S.charge_at(x, y, z);
    });
}

// Whole run (src/pic.cu:359-599 without printing); returns wall ms of the Poisson loop --
// the same span the reference brackets with its cudaEvents (:476, :563).
double orc_run(OrcHandle* h, int poisson_steps, int stepwise) {
    auto t0 = std::chrono::high_resolution_clock::now();
    for (int t = 0; t < poisson_steps; ++t) {
        orc_poisson_step(h, stepwise);
        if (orc_n(h) == 0) break;                 // "Hit 0" src/pic.cu:556-559
    }
    auto t1 = std::chrono::high_resolution_clock::now();
    return std::chrono::duration<double, std::milli>(t1 - t0).count();
}
int orc_max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}  // extern "C"

// =====================================================================================
// Optional CLI with the reference's 8 positional arguments and stdout format
// (src/main.cu:8-46, src/pic.cu:360, :457-474, :573-576, src/utility.cu:124-137,
// src/electron.h:19-26) so printed state can be diffed against the engine's `run`.
// =====================================================================================
#ifdef ORACLE_MAIN
template <class Real>
static void print_state(const orc::Sim<Real>& S, int t) {
    printf("Time %d, amount %d\n", t, (int)S.n_virtual);
    for (size_t i = 0; i < S.el.size(); ++i) {
        const auto& e = S.el[i];
        printf("%d: (%.15f, %.15f, %.15f) (%.15f, %.15f, %.15f) ((%.7f, %.7f, %.7f)) [%d]\n", (int)i,
               (double)e.x, (double)e.y, (double)e.z, (double)e.vx, (double)e.vy, (double)e.vz,
               e.ax, e.ay, e.az, e.ts);
    }
    printf("\n");
}

template <class Real>
static int run_cli(int mode, int verbose, int init_n, int max_t, int block, int max_n, int sleep_ns, int pt,
                   const char* table_path, int threads) {
    printf("PIC with\ninit n: %d\ncapacity: %d\npoisson steps: %d\npoisson_timestep: %d\nblock size: %d\nsleep time: %d\n",
           init_n, max_n, max_t, pt, block, sleep_ns);
    orc::Sim<Real> S(max_n, pt);
    S.threads = threads;
    if (orc::load_cross_sections(table_path, S.table) < 0) {
        fprintf(stderr, "Failed to open file: %s\n", table_path);
        return 2;
    }
    S.init_uniform(init_n, 0);
    printf("Multiprocessor count: %d\nBlocks per multiprocessor: %d\n", threads, 1);
    static const char* names[] = {"Dynamic", "CPU Sync", "Naive", "Dynamic Old"};
    printf("%s\n", names[mode]);
    auto t0 = std::chrono::high_resolution_clock::now();
    for (int t = 0; t < max_t; ++t) {
        if (verbose != 0 && t % verbose == 0) print_state(S, t);
        S.poisson_step(mode == 2);
        if (S.last_n_after_mobility >= max_n) printf("\n\nOVERFLOW FROM ADDING PARTICLES\n\n\n");
        if (S.el.empty()) { printf("Hit 0\n"); break; }
    }
    if (verbose != 0 && max_t % verbose == 0) print_state(S, max_t);
    auto t1 = std::chrono::high_resolution_clock::now();
    printf("Final amount of particles: %d\n", (int)std::min<int64_t>(S.n_virtual, max_n));
    printf("Particles added: %lld\n", (long long)S.total_added);
    printf("Particles removed: %lld\n", (long long)S.total_removed);
    printf("GPU time of program: %f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count());
    fprintf(stderr, "pushes: %lld\n", (long long)S.pushes);
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 9) {
        fprintf(stderr, "usage: %s <30|31|32|33> verbose init_n max_t block max_n sleep poisson_timestep\n", argv[0]);
        return 64;
    }
    int mode = atoi(argv[1]) - 30;
    if (mode < 0 || mode > 3) return 0;
    auto t0 = std::chrono::high_resolution_clock::now();
    const char* table = getenv("PICB200_TABLE") ? getenv("PICB200_TABLE") : "./src/cross_section.txt";
    int threads = getenv("ORACLE_THREADS") ? atoi(getenv("ORACLE_THREADS")) : 1;
    int fp64 = getenv("ORACLE_FP64") ? atoi(getenv("ORACLE_FP64")) : 0;
    int rc = fp64 ? run_cli<double>(mode, atoi(argv[2]), atoi(argv[3]), atoi(argv[4]), atoi(argv[5]), atoi(argv[6]), atoi(argv[7]), atoi(argv[8]), table, threads)
                  : run_cli<float>(mode, atoi(argv[2]), atoi(argv[3]), atoi(argv[4]), atoi(argv[5]), atoi(argv[6]), atoi(argv[7]), atoi(argv[8]), table, threads);
    auto t1 = std::chrono::high_resolution_clock::now();
    printf("CPU time of program: %f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count());
    return rc;
}
#endif
