/*
 * bd_device.cuh -- device-side data layout and arithmetic of the B200 Brownian-dynamics engine.
 *
 * Reference algorithm (micromd, header-only C++; MD/ = submodules/github.com/snsinfu/micromd/
 * include/md/).  Nothing here is translated from it: the reference walks a std::vector pair list
 * with Newton's-third-law scatter on one CPU thread; this engine keeps beads sorted by cell in
 * 32-byte records, evaluates every bead's full neighbour row without atomics, and updates the
 * position in the same pass.
 *
 * HBM layout (per engine, R replicas of N beads, Ntot = R*N, "slot" = index in cell-sorted order,
 * slots [r*N, (r+1)*N) belong to replica r):
 *   pos[2][Ntot]   bead4 {x, y, z, w}  ping-pong; w packs original index, class, target masks
 *   bpos[Ntot]     bead4               positions at the last Verlet-list build (same slots)
 *   slot_of[Ntot]  u32                 original global index -> slot
 *   nbr[cap][Ntot] u32, nnbr[Ntot]     ELL Verlet rows (both directions), column-major
 *   bond/bend CSR  by original index   topology, uploaded once
 */
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/bd_types.h"

namespace bd {

// ---- records -----------------------------------------------------------------------------------

/* Slot Ntot of every per-bead array is a sentinel: NaN coordinates that fail every distance test.
 * Verlet rows are padded with it to a multiple of ROW_PAD entries. */
constexpr uint32_t ROW_PAD = 4;

struct __align__(32) bead4 {
    double x, y, z;
    uint64_t w;  // bits 0..31 original index within the replica, 32..39 class,
                 // 40..47 pair-forcefield target mask, 48..55 point-forcefield target mask
};

__host__ __device__ inline uint64_t pack_w(uint32_t orig, uint32_t type, uint32_t pair_mask,
                                           uint32_t point_mask)
{
    return (uint64_t) orig | ((uint64_t) (type & 0xff) << 32) | ((uint64_t) (pair_mask & 0xff) << 40)
        | ((uint64_t) (point_mask & 0xff) << 48);
}
__host__ __device__ inline uint32_t w_orig(uint64_t w) { return (uint32_t) w; }
__host__ __device__ inline uint32_t w_type(uint64_t w) { return (uint32_t) (w >> 32) & 0xff; }
__host__ __device__ inline uint32_t w_pair_mask(uint64_t w) { return (uint32_t) (w >> 40) & 0xff; }
__host__ __device__ inline uint32_t w_point_mask(uint64_t w) { return (uint32_t) (w >> 48) & 0xff; }
/* Bit 7 of the point-mask byte marks a GHOST: a bead another engine (another GPU's slab) advances.  Its
 * position only changes through bd_scatter_positions_device; no force is evaluated for it, but it is a
 * neighbour and a bonded partner like any other bead. */
constexpr uint32_t GHOST_BIT = 0x80u;
__host__ __device__ inline bool w_ghost(uint64_t w) { return ((uint32_t) (w >> 48) & GHOST_BIT) != 0; }

/* flattened potential term with the constants the kernels need precomputed on the host */
struct __align__(16) DTerm {
    int32_t kind, p, q, pad;
    double factor, energy, length;
    double cutoff2;  // cutoff^2, 0 = none
    double k2;       // 1 / length^2
    double pad2;
};
static_assert(sizeof(DTerm) == 64, "DTerm is one 64-byte record");

struct DPair { int32_t enabled, nterms, term_off, has_targets; double dcut, dcut2; };
struct DBond { int32_t enabled, nterms, term_off, min_image; };
struct DBend { int32_t enabled, pad[3]; };
struct DWall {
    int32_t enabled, nterms, term_off, shape, direction, pad[3];
    double center[3];
    double semi[3];
    double quad[3];  // 1 / (semi * semi), rounded exactly as ellipsoid_surface_forcefield.hpp:72-76 does
    double range;    // beyond this distance from the surface every row of the table is exactly zero
                     // (softcore diameters, cutoffs); 0 = some term has unbounded range.  Host-computed.
};
struct DPoint { int32_t enabled, nterms, term_off, has_targets; double source[3]; double pad; };

/* Geometry of one wall forcefield of ONE replica.  The interphase driver moves the wall semiaxes of every
 * system from that system's own wall reaction after each step
 * (5-sim-genome/src/simulation_interphase/simulation_driver_interphase.cc:70-80), so in an ensemble the
 * geometry is per replica: wall_geom[replica][BD_MAX_WALL_FF], device-resident and the master copy
 * (bd_update_wall_ff[_replica] writes it, k_wall_relax advances it).  The first nine doubles have the
 * layout of DWall::center/semi/quad, which the kernels overwrite in their shared copy of DParams. */
struct __align__(16) DWallGeom {
    double center[3];
    double semi[3];
    double quad[3];
    double pad[3];
};
static_assert(sizeof(DWallGeom) == 96, "DWallGeom is 96 bytes");
constexpr int WALL_GEOM_DOUBLES = 9;

/* everything the host may change between steps; uploaded as one block when dirty */
struct DParams {
    int32_t ntypes, box_kind;
    int32_t pad0[2];
    double period[3], inv_period[3];
    double mobility[BD_MAX_TYPES];
    double dcut_max;       // max dcut over enabled pair forcefields (0 = none)
    double verlet_factor;  // list radius / dcut_max
    DPair pair[BD_MAX_PAIR_FF];
    DBond bond[BD_MAX_BOND_FF];
    DBend bend[BD_MAX_BEND_FF];
    DWall wall[BD_MAX_WALL_FF];
    DPoint point[BD_MAX_POINT_FF];
};

/* cell grid of the current Verlet list; written on the device at every rebuild */
struct DGrid {
    double origin[3];
    double inv_cell[3];
    int32_t n[3];
    int32_t bits[3];
    int32_t totalbits;
    int32_t pad;
    double radius;  // list radius the rows were built with
    // single-precision shadow coordinates of the pre-filter (bd_step.cuh): |r_shadow - r| <= shadow_eps / 4
    double shadow_eps;
    float shadow_scale[3];  // periodic: period * 2^-32 (fixed-point fraction -> length)
    float pad2;
};

/* device-resident control block: kernels decide on rebuilds without returning to the host */
struct DControl {
    uint32_t need_rebuild;   // host request (positions replaced, capacity changed)
    uint32_t doing_rebuild;  // latched copy for the kernels of one rebuild chain
    uint32_t overflow;       // beads whose ELL row overflowed (sticky)
    uint32_t max_row;        // longest ELL row at the last build
    uint32_t stage_overflow; // tiles whose neighbourhood did not fit the staging tables (sticky; bd_nlist.cuh)
    uint32_t max_stage;      // largest staged neighbourhood (beads) of a tile at the last build
    unsigned long long rebuilds;
    unsigned long long graph_rebuilds;  // ... of which inside a step graph's conditional node (bd_engine.cu)
    unsigned long long max_disp2_bits;  // bits of max_i |x_i - x_i(build)|^2, kept by the step kernel
    unsigned long long listed_entries;  // sum of row lengths at the last build (2 x listed pairs)
    unsigned long long bbox_lo[3], bbox_hi[3];  // order-preserving encodings of the bounding box
    alignas(16) DGrid grid;                     // copied to shared memory in 16-byte pieces
};

/* forcefield selection for compute_forces / compute_energy of a single forcefield */
struct Select {
    int32_t ff_class;  // BD_FF_ALL or one class
    int32_t ff;
};

__device__ inline bool ff_on(Select sel, int cls, int k, int enabled)
{
    return sel.ff_class == BD_FF_ALL ? enabled != 0 : (sel.ff_class == cls && sel.ff == k);
}

// ---- exact fp64 helpers (never contracted into FMA: bucket ids, pair sets and the rebuild
//      criterion must match the reference bit for bit) -------------------------------------------

struct d3 { double x, y, z; };

__device__ inline double sqnorm_exact(double x, double y, double z)
{
    // MD/basic_types/point.hpp:120-127: x*x + y*y + z*z, left to right
    return __dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z));
}

__device__ inline double round_mod_exact(double x, double m, double inv_m)
{
    // MD/misc/nsearch_detail/math.hpp:34-38: x - nearbyint(x * (1/m)) * m
    return __dsub_rn(x, __dmul_rn(nearbyint(__dmul_rn(x, inv_m)), m));
}

__device__ inline double floor_mod_exact(double x, double m, double inv_m)
{
    // MD/misc/nsearch_detail/math.hpp:27-31
    return __dsub_rn(x, __dmul_rn(floor(__dmul_rn(x, inv_m)), m));
}

/* MD/misc/box.hpp:25-28 (open), :45-51 (periodic): shortest displacement p1 - p2 */
__device__ inline d3 displacement(const DParams &P, double ax, double ay, double az, double bx,
                                  double by, double bz)
{
    d3 r;
    r.x = __dsub_rn(ax, bx);
    r.y = __dsub_rn(ay, by);
    r.z = __dsub_rn(az, bz);
    if (P.box_kind == BD_BOX_PERIODIC) {
        r.x = round_mod_exact(r.x, P.period[0], P.inv_period[0]);
        r.y = round_mod_exact(r.y, P.period[1], P.inv_period[1]);
        r.z = round_mod_exact(r.z, P.period[2], P.inv_period[2]);
    }
    return r;
}

// ---- potentials (MD/potential/*.hpp) ------------------------------------------------------------

/* MD/misc/math.hpp:18-33: power by repeated squaring */
__device__ inline double power_n(int n, double x)
{
    if (n == 0) return 1.0;
    double pw = x;
    for (int k = n - 1; k > 0; k >>= 1) {
        if (k & 1) pw *= x;
        x *= x;
    }
    return pw;
}

/* MD/misc/math.hpp:37-44: x^(n/2) from x^2 */
__device__ inline double power_sqrt_n(int n, double x2)
{
    return (n & 1) ? power_n(n, sqrt(x2)) : power_n(n >> 1, x2);
}

/* force of one term = scalar(r^2) * r for every kind; returns the scalar times the factor */
__device__ inline double term_force_scalar(const DTerm &t, double r2)
{
    if (t.cutoff2 != 0.0 && r2 >= t.cutoff2) return 0.0;  // MD/potential/cutoff_potential.hpp:42-46
    double s = 0.0;
    switch (t.kind) {
    case BD_TERM_SOFTCORE: {  // MD/potential/softcore_potential.hpp:49-61
        double const u2 = t.k2 * r2;
        double const v = power_sqrt_n(t.p - 2, u2);
        double const g = 1.0 - v * u2;
        if (g >= 0.0) s = (double) (t.p * t.q) * t.energy * t.k2 * power_n(t.q - 1, g) * v;
        break;
    }
    case BD_TERM_SOFTWELL: {  // MD/potential/softwell_potential.hpp:44-52
        double const u2 = t.k2 * r2;
        double const up_2 = power_sqrt_n(t.p - 2, u2);
        double const v1 = 1.0 + up_2 * u2;
        s = ((double) (-t.p) * t.energy * t.k2) * up_2 / (v1 * v1);
        break;
    }
    case BD_TERM_HARMONIC:  // MD/potential/harmonic_potential.hpp:29-32
        s = -t.energy;
        break;
    case BD_TERM_SPRING: {  // MD/potential/spring_potential.hpp:32-39
        double const r1 = sqrt(r2);
        if (r1 != 0.0) s = t.energy * t.length / r1 - t.energy;
        break;
    }
    case BD_TERM_SEMISPRING: {  // MD/potential/semispring_potential.hpp:37-44
        double const r1 = sqrt(r2);
        if (r1 > t.length) s = t.energy * t.length / r1 - t.energy;
        break;
    }
    default:
        break;
    }
    return t.factor * s;
}

__device__ inline double term_energy(const DTerm &t, double r2)
{
    if (t.cutoff2 != 0.0 && r2 >= t.cutoff2) return 0.0;
    double e = 0.0;
    switch (t.kind) {
    case BD_TERM_SOFTCORE: {  // MD/potential/softcore_potential.hpp:35-47
        double const g = 1.0 - power_sqrt_n(t.p, t.k2 * r2);
        if (g >= 0.0) e = t.energy * power_n(t.q, g);
        break;
    }
    case BD_TERM_SOFTWELL:  // MD/potential/softwell_potential.hpp:35-42
        e = -t.energy / (1.0 + power_sqrt_n(t.p, t.k2 * r2));
        break;
    case BD_TERM_HARMONIC:  // MD/potential/harmonic_potential.hpp:24-27
        e = 0.5 * t.energy * r2;
        break;
    case BD_TERM_SPRING: {  // MD/potential/spring_potential.hpp:26-30
        double const u = sqrt(r2) - t.length;
        e = 0.5 * t.energy * u * u;
        break;
    }
    case BD_TERM_SEMISPRING: {  // MD/potential/semispring_potential.hpp:28-35
        double const u = sqrt(r2) - t.length;
        if (u > 0.0) e = 0.5 * t.energy * u * u;
        break;
    }
    default:
        break;
    }
    return t.factor * e;
}

/* sum_potential / scaled_potential (MD/potential/sum_potential.hpp:32-40) */
__device__ inline double pot_force_scalar(const DTerm *__restrict__ row, int nterms, double r2)
{
    double s = 0.0;
    for (int k = 0; k < nterms; k++) s += term_force_scalar(row[k], r2);
    return s;
}

__device__ inline double pot_energy(const DTerm *__restrict__ row, int nterms, double r2)
{
    double e = 0.0;
    for (int k = 0; k < nterms; k++) e += term_energy(row[k], r2);
    return e;
}

// ---- tile staging ---------------------------------------------------------------------------------
//
// The step kernel advances tiles of 128 slot-consecutive beads.  After the Morton sort the neighbourhood
// of a tile -- every cell any of its beads lists from -- is a few dozen contiguous slot RUNS.  At a list
// build the runs of every tile are written down and the Verlet rows store 16-bit positions inside the
// concatenated runs; at every step a block copies the 16-byte shadows of its runs into shared memory with
// bulk asynchronous copies (no LSU traffic) and the pair walk gathers from there.  A gather of 32
// scattered shadows costs the L1 one wavefront per touched line when it goes to global memory, and a
// quarter of that from shared memory; that gather was what bound the round-1 kernel (75 % of peak
// L1 wavefronts).
constexpr uint32_t STAGE_RUNS = 254;          // runs per tile (adjacent cells merge)
constexpr uint32_t STAGE_CELLS = 512;         // distinct cells in the neighbourhood of a tile
constexpr uint32_t STAGE_HASH = 1024;         // open-addressing table that deduplicates them
constexpr uint32_t STAGE_MAX_BEADS = 65000;   // 16-bit row entries

struct TileRuns {
    uint32_t nruns, total;                    // total = staged beads; row entries are in [0, total], total = padding
    uint32_t start[STAGE_RUNS];               // first slot of run r
    uint32_t off[STAGE_RUNS];                 // its position in the staged array
};
static_assert(sizeof(TileRuns) == 2040, "TileRuns layout");

/* staged position -> slot (the rare consumers of the rows that do not stage: energy, list export) */
__device__ inline uint32_t staged_to_slot(const TileRuns &t, uint32_t local)
{
    uint32_t lo = 0, hi = t.nruns;  // last run with off <= local
    while (hi - lo > 1) {
        uint32_t const mid = (lo + hi) >> 1;
        if (t.off[mid] <= local) lo = mid; else hi = mid;
    }
    return t.start[lo] + (local - t.off[lo]);
}

/* entry k of the row at thread position p, as a slot, in either row format (for the consumers that do not
 * stage: list export, energy).  tile = the tile the position belongs to. */
__device__ inline uint32_t row_entry_slot(const uint4 *nbr4, const TileRuns *tile_runs, uint32_t staged, size_t pstride,
                                          uint32_t p, uint32_t tile, uint32_t k)
{
    if (!staged) {
        uint4 const grp = nbr4[(size_t) (k / ROW_PAD) * pstride + p];
        return k % ROW_PAD == 0 ? grp.x : k % ROW_PAD == 1 ? grp.y : k % ROW_PAD == 2 ? grp.z : grp.w;
    }
    uint2 const grp = reinterpret_cast<const uint2 *>(nbr4)[(size_t) (k / ROW_PAD) * pstride + p];
    uint32_t const word = k % ROW_PAD < 2 ? grp.x : grp.y;
    return staged_to_slot(tile_runs[tile], (k & 1u) ? word >> 16 : word & 0xffffu);
}

// ---- bulk asynchronous copies global -> shared (sm_90+: cp.async.bulk, completion on an mbarrier) ------

__device__ __forceinline__ uint32_t shared_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t arrivals)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(shared_u32(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(shared_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" :: "r"(shared_u32(bar)), "r"(parity) : "memory");
}

/* bytes is a positive multiple of 16; both addresses are 16-byte aligned */
__device__ __forceinline__ void bulk_g2s(void *dst_shared, const void *src_global, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(shared_u32(dst_shared)), "l"(src_global), "r"(bytes), "r"(shared_u32(bar)) : "memory");
}

/* Copy the shadows of a tile's runs into `stage` (all threads of the block call this; `runs` is the
 * shared-memory copy of the tile's table).  One thread announces the byte count, the runs are issued
 * round-robin by the threads, everybody waits for the data.  stage[total] gets the padding entry. */
__device__ __forceinline__ void stage_runs(float4 *stage, const float4 *shadow, const TileRuns &runs, uint64_t *bar,
                                           uint32_t parity, uint32_t pad_slot)
{
    uint32_t const n = runs.nruns, total = runs.total;
    if (threadIdx.x == 0) {
        mbar_expect(bar, total * 16u);
        float const fnan = __uint_as_float(0x7fc00000u);
        stage[total] = make_float4(fnan, fnan, fnan, __uint_as_float(pad_slot));
    }
    for (uint32_t r = threadIdx.x; r < n; r += blockDim.x) {
        uint32_t const end = r + 1 < n ? runs.off[r + 1] : total;
        bulk_g2s(stage + runs.off[r], shadow + runs.start[r], (end - runs.off[r]) * 16u, bar);
    }
    mbar_wait(bar, parity);
}

// ---- cell keys ----------------------------------------------------------------------------------

/* Morton interleave of cell coordinates with per-axis bit counts */
__device__ inline uint32_t morton_key(uint32_t cx, uint32_t cy, uint32_t cz, const int32_t bits[3])
{
    uint32_t key = 0;
    int pos = 0;
    int const maxb = max(bits[0], max(bits[1], bits[2]));
    for (int i = 0; i < maxb; i++) {
        if (i < bits[0]) key |= ((cx >> i) & 1u) << pos++;
        if (i < bits[1]) key |= ((cy >> i) & 1u) << pos++;
        if (i < bits[2]) key |= ((cz >> i) & 1u) << pos++;
    }
    return key;
}

__device__ inline void cell_coords(const DParams &P, const DGrid &G, double x, double y, double z,
                                   int c[3])
{
    double const v[3] = {x, y, z};
#pragma unroll
    for (int d = 0; d < 3; d++) {
        double u;
        if (P.box_kind == BD_BOX_PERIODIC) {
            u = floor_mod_exact(v[d], P.period[d], P.inv_period[d]) * G.inv_cell[d];
        } else {
            u = (v[d] - G.origin[d]) * G.inv_cell[d];
        }
        int ci = (int) u;
        ci = ci < 0 ? 0 : ci;
        ci = ci >= G.n[d] ? G.n[d] - 1 : ci;
        c[d] = ci;
    }
}

// ---- single-precision shadow coordinates --------------------------------------------------------
//
// The step kernel walks ~26 listed neighbours per bead of which ~8 interact.  The walk tests a
// 16-byte single-precision shadow of each neighbour against a widened cutoff; only survivors are
// loaded as 32-byte fp64 records and evaluated exactly, so the fp64 result does not depend on the
// shadow.  Open box: float offsets from the grid origin.  Periodic box: 32-bit fixed-point
// fractions of the period, whose wrapping difference IS the minimum image.

/* .w carries a tag (as bits).  Staged engines: the bead's own slot -- a kernel that reads a shadow from a staged
 * shared-memory copy (bd_step.cuh: tile staging) learns from it where the bead's fp64 record lives.  All other
 * engines: the bead's class, which is all the single-precision pair walk (bd_step.cuh: WALK_F32) needs besides
 * the coordinates. */
__device__ __forceinline__ uint32_t shadow_tag(uint32_t staged, uint32_t slot, uint64_t w)
{
    return staged ? slot : w_type(w);
}

template <int BOX>
__device__ __forceinline__ float4 make_shadow(const DParams &P, const DGrid &G, double x, double y, double z, uint32_t slot)
{
    float4 q;
    if (BOX == BD_BOX_PERIODIC) {
        double const v[3] = {x, y, z};
        uint32_t u[3];
#pragma unroll
        for (int d = 0; d < 3; d++) {
            double t = v[d] * P.inv_period[d];
            t -= floor(t);
            u[d] = (uint32_t) (unsigned long long) (t * 4294967296.0);
        }
        q.x = __uint_as_float(u[0]);
        q.y = __uint_as_float(u[1]);
        q.z = __uint_as_float(u[2]);
    } else {
        q.x = (float) (x - G.origin[0]);
        q.y = (float) (y - G.origin[1]);
        q.z = (float) (z - G.origin[2]);
    }
    q.w = __uint_as_float(slot);
    return q;
}

/* squared shadow distance between two beads */
template <int BOX>
__device__ __forceinline__ float shadow_dist2(const float4 &a, const float4 &b, const float scale[3])
{
    float dx, dy, dz;
    if (BOX == BD_BOX_PERIODIC) {
        dx = (float) (int32_t) (__float_as_uint(a.x) - __float_as_uint(b.x)) * scale[0];
        dy = (float) (int32_t) (__float_as_uint(a.y) - __float_as_uint(b.y)) * scale[1];
        dz = (float) (int32_t) (__float_as_uint(a.z) - __float_as_uint(b.z)) * scale[2];
    } else {
        dx = a.x - b.x;
        dy = a.y - b.y;
        dz = a.z - b.z;
    }
    return fmaf(dx, dx, fmaf(dy, dy, dz * dz));
}

/* the displacement a - b itself, in single precision (WALK_F32) */
template <int BOX>
__device__ __forceinline__ void shadow_delta(const float4 &a, const float4 &b, const float scale[3], float &dx, float &dy, float &dz)
{
    if (BOX == BD_BOX_PERIODIC) {
        dx = (float) (int32_t) (__float_as_uint(a.x) - __float_as_uint(b.x)) * scale[0];
        dy = (float) (int32_t) (__float_as_uint(a.y) - __float_as_uint(b.y)) * scale[1];
        dz = (float) (int32_t) (__float_as_uint(a.z) - __float_as_uint(b.z)) * scale[2];
    } else {
        dx = a.x - b.x;
        dy = a.y - b.y;
        dz = a.z - b.z;
    }
}

/* order-preserving map double -> u64 (for atomicMin/Max bounding boxes) */
__device__ inline unsigned long long encode_ordered(double v)
{
    unsigned long long b = (unsigned long long) __double_as_longlong(v);
    return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ inline double decode_ordered(unsigned long long e)
{
    unsigned long long b = (e & 0x8000000000000000ull) ? (e & 0x7fffffffffffffffull) : ~e;
    return __longlong_as_double((long long) b);
}

// ---- Philox4x32-10 counter-based noise ----------------------------------------------------------

__device__ inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t const hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
        uint32_t const hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
        uint32_t const n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
        c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

/* three standard normals for (seed, step, replica, original bead index): the stream does not
 * depend on the sort order, the launch shape or the number of GPUs.
 *
 * Box-Muller in single precision on the four 32-bit outputs of one Philox call: radii from c0, c1, angles
 * from c2, c3.  What that means against the reference's double ziggurat (MD/basic_types/ziggurat.hpp:262-322),
 * stated here and measured by tests/test_gpu_noise.py (bd_noise_statistics, 3e8 draws):
 *   - values carry a relative rounding of ~1e-7 (a displacement sigma*z of 1e-3 is resolved to 1e-10; the
 *     fp64 positions to 1e-16): invisible to any moment or histogram at any sample size a run can reach;
 *   - the radius uniform has 40 bits (the low byte of the angle words, which the float angle cannot see,
 *     refines c0 / c1), so |z| reaches sqrt(2 ln 2^41) = 7.54; the mass beyond is 5e-14 per draw, less than
 *     one draw in the 1.2e13 of a 10^6-step run of the 64-replica ensemble. */
__device__ inline void philox_normals(uint64_t seed, uint64_t step, uint32_t replica, uint32_t orig,
                                      double z[3])
{
    uint32_t c[4] = {orig, replica, (uint32_t) step, (uint32_t) (step >> 32)};
    philox4x32_10(c, (uint32_t) seed, (uint32_t) (seed >> 32));
    // u = (c + (low byte of the angle word + 1/2) / 256) * 2^-32 in (0, 1): exact in float for small c,
    // which is where the tail lives
    float const f1 = __uint2float_rn(c[2] & 0xffu) * 3.90625e-3f + 1.953125e-3f;
    float const f2 = __uint2float_rn(c[3] & 0xffu) * 3.90625e-3f + 1.953125e-3f;
    float const u1 = (__uint2float_rn(c[0]) + f1) * 2.3283064365386963e-10f;
    float const u2 = (__uint2float_rn(c[1]) + f2) * 2.3283064365386963e-10f;
    // -2 ln u = (-2 ln 2) log2 u on the special-function unit (MUFU.LG2: relative error 2^-22 away from 1,
    // absolute 2^-22 near 1, i.e. radii below 1e-3 are resolved to ~1e-4 -- mass 1e-6, no visible moment);
    // angles from the signed words, theta in [-pi, pi) where MUFU.SIN / COS are accurate to 2^-21
    float const r1 = sqrtf(-1.3862943611198906f * __log2f(fminf(u1, 1.0f)));
    float const r2 = sqrtf(-1.3862943611198906f * __log2f(fminf(u2, 1.0f)));
    float const t1 = (float) (int32_t) c[2] * 1.4629180792671596e-9f;  // pi / 2^31
    float const t2 = (float) (int32_t) c[3] * 1.4629180792671596e-9f;
    z[0] = (double) (r1 * __cosf(t1));
    z[1] = (double) (r1 * __sinf(t1));
    z[2] = (double) (r2 * __cosf(t2));
}

}  // namespace bd
