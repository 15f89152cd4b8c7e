/*
 * bd_step.cuh -- force evaluation and the Brownian update.
 *
 * Replaces, in one pass over cell-sorted beads:
 *   md::neighbor_pairwise_forcefield::compute_force   MD/forcefield/neighbor_pairwise_forcefield.hpp:68-84
 *   md::bonded_pairwise_forcefield::compute_force     MD/forcefield/bonded_pairwise_forcefield.hpp:78-93
 *   md::bonded_triplewise_forcefield::compute_force   MD/forcefield/bonded_triplewise_forcefield.hpp:82-103
 *   sphere / ellipsoid surface forcefields            MD/forcefield/sphere_surface_forcefield.hpp:108-143,
 *                                                     MD/forcefield/ellipsoid_surface_forcefield.hpp:167-211
 *   md::point_source_forcefield::compute_force        MD/forcefield/point_source_forcefield.hpp:78-103
 *   brownian_simulator::simulate_step                 MD/simulation/detail/brownian_simulator.hpp:44-76
 *   neighbor_list::check_consistency                  MD/forcefield/detail/neighbor_list.hpp:101-159
 *
 * One thread owns one bead: it gathers every partner itself (each pair is evaluated from both
 * sides), so no force is ever scattered and nothing is atomic except block-level statistics.
 * Positions are double-buffered: the fused kernel reads pos[cur] and writes pos[1-cur], which is
 * the single HBM round trip of a step.
 */
#pragma once

#include "bd_device.cuh"

namespace bd {

constexpr int STEP_BLOCK = 128;
#ifndef STEP_MIN_BLOCKS
#define STEP_MIN_BLOCKS 8  // resident CTAs per SM the fused kernel is compiled for (register budget)
#endif
constexpr int WALL_STATS = 4;

struct StepArrays {
    const bead4 *pos;
    bead4 *pos_out;
    const bead4 *bpos;
    const float4 *shadow;         // single-precision shadow of pos (pre-filter of the pair walk)
    float4 *shadow_out;           // ... of pos_out
    const uint32_t *slot_of;
    const uint4 *nbr4;            // [cap / ROW_PAD][pstride] Verlet rows by thread position (bd_nlist.cuh)
    const TileRuns *tile_runs;    // [replicas * chunks] staged neighbourhoods (staged rows: uint2 groups of 16-bit positions)
    uint32_t staged, stage_cap;   // row format; shadows a block can stage (dynamic shared memory)
    uint32_t pair_f32;            // pair forces from the single-precision shadows (option pair_precision = 32)
    uint32_t stage_off;           // byte offset of the staged shadows in the dynamic shared memory
    const uint32_t *plen;         // [pstride] row length at a position
    const uint2 *bondrow;         // [bond_deg][pstride] {partner slot, ff | second << 8}
    uint32_t bond_deg, npad, pstride;
    const uint8_t *order;         // [Ntot] thread -> bead offset within its 128-slot tile (bd_nlist.cuh: k_balance)
    const uint32_t *bond_start;   // [n+1] CSR by original index
    const uint2 *bond_adj;        // {partner original index, ff | second<<8}
    const uint2 *bend_pq;         // [bend_deg][pstride] slots of the other two beads of the thread's k-th triple (i < j < k order)
    const double *bend_e;         // [bend_deg][pstride] its bending energy
    const uint8_t *bend_meta;     // [bend_deg][pstride] 1 | role << 1 | ff << 3, 0 = no more triples
    uint32_t bend_deg;
    const uint32_t *bend_start;   // [n+1]
    const uint32_t *bend_adj;     // triple id << 2 | role
    const uint32_t *triples;      // [m][3] original indices
    const double *bend_energy;    // [m]
    const uint8_t *bend_ff;       // [m]
    const DTerm *terms;
    const DParams *params;
    const DWallGeom *wall_geom;   // [replicas][BD_MAX_WALL_FF] per-replica wall geometry (bd_device.cuh)
    DControl *ctl;
    double *wall_partial;         // [replicas][chunks][BD_MAX_WALL_FF][4]
    unsigned long long *inter_partial;  // [replicas*chunks*warps] interacting (bead, neighbour) entries so far
    double *force;                // [Ntot][3] by slot (two-pass path and compute_forces)
    double *wprev;                // [Ntot][3] by original global index (replay mode)
    const double *ext;            // [Ntot][3] by original global index: host-evaluated user forcefields, or null
    unsigned long long *dt_bits;  // [replicas] adaptive: min timestep of the current step
    const double *dt_prev;        // [replicas] adaptive: timestep of the previous step
    const struct PairRow *pair_rows;    // compiled table of the specialised pair program
    const struct BondRow *bond_rows;    // compiled bond tables [used bond ff][ntypes^2], or unused
    uint8_t bond_slot[BD_MAX_BOND_FF];  // bond forcefield -> block of bond_rows
    int32_t bonds_compiled, nbond_rows; // all enabled bond tables are single simple terms
    int32_t any_bend;                   // some bending forcefield exists
    uint32_t ntypes2;             // rows in pair_rows
    int32_t fast_ff;              // the pair forcefield the specialised program evaluates
    int32_t k2_uniform;           // every row of pair_rows has the same k2a and k2b: taken from here
    double k2a, k2b;
    uint32_t n, replicas, ntot, chunks;
};

struct StepScalars {
    double dt, dt_prev;      // fixed-timestep path
    double kT, kT_prev;
    double spacestep;
    unsigned long long seed;
    long long step;          // 1-based index of this step within the run
    const double *replay;    // device, [replicas][n][3] standard normals of this step, or null
    int adaptive;
    int want_stats;          // reduce the wall reaction statistics in this step
    // noise key of local replica r is replica_base + r * replica_stride: an ensemble sharded over
    // GPUs draws the same noise for a replica wherever it runs (ensemble.py)
    uint32_t replica_base, replica_stride;
};

template <typename T>
__device__ inline T ldg(const T *p) { return __ldg(p); }

/* one 32-byte record = one 256-bit access (sm_100: LDG.E.256 / STG.E.256) */
__device__ __forceinline__ bead4 load_bead(const bead4 *p)
{
    bead4 b;
    // .nc: the step kernels only read the buffer they gather from
    asm("ld.global.nc.v4.b64 {%0, %1, %2, %3}, [%4];" : "=d"(b.x), "=d"(b.y), "=d"(b.z), "=l"(b.w) : "l"(p));
    return b;
}

#ifdef STEP_NO_STREAM_LOADS
__device__ __forceinline__ bead4 stream_bead(const bead4 *p) { return load_bead(p); }
#endif

/* read-once streams (Verlet rows, bond rows, build positions): keep them out of the L1, which the gathers need */
#ifndef STEP_NO_STREAM_LOADS
__device__ __forceinline__ uint4 stream_u4(const uint4 *p)
{
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ uint2 stream_u2(const uint2 *p)
{
    uint2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ bead4 stream_bead(const bead4 *p)
{
    bead4 b;
    asm volatile("ld.global.nc.L1::no_allocate.v4.b64 {%0, %1, %2, %3}, [%4];" : "=d"(b.x), "=d"(b.y), "=d"(b.z), "=l"(b.w) : "l"(p));
    return b;
}
#else
__device__ __forceinline__ uint4 stream_u4(const uint4 *p) { return *p; }
__device__ __forceinline__ uint2 stream_u2(const uint2 *p) { return *p; }
#endif

__device__ __forceinline__ void store_bead(bead4 *p, double x, double y, double z, uint64_t w)
{
    asm volatile("st.global.v4.b64 [%0], {%1, %2, %3, %4};" :: "l"(p), "d"(x), "d"(y), "d"(z), "l"(w) : "memory");
}

// ---- compile-time specialisations -----------------------------------------------------------------
//
// The reference builds its pair potential from expression templates, so the compiler sees
// `a*softcore<2,3> + b*softcore<8,3>` as straight-line code (MD/potential/sum_potential.hpp:32-40,
// softcore_potential.hpp:49-61).  The engine gets the same effect by recognising the flattened
// table's program on the host and launching a kernel instantiated for it; every other table runs
// through the run-time interpreter (pot_force_scalar).

enum PairProgram {
    PROG_GENERIC = 0,     // interpreter: any terms, several pair forcefields, target subsets
    PROG_SC23_SC83 = 1,   // f0*softcore<2,3> + f1*softcore<8,3>   (4-sim-ab, 5-sim-genome 100kb, 3-sim-1kb)
    PROG_SC23 = 2,        // f0*softcore<2,3>                      (5-sim-genome 10Mb)
    PROG_COUNT = 3
};

/* one row of a compiled pair table: c = factor * P * Q * energy * k2, k2 = 1 / diameter^2.  In every
 * reference driver the diameters are global (only the factors depend on the bead classes), so the k2 of
 * all rows are equal: the kernel then takes them from the launch parameters and reads 16 bytes per
 * evaluated pair from shared memory instead of 32 (StepArrays::k2_uniform). */
struct __align__(16) PairRow {
    double ca, cb, k2a, k2b;
};

/* one row of a compiled bond table: a single harmonic / spring / semispring term (or none), in the form
 * the kernel evaluates without looking at the kind:
 *     scalar = factor * (r2 > b2 ? kb / |r| - k : 0)
 * spring (MD/potential/spring_potential.hpp:32-39):      kb = K b, b2 = 0    (acts unless r = 0)
 * semispring (semispring_potential.hpp:37-44):           kb = K b, b2 = b^2  (acts beyond b)
 * harmonic (harmonic_potential.hpp:29-32):               kb = 0,   b2 = -1   (always: scalar = -K factor)
 * none:                                                  factor = 0 */
struct __align__(16) BondRow {
    double kb, k, b2, factor;
};

/* bondrow[].y: row of the compiled bond table (bits 0-11, resolved at the list build: bd_nlist.cuh),
 * forcefield (bits 12-14), "this bead is the second of the pair" (bit 15) */
constexpr uint32_t BOND_ROW_MASK = 0xfffu;
__host__ __device__ inline uint32_t bond_word(uint32_t row, uint32_t ff, uint32_t second)
{
    return (row & BOND_ROW_MASK) | (ff << 12) | (second << 15);
}

/* shared memory through 32-bit shared-space addresses: the generic-pointer forms make the compiler rebuild
 * the shared window base inside the pair loops (S2R SR_CgaCtaId + LEA per access in the round-1 SASS) */
__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v)
{
    asm volatile("st.shared.u32 [%0], %1;" :: "r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ double lds_f64(uint32_t addr)  // read-only tables: free to be scheduled
{
    double v;
    asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}

template <int BOX>
__device__ __forceinline__ d3 displacement_t(const DParams &P, double ax, double ay, double az, double bx,
                                             double by, double bz)
{
    d3 r;
    r.x = __dsub_rn(ax, bx);
    r.y = __dsub_rn(ay, by);
    r.z = __dsub_rn(az, bz);
    if (BOX == BD_BOX_PERIODIC) {
        r.x = round_mod_exact(r.x, P.period[0], P.inv_period[0]);
        r.y = round_mod_exact(r.y, P.period[1], P.inv_period[1]);
        r.z = round_mod_exact(r.z, P.period[2], P.inv_period[2]);
    }
    return r;
}

/* softcore<P,3> force scalar without the factor: c * g^2 * (u2)^((P-2)/2), g = 1 - (u2)^(P/2) */
template <int PROG>
__device__ __forceinline__ double compiled_pair_scalar(double ca, double cb, double k2a, double k2b, double r2)
{
    double const ua = k2a * r2;
    double const ga = 1.0 - ua;
    double s = ga >= 0.0 ? ca * ga * ga : 0.0;
    if (PROG == PROG_SC23_SC83) {
        double const ub = k2b * r2;
        double const vb = ub * ub * ub;
        double const gb = 1.0 - vb * ub;
        s += gb >= 0.0 ? cb * gb * gb * vb : 0.0;
    }
    return s;
}

#ifndef STEP_HIT_CAP
#define STEP_HIT_CAP 24
#endif
constexpr int HIT_CAP = STEP_HIT_CAP;  // near-hits buffered per bead between the two phases of the pair walk

extern __shared__ __align__(16) unsigned char dyn_smem[];  // PairRow[ntypes^2] of the compiled program

struct StepShared {
    alignas(16) DParams P;
    alignas(16) DGrid G;
    float filter2;  // widened cutoff^2 of the shadow pre-filter
    uint32_t bond_on, bond_image;  // per bond forcefield (bit f): selected and enabled / uses the box
    // per bead class, for the timestep of this launch (fused kernel): mobility * dt, noise amplitudes
    double mu_dt[BD_MAX_TYPES], sigma[BD_MAX_TYPES], sigma_prev[BD_MAX_TYPES];
    uint32_t hits[HIT_CAP][STEP_BLOCK];
    double wall_red[STEP_BLOCK / 32][BD_MAX_WALL_FF][WALL_STATS];
    // tile staging (bd_device.cuh; staged engines only): whether the tile's shadows are in shared memory.
    // The run table and the staged shadows live in the dynamic shared memory, so that engines that do not
    // stage do not pay for them (the L1 gets what shared memory does not take).
    alignas(8) uint64_t bar;
    uint32_t stage_ok;
};

/* dynamic shared memory of the step kernels: [compiled pair rows][compiled bond rows] then, staged engines only,
 * [TileRuns][staged shadows (stage_cap + 1)] */
__device__ __forceinline__ TileRuns &tile_table(const StepArrays &A)
{
    return *reinterpret_cast<TileRuns *>(dyn_smem + A.stage_off);
}
constexpr uint32_t STAGE_TABLE_BYTES = (uint32_t) ((sizeof(TileRuns) + 15) & ~(size_t) 15);


/* where phase 1 of the pair walk finds the shadows of the listed neighbours */
enum WalkMode {
    WALK_GLOBAL = 0,   // rows of slots, one 16-byte global gather per entry (engines that do not stage)
    WALK_STAGED = 1,   // rows of 16-bit staged positions, gathers from the tile's shared-memory copy
    WALK_SPILLED = 2,  // staged rows of a tile whose neighbourhood does not fit the block's shared memory:
                       // positions are translated back to slots through the run table (rare, slow, correct)
    WALK_F32 = 3       // the mixed-precision production lane (north_star: "SoA float4 positions"): ONE phase, the
                       // displacement is the difference of the 16-byte single-precision shadows (coordinates as
                       // float offsets from the grid origin / 32-bit fractions of the period, class in .w), the
                       // potential is evaluated in double on it and the force accumulated in double; no fp64
                       // record of a neighbour is read.  Not the reference's arithmetic: forces agree to the
                       // rounding of a float coordinate (stated and tested: tests/test_gpu_parity.py)
};

__device__ __forceinline__ float4 lds_f4(uint32_t addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}

/* Non-bonded forces on the bead at slot s from its Verlet row, in two phases:
 *   1. walk the row (groups of ROW_PAD coalesced entries), test the 16-byte shadow of each listed
 *      neighbour against the widened cutoff in single precision and queue the survivors' slots;
 *   2. load the queued neighbours' fp64 records and evaluate them exactly -- the reference's
 *      arithmetic, in row order, so the result does not depend on the shadows at all.
 * About 26 neighbours are listed per bead and 8 interact: phase 1 is cheap per entry and fully
 * convergent, phase 2 runs on a third of the entries. */
template <int BOX, int PROG, int MODE>
__device__ __forceinline__ void pair_walk(const StepArrays &A, StepShared &sh, Select sel, uint32_t s,
                                          uint32_t pos, const bead4 &me, d3 &F, uint32_t &interactions)
{
    const DParams &P = sh.P;
    uint32_t const oi = w_orig(me.w), ti = w_type(me.w);
    int const T = P.ntypes;
    (void) oi;
    const PairRow *mine = reinterpret_cast<const PairRow *>(dyn_smem) + ti * T;
    double const dcut2 = PROG != PROG_GENERIC ? P.pair[A.fast_ff].dcut2 : 0.0;
    float const filter2 = sh.filter2;
    float const scale[3] = {sh.G.shadow_scale[0], sh.G.shadow_scale[1], sh.G.shadow_scale[2]};
    size_t const stride = A.pstride;
    uint32_t const m = (A.plen[pos] + ROW_PAD - 1) & ~(ROW_PAD - 1);  // rows are padded
    float4 const ms = A.shadow[s];
    const uint4 *col = A.nbr4 + pos;                                        // WALK_GLOBAL: four slots per group
    const uint2 *col2 = reinterpret_cast<const uint2 *>(A.nbr4) + pos;      // staged: four 16-bit positions
    uint32_t const stage = smem_addr(dyn_smem) + A.stage_off + STAGE_TABLE_BYTES;  // the tile's staged shadows
    constexpr bool SLOT_ROWS = MODE == WALK_GLOBAL || MODE == WALK_F32;             // rows of slots (uint4 groups)
    uint32_t const pad = SLOT_ROWS ? 0u : tile_table(A).total;                      // staged rows are padded with this position
    // the thread's queue column and its row of the compiled pair table, as shared-space addresses
    constexpr uint32_t QSTEP = STEP_BLOCK * 4u;  // bytes between consecutive entries of one thread's queue
    uint32_t const queue = smem_addr(&sh.hits[0][threadIdx.x]);
    uint32_t const qfull = queue + (HIT_CAP - ROW_PAD) * QSTEP;
    uint32_t const cav = smem_addr(dyn_smem) + ti * (uint32_t) T * 8u;  // uniform tables: ca[T*T] then cb[T*T]

    auto fetch = [&](uint32_t out[ROW_PAD]) {
        if (SLOT_ROWS) {
            uint4 const g = stream_u4(col);
            out[0] = g.x; out[1] = g.y; out[2] = g.z; out[3] = g.w;
        } else {
            uint2 const g = stream_u2(col2);
            out[0] = g.x & 0xffffu; out[1] = g.x >> 16; out[2] = g.y & 0xffffu; out[3] = g.y >> 16;
        }
    };
    if (MODE == WALK_F32) {
        float const dcut2f = __double2float_ru(dcut2 * (1.0 + 1e-6));
        uint32_t jj[ROW_PAD], jjn[ROW_PAD];
        if (m) fetch(jj);
        for (uint32_t kk = 0; kk < m; kk += ROW_PAD) {
            float4 q[ROW_PAD];
#pragma unroll
            for (int u = 0; u < (int) ROW_PAD; u++) q[u] = __ldg(A.shadow + jj[u]);
            col += stride;
            if (kk + ROW_PAD < m) fetch(jjn);
#pragma unroll
            for (int u = 0; u < (int) ROW_PAD; u++) {
                float dx, dy, dz;
                shadow_delta<BOX>(ms, q[u], scale, dx, dy, dz);
                // the cut in single precision first (26 of 34 listed neighbours are out of range): NaN for the
                // sentinel of the open box fails it
                bool in = fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= dcut2f;
                if (BOX == BD_BOX_PERIODIC) in = in && jj[u] < A.ntot;
                double const rx = (double) dx, ry = (double) dy, rz = (double) dz;
                double const r2 = rx * rx + ry * ry + rz * rz;
                if (in && r2 <= dcut2) {
                    uint32_t const at = cav + (__float_as_uint(q[u].w) & 0xffu) * 8u;
                    double sc;
                    if (A.k2_uniform) {
                        sc = compiled_pair_scalar<PROG>(lds_f64(at), PROG == PROG_SC23_SC83 ? lds_f64(at + A.ntypes2 * 8u) : 0.0,
                                                        A.k2a, A.k2b, r2);
                    } else {
                        const PairRow &row = mine[__float_as_uint(q[u].w) & 0xffu];
                        sc = compiled_pair_scalar<PROG>(row.ca, row.cb, row.k2a, row.k2b, r2);
                    }
                    F.x += sc * rx; F.y += sc * ry; F.z += sc * rz;
                    interactions += 1u;
                }
            }
#pragma unroll
            for (int u = 0; u < (int) ROW_PAD; u++) jj[u] = jjn[u];
        }
        return;
    }
    uint32_t k = 0;
    uint32_t j[ROW_PAD], jn[ROW_PAD];
    if (m) fetch(j);
    while (k < m) {
        // ---- phase 1: single-precision filter ------------------------------------------------
        // the entries of the next group are requested together with the shadows of this one, so
        // only the gather itself is on the dependent path
        uint32_t qp = queue;
        while (k < m && qp <= qfull) {
            float4 q[ROW_PAD];
#pragma unroll
            for (int u = 0; u < (int) ROW_PAD; u++) {
                if (MODE == WALK_GLOBAL) q[u] = __ldg(A.shadow + j[u]);
                else if (MODE == WALK_STAGED) q[u] = lds_f4(stage + j[u] * 16u);
                else q[u] = __ldg(A.shadow + (j[u] == pad ? A.ntot : staged_to_slot(tile_table(A), j[u])));
            }
            col += stride;
            col2 += stride;
            k += ROW_PAD;
            if (k < m) fetch(jn);
#pragma unroll
            for (int u = 0; u < (int) ROW_PAD; u++) {
                bool near = shadow_dist2<BOX>(ms, q[u], scale) <= filter2;
                if (BOX == BD_BOX_PERIODIC) near = near && (MODE == WALK_GLOBAL ? j[u] < A.ntot : j[u] != pad);  // fixed point has no NaN
                if (near) {
                    sts_u32(qp, MODE == WALK_GLOBAL ? j[u] : __float_as_uint(q[u].w));  // the neighbour's slot
                    qp += QSTEP;
                }
            }
#pragma unroll
            for (int u = 0; u < (int) ROW_PAD; u++) j[u] = jn[u];
        }
        if (((qp - queue) / QSTEP) & 1u) sts_u32(qp, A.ntot);  // phase 2 takes the queue two at a time

        // ---- phase 2: exact evaluation -------------------------------------------------------
        for (uint32_t qa = queue; qa < qp; qa += 2 * QSTEP) {
            uint32_t const j0 = lds_u32(qa), j1 = lds_u32(qa + QSTEP);
            bead4 const o0 = load_bead(A.pos + j0);
            bead4 const o1 = load_bead(A.pos + j1);
#pragma unroll
            for (int u = 0; u < 2; u++) {
                const bead4 &o = u == 0 ? o0 : o1;
                d3 const r = displacement_t<BOX>(P, me.x, me.y, me.z, o.x, o.y, o.z);
                double const r2 = sqnorm_exact(r.x, r.y, r.z);  // NaN for the sentinel: every test fails
                if (PROG != PROG_GENERIC) {
                    if (r2 <= dcut2) {
                        double sc;
                        if (A.k2_uniform) {
                            // uniform tables are stored as two arrays of doubles, ca[T*T] then cb[T*T]: 8-byte
                            // reads of 16 or fewer distinct rows never collide on a shared-memory bank
                            uint32_t const at = cav + w_type(o.w) * 8u;
                            sc = compiled_pair_scalar<PROG>(lds_f64(at), PROG == PROG_SC23_SC83 ? lds_f64(at + A.ntypes2 * 8u) : 0.0,
                                                            A.k2a, A.k2b, r2);
                        } else {
                            const PairRow &row = mine[w_type(o.w)];
                            sc = compiled_pair_scalar<PROG>(row.ca, row.cb, row.k2a, row.k2b, r2);
                        }
                        F.x += sc * r.x; F.y += sc * r.y; F.z += sc * r.z;
                        interactions += 1u;  // every pair is seen from both ends: halved when read out
                    }
                } else {
                    uint32_t const oj = w_orig(o.w), tj = w_type(o.w);
                    int const cell = oi < oj ? (int) (ti * T + tj) : (int) (tj * T + ti);
                    bool hit = false;
#pragma unroll
                    for (int f = 0; f < BD_MAX_PAIR_FF; f++) {
                        const DPair &pf = P.pair[f];
                        if (!ff_on(sel, BD_FF_PAIR, f, pf.enabled) || !(r2 <= pf.dcut2)) continue;
                        if (pf.has_targets && !((w_pair_mask(me.w) >> f) & (w_pair_mask(o.w) >> f) & 1u)) continue;
                        double const sc = pot_force_scalar(A.terms + pf.term_off + cell * pf.nterms, pf.nterms, r2);
                        F.x += sc * r.x; F.y += sc * r.y; F.z += sc * r.z;
                        hit = true;
                    }
                    interactions += hit ? 1u : 0u;
                }
            }
        }
    }
}

template <int BOX, int PROG>
__device__ __forceinline__ void pair_forces(const StepArrays &A, StepShared &sh, Select sel, uint32_t s,
                                            uint32_t pos, const bead4 &me, d3 &F, uint32_t &interactions)
{
    if (PROG == PROG_GENERIC) {
        bool any_pair = false;
        for (int f = 0; f < BD_MAX_PAIR_FF; f++) any_pair |= ff_on(sel, BD_FF_PAIR, f, sh.P.pair[f].enabled);
        if (!any_pair) return;
    }
    // block-uniform choice
    if (PROG != PROG_GENERIC && A.pair_f32) pair_walk<BOX, PROG, WALK_F32>(A, sh, sel, s, pos, me, F, interactions);
    else if (!A.staged) pair_walk<BOX, PROG, WALK_GLOBAL>(A, sh, sel, s, pos, me, F, interactions);
    else if (sh.stage_ok) pair_walk<BOX, PROG, WALK_STAGED>(A, sh, sel, s, pos, me, F, interactions);
    else pair_walk<BOX, PROG, WALK_SPILLED>(A, sh, sel, s, pos, me, F, interactions);
}

/* Bring the tile's neighbourhood into shared memory (staged engines; every thread of the block calls this
 * before block_forces).  `parity` is the block's count of stagings so far, modulo 2. */
__device__ __forceinline__ void stage_tile(const StepArrays &A, StepShared &sh, uint32_t replica, uint32_t chunk,
                                           uint32_t &parity)
{
    if (!A.staged) return;
    __syncthreads();  // the block's previous tile is done with the table and the staged copy
    const uint32_t *src = reinterpret_cast<const uint32_t *>(A.tile_runs + (size_t) replica * A.chunks + chunk);
    TileRuns &runs = tile_table(A);
    uint32_t *dst = reinterpret_cast<uint32_t *>(&runs);
    uint32_t const nruns = src[0];
    for (uint32_t i = threadIdx.x; i < 2u + STAGE_RUNS + nruns; i += blockDim.x)
        if (i < 2u + nruns || i >= 2u + STAGE_RUNS) dst[i] = src[i];
    __syncthreads();
    bool const ok = runs.total <= A.stage_cap;
    if (threadIdx.x == 0) sh.stage_ok = ok;
    if (ok) {
        stage_runs(reinterpret_cast<float4 *>(dyn_smem + A.stage_off + STAGE_TABLE_BYTES), A.shadow, runs, &sh.bar, parity, A.ntot);
        parity ^= 1u;
    }
    __syncthreads();
}

/* bonded pairs, bending triples and point sources acting on one bead */
__device__ __forceinline__ void bonded_forces(const StepArrays &A, const StepShared &sh, Select sel,
                                              uint32_t replica, uint32_t pos, const bead4 &me, d3 &F)
{
    const DParams &P = sh.P;
    uint32_t const oi = w_orig(me.w), ti = w_type(me.w);
    int const T = P.ntypes;
    uint32_t const base = replica * A.n;

    // bonded pairs (no minimum image: bonded_pairwise_forcefield.hpp:64); the partners were resolved
    // to slots, and their rows of the compiled table chosen, at the last rebuild (bd_nlist.cuh: k_balance)
    {
        uint32_t const brows = smem_addr(dyn_smem) + A.ntypes2 * (uint32_t) sizeof(PairRow);
        uint32_t const on_mask = sh.bond_on, image_mask = sh.bond_image;
        // software pipeline: the next bond's row entry and partner record are requested while this
        // bond is evaluated (the loop was bound by these two dependent loads)
        const uint2 *col = A.bondrow + pos;
        uint2 adj_next = A.bond_deg ? stream_u2(col) : make_uint2(0xffffffffu, 0u);
        bead4 o_next = me;
        if (adj_next.x != 0xffffffffu) o_next = load_bead(A.pos + adj_next.x);
        for (uint32_t k = 0; k < A.bond_deg; k++) {
            uint2 const adj = adj_next;
            if (adj.x == 0xffffffffu) break;
            bead4 const o = o_next;
            col += A.pstride;
            adj_next = k + 1 < A.bond_deg ? stream_u2(col) : make_uint2(0xffffffffu, 0u);
            if (adj_next.x != 0xffffffffu) o_next = load_bead(A.pos + adj_next.x);
            uint32_t const f = (adj.y >> 12) & 7u;
            if (!((on_mask >> f) & 1u)) continue;
            // the force on this bead is scalar * (x_me - x_partner) whichever end it is; only the
            // table row depends on the order (type_first, type_second)
            double rx = me.x - o.x, ry = me.y - o.y, rz = me.z - o.z;
            if ((image_mask >> f) & 1u) {  // a dynamic pair list of glue_forcefield's kind (glue_forcefield.cpp:40)
                d3 const r = displacement(P, me.x, me.y, me.z, o.x, o.y, o.z);
                rx = r.x; ry = r.y; rz = r.z;
            }
            double const r2 = rx * rx + ry * ry + rz * rz;
            double sc;
            if (A.bonds_compiled) {
                // single-term tables compiled on the host (BondRow): K b / |r| - K through rsqrt
                uint32_t const at = brows + (adj.y & BOND_ROW_MASK) * (uint32_t) sizeof(BondRow);
                double const kb = lds_f64(at), kk = lds_f64(at + 8), b2 = lds_f64(at + 16);
                double v = 0.0;
                if (r2 > b2) {
                    v = -kk;
                    if (kb != 0.0) v = kb * rsqrt(r2) - kk;
                }
                sc = lds_f64(at + 24) * v;
            } else {
                const DBond &bf = P.bond[f];
                uint32_t const tj = w_type(o.w);
                bool const second = (adj.y >> 15) & 1u;
                int const cell = second ? (int) (tj * T + ti) : (int) (ti * T + tj);
                sc = pot_force_scalar(A.terms + bf.term_off + cell * bf.nterms, bf.nterms, r2);
            }
            F.x += sc * rx; F.y += sc * ry; F.z += sc * rz;
        }
    }

    // bending triples (cosine_bending_potential.hpp:43-60)
    // The triples of the bead at this thread position were resolved to slots at the last rebuild (bd_nlist.cuh:
    // fill_bend_rows): three coalesced row reads and two record gathers per triple, where walking the CSR by
    // original index cost ten dependent gathers (adjacency, forcefield, energy, three indices, two slot lookups,
    // two records) -- 40 % of the gathers of the 1kb chromatin model.
    if (A.any_bend) {
        for (uint32_t k = 0; k < A.bend_deg; k++) {
            size_t const at = (size_t) k * A.pstride + pos;
            uint32_t const meta = A.bend_meta[at];
            if (!(meta & 1u)) break;
            uint32_t const role = (meta >> 1) & 3u;
            int const f = (int) (meta >> 3);
            if (!ff_on(sel, BD_FF_BEND, f, P.bend[f].enabled)) continue;
            double const E = A.bend_e[at];
            uint2 const pq = stream_u2(A.bend_pq + at);
            bead4 const rp = load_bead(A.pos + pq.x), rq = load_bead(A.pos + pq.y);
            bead4 const pi = role == 0 ? me : rp;
            bead4 const pj = role == 1 ? me : (role == 0 ? rp : rq);
            bead4 const pk = role == 2 ? me : rq;
            double const ax = pi.x - pj.x, ay = pi.y - pj.y, az = pi.z - pj.z;  // rij
            double const bx = pj.x - pk.x, by = pj.y - pk.y, bz = pj.z - pk.z;  // rjk
            double const dij = ax * ax + ay * ay + az * az, djk = bx * bx + by * by + bz * bz;
            double const dd = sqrt(dij * djk);
            if (dd == 0.0) continue;
            double const dot = ax * bx + ay * by + az * bz;
            double const ed = E / dd, ca = dot / dij, cb = dot / djk;
            double const fijx = ed * (bx - ca * ax), fijy = ed * (by - ca * ay), fijz = ed * (bz - ca * az);
            double const fjkx = ed * (ax - cb * bx), fjky = ed * (ay - cb * by), fjkz = ed * (az - cb * bz);
            if (role == 0) { F.x += fijx; F.y += fijy; F.z += fijz; }
            else if (role == 1) { F.x += fjkx - fijx; F.y += fjky - fijy; F.z += fjkz - fijz; }
            else { F.x -= fjkx; F.y -= fjky; F.z -= fjkz; }
        }
    }

    // point sources (point_source_forcefield.hpp:78-103)
#pragma unroll 1
    for (int f = 0; f < BD_MAX_POINT_FF; f++) {
        const DPoint &pf = P.point[f];
        if (!ff_on(sel, BD_FF_POINT, f, pf.enabled)) continue;
        if (pf.has_targets && !((w_point_mask(me.w) >> f) & 1u)) continue;
        double const rx = me.x - pf.source[0], ry = me.y - pf.source[1], rz = me.z - pf.source[2];
        double const sc = pot_force_scalar(A.terms + pf.term_off + ti * pf.nterms, pf.nterms,
                                           rx * rx + ry * ry + rz * rz);
        F.x += sc * rx; F.y += sc * ry; F.z += sc * rz;
    }
}

/* one wall forcefield on one bead: adds the force, returns the reaction statistics in st[4] */
__device__ __forceinline__ void wall_force(const StepArrays &A, const DWall &wf, const bead4 &me, d3 &F,
                                           double st[WALL_STATS])
{
    st[0] = st[1] = st[2] = st[3] = 0.0;
    double const rx = me.x - wf.center[0], ry = me.y - wf.center[1], rz = me.z - wf.center[2];
    const DTerm *row = A.terms + wf.term_off + w_type(me.w) * wf.nterms;
    if (wf.shape == BD_WALL_SPHERE) {
        // sphere_surface_forcefield.hpp:108-143
        double const radius = wf.semi[0];
        double const r2 = rx * rx + ry * ry + rz * rz;
        if (r2 == 0.0) return;
        bool const inside = r2 < radius * radius;
        if (inside ? wf.direction != BD_WALL_INWARD : wf.direction != BD_WALL_OUTWARD) return;
        if (wf.range > 0.0) {
            // farther from the surface than any term reaches: force and statistics are exactly zero
            double const reach = wf.range * (1.0 + 1e-6);
            double const lim = inside ? radius - reach : radius + reach;
            if (inside ? (lim > 0.0 && r2 < lim * lim) : r2 > lim * lim) return;
        }
        double const r1 = sqrt(r2);
        double const scale = radius / r1;
        double const sx = rx - scale * rx, sy = ry - scale * ry, sz = rz - scale * rz;
        double const sc = pot_force_scalar(row, wf.nterms, sx * sx + sy * sy + sz * sz);
        double const fx = sc * sx, fy = sc * sy, fz = sc * sz;
        double const fr = fx * rx + fy * ry + fz * rz;
        double const pr = fr / r2;  // project(force, r)
        F.x += fx + scale * (pr * rx - fx);
        F.y += fy + scale * (pr * ry - fy);
        F.z += fz + scale * (pr * rz - fz);
        st[0] = -fr / r1;
    } else {
        // ellipsoid_surface_forcefield.hpp:54-99 (evaluate_point), :167-211
        double const qx = wf.quad[0], qy = wf.quad[1], qz = wf.quad[2];  // 1 / semiaxis^2 (host)
        double const dx = qx * rx, dy = qy * ry, dz = qz * rz;  // dual
        double const d2 = dx * dx + dy * dy + dz * dz;
        if (d2 == 0.0) return;
        double const implicit = dx * rx + dy * ry + dz * rz - 1.0;
        bool const inside = implicit < 0.0;
        if (inside ? wf.direction != BD_WALL_INWARD : wf.direction != BD_WALL_OUTWARD) return;
        // delta^2 = implicit^2 / (4 d2): farther from the surface than any term reaches (with a margin for
        // the rounding of this estimate), force and statistics are exactly zero
        if (wf.range > 0.0 && implicit * implicit > 4.0 * d2 * wf.range * wf.range * (1.0 + 1e-6)) return;
        double const scale = implicit / (2.0 * d2);
        double const ex = scale * dx, ey = scale * dy, ez = scale * dz;  // delta
        double const e2 = ex * ex + ey * ey + ez * ez;
        double const sc = pot_force_scalar(row, wf.nterms, e2);
        double const bx = sc * ex, by = sc * ey, bz = sc * ez;  // basic force
        double const pr = e2 != 0.0 ? (bx * ex + by * ey + bz * ez) / e2 : 0.0;
        double const ix = pr * ex, iy = pr * ey, iz = pr * ez;  // iso = project(basic, delta)
        double const fx = ix + ((bx - ix) - ix) * (scale * qx);
        double const fy = iy + ((by - iy) - iy) * (scale * qy);
        double const fz = iz + ((bz - iz) - iz) * (scale * qz);
        F.x += fx; F.y += fy; F.z += fz;
        double const en = e2 != 0.0 ? (inside ? -1.0 : 1.0) / sqrt(e2) : 0.0;
        double const nx = en * ex, ny = en * ey, nz = en * ez;  // outward normal
        st[0] = -(fx * nx + fy * ny + fz * nz);
        st[1] = -(nx > 0 ? fx : -fx);
        st[2] = -(ny > 0 ? fy : -fy);
        st[3] = -(nz > 0 ? fz : -fz);
    }
}

/* block-wide copy in 16-byte pieces (both sides 16-byte aligned, size a multiple of 16) */
__device__ __forceinline__ void copy16(void *dst, const void *src, uint32_t bytes)
{
    for (uint32_t i = threadIdx.x; i < bytes / 16u; i += blockDim.x) ((uint4 *) dst)[i] = ((const uint4 *) src)[i];
}

/* this replica's wall geometry over the shared copy of the parameters (call between two barriers) */
__device__ __forceinline__ void load_replica_walls(DParams &P, const DWallGeom *geom, uint32_t replica)
{
    for (uint32_t i = threadIdx.x; i < BD_MAX_WALL_FF * WALL_GEOM_DOUBLES; i += blockDim.x) {
        uint32_t const f = i / WALL_GEOM_DOUBLES, c = i % WALL_GEOM_DOUBLES;
        if (!P.wall[f].enabled) continue;
        (&P.wall[f].center[0])[c] = reinterpret_cast<const double *>(geom + (size_t) replica * BD_MAX_WALL_FF + f)[c];
    }
}

template <int PROG>
__device__ __forceinline__ void load_shared(StepShared &sh, const StepArrays &A, Select sel)
{
    static_assert(sizeof(DParams) % 16 == 0 && sizeof(DGrid) % 16 == 0, "copied in 16-byte pieces");
    if (threadIdx.x == 0) mbar_init(&sh.bar, 1);
    copy16(&sh.P, A.params, sizeof(DParams));
    copy16(&sh.G, &A.ctl->grid, sizeof(DGrid));
    if (PROG != PROG_GENERIC) copy16(dyn_smem, A.pair_rows, A.ntypes2 * (uint32_t) sizeof(PairRow));
    if (A.bonds_compiled)
        copy16(dyn_smem + A.ntypes2 * sizeof(PairRow), A.bond_rows, (uint32_t) A.nbond_rows * (uint32_t) sizeof(BondRow));
    __syncthreads();
    load_replica_walls(sh.P, A.wall_geom, blockIdx.y);
    if (threadIdx.x == 0) {
        // widened cutoff of the pre-filter: r <= dcut implies r_shadow <= dcut + shadow_eps / 4, and the
        // factor covers the single-precision rounding of the squared distance itself
        double dcut = 0.0;
        for (int f = 0; f < BD_MAX_PAIR_FF; f++)
            if (ff_on(sel, BD_FF_PAIR, f, sh.P.pair[f].enabled)) dcut = fmax(dcut, sh.P.pair[f].dcut);
        double const wide = dcut + sh.G.shadow_eps;
        sh.filter2 = __double2float_ru(wide * wide * (1.0 + 4e-6));
        uint32_t on = 0, image = 0;
        for (int f = 0; f < BD_MAX_BOND_FF; f++) {
            if (ff_on(sel, BD_FF_BOND, f, sh.P.bond[f].enabled)) on |= 1u << f;
            if (sh.P.bond[f].min_image) image |= 1u << f;
        }
        sh.bond_on = on;
        sh.bond_image = image;
    }
    __syncthreads();
}

/* all selected forces on one bead; every thread of the block must call this (walls reduce their
 * statistics across the block), `live` masks the threads past the end of the replica */
template <int BOX, int PROG>
__device__ __forceinline__ d3 block_forces(const StepArrays &A, StepShared &sh, Select sel, bool live,
                                           bool want_stats, uint32_t s, uint32_t replica, uint32_t chunk,
                                           const bead4 &me, uint32_t &interactions)
{
    const DParams &P = sh.P;
    d3 F = {0.0, 0.0, 0.0};
    uint32_t const pos = replica * A.npad + chunk * STEP_BLOCK + threadIdx.x;
    if (live) {
        pair_forces<BOX, PROG>(A, sh, sel, s, pos, me, F, interactions);
        bonded_forces(A, sh, sel, replica, pos, me, F);
        if (A.ext && sel.ff_class == BD_FF_ALL) {
            const double *e = A.ext + 3 * ((size_t) replica * A.n + w_orig(me.w));
            F.x += e[0]; F.y += e[1]; F.z += e[2];
        }
    }
    bool any_wall = false;
#pragma unroll 1
    for (int f = 0; f < BD_MAX_WALL_FF; f++) {
        const DWall &wf = P.wall[f];
        if (!ff_on(sel, BD_FF_WALL, f, wf.enabled)) continue;  // block-uniform
        any_wall = true;
        double st[WALL_STATS] = {0.0, 0.0, 0.0, 0.0};
        if (live) wall_force(A, wf, me, F, st);
        if (!want_stats) continue;  // block-uniform: the last step of a call, or every step with wall relaxation
        // most warps are nowhere near the wall: all their statistics are exactly zero and need no reduction
        bool const touched = st[0] != 0.0 || st[1] != 0.0 || st[2] != 0.0 || st[3] != 0.0;
        if (__any_sync(0xffffffffu, touched)) {
#pragma unroll
            for (int c = 0; c < WALL_STATS; c++)
                for (int o = 16; o > 0; o >>= 1) st[c] += __shfl_xor_sync(0xffffffffu, st[c], o);
        }
        if ((threadIdx.x & 31) == 0) {
#pragma unroll
            for (int c = 0; c < WALL_STATS; c++) sh.wall_red[threadIdx.x >> 5][f][c] = st[c];
        }
    }
    if (any_wall && want_stats) {
        __syncthreads();
        if (threadIdx.x < BD_MAX_WALL_FF * WALL_STATS) {
            int const f = threadIdx.x / WALL_STATS, c = threadIdx.x % WALL_STATS;
            if (ff_on(sel, BD_FF_WALL, f, P.wall[f].enabled)) {
                double t = 0.0;
                for (int w = 0; w < STEP_BLOCK / 32; w++) t += sh.wall_red[w][f][c];
                A.wall_partial[(((size_t) replica * A.chunks + chunk) * BD_MAX_WALL_FF + f) * WALL_STATS + c] = t;
            }
        }
        __syncthreads();  // the fused kernel comes back with the block's next tile
    }
    return F;
}

__device__ inline void load_params(DParams &P, const DParams *src)
{
    for (uint32_t i = threadIdx.x; i < sizeof(DParams) / 4; i += blockDim.x)
        ((uint32_t *) &P)[i] = ((const uint32_t *) src)[i];
    __syncthreads();
}


/* MD/simulation/detail/brownian_simulator.hpp:57-75 for one bead; returns the squared distance
 * from the bead's position at the last list build (neighbor_list.hpp:138-157).
 * mu_dt = timestep * mobility, sigma = sqrt(2 kT mu_dt), sigma_prev the same for the previous step
 * (brownian_simulator.hpp:57-62): per bead class, computed by the caller. */
template <int BOX>
__device__ __forceinline__ double integrate_bead(const StepArrays &A, const DParams &P, const DGrid &G,
                                                 const StepScalars &S,
                                                 uint32_t s, uint32_t replica, const bead4 &me, d3 F,
                                                 double mu_dt, double sigma, double sigma_prev)
{
    uint32_t const oi = w_orig(me.w);
    double wx, wy, wz, px, py, pz;
    size_t const g = (size_t) replica * A.n + oi;
    if (S.replay) {
        wx = sigma * S.replay[3 * g]; wy = sigma * S.replay[3 * g + 1]; wz = sigma * S.replay[3 * g + 2];
        px = A.wprev[3 * g]; py = A.wprev[3 * g + 1]; pz = A.wprev[3 * g + 2];
        A.wprev[3 * g] = wx; A.wprev[3 * g + 1] = wy; A.wprev[3 * g + 2] = wz;
    } else {
        double z[3];
        philox_normals(S.seed, (uint64_t) S.step, S.replica_base + replica * S.replica_stride, oi, z);
        wx = sigma * z[0]; wy = sigma * z[1]; wz = sigma * z[2];
        if (S.step > 1) {
            philox_normals(S.seed, (uint64_t) (S.step - 1), S.replica_base + replica * S.replica_stride, oi, z);
            px = sigma_prev * z[0]; py = sigma_prev * z[1]; pz = sigma_prev * z[2];
        } else {
            px = py = pz = 0.0;
        }
    }
    double x = me.x + mu_dt * F.x, y = me.y + mu_dt * F.y, z2 = me.z + mu_dt * F.z;
    x += 0.5 * (wx + px); y += 0.5 * (wy + py); z2 += 0.5 * (wz + pz);
    store_bead(A.pos_out + s, x, y, z2, me.w);
    A.shadow_out[s] = make_shadow<BOX>(P, G, x, y, z2, shadow_tag(A.staged, s, me.w));

    bead4 const b0 = stream_bead(A.bpos + s);
    d3 const d = displacement_t<BOX>(P, x, y, z2, b0.x, b0.y, b0.z);
    return sqnorm_exact(d.x, d.y, d.z);
}

/* a ghost keeps the position its owner sent: copy it to the other buffer and report how far it is
 * from where it was at the last list build */
template <int BOX>
__device__ __forceinline__ double carry_ghost(const StepArrays &A, const DParams &P, const DGrid &G, uint32_t s,
                                              const bead4 &me)
{
    store_bead(A.pos_out + s, me.x, me.y, me.z, me.w);
    A.shadow_out[s] = make_shadow<BOX>(P, G, me.x, me.y, me.z, shadow_tag(A.staged, s, me.w));
    bead4 const b0 = stream_bead(A.bpos + s);
    d3 const d = displacement_t<BOX>(P, me.x, me.y, me.z, b0.x, b0.y, b0.z);
    return sqnorm_exact(d.x, d.y, d.z);
}

/* fold the block's largest displacement^2 into the control block; the rebuild decision itself is
 * taken by k_rebuild_begin before the next force evaluation (neighbor_list.hpp:132-158) */
__device__ __forceinline__ void publish_displacement(const StepArrays &A, double moved2, double known)
{
    for (int o = 16; o > 0; o >>= 1) moved2 = fmax(moved2, __shfl_xor_sync(0xffffffffu, moved2, o));
    if ((threadIdx.x & 31) == 0 && moved2 > known)
        atomicMax(&A.ctl->max_disp2_bits, (unsigned long long) __double_as_longlong(moved2));
}

/* the hot kernel: forces + Brownian update + skin test, fixed timestep.
 * One block advances STEP_TILES consecutive 128-bead tiles of one replica, so that the per-block set-up
 * (parameter block, this replica's walls, compiled tables, noise amplitudes per class: 7 % of the
 * instructions of the one-tile-per-block kernel) is paid once per STEP_TILES * 128 beads. */
#ifndef STEP_TILES_PER_BLOCK
#define STEP_TILES_PER_BLOCK 1
#endif
constexpr int STEP_TILES = STEP_TILES_PER_BLOCK;

template <int BOX, int PROG>
__global__ void __launch_bounds__(STEP_BLOCK, STEP_MIN_BLOCKS) k_step_fused(StepArrays A, StepScalars S)
{
    __shared__ StepShared sh;
    Select const all = {BD_FF_ALL, -1};
    load_shared<PROG>(sh, A, all);
    uint32_t const replica = blockIdx.y;
    if (threadIdx.x < (uint32_t) sh.P.ntypes) {
        // brownian_simulator.hpp:57-62, once per bead class instead of once per bead
        double const mu = sh.P.mobility[threadIdx.x];
        double const mu_dt = S.dt * mu;
        sh.mu_dt[threadIdx.x] = mu_dt;
        sh.sigma[threadIdx.x] = sqrt(2.0 * S.kT * mu_dt);
        sh.sigma_prev[threadIdx.x] = sqrt(2.0 * S.kT_prev * (S.dt_prev * mu));
    }
    __syncthreads();
    double const known = __longlong_as_double((long long) A.ctl->max_disp2_bits);
    double moved2 = 0.0;
    uint32_t parity = 0;
    uint32_t const last = min((blockIdx.x + 1) * STEP_TILES, A.chunks);
    for (uint32_t chunk = blockIdx.x * STEP_TILES; chunk < last; chunk++) {
        stage_tile(A, sh, replica, chunk, parity);
        bool const live = chunk * STEP_BLOCK + threadIdx.x < A.n;
        // threads take the tile's beads in order of row length, so that the rows of a warp end together
        uint32_t const s = replica * A.n
            + (live ? chunk * STEP_BLOCK + A.order[replica * A.n + chunk * STEP_BLOCK + threadIdx.x] : 0u);
        uint32_t inter = 0;
        bead4 const me = load_bead(A.pos + s);
        bool const active = live && !w_ghost(me.w);
        d3 const F = block_forces<BOX, PROG>(A, sh, all, active, S.want_stats != 0, s, replica, chunk, me, inter);
        uint32_t const t = w_type(me.w);
        if (active) moved2 = fmax(moved2, integrate_bead<BOX>(A, sh.P, sh.G, S, s, replica, me, F, sh.mu_dt[t], sh.sigma[t],
                                                             sh.sigma_prev[t]));
        else if (live) moved2 = fmax(moved2, carry_ghost<BOX>(A, sh.P, sh.G, s, me));
        for (int o = 16; o > 0; o >>= 1) inter += __shfl_xor_sync(0xffffffffu, inter, o);
        if ((threadIdx.x & 31) == 0)  // one private counter per warp: no barrier, no atomics
            A.inter_partial[((size_t) replica * A.chunks + chunk) * (STEP_BLOCK / 32) + (threadIdx.x >> 5)] += inter;
    }
    publish_displacement(A, moved2, known);
}

/* forces only (compute_forces, and the first pass of an adaptive-timestep step) */
template <int BOX, int PROG>
__global__ void __launch_bounds__(STEP_BLOCK) k_force(StepArrays A, StepScalars S, Select sel)
{
    __shared__ StepShared sh;
    load_shared<PROG>(sh, A, sel);
    uint32_t const replica = blockIdx.y, chunk = blockIdx.x;
    uint32_t parity = 0;
    stage_tile(A, sh, replica, chunk, parity);
    bool const live = chunk * STEP_BLOCK + threadIdx.x < A.n;
    uint32_t const s = replica * A.n
        + (live ? chunk * STEP_BLOCK + A.order[replica * A.n + chunk * STEP_BLOCK + threadIdx.x] : 0u);
    uint32_t inter = 0;
    double dt = S.dt;
    bead4 const me = load_bead(A.pos + s);
    bool const active = live && !w_ghost(me.w);
    d3 const F = block_forces<BOX, PROG>(A, sh, sel, active, true, s, replica, chunk, me, inter);
    if (live) {
        A.force[3 * (size_t) s] = F.x;
        A.force[3 * (size_t) s + 1] = F.y;
        A.force[3 * (size_t) s + 2] = F.z;
        if (S.adaptive && active) {
            // brownian_timestepper.hpp:72-91,109-136
            double const mu = sh.P.mobility[w_type(me.w)];
            double const f2 = F.x * F.x + F.y * F.y + F.z * F.z;
            if (mu != 0.0 || f2 != 0.0) {
                double const a = mu * mu * f2, b = 2.55 * mu * S.kT, c = S.spacestep * S.spacestep;
                dt = (a * c < 1e-6 * b * b) ? c / b : (-b + sqrt(b * b + 4.0 * a * c)) / (2.0 * a);
            }
            dt = fmin(dt, S.dt);
        }
    }
    if (S.adaptive) {
        for (int o = 16; o > 0; o >>= 1) dt = fmin(dt, __shfl_xor_sync(0xffffffffu, dt, o));
        if ((threadIdx.x & 31) == 0)
            atomicMin(A.dt_bits + replica, (unsigned long long) __double_as_longlong(dt));
    }
}

/* second pass of an adaptive step: update from the stored forces */
template <int BOX>
__global__ void __launch_bounds__(STEP_BLOCK) k_integrate(StepArrays A, StepScalars S)
{
    __shared__ DParams P;
    __shared__ DGrid G;
    for (uint32_t i = threadIdx.x; i < sizeof(DGrid) / 4; i += blockDim.x)
        ((uint32_t *) &G)[i] = ((const uint32_t *) &A.ctl->grid)[i];
    load_params(P, A.params);
    uint32_t const replica = blockIdx.y;
    uint32_t const local = blockIdx.x * STEP_BLOCK + threadIdx.x;
    double const known = __longlong_as_double((long long) A.ctl->max_disp2_bits);
    double moved2 = 0.0;
    if (local < A.n) {
        uint32_t const s = replica * A.n + local;
        bead4 const me = load_bead(A.pos + s);
        d3 const F = {A.force[3 * (size_t) s], A.force[3 * (size_t) s + 1], A.force[3 * (size_t) s + 2]};
        double const dt = __longlong_as_double((long long) A.dt_bits[replica]);
        if (w_ghost(me.w)) {
            moved2 = carry_ghost<BOX>(A, P, G, s, me);
        } else {
            double const mu = P.mobility[w_type(me.w)];
            double const mu_dt = dt * mu;
            moved2 = integrate_bead<BOX>(A, P, G, S, s, replica, me, F, mu_dt, sqrt(2.0 * S.kT * mu_dt),
                                         sqrt(2.0 * S.kT_prev * (A.dt_prev[replica] * mu)));
        }
    }
    publish_displacement(A, moved2, known);
}

/* adaptive bookkeeping between steps: dt_prev <- dt, dt <- max timestep */
__global__ void k_dt_rotate(unsigned long long *dt_bits, double *dt_prev, double *dt_last,
                            double max_dt, uint32_t replicas)
{
    uint32_t const r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= replicas) return;
    double const dt = __longlong_as_double((long long) dt_bits[r]);
    dt_prev[r] = dt;
    dt_last[r] = dt;
    dt_bits[r] = (unsigned long long) __double_as_longlong(max_dt);
}

/* potential energy of the selected forcefields: per-block partial sums, fixed order */
__global__ void __launch_bounds__(STEP_BLOCK) k_energy(StepArrays A, Select sel, double *partial)
{
    __shared__ DParams P;
    load_params(P, A.params);
    load_replica_walls(P, A.wall_geom, blockIdx.y);
    __syncthreads();
    uint32_t const replica = blockIdx.y, chunk = blockIdx.x;
    uint32_t const local = chunk * STEP_BLOCK + threadIdx.x;
    uint32_t const pos = replica * A.npad + local;
    uint32_t const s = replica * A.n + (local < A.n ? chunk * STEP_BLOCK + A.order[replica * A.n + local] : 0u);
    double E = 0.0;
    if (local < A.n && !w_ghost(load_bead(A.pos + s).w)) {
        bead4 const me = load_bead(A.pos + s);
        uint32_t const oi = w_orig(me.w), ti = w_type(me.w);
        int const T = P.ntypes;
        uint32_t const base = replica * A.n;
        uint32_t const m = A.plen[pos];
        for (uint32_t k = 0; k < m; k++) {
            uint32_t const j = row_entry_slot(A.nbr4, A.tile_runs, A.staged, A.pstride, pos, replica * A.chunks + chunk, k);
            bead4 const o = load_bead(A.pos + j);
            uint32_t const oj = w_orig(o.w), tj = w_type(o.w);
            if (oi >= oj) continue;
            d3 const r = displacement(P, me.x, me.y, me.z, o.x, o.y, o.z);
            double const r2 = sqnorm_exact(r.x, r.y, r.z);
            for (int f = 0; f < BD_MAX_PAIR_FF; f++) {
                const DPair &pf = P.pair[f];
                if (!ff_on(sel, BD_FF_PAIR, f, pf.enabled) || r2 > pf.dcut2) continue;
                if (pf.has_targets && !((w_pair_mask(me.w) >> f) & (w_pair_mask(o.w) >> f) & 1u)) continue;
                E += pot_energy(A.terms + pf.term_off + (ti * T + tj) * pf.nterms, pf.nterms, r2);
            }
        }
        for (uint32_t k = A.bond_start[oi]; k < A.bond_start[oi + 1]; k++) {
            uint2 const adj = A.bond_adj[k];
            int const f = adj.y & 0xff;
            if ((adj.y >> 8) & 1u) continue;  // counted from the first bead
            const DBond &bf = P.bond[f];
            if (!ff_on(sel, BD_FF_BOND, f, bf.enabled)) continue;
            bead4 const o = load_bead(A.pos + A.slot_of[base + adj.x]);
            double rx = me.x - o.x, ry = me.y - o.y, rz = me.z - o.z;
            if (bf.min_image) {
                d3 const r = displacement(P, me.x, me.y, me.z, o.x, o.y, o.z);
                rx = r.x; ry = r.y; rz = r.z;
            }
            E += pot_energy(A.terms + bf.term_off + (ti * T + w_type(o.w)) * bf.nterms, bf.nterms,
                            rx * rx + ry * ry + rz * rz);
        }
        for (uint32_t k = A.bend_start[oi]; k < A.bend_start[oi + 1]; k++) {
            uint32_t const adj = A.bend_adj[k];
            if ((adj & 3u) != 0) continue;  // counted from the first bead
            uint32_t const t = adj >> 2;
            int const f = A.bend_ff[t];
            if (!ff_on(sel, BD_FF_BEND, f, P.bend[f].enabled)) continue;
            bead4 const pj = load_bead(A.pos + A.slot_of[base + A.triples[3 * t + 1]]);
            bead4 const pk = load_bead(A.pos + A.slot_of[base + A.triples[3 * t + 2]]);
            double const ax = me.x - pj.x, ay = me.y - pj.y, az = me.z - pj.z;
            double const bx = pj.x - pk.x, by = pj.y - pk.y, bz = pj.z - pk.z;
            double const dij = ax * ax + ay * ay + az * az, djk = bx * bx + by * by + bz * bz;
            if (dij * djk == 0.0) continue;
            double const e = A.bend_energy[t];
            E += e - e * ((ax * bx + ay * by + az * bz) / sqrt(dij * djk));
        }
        for (int f = 0; f < BD_MAX_WALL_FF; f++) {
            const DWall &wf = P.wall[f];
            if (!ff_on(sel, BD_FF_WALL, f, wf.enabled)) continue;
            double const rx = me.x - wf.center[0], ry = me.y - wf.center[1], rz = me.z - wf.center[2];
            const DTerm *row = A.terms + wf.term_off + ti * wf.nterms;
            if (wf.shape == BD_WALL_SPHERE) {
                double const radius = wf.semi[0];
                double const r2 = rx * rx + ry * ry + rz * rz;
                if (r2 == 0.0) continue;
                double const scale = radius / sqrt(r2);
                double const sx = rx - scale * rx, sy = ry - scale * ry, sz = rz - scale * rz;
                bool const inside = r2 < radius * radius;
                if (inside ? wf.direction == BD_WALL_INWARD : wf.direction == BD_WALL_OUTWARD)
                    E += pot_energy(row, wf.nterms, sx * sx + sy * sy + sz * sz);
            } else {
                double const qx = wf.quad[0], qy = wf.quad[1], qz = wf.quad[2];
                double const dx = qx * rx, dy = qy * ry, dz = qz * rz;
                double const d2 = dx * dx + dy * dy + dz * dz;
                if (d2 == 0.0) continue;
                double const implicit = dx * rx + dy * ry + dz * rz - 1.0;
                double const scale = implicit / (2.0 * d2);
                bool const inside = implicit < 0.0;
                if (inside ? wf.direction == BD_WALL_INWARD : wf.direction == BD_WALL_OUTWARD)
                    E += pot_energy(row, wf.nterms, scale * scale * d2);
            }
        }
        for (int f = 0; f < BD_MAX_POINT_FF; f++) {
            const DPoint &pf = P.point[f];
            if (!ff_on(sel, BD_FF_POINT, f, pf.enabled)) continue;
            if (pf.has_targets && !((w_point_mask(me.w) >> f) & 1u)) continue;
            double const rx = me.x - pf.source[0], ry = me.y - pf.source[1], rz = me.z - pf.source[2];
            E += pot_energy(A.terms + pf.term_off + ti * pf.nterms, pf.nterms, rx * rx + ry * ry + rz * rz);
        }
    }
    for (int o = 16; o > 0; o >>= 1) E += __shfl_xor_sync(0xffffffffu, E, o);
    __shared__ double red[STEP_BLOCK / 32];
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = E;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < STEP_BLOCK / 32; w++) t += red[w];
        partial[(size_t) replica * A.chunks + chunk] = t;
    }
}

/* out[replica][c] = sum over chunks of partial[replica][chunk][c], in chunk order */
__global__ void k_reduce_partials(const double *partial, double *out, uint32_t chunks, uint32_t width)
{
    uint32_t const replica = blockIdx.x;
    for (uint32_t c = threadIdx.x; c < width; c += blockDim.x) {
        double t = 0.0;
        for (uint32_t k = 0; k < chunks; k++) t += partial[((size_t) replica * chunks + k) * width + c];
        out[(size_t) replica * width + c] = t;
    }
}

/* Device-resident wall relaxation: what the interphase driver's per-step callback does on the host for
 * its one system (5-sim-genome/src/simulation_interphase/simulation_driver_interphase.cc:70-80,
 * simulation_driver_forcefield.cc:230-234), done here for every replica of an ensemble without leaving
 * the GPU.  After a step, per replica:
 *     net   = sum over the walls of the group of stats.axial_reaction  -  spring o semiaxes
 *     semiaxes += dt * mobility * net            (all walls of the group share the semiaxes)
 * One block per replica: reduces the per-chunk statistics in a fixed order (also the answer of
 * bd_get_wall_stats[_all]) and rewrites the replica's geometry for the next step. */
struct WallRelax {
    uint32_t mask;       // walls of the group (bit f), 0 = no relaxation: the kernel only reduces
    uint32_t pad;
    double spring[3];
    double mobility;
};

constexpr int RELAX_LANES = 32;  // threads per statistic: the chunk sums are a chain of dependent loads
constexpr int RELAX_WIDTH = BD_MAX_WALL_FF * WALL_STATS;

__global__ void __launch_bounds__(RELAX_WIDTH * RELAX_LANES)
k_wall_relax(const double *partial, double *out, DWallGeom *geom, WallRelax law, uint32_t chunks, double dt)
{
    __shared__ double part[RELAX_WIDTH][RELAX_LANES];
    __shared__ double total[RELAX_WIDTH];
    uint32_t const replica = blockIdx.x;
    uint32_t const c = threadIdx.x / RELAX_LANES, lane = threadIdx.x % RELAX_LANES;
    double t = 0.0;
    for (uint32_t k = lane; k < chunks; k += RELAX_LANES) t += partial[((size_t) replica * chunks + k) * RELAX_WIDTH + c];
    part[c][lane] = t;
    __syncthreads();
    if (lane == 0) {
        double sum = 0.0;
        for (int l = 0; l < RELAX_LANES; l++) sum += part[c][l];
        total[c] = sum;
        out[(size_t) replica * RELAX_WIDTH + c] = sum;
    }
    __syncthreads();
    if (law.mask == 0 || threadIdx.x >= 3) return;
    int const d = threadIdx.x;
    DWallGeom *g = geom + (size_t) replica * BD_MAX_WALL_FF;
    double semi = 0.0, net = 0.0;
    bool first = true;
    for (int f = 0; f < BD_MAX_WALL_FF; f++) {
        if (!((law.mask >> f) & 1u)) continue;
        if (first) semi = g[f].semi[d];
        first = false;
        net += total[f * WALL_STATS + 1 + d];
    }
    net = __dsub_rn(net, __dmul_rn(law.spring[d], semi));
    semi = __dadd_rn(semi, __dmul_rn(__dmul_rn(dt, law.mobility), net));
    for (int f = 0; f < BD_MAX_WALL_FF; f++) {
        if (!((law.mask >> f) & 1u)) continue;
        g[f].semi[d] = semi;
        g[f].quad[d] = 1.0 / __dmul_rn(semi, semi);
    }
}

// ---- original-order <-> slot-order movement -----------------------------------------------------

/* xyz[replica][n][3] (original order) -> pos[slot] */
__global__ void k_scatter_positions(bead4 *pos, const uint32_t *slot_of, const double *xyz,
                                    uint32_t first, uint32_t count)
{
    uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    uint32_t const g = first + i;
    bead4 *p = pos + slot_of[g];
    p->x = xyz[3 * (size_t) i];
    p->y = xyz[3 * (size_t) i + 1];
    p->z = xyz[3 * (size_t) i + 2];
}

/* positions replaced by the caller while a list exists: the rows stay valid as long as no bead is
 * further than the skin allows from where it was at the last build, which is the reference's own
 * criterion (neighbor_list.hpp:132-158 looks at the positions, not at who wrote them).  Write the
 * record and its shadow, and fold the displacement into the control block; k_rebuild_begin decides. */
__global__ void k_scatter_tracked(bead4 *pos, float4 *shadow, const bead4 *bpos, const uint32_t *slot_of,
                                  const double *xyz, uint32_t first, uint32_t count, const DParams *params,
                                  DControl *ctl, uint32_t staged)
{
    uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
    double moved2 = 0.0;
    if (i < count) {
        uint32_t const s = slot_of[first + i];
        double const x = xyz[3 * (size_t) i], y = xyz[3 * (size_t) i + 1], z = xyz[3 * (size_t) i + 2];
        pos[s].x = x;
        pos[s].y = y;
        pos[s].z = z;
        bead4 const b0 = bpos[s];
        d3 d;
        if (params->box_kind == BD_BOX_PERIODIC) {
            shadow[s] = make_shadow<BD_BOX_PERIODIC>(*params, ctl->grid, x, y, z, shadow_tag(staged, s, b0.w));
            d = displacement_t<BD_BOX_PERIODIC>(*params, x, y, z, b0.x, b0.y, b0.z);
        } else {
            shadow[s] = make_shadow<BD_BOX_OPEN>(*params, ctl->grid, x, y, z, shadow_tag(staged, s, b0.w));
            d = displacement_t<BD_BOX_OPEN>(*params, x, y, z, b0.x, b0.y, b0.z);
        }
        moved2 = sqnorm_exact(d.x, d.y, d.z);
        if (!(moved2 == moved2)) moved2 = __longlong_as_double(0x7ff0000000000000ll);  // NaN: force a rebuild
    }
    for (int o = 16; o > 0; o >>= 1) moved2 = fmax(moved2, __shfl_xor_sync(0xffffffffu, moved2, o));
    if ((threadIdx.x & 31) == 0 && moved2 > 0.0)
        atomicMax(&ctl->max_disp2_bits, (unsigned long long) __double_as_longlong(moved2));
}

__global__ void k_gather_positions(const bead4 *pos, const uint32_t *slot_of, double *xyz,
                                   uint32_t first, uint32_t count)
{
    uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const bead4 *p = pos + slot_of[first + i];
    xyz[3 * (size_t) i] = p->x;
    xyz[3 * (size_t) i + 1] = p->y;
    xyz[3 * (size_t) i + 2] = p->z;
}

__global__ void k_gather_forces(const double *force, const uint32_t *slot_of, double *out,
                                uint32_t first, uint32_t count)
{
    uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    size_t const s = slot_of[first + i];
    out[3 * (size_t) i] = force[3 * s];
    out[3 * (size_t) i + 1] = force[3 * s + 1];
    out[3 * (size_t) i + 2] = force[3 * s + 2];
}

/* positions of a list of original indices (replica 0) -> xyz[count][3]: what a slab sends to its
 * neighbours every step (slab.py) */
__global__ void k_gather_indexed(const bead4 *pos, const uint32_t *slot_of, const uint32_t *idx, double *xyz,
                                 uint32_t count)
{
    uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    bead4 const b = load_bead(pos + slot_of[idx[i]]);
    xyz[3 * (size_t) i] = b.x;
    xyz[3 * (size_t) i + 1] = b.y;
    xyz[3 * (size_t) i + 2] = b.z;
}

/* xyz[count][3] -> positions (and shadows) of a list of original indices: the ghosts a slab receives.
 * Like every other writer of positions it folds the displacement from the bead's position at the last list
 * build into the control block, so that k_rebuild_begin sees a ghost that left its skin BEFORE the next
 * force evaluation (neighbor_list.hpp:132-158 looks at all points, whoever moved them). */
__global__ void k_scatter_indexed(bead4 *pos, float4 *shadow, const bead4 *bpos, const uint32_t *slot_of,
                                  const uint32_t *idx, const double *xyz, uint32_t count, const DParams *params,
                                  DControl *ctl, uint32_t staged)
{
    uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
    double moved2 = 0.0;
    if (i < count) {
        uint32_t const s = slot_of[idx[i]];
        double const x = xyz[3 * (size_t) i], y = xyz[3 * (size_t) i + 1], z = xyz[3 * (size_t) i + 2];
        pos[s].x = x;
        pos[s].y = y;
        pos[s].z = z;
        bead4 const b0 = bpos[s];
        d3 d;
        if (params->box_kind == BD_BOX_PERIODIC) {
            shadow[s] = make_shadow<BD_BOX_PERIODIC>(*params, ctl->grid, x, y, z, shadow_tag(staged, s, b0.w));
            d = displacement_t<BD_BOX_PERIODIC>(*params, x, y, z, b0.x, b0.y, b0.z);
        } else {
            shadow[s] = make_shadow<BD_BOX_OPEN>(*params, ctl->grid, x, y, z, shadow_tag(staged, s, b0.w));
            d = displacement_t<BD_BOX_OPEN>(*params, x, y, z, b0.x, b0.y, b0.z);
        }
        moved2 = sqnorm_exact(d.x, d.y, d.z);
        if (!(moved2 == moved2)) moved2 = __longlong_as_double(0x7ff0000000000000ll);  // NaN: force a rebuild
    }
    for (int o = 16; o > 0; o >>= 1) moved2 = fmax(moved2, __shfl_xor_sync(0xffffffffu, moved2, o));
    if ((threadIdx.x & 31) == 0 && moved2 > 0.0)
        atomicMax(&ctl->max_disp2_bits, (unsigned long long) __double_as_longlong(moved2));
}

/* (re)write the packed word of every bead: original index, class, target masks */
__global__ void k_set_words(bead4 *pos0, bead4 *pos1, bead4 *bpos, const uint32_t *slot_of,
                            const uint8_t *type, const uint8_t *pair_mask, const uint8_t *point_mask,
                            uint32_t n, uint32_t ntot)
{
    uint32_t const g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ntot) return;
    uint32_t const o = g % n;
    uint32_t const s = slot_of[g];
    // ownership (the ghost bit) is device state: bd_set_owned / bd_set_owned_device
    uint64_t const w = pack_w(o, type[o], pair_mask[o], point_mask[o]) | (pos0[s].w & ((uint64_t) GHOST_BIT << 48));
    pos0[s].w = w;
    pos1[s].w = w;
    bpos[s].w = w;
}

/* ownership of every bead (replica 0): owned[original index] != 0, or everything when owned is null */
__global__ void k_set_ghost(bead4 *pos0, bead4 *pos1, bead4 *bpos, const uint32_t *slot_of, const uint8_t *owned,
                            uint32_t n)
{
    uint32_t const g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    uint32_t const s = slot_of[g];
    uint64_t const bit = (uint64_t) GHOST_BIT << 48;
    uint64_t const w = (pos0[s].w & ~bit) | ((owned && !owned[g]) ? bit : 0ull);
    pos0[s].w = w;
    pos1[s].w = w;
    bpos[s].w = w;
}

__global__ void k_init_identity(uint32_t *slot_of, bead4 *pos0, bead4 *pos1, bead4 *bpos, float4 *shadow0,
                                float4 *shadow1, uint8_t *order, uint32_t n, uint32_t ntot)
{
    uint32_t const g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g > ntot) return;
    if (g == ntot) {
        // the sentinel slot: NaN coordinates fail every distance test
        double const qnan = __longlong_as_double(0x7ff8000000000000ll);
        float const fnan = __uint_as_float(0x7fc00000u);
        bead4 b;
        b.x = b.y = b.z = qnan;
        b.w = pack_w(0xffffffffu, 0, 0, 0);
        pos0[g] = b;
        pos1[g] = b;
        shadow0[g] = make_float4(fnan, fnan, fnan, __uint_as_float(g));
        shadow1[g] = make_float4(fnan, fnan, fnan, __uint_as_float(g));
        return;
    }
    slot_of[g] = g;
    order[g] = (uint8_t) ((g % n) % STEP_BLOCK);
    bead4 b;
    b.x = b.y = b.z = 0.0;
    b.w = pack_w(g % n, 0, 0, 0);
    pos0[g] = b;
    pos1[g] = b;
    bpos[g] = b;
    shadow0[g] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(g));
    shadow1[g] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(g));
}

}  // namespace bd
