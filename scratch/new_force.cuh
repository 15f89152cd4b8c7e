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

/* one row of a compiled pair table: c = factor * P * Q * energy * k2, k2 = 1 / diameter^2 */
struct __align__(16) PairRow {
    double k2a, ca, k2b, cb;
};

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
__device__ __forceinline__ double compiled_pair_scalar(const PairRow &row, double r2)
{
    double const ua = row.k2a * r2;
    double const ga = 1.0 - ua;
    double s = ga >= 0.0 ? row.ca * ga * ga : 0.0;
    if (PROG == PROG_SC23_SC83) {
        double const ub = row.k2b * r2;
        double const vb = ub * ub * ub;
        double const gb = 1.0 - vb * ub;
        s += gb >= 0.0 ? row.cb * gb * gb * vb : 0.0;
    }
    return s;
}

/* non-bonded forces on the bead at slot s from its Verlet row */
template <int BOX, int PROG>
__device__ __forceinline__ void pair_forces(const StepArrays &A, const DParams &P, const PairRow *rows,
                                            Select sel, uint32_t s, const bead4 &me, d3 &F,
                                            uint32_t &interactions)
{
    uint32_t const oi = w_orig(me.w), ti = w_type(me.w);
    int const T = P.ntypes;
    uint32_t const m = A.nnbr[s];
    const uint32_t *col = A.nbr + s;
    if (PROG != PROG_GENERIC) {
        double const dcut2 = P.pair[A.fast_ff].dcut2;
        const PairRow *mine = rows + ti * T;
        for (uint32_t k = 0; k < m; k++) {
            uint32_t const j = col[(size_t) k * A.ntot];
            bead4 const o = load_bead(A.pos + j);
            d3 const r = displacement_t<BOX>(P, me.x, me.y, me.z, o.x, o.y, o.z);
            double const r2 = sqnorm_exact(r.x, r.y, r.z);
            if (r2 <= dcut2) {
                double const sc = compiled_pair_scalar<PROG>(mine[w_type(o.w)], r2);
                F.x += sc * r.x; F.y += sc * r.y; F.z += sc * r.z;
                interactions += oi < w_orig(o.w) ? 1u : 0u;
            }
        }
        return;
    }
    bool any_pair = false;
    for (int f = 0; f < BD_MAX_PAIR_FF; f++) any_pair |= ff_on(sel, BD_FF_PAIR, f, P.pair[f].enabled);
    if (!any_pair) return;
    for (uint32_t k = 0; k < m; k++) {
        uint32_t const j = col[(size_t) k * A.ntot];
        bead4 const o = load_bead(A.pos + j);
        d3 const r = displacement_t<BOX>(P, me.x, me.y, me.z, o.x, o.y, o.z);
        double const r2 = sqnorm_exact(r.x, r.y, r.z);
        uint32_t const oj = w_orig(o.w), tj = w_type(o.w);
        int const cell = oi < oj ? (int) (ti * T + tj) : (int) (tj * T + ti);
        bool hit = false;
#pragma unroll
        for (int f = 0; f < BD_MAX_PAIR_FF; f++) {
            const DPair &pf = P.pair[f];
            if (!ff_on(sel, BD_FF_PAIR, f, pf.enabled) || r2 > pf.dcut2) continue;
            if (pf.has_targets && !((w_pair_mask(me.w) >> f) & (w_pair_mask(o.w) >> f) & 1u)) continue;
            double const sc = pot_force_scalar(A.terms + pf.term_off + cell * pf.nterms, pf.nterms, r2);
            F.x += sc * r.x; F.y += sc * r.y; F.z += sc * r.z;
            hit = true;
        }
        interactions += (hit && oi < oj) ? 1u : 0u;
    }
}

/* bonded pairs, bending triples and point sources acting on one bead */
__device__ __forceinline__ void bonded_forces(const StepArrays &A, const DParams &P, Select sel,
                                              uint32_t replica, const bead4 &me, d3 &F)
{
    uint32_t const oi = w_orig(me.w), ti = w_type(me.w);
    int const T = P.ntypes;
    uint32_t const base = replica * A.n;

    // bonded pairs (no minimum image: bonded_pairwise_forcefield.hpp:64)
    {
        uint32_t const b = A.bond_start[oi], e = A.bond_start[oi + 1];
        for (uint32_t k = b; k < e; k++) {
            uint2 const adj = A.bond_adj[k];
            int const f = adj.y & 0xff;
            const DBond &bf = P.bond[f];
            if (!ff_on(sel, BD_FF_BOND, f, bf.enabled)) continue;
            bead4 const o = load_bead(A.pos + A.slot_of[base + adj.x]);
            uint32_t const tj = w_type(o.w);
            bool const second = (adj.y >> 8) & 1u;
            // the force on this bead is scalar * (x_me - x_partner) whichever end it is; only the
            // table row depends on the order (type_first, type_second)
            double const rx = me.x - o.x, ry = me.y - o.y, rz = me.z - o.z;
            int const cell = second ? (int) (tj * T + ti) : (int) (ti * T + tj);
            double const sc = pot_force_scalar(A.terms + bf.term_off + cell * bf.nterms, bf.nterms,
                                               rx * rx + ry * ry + rz * rz);
            F.x += sc * rx; F.y += sc * ry; F.z += sc * rz;
        }
    }

    // bending triples (cosine_bending_potential.hpp:43-60)
    {
        uint32_t const b = A.bend_start[oi], e = A.bend_start[oi + 1];
        for (uint32_t k = b; k < e; k++) {
            uint32_t const adj = A.bend_adj[k];
            uint32_t const t = adj >> 2, role = adj & 3u;
            int const f = A.bend_ff[t];
            if (!ff_on(sel, BD_FF_BEND, f, P.bend[f].enabled)) continue;
            double const E = A.bend_energy[t];
            uint32_t const i0 = A.triples[3 * t], i1 = A.triples[3 * t + 1], i2 = A.triples[3 * t + 2];
            bead4 const pi = role == 0 ? me : load_bead(A.pos + A.slot_of[base + i0]);
            bead4 const pj = role == 1 ? me : load_bead(A.pos + A.slot_of[base + i1]);
            bead4 const pk = role == 2 ? me : load_bead(A.pos + A.slot_of[base + i2]);
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
        double const qx = 1.0 / (wf.semi[0] * wf.semi[0]), qy = 1.0 / (wf.semi[1] * wf.semi[1]),
                     qz = 1.0 / (wf.semi[2] * wf.semi[2]);
        double const dx = qx * rx, dy = qy * ry, dz = qz * rz;  // dual
        double const d2 = dx * dx + dy * dy + dz * dz;
        if (d2 == 0.0) return;
        double const implicit = dx * rx + dy * ry + dz * rz - 1.0;
        bool const inside = implicit < 0.0;
        if (inside ? wf.direction != BD_WALL_INWARD : wf.direction != BD_WALL_OUTWARD) return;
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

struct StepShared {
    DParams P;
    PairRow rows[BD_MAX_TYPES * BD_MAX_TYPES];
    double wall_red[STEP_BLOCK / 32][BD_MAX_WALL_FF][WALL_STATS];
    uint32_t inter[STEP_BLOCK / 32];
};

template <int PROG>
__device__ __forceinline__ void load_shared(StepShared &sh, const StepArrays &A)
{
    for (uint32_t i = threadIdx.x; i < sizeof(DParams) / 4; i += blockDim.x)
        ((uint32_t *) &sh.P)[i] = ((const uint32_t *) A.params)[i];
    if (PROG != PROG_GENERIC) {
        uint32_t const words = A.ntypes2 * (sizeof(PairRow) / 4);
        for (uint32_t i = threadIdx.x; i < words; i += blockDim.x)
            ((uint32_t *) sh.rows)[i] = ((const uint32_t *) A.pair_rows)[i];
    }
    __syncthreads();
}

/* all selected forces on one bead; every thread of the block must call this (walls reduce their
 * statistics across the block), `live` masks the threads past the end of the replica */
template <int BOX, int PROG>
__device__ __forceinline__ d3 block_forces(const StepArrays &A, StepShared &sh, Select sel, bool live,
                                           uint32_t s, uint32_t replica, uint32_t chunk, const bead4 &me,
                                           uint32_t &interactions)
{
    const DParams &P = sh.P;
    d3 F = {0.0, 0.0, 0.0};
    if (live) {
        pair_forces<BOX, PROG>(A, P, sh.rows, sel, s, me, F, interactions);
        bonded_forces(A, P, sel, replica, me, F);
    }
    bool any_wall = false;
#pragma unroll 1
    for (int f = 0; f < BD_MAX_WALL_FF; f++) {
        const DWall &wf = P.wall[f];
        if (!ff_on(sel, BD_FF_WALL, f, wf.enabled)) continue;  // block-uniform
        any_wall = true;
        double st[WALL_STATS] = {0.0, 0.0, 0.0, 0.0};
        if (live) wall_force(A, wf, me, F, st);
#pragma unroll
        for (int c = 0; c < WALL_STATS; c++) {
            for (int o = 16; o > 0; o >>= 1) st[c] += __shfl_xor_sync(0xffffffffu, st[c], o);
            if ((threadIdx.x & 31) == 0) sh.wall_red[threadIdx.x >> 5][f][c] = st[c];
        }
    }
    if (any_wall) {
        __syncthreads();
        if (threadIdx.x < BD_MAX_WALL_FF * WALL_STATS) {
            int const f = threadIdx.x / WALL_STATS, c = threadIdx.x % WALL_STATS;
            if (ff_on(sel, BD_FF_WALL, f, P.wall[f].enabled)) {
                double t = 0.0;
                for (int w = 0; w < STEP_BLOCK / 32; w++) t += sh.wall_red[w][f][c];
                A.wall_partial[(((size_t) replica * A.chunks + chunk) * BD_MAX_WALL_FF + f) * WALL_STATS + c] = t;
            }
        }
    }
    return F;
}
