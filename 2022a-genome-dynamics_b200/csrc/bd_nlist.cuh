/*
 * bd_nlist.cuh -- GPU cell list and Verlet neighbour rows.
 *
 * Replaces md::neighbor_list<Box> + md::neighbor_searcher<Box> + search_grid
 * (MD/forcefield/detail/neighbor_list.hpp:63-239, MD/misc/neighbor_searcher.hpp:23-131,
 * MD/misc/nsearch_detail/search_grid.hpp:67-295).  The reference hashes beads into
 * std::vector buckets and emits a std::pair list on the host; here a rebuild is a chain of kernels
 * that never returns to the host: bounding box -> grid -> counting sort by (replica, Morton cell
 * key) with a deterministic in-cell order -> ELL rows.  Every kernel of the chain is gated on a
 * device-side flag, so the step loop can enqueue the chain unconditionally.
 *
 * What is kept bit-for-bit from the reference: the list radius 1.5*dcut (neighbor_list.hpp:182),
 * the fp64 distance test r^2 <= radius^2 on the build positions (neighbor_searcher.hpp:81-92) and
 * the rebuild criterion (neighbor_list.hpp:132-158), so the listed pair SET equals the reference's.
 */
#pragma once

#include "bd_device.cuh"

namespace bd {

constexpr int NL_BLOCK = 256;
constexpr int BAL_TILE = 128;                       // == STEP_BLOCK (checked in bd_engine.cu)
static_assert(ROW_PAD == 4, "Verlet rows are stored in uint4 groups");
constexpr int SCAN_ITEMS = 8;                       // counts per thread in the scan kernels
constexpr int SCAN_TILE = NL_BLOCK * SCAN_ITEMS;    // 2048 counts per block

struct NlistArrays {
    const bead4 *pos;        // current positions (slot order of the previous build)
    bead4 *pos_out;          // == pos buffer, rewritten in the new slot order by k_rows
    bead4 *bpos;             // build positions, new slot order
    float4 *shadow;          // single-precision shadow of pos_out (bd_device.cuh: make_shadow)
    uint32_t *slot_of;       // [Ntot]
    uint32_t *cellkey;       // [Ntot] key of the bead at OLD slot
    uint32_t *rank;          // [Ntot] arrival rank within the cell
    uint32_t *stage;         // [Ntot] old slot, grouped by cell
    uint32_t *cell_count;    // [cell_cap]
    uint32_t *cell_start;    // [cell_cap + 1]
    uint32_t *block_sum;     // [cell_cap / SCAN_TILE + 1]
    // Verlet rows are stored by THREAD POSITION P = replica * npad + chunk * BAL_TILE + t (npad = N
    // rounded up to whole tiles), in groups of ROW_PAD indices: nbr4[g * pstride + P]
    uint4 *nbr4;             // [cap / ROW_PAD][pstride]
    uint32_t *plen;          // [pstride] row length of the bead at position P
    uint32_t *nnbr;          // [Ntot] row length by slot
    uint8_t *order;          // [Ntot] position -> bead offset within its BAL_TILE-slot tile, by row length
    uint2 *bondrow;          // [bond_deg][pstride] {partner slot, ff | second << 8}, 0xffffffff slot = none
    const uint32_t *bond_start;  // CSR of all bond forcefields by original index (bd_engine.cu: upload_topology)
    const uint2 *bond_adj;
    const uint8_t *type;     // [n] bead class by original index
    // bending triples by original index (CSR) and, resolved to slots per thread position, the rows the step kernel reads
    const uint32_t *bend_start, *bend_adj, *triples;
    const double *bend_energy;
    const uint8_t *bend_ff;
    uint2 *bend_pq;
    double *bend_e;
    uint8_t *bend_meta;
    uint32_t bend_deg;
    uint8_t bond_slot[BD_MAX_BOND_FF];  // bond forcefield -> block of the compiled bond table (bd_step.cuh: BondRow)
    uint32_t ntypes;
    uint32_t bond_deg;       // largest number of bonds on one bead
    TileRuns *tile_runs;     // [replicas * chunks] staged neighbourhood of every tile (bd_device.cuh), or null
    uint32_t staged;         // rows hold 16-bit staged positions (uint2 groups) instead of slots (uint4 groups)
    uint32_t balance;        // k_balance ranks the beads of a tile by row length (option "balance_rows")
    uint32_t npad, pstride;
    DControl *ctl;
    const DParams *params;
    uint32_t n;              // beads per replica
    uint32_t replicas;
    uint32_t ntot;
    uint32_t cell_cap;       // capacity of cell arrays (power of two)
    uint32_t max_bits;       // Morton bits available per replica
    uint32_t cap;            // ELL row capacity
};

/* decide whether this chain rebuilds the list (neighbor_list.hpp:101-159) and latch the answer:
 * a host request (new positions), a list radius that no longer covers dcut, or a bead that has
 * moved further than (radius - dcut)/2 from its build position */
/* `gate`: inside a step graph (bd_engine.cu) the nine kernels of the chain sit in a conditional node that this
 * kernel opens only when a rebuild is due -- a step that keeps its list launches nothing but this kernel and
 * the step kernel.  Outside a graph (use_gate = 0) the chain kernels are launched anyway and return at once. */
__global__ void k_rebuild_begin(DControl *ctl, const DParams *params, cudaGraphConditionalHandle gate, uint32_t use_gate)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    uint32_t need = ctl->need_rebuild;
    double const dcut = params->dcut_max;
    if (dcut > 0.0) {
        double const thr = __dmul_rn(__dsub_rn(ctl->grid.radius, dcut), 0.5);
        double const moved2 = __longlong_as_double((long long) ctl->max_disp2_bits);
        if (thr <= 0.0 || moved2 > __dmul_rn(thr, thr)) need = 1;
    }
    ctl->doing_rebuild = need;
    ctl->need_rebuild = 0;
    if (use_gate) cudaGraphSetConditional(gate, need);
    if (need) {
        ctl->rebuilds += 1;
        ctl->graph_rebuilds += use_gate;
        ctl->listed_entries = 0;
        ctl->max_row = 0;
        ctl->max_stage = 0;
        ctl->max_disp2_bits = 0;
        for (int d = 0; d < 3; d++) {
            ctl->bbox_lo[d] = ~0ull;
            ctl->bbox_hi[d] = 0ull;
        }
    }
}

/* bounding box of all replicas (open box) and zero-fill of the cell counters */
__global__ void __launch_bounds__(NL_BLOCK) k_bbox_zero(NlistArrays A)
{
    if (!A.ctl->doing_rebuild) return;
    uint32_t const tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t const stride = gridDim.x * blockDim.x;
    for (uint32_t c = tid; c < A.cell_cap; c += stride) A.cell_count[c] = 0;

    unsigned long long lo[3] = {~0ull, ~0ull, ~0ull}, hi[3] = {0, 0, 0};
    for (uint32_t s = tid; s < A.ntot; s += stride) {
        bead4 const b = A.pos[s];
        double const v[3] = {b.x, b.y, b.z};
#pragma unroll
        for (int d = 0; d < 3; d++) {
            unsigned long long const e = encode_ordered(v[d]);
            lo[d] = e < lo[d] ? e : lo[d];
            hi[d] = e > hi[d] ? e : hi[d];
        }
    }
#pragma unroll
    for (int d = 0; d < 3; d++) {
        for (int o = 16; o > 0; o >>= 1) {
            unsigned long long const l2 = __shfl_xor_sync(0xffffffffu, lo[d], o);
            unsigned long long const h2 = __shfl_xor_sync(0xffffffffu, hi[d], o);
            lo[d] = l2 < lo[d] ? l2 : lo[d];
            hi[d] = h2 > hi[d] ? h2 : hi[d];
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int d = 0; d < 3; d++) {
            atomicMin(&A.ctl->bbox_lo[d], lo[d]);
            atomicMax(&A.ctl->bbox_hi[d], hi[d]);
        }
    }
}

/* choose the grid: cell edge >= list radius, Morton bits within the cell-array capacity */
__global__ void k_make_grid(NlistArrays A)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    DControl *ctl = A.ctl;
    if (!ctl->doing_rebuild) return;
    const DParams &P = *A.params;
    DGrid &G = ctl->grid;
    double radius = __dmul_rn(P.verlet_factor, P.dcut_max);  // neighbor_list.hpp:182
    G.radius = radius;
    double const edge = radius > 0 ? radius * (1.0 + 1e-9) : 1.0;
    double extent[3];
    for (int d = 0; d < 3; d++) {
        if (P.box_kind == BD_BOX_PERIODIC) {
            G.origin[d] = 0;
            extent[d] = P.period[d];
        } else {
            double const lo = decode_ordered(ctl->bbox_lo[d]), hi = decode_ordered(ctl->bbox_hi[d]);
            G.origin[d] = lo;
            extent[d] = (hi - lo) * (1.0 + 1e-12) + 1e-12;
        }
        double const cells = floor(extent[d] / edge);
        int n = cells < 1.0 ? 1 : (cells > 1048576.0 ? 1048576 : (int) cells);
        int b = 0;
        while ((1 << b) < n) b++;
        G.n[d] = n;
        G.bits[d] = b;
    }
    // shed bits from the finest axis until the key fits
    while (G.bits[0] + G.bits[1] + G.bits[2] > (int) A.max_bits) {
        int d = 0;
        if (G.bits[1] > G.bits[d]) d = 1;
        if (G.bits[2] > G.bits[d]) d = 2;
        G.bits[d] -= 1;
        if (G.n[d] > (1 << G.bits[d])) G.n[d] = 1 << G.bits[d];
    }
    G.totalbits = G.bits[0] + G.bits[1] + G.bits[2];
    for (int d = 0; d < 3; d++) G.inv_cell[d] = (double) G.n[d] / extent[d];
    // error bound of the shadow coordinates (bd_device.cuh: make_shadow).  Open box: two float
    // roundings of offsets no larger than E = extent + 2 radius (a bead further than (radius-dcut)/2
    // from its build position forces a rebuild before the next force evaluation) give
    // |r_shadow - r| <= sqrt(3) 2^-23 E < 2^-22 E.  Periodic: two 2^-32 L quantisations.
    double const big = fmax(extent[0], fmax(extent[1], extent[2]));
    if (P.box_kind == BD_BOX_PERIODIC) {
        G.shadow_eps = big * 3.725290298461914e-09;  // 2^-28
        for (int d = 0; d < 3; d++) G.shadow_scale[d] = (float) (P.period[d] * 2.3283064365386963e-10);
    } else {
        G.shadow_eps = (big + 2.0 * radius) * 9.5367431640625e-07;  // 2^-20
        for (int d = 0; d < 3; d++) G.shadow_scale[d] = 1.0f;
    }
}

__device__ inline uint32_t bead_cell_key(const DParams &P, const DGrid &G, const bead4 &b,
                                         uint32_t replica)
{
    int c[3];
    cell_coords(P, G, b.x, b.y, b.z, c);
    return (replica << G.totalbits) | morton_key(c[0], c[1], c[2], G.bits);
}

/* The per-bead kernels of the chain run grid-stride on a fixed, small grid (bd_engine.cu:
 * chain_blocks): when no rebuild is due -- nine steps out of ten -- each costs one wave of
 * immediately returning blocks instead of Ntot/256 of them. */
__global__ void __launch_bounds__(NL_BLOCK) k_cell_count(NlistArrays A)
{
    if (!A.ctl->doing_rebuild) return;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < A.ntot; s += gridDim.x * blockDim.x) {
        uint32_t const key = bead_cell_key(*A.params, A.ctl->grid, A.pos[s], s / A.n);
        A.cellkey[s] = key;
        A.rank[s] = atomicAdd(&A.cell_count[key], 1u);
    }
}

/* ---- exclusive scan of cell_count[0 .. replicas << totalbits) -> cell_start --------------- */

__device__ inline uint32_t ncells_total(const NlistArrays &A)
{
    return A.replicas << A.ctl->grid.totalbits;
}

__global__ void __launch_bounds__(NL_BLOCK) k_scan_reduce(NlistArrays A)
{
    if (!A.ctl->doing_rebuild) return;
    uint32_t const ncell = ncells_total(A);
    uint32_t const base = blockIdx.x * SCAN_TILE;
    if (base >= ncell) return;
    uint32_t sum = 0;
    for (int k = 0; k < SCAN_ITEMS; k++) {
        uint32_t const i = base + k * NL_BLOCK + threadIdx.x;
        sum += i < ncell ? A.cell_count[i] : 0u;
    }
    __shared__ uint32_t warp_sum[NL_BLOCK / 32];
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < NL_BLOCK / 32; w++) t += warp_sum[w];
        A.block_sum[blockIdx.x] = t;
    }
}

/* one block: exclusive scan of the per-tile sums */
__global__ void __launch_bounds__(1024) k_scan_top(NlistArrays A)
{
    if (!A.ctl->doing_rebuild) return;
    uint32_t const ncell = ncells_total(A);
    uint32_t const ntile = (ncell + SCAN_TILE - 1) / SCAN_TILE;
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t base = 0; base < ntile; base += 1024) {
        uint32_t const i = base + threadIdx.x;
        uint32_t const v = i < ntile ? A.block_sum[i] : 0u;
        uint32_t incl = v;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t const t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((threadIdx.x & 31) >= o) incl += t;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_tot[threadIdx.x];
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t const t = __shfl_up_sync(0xffffffffu, w, o);
                if (threadIdx.x >= o) w += t;
            }
            warp_tot[threadIdx.x] = w;  // inclusive totals of warps
        }
        __syncthreads();
        uint32_t const warp_off = (threadIdx.x >> 5) ? warp_tot[(threadIdx.x >> 5) - 1] : 0u;
        uint32_t const carry = carry_s;
        if (i < ntile) A.block_sum[i] = carry + warp_off + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + warp_off + incl;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(NL_BLOCK) k_scan_down(NlistArrays A)
{
    if (!A.ctl->doing_rebuild) return;
    uint32_t const ncell = ncells_total(A);
    uint32_t const base = blockIdx.x * SCAN_TILE;
    if (base >= ncell) return;
    // thread t owns SCAN_ITEMS consecutive counts
    uint32_t v[SCAN_ITEMS];
    uint32_t sum = 0;
    uint32_t const first = base + threadIdx.x * SCAN_ITEMS;
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = first + k < ncell ? A.cell_count[first + k] : 0u;
        sum += v[k];
    }
    uint32_t incl = sum;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t const t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += t;
    }
    __shared__ uint32_t warp_tot[NL_BLOCK / 32];
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t off = A.block_sum[blockIdx.x];
    for (int w = 0; w < (int) (threadIdx.x >> 5); w++) off += warp_tot[w];
    off += incl - sum;
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (first + k < ncell) A.cell_start[first + k] = off;
        off += v[k];
    }
    if (first < ncell && first + SCAN_ITEMS >= ncell) A.cell_start[ncell] = A.ntot;
}

/* group old slots by cell (arrival order) */
__global__ void __launch_bounds__(NL_BLOCK) k_group(NlistArrays A)
{
    if (!A.ctl->doing_rebuild) return;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < A.ntot; s += gridDim.x * blockDim.x)
        A.stage[A.cell_start[A.cellkey[s]] + A.rank[s]] = s;
}

/* deterministic order inside each cell (ascending original index), then move the records */
__global__ void __launch_bounds__(NL_BLOCK) k_place(NlistArrays A)
{
    if (!A.ctl->doing_rebuild) return;
    const DParams &P = *A.params;
    const DGrid &G = A.ctl->grid;
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < A.ntot; t += gridDim.x * blockDim.x) {
        uint32_t const old = A.stage[t];
        uint32_t const key = A.cellkey[old];
        uint32_t const b = A.cell_start[key], e = A.cell_start[key + 1];
        bead4 const rec = A.pos[old];
        uint32_t const mine = w_orig(rec.w);
        uint32_t rnk = 0;
        for (uint32_t u = b; u < e; u++) {
            uint32_t const other = w_orig(A.pos[A.stage[u]].w);
            rnk += other < mine;
        }
        uint32_t const slot = b + rnk;
        A.bpos[slot] = rec;
        // nobody reads the shadow buffer during a rebuild: it can be rewritten in the new order here
        A.shadow[slot] = P.box_kind == BD_BOX_PERIODIC ? make_shadow<BD_BOX_PERIODIC>(P, G, rec.x, rec.y, rec.z, shadow_tag(A.staged, slot, rec.w))
                                                       : make_shadow<BD_BOX_OPEN>(P, G, rec.x, rec.y, rec.z, shadow_tag(A.staged, slot, rec.w));
        A.slot_of[(old / A.n) * A.n + mine] = slot;
    }
}

constexpr int ROWS_BLOCK = 128;
constexpr uint32_t ROWS_QUEUE = 16;  // queued hits per thread between write-outs

/* ELL rows: every bead lists the beads of its replica within the list radius at build time.
 *
 * One thread per bead.  The (at most) 27 surrounding cells are looked up first, in lockstep, into
 * shared memory; the candidates are then walked in ONE flat loop per thread, so a lane does not wait
 * at every cell boundary for the fullest cell of its warp.  Candidates are tested on their 16-byte
 * single-precision shadows against the widened radius: the stored rows are a superset of the
 * reference's list (neighbor_searcher.hpp:81-92) by at most the shadow error bound.  That changes no
 * result -- the step kernel decides every interaction with the exact fp64 test -- and the exported
 * list (k_export_pairs) applies the reference's fp64 test to the build positions. */
__global__ void __launch_bounds__(ROWS_BLOCK) k_rows(NlistArrays A)
{
    if (!A.ctl->doing_rebuild) return;
    __shared__ DParams P;
    __shared__ DGrid G;
    __shared__ uint32_t runs[27][ROWS_BLOCK];   // [cell][thread] keys of the non-empty surrounding cells
    __shared__ uint32_t hits[ROWS_QUEUE][ROWS_BLOCK];  // [entry][thread] listed slots not yet written out
    __shared__ uint8_t bitpos[3][32];            // position of bit i of axis d inside morton_key()
    for (uint32_t i = threadIdx.x; i < sizeof(DParams) / 4; i += blockDim.x)
        ((uint32_t *) &P)[i] = ((const uint32_t *) A.params)[i];
    for (uint32_t i = threadIdx.x; i < sizeof(DGrid) / 4; i += blockDim.x)
        ((uint32_t *) &G)[i] = ((const uint32_t *) &A.ctl->grid)[i];
    if (threadIdx.x == 0) {
        const int32_t *bits = A.ctl->grid.bits;
        int pos = 0;
        int const maxb = max(bits[0], max(bits[1], bits[2]));
        for (int i = 0; i < maxb; i++)
            for (int d = 0; d < 3; d++)
                if (i < bits[d]) bitpos[d][i] = (uint8_t) pos++;
    }
    __syncthreads();
    double const wide = G.radius + G.shadow_eps;
    float const filter2 = __double2float_ru(wide * wide * (1.0 + 4e-6));
    float const scale[3] = {G.shadow_scale[0], G.shadow_scale[1], G.shadow_scale[2]};
    bool const periodic = P.box_kind == BD_BOX_PERIODIC;
    uint32_t const tid = threadIdx.x;

    unsigned long long tot = 0;
    uint32_t longest = 0, overflowed = 0;
    // every thread of the block runs the same number of outer iterations (the loop has no barrier,
    // but the warp reductions below want whole warps)
    for (uint32_t s0 = blockIdx.x * blockDim.x; s0 < A.ntot; s0 += gridDim.x * blockDim.x) {
        uint32_t const s = s0 + tid;
        if (s >= A.ntot) continue;
        bead4 const me = A.bpos[s];
        float4 const ms = A.shadow[s];
        A.pos_out[s] = me;
        uint32_t const replica = s / A.n;
        int c[3];
        cell_coords(P, G, me.x, me.y, me.z, c);

        // per axis: up to three neighbouring cell coordinates, already dilated to Morton bits
        uint32_t dil[3][3];
        int cnt[3];
#pragma unroll
        for (int d = 0; d < 3; d++) {
            int const n = G.n[d];
            int v[3];
            int m = 0;
            if (periodic) {
                if (n == 1) {
                    v[m++] = c[d];
                } else if (n == 2) {
                    v[m++] = c[d];
                    v[m++] = c[d] ^ 1;
                } else {
                    v[m++] = c[d] == 0 ? n - 1 : c[d] - 1;
                    v[m++] = c[d];
                    v[m++] = c[d] == n - 1 ? 0 : c[d] + 1;
                }
            } else {
                if (c[d] > 0) v[m++] = c[d] - 1;
                v[m++] = c[d];
                if (c[d] < n - 1) v[m++] = c[d] + 1;
            }
            cnt[d] = m;
#pragma unroll
            for (int i = 0; i < 3; i++) {
                uint32_t key = 0;
                if (i < m)
                    for (int bit = 0; bit < G.bits[d]; bit++) key |= (((uint32_t) v[i] >> bit) & 1u) << bitpos[d][bit];
                dil[d][i] = key;
            }
        }
        // look the cells up (lockstep: no divergence), keep the non-empty ones
        uint32_t nruns = 0;
#pragma unroll
        for (int iz = 0; iz < 3; iz++)
#pragma unroll
            for (int iy = 0; iy < 3; iy++)
#pragma unroll
                for (int ix = 0; ix < 3; ix++) {
                    if (ix < cnt[0] && iy < cnt[1] && iz < cnt[2]) {
                        uint32_t const key = (replica << G.totalbits) | dil[0][ix] | dil[1][iy] | dil[2][iz];
                        uint32_t const b = A.cell_start[key], e = A.cell_start[key + 1];
                        if (b < e) runs[nruns++][tid] = key;
                    }
                }

        // one flat walk over all candidates
        uint4 *dst = A.nbr4 + (size_t) replica * A.npad + (s - replica * A.n);  // natural position; k_balance permutes
        uint32_t count = 0, run = 0, j = 0, e = 0;
        bool over = false;
        if (w_ghost(me.w)) nruns = 0;  // no force is evaluated for a ghost: it needs no row
        // A hit is one predicated store into the thread's queue column; whole groups of ROW_PAD leave the
        // queue together when it fills up.  (Writing each group as it completes costs a divergent branch
        // per candidate -- some lane of the warp nearly always has a hit -- and was 40 % of this kernel.)
        uint32_t qn = 0;
        uint32_t *const queue = &hits[0][tid];
        auto flush = [&](bool last) {
            uint32_t const groups = qn / ROW_PAD;
            for (uint32_t g = 0; g < groups; g++) {
                const uint32_t *q4 = queue + g * ROW_PAD * ROWS_BLOCK;
                if (count < A.cap) {
                    *dst = make_uint4(q4[0], q4[ROWS_BLOCK], q4[2 * ROWS_BLOCK], q4[3 * ROWS_BLOCK]);
                    dst += A.pstride;
                    count += ROW_PAD;
                } else {
                    over = true;
                }
            }
            uint32_t const rem = qn - groups * ROW_PAD;
            if (!last) {
                if (groups)
                    for (uint32_t k = 0; k < rem; k++) queue[k * ROWS_BLOCK] = queue[(groups * ROW_PAD + k) * ROWS_BLOCK];
            } else if (rem) {
                // the last group is padded with the sentinel slot: the step kernel reads whole groups
                const uint32_t *q4 = queue + groups * ROW_PAD * ROWS_BLOCK;
                if (count < A.cap) {
                    *dst = make_uint4(q4[0], rem > 1 ? q4[ROWS_BLOCK] : A.ntot, rem > 2 ? q4[2 * ROWS_BLOCK] : A.ntot, A.ntot);
                    count += rem;
                } else {
                    over = true;
                }
            }
            qn = rem;
        };
        auto consider = [&](uint32_t jj, const float4 &q, bool on) {
            float const d2 = periodic ? shadow_dist2<BD_BOX_PERIODIC>(ms, q, scale)
                                      : shadow_dist2<BD_BOX_OPEN>(ms, q, scale);
            if (on && d2 <= filter2 && jj != s) {
                queue[qn * ROWS_BLOCK] = jj;
                qn++;
            }
        };
        // the bounds of the next run are requested one run ahead (they were just read: cache hits)
        uint32_t nb = 0, ne = 0;
        if (nruns) {
            uint32_t const key = runs[0][tid];
            nb = A.cell_start[key];
            ne = A.cell_start[key + 1];
        }
        while (true) {
            if (j >= e) {
                if (run == nruns) break;
                j = nb;
                e = ne;
                run++;
                if (run < nruns) {
                    uint32_t const key = runs[run][tid];
                    nb = A.cell_start[key];
                    ne = A.cell_start[key + 1];
                }
            }
            // Up to four candidates of the run per iteration, requested together (the walk is bound by
            // the latency of these loads).  One path for every lane: a run with fewer than four beads left
            // repeats its first candidate with the test switched off -- two code paths here had the lanes
            // of a warp take turns (15 of 32 active).
            uint32_t const left = e - j;
            uint32_t const j1 = left > 1u ? j + 1 : j, j2 = left > 2u ? j + 2 : j, j3 = left > 3u ? j + 3 : j;
            float4 const q0 = A.shadow[j], q1 = A.shadow[j1], q2 = A.shadow[j2], q3 = A.shadow[j3];
            consider(j, q0, true);
            consider(j1, q1, left > 1u);
            consider(j2, q2, left > 2u);
            consider(j3, q3, left > 3u);
            j += left < 4u ? left : 4u;
            // write-outs are taken together by the lanes that are here: whoever has whole groups queued
            // writes them when any lane must (early write-outs are always safe)
            if (__any_sync(__activemask(), qn > ROWS_QUEUE - 4)) flush(false);
        }
        flush(true);
        A.nnbr[s] = count;
        overflowed += over;
        tot += count;
        longest = max(longest, count);
    }
    // listed entries (2 x listed pairs), longest row, overflowed rows
    for (int o = 16; o > 0; o >>= 1) {
        tot += __shfl_xor_sync(0xffffffffu, tot, o);
        longest = max(longest, __shfl_xor_sync(0xffffffffu, longest, o));
        overflowed += __shfl_xor_sync(0xffffffffu, overflowed, o);
    }
    if ((threadIdx.x & 31) == 0 && tot) {
        atomicAdd(&A.ctl->listed_entries, tot);
        atomicMax(&A.ctl->max_row, longest);
        if (overflowed) atomicAdd(&A.ctl->overflow, overflowed);
    }
}


// ---- staged rows ------------------------------------------------------------------------------------

constexpr uint32_t STAGE_GAP = 8;           // slots between two cells that one run bridges
constexpr uint32_t ROWS_STAGE_CAP = 2048;   // shadows k_rows_staged itself can hold (32 KB); larger tiles read global

struct RowsStagedShared {
    alignas(16) float4 stage[ROWS_STAGE_CAP + 1];
    alignas(16) DParams P;
    alignas(16) DGrid G;
    alignas(16) TileRuns runs;
    uint32_t hkey[STAGE_HASH];               // cell key, 0xffffffff = empty
    uint32_t hval[STAGE_HASH];               // staged offset of the cell | its bead count << 16
    uint16_t uniq[STAGE_CELLS];              // hash slots of the distinct cells, in arrival order
    uint16_t sorted[STAGE_CELLS];            // ... in ascending key order
    uint32_t cbeg[STAGE_CELLS], cend[STAGE_CELLS];  // slot range of the sorted cells
    uint16_t cells[27][ROWS_BLOCK];          // [k][thread] hash slots of the thread's non-empty surrounding cells
    uint16_t hits[ROWS_QUEUE][ROWS_BLOCK];   // [entry][thread] listed staged positions not yet written out
    uint8_t bitpos[3][32];
    uint32_t nuniq, bad;
    alignas(8) uint64_t bar;
};

/* ELL rows of 16-bit staged positions + the run table of every tile (bd_device.cuh: tile staging).
 * One block per tile of ROWS_BLOCK slot-consecutive beads of one replica:
 *   A. every thread looks up the (at most 27) cells around its bead and the block collects the distinct
 *      non-empty ones in a shared hash table;
 *   B. the distinct cells are sorted by key -- which is slot order -- their slot ranges concatenated into the
 *      staged array, adjacent ranges merged into runs, and the run table written out for the step kernel;
 *   C. the shadows of the runs are copied into shared memory (bulk asynchronous copies);
 *   D. every thread walks the candidates of its cells in the staged copy and lists those within the list
 *      radius, as positions in the staged array.
 * The listed SET is what k_rows lists: the same candidates, the same single-precision test. */
__global__ void __launch_bounds__(ROWS_BLOCK) k_rows_staged(NlistArrays A)
{
    if (!A.ctl->doing_rebuild) return;
    extern __shared__ __align__(16) unsigned char rows_smem[];
    RowsStagedShared &sh = *reinterpret_cast<RowsStagedShared *>(rows_smem);
    uint32_t const tid = threadIdx.x;
    for (uint32_t i = tid; i < sizeof(DParams) / 4; i += blockDim.x) ((uint32_t *) &sh.P)[i] = ((const uint32_t *) A.params)[i];
    for (uint32_t i = tid; i < sizeof(DGrid) / 4; i += blockDim.x) ((uint32_t *) &sh.G)[i] = ((const uint32_t *) &A.ctl->grid)[i];
    if (tid == 0) {
        const int32_t *bits = A.ctl->grid.bits;
        int pos = 0;
        int const maxb = max(bits[0], max(bits[1], bits[2]));
        for (int i = 0; i < maxb; i++)
            for (int d = 0; d < 3; d++)
                if (i < bits[d]) sh.bitpos[d][i] = (uint8_t) pos++;
        mbar_init(&sh.bar, 1);
    }
    __syncthreads();
    const DParams &P = sh.P;
    const DGrid &G = sh.G;
    double const wide = G.radius + G.shadow_eps;
    float const filter2 = __double2float_ru(wide * wide * (1.0 + 4e-6));
    float const scale[3] = {G.shadow_scale[0], G.shadow_scale[1], G.shadow_scale[2]};
    bool const periodic = P.box_kind == BD_BOX_PERIODIC;
    uint32_t const chunks = A.npad / ROWS_BLOCK;
    uint32_t parity = 0;

    unsigned long long tot = 0;
    uint32_t longest = 0, overflowed = 0;
    for (uint32_t tile = blockIdx.x; tile < chunks * A.replicas; tile += gridDim.x) {
        uint32_t const chunk = tile % chunks, replica = tile / chunks;
        uint32_t const local = chunk * ROWS_BLOCK + tid;
        bool const live = local < A.n;
        uint32_t const s = replica * A.n + (live ? local : 0u);
        __syncthreads();  // the previous tile is done with the tables
        for (uint32_t i = tid; i < STAGE_HASH; i += blockDim.x) sh.hkey[i] = 0xffffffffu;
        if (tid == 0) {
            sh.nuniq = 0;
            sh.bad = 0;
        }
        __syncthreads();

        // ---- A: the thread's cells, collected into the block's set ----------------------------------
        bead4 me;
        float4 ms = make_float4(0.f, 0.f, 0.f, 0.f);
        uint32_t ncells = 0, self_key = 0;
        if (live) {
            me = A.bpos[s];
            ms = A.shadow[s];
            A.pos_out[s] = me;
            int c[3];
            cell_coords(P, G, me.x, me.y, me.z, c);
            uint32_t dil[3][3];
            int cnt[3];
#pragma unroll
            for (int d = 0; d < 3; d++) {
                int const n = G.n[d];
                int v[3];
                int m = 0;
                if (periodic) {
                    if (n == 1) {
                        v[m++] = c[d];
                    } else if (n == 2) {
                        v[m++] = c[d];
                        v[m++] = c[d] ^ 1;
                    } else {
                        v[m++] = c[d] == 0 ? n - 1 : c[d] - 1;
                        v[m++] = c[d];
                        v[m++] = c[d] == n - 1 ? 0 : c[d] + 1;
                    }
                } else {
                    if (c[d] > 0) v[m++] = c[d] - 1;
                    v[m++] = c[d];
                    if (c[d] < n - 1) v[m++] = c[d] + 1;
                }
                cnt[d] = m;
#pragma unroll
                for (int i = 0; i < 3; i++) {
                    uint32_t key = 0;
                    if (i < m)
                        for (int bit = 0; bit < G.bits[d]; bit++) key |= (((uint32_t) v[i] >> bit) & 1u) << sh.bitpos[d][bit];
                    dil[d][i] = key;
                }
            }
            {
                uint32_t key = 0;
                for (int d = 0; d < 3; d++)
                    for (int bit = 0; bit < G.bits[d]; bit++) key |= (((uint32_t) c[d] >> bit) & 1u) << sh.bitpos[d][bit];
                self_key = (replica << G.totalbits) | key;
            }
            bool const ghost = w_ghost(me.w);  // no force is evaluated for a ghost: it needs no row
#pragma unroll
            for (int iz = 0; iz < 3; iz++)
#pragma unroll
                for (int iy = 0; iy < 3; iy++)
#pragma unroll
                    for (int ix = 0; ix < 3; ix++) {
                        if (!(ix < cnt[0] && iy < cnt[1] && iz < cnt[2]) || ghost) continue;
                        uint32_t const key = (replica << G.totalbits) | dil[0][ix] | dil[1][iy] | dil[2][iz];
                        if (A.cell_start[key] == A.cell_start[key + 1]) continue;
                        uint32_t h = (key * 2654435761u) >> 22;
                        bool placed = false;
                        for (uint32_t probe = 0; probe < STAGE_HASH; probe++) {
                            uint32_t const old = atomicCAS(&sh.hkey[h], 0xffffffffu, key);
                            if (old == 0xffffffffu) {
                                uint32_t const at = atomicAdd(&sh.nuniq, 1u);
                                if (at < STAGE_CELLS) sh.uniq[at] = (uint16_t) h;
                                placed = true;
                                break;
                            }
                            if (old == key) {
                                placed = true;
                                break;
                            }
                            h = (h + 1) & (STAGE_HASH - 1);
                        }
                        if (!placed) sh.bad = 1;
                        sh.cells[ncells++][tid] = (uint16_t) h;
                    }
        }
        __syncthreads();

        // ---- B: sort the distinct cells by key, lay them out, merge adjacent ones into runs -----------
        uint32_t const nu = min(sh.nuniq, STAGE_CELLS);
        bool bad = sh.bad != 0 || sh.nuniq > STAGE_CELLS;
        for (uint32_t i = tid; i < nu; i += blockDim.x) {
            uint32_t const mine = sh.hkey[sh.uniq[i]];
            uint32_t rnk = 0;
            for (uint32_t u = 0; u < nu; u++) rnk += sh.hkey[sh.uniq[u]] < mine;
            sh.sorted[rnk] = sh.uniq[i];
            sh.cbeg[rnk] = A.cell_start[mine];
            sh.cend[rnk] = A.cell_start[mine + 1];
        }
        __syncthreads();
        if (tid < 32) {
            uint32_t off = 0, nruns = 0;
            for (uint32_t base = 0; base < nu; base += 32) {
                uint32_t const i = base + tid;
                bool const in = i < nu;
                uint32_t const b = in ? sh.cbeg[i] : 0u, e = in ? sh.cend[i] : 0u;
                // a cell opens a new run only when it begins more than STAGE_GAP slots after the previous cell:
                // the few beads in between are staged along (cheaper than one more bulk copy)
                uint32_t const gap = (in && i > 0) ? b - sh.cend[i - 1] : 0u;
                bool const starts = in && (i == 0 || gap > STAGE_GAP);
                uint32_t const count = (e - b) + (starts ? 0u : gap);
                uint32_t incl = count;
                for (int o = 1; o < 32; o <<= 1) {
                    uint32_t const t = __shfl_up_sync(0xffffffffu, incl, o);
                    if ((int) tid >= o) incl += t;
                }
                uint32_t const mask = __ballot_sync(0xffffffffu, starts);
                uint32_t const my_off = off + incl - (e - b);  // the cell itself sits after the bridged gap
                if (in) {
                    sh.hval[sh.sorted[i]] = (my_off & 0xffffu) | ((e - b) << 16);
                    if (starts) {
                        uint32_t const r = nruns + __popc(mask & ((1u << tid) - 1u));
                        if (r < STAGE_RUNS) {
                            sh.runs.start[r] = b;
                            sh.runs.off[r] = my_off;
                        }
                    }
                }
                off += __shfl_sync(0xffffffffu, incl, 31);
                nruns += __popc(mask);
            }
            if (tid == 0) {
                if (nruns > STAGE_RUNS || off > STAGE_MAX_BEADS) sh.bad = 1;
                sh.runs.nruns = min(nruns, STAGE_RUNS);
                sh.runs.total = off;
            }
        }
        __syncthreads();
        bad = bad || sh.bad != 0;
        uint32_t const total = bad ? 0u : sh.runs.total;
        if (bad && tid == 0) {
            sh.runs.nruns = 0;
            sh.runs.total = 0;
            atomicAdd(&A.ctl->stage_overflow, 1u);
        }
        __syncthreads();
        {
            const uint32_t *src = reinterpret_cast<const uint32_t *>(&sh.runs);
            uint32_t *dst = reinterpret_cast<uint32_t *>(A.tile_runs + tile);
            uint32_t const words = 2 + STAGE_RUNS;  // header + starts; the offsets follow
            for (uint32_t i = tid; i < sizeof(TileRuns) / 4; i += blockDim.x)
                if (i < 2 + sh.runs.nruns || (i >= words && i < words + sh.runs.nruns)) dst[i] = src[i];
        }

        // ---- C: the shadows of the runs into shared memory ------------------------------------------
        bool const in_smem = total <= ROWS_STAGE_CAP;
        if (in_smem) {
            stage_runs(sh.stage, A.shadow, sh.runs, &sh.bar, parity, A.ntot);
            parity ^= 1u;
        }
        __syncthreads();

        // ---- D: one flat walk over the candidates of the thread's cells -------------------------------
        uint32_t count = 0;
        bool over = false;
        if (live && !bad) {
            uint2 *dst = reinterpret_cast<uint2 *>(A.nbr4) + (size_t) replica * A.npad + local;  // natural position; k_balance permutes
            uint32_t qn = 0;
            uint16_t *const queue = &sh.hits[0][tid];
            auto flush = [&](bool last) {
                uint32_t const groups = qn / ROW_PAD;
                for (uint32_t g = 0; g < groups; g++) {
                    const uint16_t *q4 = queue + g * ROW_PAD * ROWS_BLOCK;
                    if (count < A.cap) {
                        *dst = make_uint2((uint32_t) q4[0] | ((uint32_t) q4[ROWS_BLOCK] << 16),
                                          (uint32_t) q4[2 * ROWS_BLOCK] | ((uint32_t) q4[3 * ROWS_BLOCK] << 16));
                        dst += A.pstride;
                        count += ROW_PAD;
                    } else {
                        over = true;
                    }
                }
                uint32_t const rem = qn - groups * ROW_PAD;
                if (!last) {
                    if (groups)
                        for (uint32_t k = 0; k < rem; k++) queue[k * ROWS_BLOCK] = queue[(groups * ROW_PAD + k) * ROWS_BLOCK];
                } else if (rem) {
                    // the last group is padded with the padding position: the step kernel reads whole groups
                    const uint16_t *q4 = queue + groups * ROW_PAD * ROWS_BLOCK;
                    uint32_t const pad = total;
                    if (count < A.cap) {
                        uint32_t const e1 = rem > 1 ? q4[ROWS_BLOCK] : pad, e2 = rem > 2 ? q4[2 * ROWS_BLOCK] : pad;
                        *dst = make_uint2((uint32_t) q4[0] | (e1 << 16), e2 | (pad << 16));
                        count += rem;
                    } else {
                        over = true;
                    }
                }
                qn = rem;
            };
            // position of this bead itself in the staged array
            uint32_t self = 0xffffffffu;
            uint32_t run = 0, j = 0, e = 0, jslot = 0;
            while (true) {
                if (j >= e) {
                    if (run == ncells) break;
                    uint32_t const hs = sh.cells[run][tid];
                    uint32_t const v = sh.hval[hs];
                    j = v & 0xffffu;
                    e = j + (v >> 16);
                    jslot = A.cell_start[sh.hkey[hs]];
                    if (sh.hkey[hs] == self_key) self = j + (s - jslot);
                    run++;
                }
                uint32_t const left = e - j;
                uint32_t const j1 = left > 1u ? j + 1 : j, j2 = left > 2u ? j + 2 : j, j3 = left > 3u ? j + 3 : j;
                float4 q0, q1, q2, q3;
                if (in_smem) {
                    q0 = sh.stage[j]; q1 = sh.stage[j1]; q2 = sh.stage[j2]; q3 = sh.stage[j3];
                } else {
                    const float4 *g = A.shadow + jslot;
                    q0 = g[0]; q1 = g[j1 - j]; q2 = g[j2 - j]; q3 = g[j3 - j];
                }
                auto consider = [&](uint32_t jj, const float4 &q, bool on) {
                    float const d2 = periodic ? shadow_dist2<BD_BOX_PERIODIC>(ms, q, scale)
                                              : shadow_dist2<BD_BOX_OPEN>(ms, q, scale);
                    if (on && d2 <= filter2 && jj != self) {
                        queue[qn * ROWS_BLOCK] = (uint16_t) jj;
                        qn++;
                    }
                };
                consider(j, q0, true);
                consider(j1, q1, left > 1u);
                consider(j2, q2, left > 2u);
                consider(j3, q3, left > 3u);
                uint32_t const adv = left < 4u ? left : 4u;
                j += adv;
                jslot += adv;
                if (__any_sync(__activemask(), qn > ROWS_QUEUE - 4)) flush(false);
            }
            flush(true);
        }
        if (live) A.nnbr[s] = count;
        overflowed += over;
        tot += count;
        longest = max(longest, count);
        if (tid == 0 && !bad) atomicMax(&A.ctl->max_stage, total);
    }
    for (int o = 16; o > 0; o >>= 1) {
        tot += __shfl_xor_sync(0xffffffffu, tot, o);
        longest = max(longest, __shfl_xor_sync(0xffffffffu, longest, o));
        overflowed += __shfl_xor_sync(0xffffffffu, overflowed, o);
    }
    if ((threadIdx.x & 31) == 0 && tot) {
        atomicAdd(&A.ctl->listed_entries, tot);
        atomicMax(&A.ctl->max_row, longest);
        if (overflowed) atomicAdd(&A.ctl->overflow, overflowed);
    }
}

/* one bond of bead `oi` (CSR entry {partner original index, ff | second << 8}) as the step kernel wants it:
 * {partner slot, row of the compiled bond table | ff << 12 | second << 15} -- classes do not change between
 * list builds, so the (class_first, class_second) cell of the table is resolved here, once */
__device__ inline uint2 resolve_bond(const NlistArrays &A, uint32_t replica, uint32_t oi, uint2 adj)
{
    uint32_t const f = adj.y & 0xffu, second = (adj.y >> 8) & 1u;
    uint32_t const ti = A.type[oi], tj = A.type[adj.x];
    uint32_t const cell = second ? tj * A.ntypes + ti : ti * A.ntypes + tj;
    uint32_t const row = A.bond_slot[f] * A.ntypes * A.ntypes + cell;
    return make_uint2(A.slot_of[replica * A.n + adj.x], (row & 0xfffu) | (f << 12) | (second << 15));
}

/* the bending triples of bead `oi`, for the thread at position P: slots of the other two beads (in triple order),
 * energy, role of this bead; what the step kernel reads instead of walking the CSR */
__device__ inline void fill_bend_rows(const NlistArrays &A, uint32_t replica, uint32_t oi, bool live, size_t P)
{
    uint32_t k = 0;
    if (live) {
        for (uint32_t c = A.bend_start[oi]; c < A.bend_start[oi + 1] && k < A.bend_deg; c++, k++) {
            uint32_t const adj = A.bend_adj[c];
            uint32_t const t = adj >> 2, role = adj & 3u;
            uint32_t const i0 = A.triples[3 * t], i1 = A.triples[3 * t + 1], i2 = A.triples[3 * t + 2];
            uint32_t const p = role == 0 ? i1 : i0, q = role == 2 ? i1 : i2;
            size_t const at = (size_t) k * A.pstride + P;
            A.bend_pq[at] = make_uint2(A.slot_of[replica * A.n + p], A.slot_of[replica * A.n + q]);
            A.bend_e[at] = A.bend_energy[t];
            A.bend_meta[at] = (uint8_t) (1u | (role << 1) | ((uint32_t) A.bend_ff[t] << 3));
        }
    }
    for (; k < A.bend_deg; k++) A.bend_meta[(size_t) k * A.pstride + P] = 0;
}

/* Thread -> bead assignment of the step kernels.  Inside every tile of BAL_TILE consecutive slots
 * the beads are ranked by row length, so that the 32 rows a warp walks in lockstep end together (row
 * lengths inside one tile of a polymer melt differ by 4x; unranked, half of the lanes idle).  The
 * tile's Verlet rows are then moved to their thread positions, so that the step kernel reads them
 * with one coalesced 16-byte load per lane and group, and the bonded partners of every bead are
 * resolved to slots (the step kernel then needs neither the CSR nor slot_of).
 * One tile per block iteration, BAL_TILE threads. */
__global__ void __launch_bounds__(BAL_TILE) k_balance(NlistArrays A)
{
    if (!A.ctl->doing_rebuild) return;
    __shared__ uint32_t len[BAL_TILE];
    __shared__ uint8_t src[BAL_TILE];
    __shared__ uint32_t longest;
    uint32_t const chunks = A.npad / BAL_TILE;
    for (uint32_t tile = blockIdx.x; tile < chunks * A.replicas; tile += gridDim.x) {
        uint32_t const chunk = tile % chunks, replica = tile / chunks;
        uint32_t const local = chunk * BAL_TILE + threadIdx.x;
        uint32_t const base = replica * A.n + chunk * BAL_TILE;     // slot of the tile's first bead
        uint32_t const pbase = replica * A.npad + chunk * BAL_TILE;  // position of the tile's first thread
        bool const live = local < A.n;
        uint32_t const mine = live ? A.nnbr[base + threadIdx.x] : 0xffffffffu;  // dead lanes rank last
        __syncthreads();
        if (threadIdx.x == 0) longest = 0;
        len[threadIdx.x] = mine;
        __syncthreads();
        // A.balance = 0 (default since round 2): threads keep slot order, rows stay where k_rows wrote them.  With
        // the pair walk bound by L1 gathers instead of issue slots, lanes that end together buy nothing
        // (measured: 0.718 vs 0.721 ms per launch), and the row move below was 0.35 ms of every rebuild.
        uint32_t rnk = threadIdx.x;
        if (A.balance) {
            rnk = 0;
            for (uint32_t t = 0; t < BAL_TILE; t++) {
                uint32_t const other = len[t];
                rnk += (other < mine) || (other == mine && t < threadIdx.x);
            }
        }
        src[rnk] = (uint8_t) threadIdx.x;
        if (live) {
            A.order[base + rnk] = (uint8_t) threadIdx.x;
            atomicMax(&longest, mine);
        }
        __syncthreads();
        // thread t now owns the bead at offset src[t]: move that bead's row to column pbase + t
        uint32_t const from = src[threadIdx.x];
        bool const owner = chunk * BAL_TILE + from < A.n;  // dead offsets were ranked last
        // (only the groups the row has: what lies beyond a row's padded length is never read)
        uint32_t const groups = (longest + ROW_PAD - 1) / ROW_PAD;
        uint32_t const mine_groups = owner ? (len[from] + ROW_PAD - 1) / ROW_PAD : 0u;
        if (!A.balance) {
            // identity: nothing to move
        } else if (A.staged) {
            uint2 *rows = reinterpret_cast<uint2 *>(A.nbr4);  // groups of four 16-bit staged positions
            for (uint32_t g = 0; g < groups; g++) {
                uint2 v = make_uint2(0u, 0u);
                if (g < mine_groups) v = rows[(size_t) g * A.pstride + pbase + from];
                __syncthreads();
                if (g < mine_groups) rows[(size_t) g * A.pstride + pbase + threadIdx.x] = v;
            }
        } else {
            for (uint32_t g = 0; g < groups; g++) {
                uint4 v = make_uint4(0u, 0u, 0u, 0u);
                if (g < mine_groups) v = A.nbr4[(size_t) g * A.pstride + pbase + from];
                __syncthreads();
                if (g < mine_groups) A.nbr4[(size_t) g * A.pstride + pbase + threadIdx.x] = v;
            }
        }
        A.plen[pbase + threadIdx.x] = owner ? len[from] : 0u;
        if (A.bond_deg) {
            uint32_t k = 0;
            if (owner) {
                uint32_t const oi = w_orig(A.bpos[base + from].w);
                for (uint32_t c = A.bond_start[oi]; c < A.bond_start[oi + 1] && k < A.bond_deg; c++, k++)
                    A.bondrow[(size_t) k * A.pstride + pbase + threadIdx.x] = resolve_bond(A, replica, oi, A.bond_adj[c]);
            }
            for (; k < A.bond_deg; k++)
                A.bondrow[(size_t) k * A.pstride + pbase + threadIdx.x] = make_uint2(0xffffffffu, 0u);
        }
        if (A.bend_deg)
            fill_bend_rows(A, replica, owner ? w_orig(A.bpos[base + from].w) : 0u, owner, (size_t) pbase + threadIdx.x);
    }
}

/* The slot-space bond rows alone, for the current assignment of beads to thread positions: what k_balance
 * writes at a list build, redone when only the bonded pair lists changed (bd_set_bond_pairs: the glue /
 * loop lists of 3-sim-1kb are edited every few steps, the Verlet rows are not affected). */
__global__ void __launch_bounds__(BAL_TILE) k_bond_rows(NlistArrays A)
{
    uint32_t const chunks = A.npad / BAL_TILE;
    for (uint32_t tile = blockIdx.x; tile < chunks * A.replicas; tile += gridDim.x) {
        uint32_t const chunk = tile % chunks, replica = tile / chunks;
        uint32_t const base = replica * A.n + chunk * BAL_TILE;
        uint32_t const pbase = replica * A.npad + chunk * BAL_TILE;
        bool const live = chunk * BAL_TILE + threadIdx.x < A.n;
        uint32_t k = 0;
        if (live) {
            uint32_t const from = A.order[base + threadIdx.x];
            uint32_t const oi = w_orig(A.bpos[base + from].w);
            for (uint32_t c = A.bond_start[oi]; c < A.bond_start[oi + 1] && k < A.bond_deg; c++, k++)
                A.bondrow[(size_t) k * A.pstride + pbase + threadIdx.x] = resolve_bond(A, replica, oi, A.bond_adj[c]);
        }
        for (; k < A.bond_deg; k++) A.bondrow[(size_t) k * A.pstride + pbase + threadIdx.x] = make_uint2(0xffffffffu, 0u);
        if (A.bend_deg)
            fill_bend_rows(A, replica, live ? w_orig(A.bpos[base + A.order[base + threadIdx.x]].w) : 0u, live,
                           (size_t) pbase + threadIdx.x);
    }
}

/* Verlet list as unordered pairs of original indices (validation, get_neighbor_list) */
__global__ void __launch_bounds__(NL_BLOCK) k_export_pairs(NlistArrays A, uint32_t replica,
                                                           uint32_t *pairs_out,
                                                           unsigned long long cap,
                                                           unsigned long long *count)
{
    uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.n) return;
    uint32_t const tile = i / BAL_TILE * BAL_TILE;
    uint32_t const s = replica * A.n + tile + A.order[replica * A.n + i];
    uint32_t const p = replica * A.npad + i;
    bead4 const me = A.bpos[s];
    uint32_t const oi = w_orig(me.w);
    double const radius = A.ctl->grid.radius;
    double const r2max = __dmul_rn(radius, radius);
    uint32_t const m = A.plen[p];
    uint32_t const tile_id = replica * (A.npad / BAL_TILE) + i / BAL_TILE;
    for (uint32_t k = 0; k < m; k++) {
        uint32_t const j = row_entry_slot(A.nbr4, A.tile_runs, A.staged, A.pstride, p, tile_id, k);
        bead4 const o = A.bpos[j];
        uint32_t const oj = w_orig(o.w);
        // the stored rows are a superset by the shadow error bound: apply the reference's test
        d3 const r = displacement(*A.params, me.x, me.y, me.z, o.x, o.y, o.z);
        if (sqnorm_exact(r.x, r.y, r.z) > r2max) continue;
        if (oi < oj) {
            unsigned long long const at = atomicAdd(count, 1ull);
            if (at < cap) {
                pairs_out[2 * at] = oi;
                pairs_out[2 * at + 1] = oj;
            }
        }
    }
}

}  // namespace bd
