/*
 * bd_engine.cu -- host side of libbd_b200.so: the C ABI of include/bd_cuda.h.
 *
 * One bd_engine is one md::system (or an ensemble of R replicas of it) whose state lives in HBM.
 * The host keeps only parameter tables and topology; every per-bead quantity is device-resident
 * and a run of K steps is K enqueues of [gated rebuild chain] -> [fused step kernel] with no
 * host synchronisation in between.  There is no CPU fallback: without a CUDA device every entry
 * point fails.
 */
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <cub/cub.cuh>

#include "../../include/bd_cuda.h"
#include "bd_device.cuh"
#include "bd_nlist.cuh"
#include "bd_step.cuh"

using namespace bd;

static_assert(BAL_TILE == STEP_BLOCK, "k_balance ranks the beads of one step-kernel block");
static_assert(ROWS_BLOCK == STEP_BLOCK, "k_rows_staged writes the run table of one step-kernel tile");
constexpr uint32_t STAGE_CAP_CEILING = 2048;  // shadows a step-kernel block stages at most (32 KB)

// ---- errors --------------------------------------------------------------------------------------

static thread_local std::string g_error;

static int fail(const std::string &msg)
{
    g_error = msg;
    return 1;
}

#define CK(call)                                                                                  \
    do {                                                                                          \
        cudaError_t const e_ = (call);                                                            \
        if (e_ != cudaSuccess) {                                                                  \
            return fail(std::string(#call) + ": " + cudaGetErrorString(e_));                      \
        }                                                                                         \
    } while (0)

extern "C" const char *bd_last_error(void)
{
    return g_error.c_str();
}

// ---- engine state --------------------------------------------------------------------------------

template <typename T>
struct DevBuf {
    T *p = nullptr;
    size_t count = 0;
    cudaError_t alloc(size_t n)
    {
        release();
        count = n;
        return cudaMalloc((void **) &p, std::max<size_t>(n, 1) * sizeof(T));
    }
    /* keep the allocation when it is already large enough (arrays that are re-uploaded often) */
    cudaError_t ensure(size_t n)
    {
        if (p && count >= n) return cudaSuccess;
        return alloc(n + n / 4);
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        count = 0;
    }
};

/* a DevBuf local to one call: released when the call returns, whichever way (engine members stay plain DevBufs,
 * released by bd_destroy) */
template <typename T>
struct ScopedBuf : DevBuf<T> {
    ScopedBuf() = default;
    ScopedBuf(const ScopedBuf &) = delete;
    ScopedBuf &operator=(const ScopedBuf &) = delete;
    ~ScopedBuf() { this->release(); }
};

struct HostFF {
    bool used = false;
    int nterms = 0;
    int term_off = 0;
    int cells = 0;  // table rows
};

struct bd_engine {
    int device = 0;
    uint32_t n = 0, replicas = 1, ntot = 0, chunks = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;

    // per-bead device state
    DevBuf<bead4> pos[2], bpos;
    DevBuf<float4> shadow[2];  // single-precision shadow of pos[] (slot Ntot of pos/shadow is the NaN sentinel)
    int cur = 0;
    DevBuf<uint32_t> slot_of, cellkey, rank, stage, cell_count, cell_start, block_sum, nnbr, plen;
    DevBuf<uint4> nbr4;       // Verlet rows by thread position, ROW_PAD indices per group
    DevBuf<TileRuns> tile_runs;  // staged neighbourhood of every tile (bd_device.cuh: tile staging)
    bool staged = false;      // rows hold 16-bit staged positions (option "staged"); false = rows of slots, global gathers
    bool pair_f32 = false;      // option "pair_precision" = 32: pair forces straight from the single-precision shadows
    bool balance_rows = false;  // tile threads ordered by row length (round 1) or in slot order (default)
    bool stage_auto = true;   // the block's staging capacity follows the largest tile (bd_set_option stage_capacity fixes it)
    uint32_t stage_cap = 512; // shadows a block of the step kernel stages
    DevBuf<uint2> bondrow;    // bonded partners by thread position, resolved to slots at every rebuild
    DevBuf<uint2> bend_pq;    // bending triples by thread position: slots of the other two beads,
    DevBuf<double> bend_e;    // ... the triple's energy,
    DevBuf<uint8_t> bend_meta;  // ... role and forcefield (bd_nlist.cuh: fill_bend_rows)
    uint32_t bend_deg = 0;
    DevBuf<BondRow> d_bond_rows;
    uint32_t npad = 0, pstride = 0, bond_deg = 0;
    bool bond_rows_stale = false;  // bonded pair lists changed since the bond rows were written
    bool bonds_compiled = false;
    int nbond_rows = 0;
    uint8_t bond_slot[BD_MAX_BOND_FF] = {};
    uint32_t cell_cap = 0, max_bits = 0, cap = 32;
    DevBuf<DControl> ctl;
    DevBuf<DParams> d_params;
    DevBuf<DWallGeom> wall_geom;  // [replicas][BD_MAX_WALL_FF], the master copy of every wall's geometry
    WallRelax relax = {};         // device-resident semiaxis relaxation (bd_set_wall_relaxation)
    bool wall_stats_fresh = false;  // wall_out holds the reduction of the latest force evaluation
    DevBuf<DTerm> d_terms;
    DevBuf<PairRow> d_pair_rows;
    int prog = PROG_GENERIC;   // specialised pair program of the current tables (upload_params)
    bool k2_uniform = false;   // ... whose rows all share k2a / k2b
    double k2a = 0, k2b = 0;
    int fast_ff = 0;
    bool specialize = true;
    DevBuf<double> force, wprev, xyz_stage, wall_partial, wall_out, energy_partial, energy_out, dt_prev, dt_last;
    DevBuf<unsigned long long> inter_partial, dt_bits, pair_count;
    DevBuf<uint8_t> d_type, d_pair_mask, d_point_mask, order, d_owned;
    DevBuf<double> replay, xyz_all, xyz_out, ext;
    bool ext_on = false;

    // contact map accumulated on the device (bd_contact_*)
    bd_engine *contact_search = nullptr;  // a second engine used as the neighbour searcher at the contact distance
    double contact_distance = 0;
    DevBuf<unsigned long long> contact_keys[2], contact_new;  // (i << 32 | j), sorted unique; double-buffered
    DevBuf<uint32_t> contact_counts[2], contact_ones, contact_pairs;
    DevBuf<unsigned char> contact_temp;
    DevBuf<unsigned long long> contact_runs;
    int contact_cur = 0;
    size_t contact_size = 0;
    // single-precision snapshots (bd_snapshot_begin / bd_snapshot_wait)
    DevBuf<float> snap_dev;
    float *snap_host = nullptr;  // pinned
    size_t snap_host_count = 0, snap_count = 0;
    cudaStream_t snap_stream = nullptr;
    cudaEvent_t snap_ready = nullptr, snap_done = nullptr;
    bool snap_pending = false;

    // topology (device CSR by original index)
    DevBuf<uint32_t> bond_start, bend_start, bend_adj, triples;
    DevBuf<uint2> bond_adj;
    DevBuf<double> bend_energy;
    DevBuf<uint8_t> bend_ff;

    // host mirrors
    DParams hp;
    std::vector<DTerm> hterms;
    bool params_dirty = true;
    HostFF pair[BD_MAX_PAIR_FF], bond[BD_MAX_BOND_FF], wall[BD_MAX_WALL_FF], point[BD_MAX_POINT_FF];
    bool bend_used[BD_MAX_BEND_FF] = {};
    std::vector<uint8_t> h_type, h_pair_mask, h_point_mask;
    std::vector<std::vector<uint32_t>> bond_pairs{BD_MAX_BOND_FF};
    std::vector<std::vector<uint32_t>> bend_triples{BD_MAX_BEND_FF};
    std::vector<std::vector<double>> bend_energies{BD_MAX_BEND_FF};
    bool has_types_locked = false;

    // run state
    double last_dt = 0, last_kT = 0;
    bool wprev_valid = false;
    int64_t steps_done = 0;
    int64_t launches = 0;
    bool stats_adaptive_last = false;
    bool list_verified = false;  // rows were checked for overflow since the last host-side change
    int64_t capacity_rebuilds = 0;  // builds repeated only because the ELL rows were re-sized

    // one BD step as a CUDA graph per ping-pong side (step_graph below)
    struct StepGraph {
        cudaGraph_t graph = nullptr;
        cudaGraphExec_t exec = nullptr;
        cudaGraphNode_t step_node = nullptr;
        cudaKernelNodeParams step_params = {};
        std::vector<unsigned char> signature;  // everything the nodes were built from
    } graphs[2];
    bool use_graph = true;
    int64_t graph_steps = 0;

    // optional CUDA-event timing of the step kernel alone (bench.py roofline)
    bool time_kernel = false;
    uint32_t replica_base = 0, replica_stride = 1;  // noise key of local replica r (bd_set_option)
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> kernel_events;
};

static int sync_and_check(bd_engine *h);
static void opt_in_shared_memory();
static void destroy_step_graph(bd_engine::StepGraph &g);
static size_t table_bytes(uint32_t ntypes2, uint32_t nbond_rows);

// ---- small helpers -------------------------------------------------------------------------------

static inline unsigned blocks_for(size_t count, unsigned block)
{
    return (unsigned) std::max<size_t>(1, (count + block - 1) / block);
}

static DTerm make_dterm(const bd_term &t)
{
    DTerm d;
    std::memset(&d, 0, sizeof d);
    d.kind = t.kind;
    d.p = t.p;
    d.q = t.q;
    d.factor = t.factor;
    d.energy = t.energy;
    d.length = t.length;
    d.cutoff2 = t.cutoff * t.cutoff;
    d.k2 = t.length != 0.0 ? 1.0 / (t.length * t.length) : 0.0;
    return d;
}

static int check_terms(const bd_term *table, size_t count)
{
    for (size_t i = 0; i < count; i++) {
        const bd_term &t = table[i];
        if (t.kind < BD_TERM_NONE || t.kind > BD_TERM_SEMISPRING) return fail("bd: unknown potential kind");
        if (t.kind == BD_TERM_SOFTCORE && (t.p < 2 || t.q < 1 || t.p > 64 || t.q > 64))
            return fail("bd: softcore exponents out of range (need p >= 2, q >= 1)");
        if (t.kind == BD_TERM_SOFTWELL && (t.p < 2 || t.p > 64)) return fail("bd: softwell exponent out of range");
        if ((t.kind == BD_TERM_SOFTCORE || t.kind == BD_TERM_SOFTWELL) && !(t.length > 0))
            return fail("bd: softcore/softwell length must be positive");
    }
    return 0;
}

static int store_table(bd_engine *h, HostFF &ff, int nterms, const bd_term *table, int cells)
{
    if (nterms < 1 || nterms > BD_MAX_TERMS) return fail("bd: nterms out of range");
    if (check_terms(table, (size_t) cells * nterms)) return 1;
    if (!ff.used) {
        ff.used = true;
        ff.nterms = nterms;
        ff.cells = cells;
        ff.term_off = (int) h->hterms.size();
        h->hterms.resize(h->hterms.size() + (size_t) cells * nterms);
    } else if (ff.nterms != nterms || ff.cells != cells) {
        return fail("bd: table shape changed");
    }
    for (size_t i = 0; i < (size_t) cells * nterms; i++) h->hterms[ff.term_off + i] = make_dterm(table[i]);
    h->params_dirty = true;
    h->has_types_locked = true;
    return 0;
}

static void refresh_dcut_max(bd_engine *h)
{
    double m = 0;
    for (int k = 0; k < BD_MAX_PAIR_FF; k++)
        if (h->pair[k].used && h->hp.pair[k].enabled) m = std::max(m, h->hp.pair[k].dcut);
    h->hp.dcut_max = m;
}

/* Recognise the pair table's program (bd_step.cuh: PairProgram) and compile its rows.  Anything
 * outside the recognised shapes keeps the interpreter. */
static int compile_pair_program(bd_engine *h, std::vector<PairRow> &rows)
{
    h->prog = PROG_GENERIC;
    if (!h->specialize) return 0;
    int used = 0, ff = -1;
    for (int k = 0; k < BD_MAX_PAIR_FF; k++)
        if (h->pair[k].used && h->hp.pair[k].enabled) {
            used++;
            ff = k;
        }
    if (used != 1 || h->hp.pair[ff].has_targets) return 0;
    int const T = h->hp.ntypes, nterms = h->pair[ff].nterms;
    if (nterms > 2) return 0;
    const DTerm *tab = h->hterms.data() + h->pair[ff].term_off;
    bool second = false;
    rows.assign((size_t) T * T, PairRow{0, 0, 0, 0});  // {ca, cb, k2a, k2b}
    for (int c = 0; c < T * T; c++) {
        for (int k = 0; k < nterms; k++) {
            const DTerm &t = tab[c * nterms + k];
            if (t.kind == BD_TERM_NONE) continue;
            if (t.kind != BD_TERM_SOFTCORE || t.q != 3 || t.cutoff2 != 0.0 || t.p != (k == 0 ? 2 : 8)) return 0;
            double const coef = t.factor * ((double) (t.p * t.q) * t.energy * t.k2);
            if (k == 0) {
                rows[c].k2a = t.k2;
                rows[c].ca = coef;
            } else {
                rows[c].k2b = t.k2;
                rows[c].cb = coef;
                second = true;
            }
        }
    }
    // the specialised kernels index rows by (own class, partner class): the table must be symmetric
    for (int a = 0; a < T; a++)
        for (int b = 0; b < a; b++)
            if (std::memcmp(&rows[a * T + b], &rows[b * T + a], sizeof(PairRow)) != 0) return 0;
    h->prog = second ? PROG_SC23_SC83 : PROG_SC23;
    h->fast_ff = ff;
    // rows with a zero coefficient do not care about their k2 (0 * g^2 is +0 for any finite g)
    h->k2_uniform = true;
    h->k2a = h->k2b = 0.0;
    bool seen_a = false, seen_b = false;
    for (const PairRow &r : rows) {
        if (r.ca != 0.0) {
            if (seen_a && r.k2a != h->k2a) h->k2_uniform = false;
            h->k2a = r.k2a;
            seen_a = true;
        }
        if (r.cb != 0.0) {
            if (seen_b && r.k2b != h->k2b) h->k2_uniform = false;
            h->k2b = r.k2b;
            seen_b = true;
        }
    }
    return 0;
}

/* Bond tables whose cells are one harmonic / spring / semispring term (every reference driver) are
 * compiled into BondRow blocks that the step kernel keeps in shared memory; anything else keeps the
 * interpreter. */
static void compile_bond_rows(bd_engine *h, std::vector<BondRow> &rows)
{
    h->bonds_compiled = false;
    h->nbond_rows = 0;
    rows.clear();
    int const T2 = h->hp.ntypes * h->hp.ntypes;
    int used = 0;
    // the block of every used forcefield is fixed by the order of registration (the list build resolves
    // bonds to rows of these blocks: bd_nlist.cuh resolve_bond), whether or not the tables compile
    for (int f = 0; f < BD_MAX_BOND_FF; f++)
        if (h->bond[f].used) h->bond_slot[f] = (uint8_t) used++;
    if (!h->specialize) return;
    for (int f = 0; f < BD_MAX_BOND_FF; f++) {
        if (!h->bond[f].used) continue;
        const DTerm *tab = h->hterms.data() + h->bond[f].term_off;
        int const nterms = h->bond[f].nterms;
        for (int c = 0; c < T2; c++) {
            BondRow row = {0.0, 0.0, 0.0, 0.0};  // none: factor 0
            int live = 0;
            for (int k = 0; k < nterms; k++) {
                const DTerm &t = tab[c * nterms + k];
                if (t.kind == BD_TERM_NONE) continue;
                if (++live > 1 || t.cutoff2 != 0.0) return;
                if (t.kind != BD_TERM_HARMONIC && t.kind != BD_TERM_SPRING && t.kind != BD_TERM_SEMISPRING) return;
                row.k = t.energy;
                row.factor = t.factor;
                if (t.kind == BD_TERM_HARMONIC) {
                    row.kb = 0.0;
                    row.b2 = -1.0;
                } else {
                    row.kb = t.energy * t.length;  // K * b, as (K * b) / r is evaluated (spring_potential.hpp:37)
                    row.b2 = t.kind == BD_TERM_SPRING ? 0.0 : t.length * t.length;
                }
            }
            rows.push_back(row);
        }
    }
    if (used == 0 || rows.size() > (size_t) BD_MAX_TYPES * BD_MAX_TYPES) return;
    h->bonds_compiled = true;
    h->nbond_rows = (int) rows.size();
}

static int upload_params(bd_engine *h)
{
    if (!h->params_dirty) return 0;
    refresh_dcut_max(h);
    std::vector<BondRow> brows;
    compile_bond_rows(h, brows);
    if (h->bonds_compiled)
        CK(cudaMemcpyAsync(h->d_bond_rows.p, brows.data(), brows.size() * sizeof(BondRow), cudaMemcpyHostToDevice,
                           h->stream));
    std::vector<PairRow> rows;
    compile_pair_program(h, rows);
    if (h->prog != PROG_GENERIC && h->k2_uniform) {
        // the kernel reads uniform tables as ca[T*T] followed by cb[T*T] (bd_step.cuh: pair_forces)
        std::vector<PairRow> packed(rows.size(), PairRow{0, 0, 0, 0});
        double *flat = reinterpret_cast<double *>(packed.data());
        for (size_t c = 0; c < rows.size(); c++) {
            flat[c] = rows[c].ca;
            flat[rows.size() + c] = rows[c].cb;
        }
        rows.swap(packed);
    }
    if (h->prog != PROG_GENERIC)
        CK(cudaMemcpyAsync(h->d_pair_rows.p, rows.data(), rows.size() * sizeof(PairRow), cudaMemcpyHostToDevice,
                           h->stream));
    if (h->d_terms.count < h->hterms.size()) {
        CK(cudaStreamSynchronize(h->stream));
        CK(h->d_terms.alloc(std::max<size_t>(64, h->hterms.size() * 2)));
    }
    if (!h->hterms.empty())
        CK(cudaMemcpyAsync(h->d_terms.p, h->hterms.data(), h->hterms.size() * sizeof(DTerm),
                           cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->d_params.p, &h->hp, sizeof(DParams), cudaMemcpyHostToDevice, h->stream));
    h->params_dirty = false;
    return 0;
}

static int request_rebuild(bd_engine *h)
{
    h->list_verified = false;
    static const uint32_t one = 1;
    CK(cudaMemcpyAsync(&h->ctl.p->need_rebuild, &one, sizeof one, cudaMemcpyHostToDevice, h->stream));
    return 0;
}

static int upload_words(bd_engine *h)
{
    if (h->n == 0) return 0;
    CK(cudaMemcpyAsync(h->d_type.p, h->h_type.data(), h->n, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->d_pair_mask.p, h->h_pair_mask.data(), h->n, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->d_point_mask.p, h->h_point_mask.data(), h->n, cudaMemcpyHostToDevice, h->stream));
    k_set_words<<<blocks_for(h->ntot, 256), 256, 0, h->stream>>>(
        h->pos[0].p, h->pos[1].p, h->bpos.p, h->slot_of.p, h->d_type.p, h->d_pair_mask.p,
        h->d_point_mask.p, h->n, h->ntot);
    h->launches++;
    CK(cudaGetLastError());
    return request_rebuild(h);  // the shadows carry the class (bd_device.cuh: shadow_tag): the next build rewrites them
}

/* CSR adjacency of all bond / bend forcefields by original bead index */
static int upload_topology(bd_engine *h)
{
    uint32_t const n = h->n;
    std::vector<uint32_t> start(n + 1, 0);
    for (int f = 0; f < BD_MAX_BOND_FF; f++)
        for (uint32_t v : h->bond_pairs[f]) start[v + 1]++;
    for (uint32_t i = 0; i < n; i++) start[i + 1] += start[i];
    std::vector<uint2> adj(start[n]);
    std::vector<uint32_t> fill(start.begin(), start.end() - 1);
    for (int f = 0; f < BD_MAX_BOND_FF; f++) {
        const std::vector<uint32_t> &p = h->bond_pairs[f];
        for (size_t k = 0; k + 1 < p.size(); k += 2) {
            uint32_t const i = p[k], j = p[k + 1];
            adj[fill[i]++] = make_uint2(j, (uint32_t) f);
            adj[fill[j]++] = make_uint2(i, (uint32_t) f | 0x100u);
        }
    }
    uint32_t deg = 0;
    for (uint32_t i = 0; i < n; i++) deg = std::max(deg, start[i + 1] - start[i]);
    CK(cudaStreamSynchronize(h->stream));
    h->bond_deg = deg;
    CK(h->bondrow.ensure((size_t) deg * h->pstride));
    h->list_verified = false;
    CK(h->bond_start.ensure(n + 1));
    CK(h->bond_adj.ensure(adj.size()));
    CK(cudaMemcpy(h->bond_start.p, start.data(), (n + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice));
    if (!adj.empty()) CK(cudaMemcpy(h->bond_adj.p, adj.data(), adj.size() * sizeof(uint2), cudaMemcpyHostToDevice));

    std::vector<uint32_t> tri;
    std::vector<double> ten;
    std::vector<uint8_t> tff;
    for (int f = 0; f < BD_MAX_BEND_FF; f++) {
        tri.insert(tri.end(), h->bend_triples[f].begin(), h->bend_triples[f].end());
        ten.insert(ten.end(), h->bend_energies[f].begin(), h->bend_energies[f].end());
        tff.insert(tff.end(), h->bend_energies[f].size(), (uint8_t) f);
    }
    std::fill(start.begin(), start.end(), 0u);
    for (uint32_t v : tri) start[v + 1]++;
    for (uint32_t i = 0; i < n; i++) start[i + 1] += start[i];
    std::vector<uint32_t> badj(start[n]);
    fill.assign(start.begin(), start.end() - 1);
    for (size_t t = 0; t < ten.size(); t++)
        for (uint32_t role = 0; role < 3; role++) badj[fill[tri[3 * t + role]]++] = (uint32_t) (t << 2) | role;
    uint32_t bdeg = 0;
    for (uint32_t i = 0; i < n; i++) bdeg = std::max(bdeg, start[i + 1] - start[i]);
    h->bend_deg = bdeg;
    CK(h->bend_pq.ensure((size_t) bdeg * h->pstride));
    CK(h->bend_e.ensure((size_t) bdeg * h->pstride));
    CK(h->bend_meta.ensure((size_t) bdeg * h->pstride));
    if (bdeg) CK(cudaMemset(h->bend_meta.p, 0, (size_t) bdeg * h->pstride));
    CK(h->bend_start.alloc(n + 1));
    CK(h->bend_adj.alloc(badj.size()));
    CK(h->triples.alloc(tri.size()));
    CK(h->bend_energy.alloc(ten.size()));
    CK(h->bend_ff.alloc(tff.size()));
    CK(cudaMemcpy(h->bend_start.p, start.data(), (n + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice));
    if (!badj.empty()) {
        CK(cudaMemcpy(h->bend_adj.p, badj.data(), badj.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(h->triples.p, tri.data(), tri.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(h->bend_energy.p, ten.data(), ten.size() * sizeof(double), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(h->bend_ff.p, tff.data(), tff.size(), cudaMemcpyHostToDevice));
    }
    // the slot-space bond rows are rewritten behind the next rebuild chain (k_bond_rows); the Verlet rows
    // do not depend on the bonded lists, so no list build is asked for
    h->bond_rows_stale = true;
    return 0;
}

// ---- lifetime ------------------------------------------------------------------------------------

static int alloc_rows(bd_engine *h)
{
    CK(h->nbr4.alloc((size_t) (h->cap / ROW_PAD) * h->pstride));
    return 0;
}

extern "C" int bd_create(int64_t n_beads, int32_t n_replicas, int32_t device, bd_engine **out)
{
    if (!out) return fail("bd_create: null output");
    *out = nullptr;
    if (n_beads < 0 || n_replicas < 1) return fail("bd_create: bad bead or replica count");
    if ((double) n_beads * n_replicas >= 4.0e9) return fail("bd_create: more than 2^32 beads");
    if (n_replicas > 65535) return fail("bd_create: at most 65535 replicas per engine");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail("bd_create: no CUDA device (this engine has no CPU fallback)");
    if (device < 0 || device >= ndev) return fail("bd_create: bad device ordinal");
    CK(cudaSetDevice(device));

    opt_in_shared_memory();
    bd_engine *h = new bd_engine();
    h->device = device;
    h->n = (uint32_t) n_beads;
    h->replicas = (uint32_t) n_replicas;
    h->ntot = h->n * h->replicas;
    h->chunks = blocks_for(h->n, STEP_BLOCK);
    h->npad = h->chunks * STEP_BLOCK;
    h->pstride = h->npad * h->replicas;
    if (cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete h;
        return fail("bd_create: cannot create a stream");
    }
    h->stream = h->own_stream;

    // Morton bits per replica: about two cells per bead, whole key space within the cell arrays
    uint32_t bits = 6;
    while ((1u << bits) < 2u * std::max<uint32_t>(h->n, 1u) && bits < 24) bits++;
    uint32_t rbits = 0;
    while ((1u << rbits) < h->replicas) rbits++;
    if (bits + rbits > 27) bits = 27 - rbits;
    h->max_bits = bits;
    h->cell_cap = h->replicas << bits;

    std::memset(&h->hp, 0, sizeof h->hp);
    h->hp.ntypes = 1;
    h->hp.box_kind = BD_BOX_OPEN;
    for (int d = 0; d < 3; d++) h->hp.period[d] = h->hp.inv_period[d] = 1.0;
    h->hp.mobility[0] = 1.0;  // md::mobility_attribute default (MD/system.hpp:41-44)
    h->hp.verlet_factor = 1.5;  // MD/forcefield/detail/neighbor_list.hpp:182
    h->h_type.assign(h->n, 0);
    h->h_pair_mask.assign(h->n, 0);
    h->h_point_mask.assign(h->n, 0);

    size_t const nt = h->ntot;
    bool ok = h->pos[0].alloc(nt + 1) == cudaSuccess && h->pos[1].alloc(nt + 1) == cudaSuccess
        && h->shadow[0].alloc(nt + 1) == cudaSuccess && h->shadow[1].alloc(nt + 1) == cudaSuccess
        && h->bpos.alloc(nt) == cudaSuccess && h->slot_of.alloc(nt) == cudaSuccess
        && h->cellkey.alloc(nt) == cudaSuccess && h->rank.alloc(nt) == cudaSuccess
        && h->stage.alloc(nt) == cudaSuccess && h->nnbr.alloc(nt) == cudaSuccess
        && h->cell_count.alloc(h->cell_cap) == cudaSuccess
        && h->cell_start.alloc((size_t) h->cell_cap + 1) == cudaSuccess
        && h->block_sum.alloc(h->cell_cap / SCAN_TILE + 2) == cudaSuccess
        && h->ctl.alloc(1) == cudaSuccess && h->d_params.alloc(1) == cudaSuccess
        && h->wall_geom.alloc((size_t) h->replicas * BD_MAX_WALL_FF) == cudaSuccess
        && h->d_terms.alloc(64) == cudaSuccess && h->d_pair_rows.alloc(BD_MAX_TYPES * BD_MAX_TYPES) == cudaSuccess && h->force.alloc(3 * nt) == cudaSuccess
        && h->xyz_stage.alloc(3 * (size_t) h->n) == cudaSuccess
        && h->wall_partial.alloc((size_t) h->replicas * h->chunks * BD_MAX_WALL_FF * WALL_STATS) == cudaSuccess
        && h->wall_out.alloc((size_t) h->replicas * BD_MAX_WALL_FF * WALL_STATS) == cudaSuccess
        && h->energy_partial.alloc((size_t) h->replicas * h->chunks) == cudaSuccess
        && h->energy_out.alloc(h->replicas) == cudaSuccess
        && h->inter_partial.alloc((size_t) h->replicas * h->chunks * (STEP_BLOCK / 32)) == cudaSuccess
        && h->dt_bits.alloc(h->replicas) == cudaSuccess && h->dt_prev.alloc(h->replicas) == cudaSuccess
        && h->dt_last.alloc(h->replicas) == cudaSuccess && h->pair_count.alloc(1) == cudaSuccess
        && h->d_type.alloc(h->n) == cudaSuccess && h->d_pair_mask.alloc(h->n) == cudaSuccess
        && h->d_point_mask.alloc(h->n) == cudaSuccess && h->order.alloc(nt) == cudaSuccess && h->plen.alloc(h->pstride) == cudaSuccess
        && h->d_bond_rows.alloc(BD_MAX_TYPES * BD_MAX_TYPES) == cudaSuccess
        && h->nbr4.alloc((size_t) (h->cap / ROW_PAD) * h->pstride) == cudaSuccess
        && h->tile_runs.alloc((size_t) h->chunks * h->replicas) == cudaSuccess;
    if (!ok) {
        bd_destroy(h);
        return fail(std::string("bd_create: device allocation failed: ") + cudaGetErrorString(cudaGetLastError()));
    }
    cudaMemsetAsync(h->ctl.p, 0, sizeof(DControl), h->stream);
    cudaMemsetAsync(h->wall_geom.p, 0, h->wall_geom.count * sizeof(DWallGeom), h->stream);
    cudaMemsetAsync(h->plen.p, 0, std::max<size_t>(h->pstride, 1) * sizeof(uint32_t), h->stream);
    cudaMemsetAsync(h->wall_partial.p, 0, h->wall_partial.count * sizeof(double), h->stream);
    cudaMemsetAsync(h->inter_partial.p, 0, h->inter_partial.count * sizeof(unsigned long long), h->stream);
    cudaMemsetAsync(h->dt_last.p, 0, h->replicas * sizeof(double), h->stream);
    cudaMemsetAsync(h->dt_prev.p, 0, h->replicas * sizeof(double), h->stream);
    k_init_identity<<<blocks_for(h->ntot + 1, 256), 256, 0, h->stream>>>(
        h->slot_of.p, h->pos[0].p, h->pos[1].p, h->bpos.p, h->shadow[0].p, h->shadow[1].p, h->order.p, h->n, h->ntot);
    h->launches++;
    if (upload_topology(h) || request_rebuild(h)) {
        bd_destroy(h);
        return 1;
    }
    if (cudaStreamSynchronize(h->stream) != cudaSuccess) {
        bd_destroy(h);
        return fail("bd_create: device initialisation failed");
    }
    *out = h;
    return 0;
}

extern "C" int bd_destroy(bd_engine *h)
{
    if (!h) return 0;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (auto *b : {&h->pos[0], &h->pos[1], &h->bpos}) b->release();
    for (auto *b : {&h->shadow[0], &h->shadow[1]}) b->release();
    for (auto *b : {&h->slot_of, &h->cellkey, &h->rank, &h->stage, &h->cell_count, &h->cell_start,
                    &h->block_sum, &h->plen, &h->nnbr, &h->bond_start, &h->bend_start, &h->bend_adj, &h->triples})
        b->release();
    for (auto *b : {&h->force, &h->wprev, &h->xyz_stage, &h->wall_partial, &h->wall_out, &h->energy_partial,
                    &h->energy_out, &h->dt_prev, &h->dt_last, &h->bend_energy, &h->replay, &h->xyz_all, &h->xyz_out, &h->ext})
        b->release();
    for (auto *b : {&h->inter_partial, &h->dt_bits, &h->pair_count}) b->release();
    for (auto *b : {&h->d_type, &h->d_pair_mask, &h->d_point_mask, &h->bend_ff, &h->order, &h->d_owned}) b->release();
    h->ctl.release();
    h->d_params.release();
    h->wall_geom.release();
    h->d_terms.release();
    h->d_pair_rows.release();
    h->bond_adj.release();
    h->nbr4.release();
    h->tile_runs.release();
    h->bondrow.release();
    h->bend_pq.release();
    h->bend_e.release();
    h->bend_meta.release();
    h->d_bond_rows.release();
    destroy_step_graph(h->graphs[0]);
    destroy_step_graph(h->graphs[1]);
    if (h->contact_search) bd_destroy(h->contact_search);
    for (int k = 0; k < 2; k++) {
        h->contact_keys[k].release();
        h->contact_counts[k].release();
    }
    h->contact_new.release();
    h->contact_ones.release();
    h->contact_pairs.release();
    h->contact_temp.release();
    h->contact_runs.release();
    h->snap_dev.release();
    if (h->snap_host) cudaFreeHost(h->snap_host);
    if (h->snap_ready) cudaEventDestroy(h->snap_ready);
    if (h->snap_done) cudaEventDestroy(h->snap_done);
    if (h->snap_stream) cudaStreamDestroy(h->snap_stream);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
    return 0;
}

// ---- system description --------------------------------------------------------------------------

#define NEED(h)                                                                                   \
    do {                                                                                          \
        if (!(h)) return fail("bd: null engine handle");                                          \
        cudaSetDevice((h)->device);                                                               \
    } while (0)

extern "C" int bd_set_box(bd_engine *h, int32_t box_kind, const double period[3])
{
    NEED(h);
    if (box_kind != BD_BOX_OPEN && box_kind != BD_BOX_PERIODIC) return fail("bd_set_box: unknown box kind");
    h->hp.box_kind = box_kind;
    for (int d = 0; d < 3; d++) {
        double const p = box_kind == BD_BOX_PERIODIC ? period[d] : 1.0;
        if (!(p > 0)) return fail("bd_set_box: periods must be positive");
        h->hp.period[d] = p;
        h->hp.inv_period[d] = 1 / p;  // MD/misc/nsearch_detail/math.hpp:34-38 multiplies by 1/m
    }
    h->params_dirty = true;
    return request_rebuild(h);
}

extern "C" int bd_set_types(bd_engine *h, const uint8_t *type, int32_t ntypes, const double *mobility)
{
    NEED(h);
    if (ntypes < 1 || ntypes > BD_MAX_TYPES) return fail("bd_set_types: ntypes out of range");
    if (h->has_types_locked && ntypes != h->hp.ntypes)
        return fail("bd_set_types: the number of classes cannot change after forcefields were added");
    for (uint32_t i = 0; i < h->n; i++)
        if (type[i] >= ntypes) return fail("bd_set_types: class id out of range");
    h->hp.ntypes = ntypes;
    for (int t = 0; t < ntypes; t++) h->hp.mobility[t] = mobility[t];
    if (h->n) std::memcpy(h->h_type.data(), type, h->n);
    h->params_dirty = true;
    return upload_words(h);
}

/* Positions replaced from outside: records and shadows are rewritten in place and the displacement
 * from the last list build is folded into the device-side rebuild criterion -- a caller that uploads
 * what it just downloaded (a per-step driver callback) does not pay for a list build. */
static int scatter_tracked(bd_engine *h, const double *d_xyz, uint32_t first, uint32_t count)
{
    if (upload_params(h)) return 1;
    k_scatter_tracked<<<blocks_for(count, 256), 256, 0, h->stream>>>(h->pos[h->cur].p, h->shadow[h->cur].p, h->bpos.p,
                                                                    h->slot_of.p, d_xyz, first, count, h->d_params.p,
                                                                    h->ctl.p, h->staged);
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

extern "C" int bd_set_positions(bd_engine *h, int32_t replica, const double *xyz)
{
    NEED(h);
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_set_positions: bad replica");
    if (h->n == 0) return 0;
    CK(cudaMemcpyAsync(h->xyz_stage.p, xyz, 3 * (size_t) h->n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (scatter_tracked(h, h->xyz_stage.p, (uint32_t) replica * h->n, h->n)) return 1;
    CK(cudaStreamSynchronize(h->stream));  // xyz may be pageable and is reused by the caller
    return 0;
}

extern "C" int bd_get_positions(bd_engine *h, int32_t replica, double *xyz)
{
    NEED(h);
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_get_positions: bad replica");
    if (h->n) {
        k_gather_positions<<<blocks_for(h->n, 256), 256, 0, h->stream>>>(h->pos[h->cur].p, h->slot_of.p,
                                                                         h->xyz_stage.p, (uint32_t) replica * h->n, h->n);
        h->launches++;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(xyz, h->xyz_stage.p, 3 * (size_t) h->n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    }
    return sync_and_check(h);
}

/* every replica at once, host buffers [replicas][n][3] (pinned buffers make the copies asynchronous) */
extern "C" int bd_set_all_positions(bd_engine *h, const double *xyz)
{
    NEED(h);
    if (h->ntot == 0) return 0;
    if (h->xyz_all.count < 3 * (size_t) h->ntot) CK(h->xyz_all.alloc(3 * (size_t) h->ntot));
    CK(cudaMemcpyAsync(h->xyz_all.p, xyz, 3 * (size_t) h->ntot * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    return scatter_tracked(h, h->xyz_all.p, 0, h->ntot);
}

/* enqueue the download and return: the data is there after bd_sync (xyz must be pinned for the copy to be
 * asynchronous).  Several engines on their own streams -- chunks of one replica ensemble -- overlap the
 * upload of one chunk, the step of another and the download of a third this way (ensemble.HostPipeline). */
extern "C" int bd_get_all_positions_async(bd_engine *h, double *xyz)
{
    NEED(h);
    if (h->ntot == 0) return 0;
    if (h->xyz_out.count < 3 * (size_t) h->ntot) CK(h->xyz_out.alloc(3 * (size_t) h->ntot));
    k_gather_positions<<<blocks_for(h->ntot, 256), 256, 0, h->stream>>>(h->pos[h->cur].p, h->slot_of.p, h->xyz_out.p, 0, h->ntot);
    h->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(xyz, h->xyz_out.p, 3 * (size_t) h->ntot * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    return 0;
}

extern "C" int bd_get_all_positions(bd_engine *h, double *xyz)
{
    if (bd_get_all_positions_async(h, xyz)) return 1;
    return sync_and_check(h);
}

extern "C" int bd_set_positions_device(bd_engine *h, const double *d_xyz)
{
    NEED(h);
    if (h->ntot == 0) return 0;
    return scatter_tracked(h, d_xyz, 0, h->ntot);
}

extern "C" int bd_get_positions_device(bd_engine *h, double *d_xyz)
{
    NEED(h);
    if (h->ntot == 0) return 0;
    k_gather_positions<<<blocks_for(h->ntot, 256), 256, 0, h->stream>>>(h->pos[h->cur].p, h->slot_of.p, d_xyz, 0, h->ntot);
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

extern "C" int bd_set_external_forces(bd_engine *h, int32_t replica, const double *xyz)
{
    NEED(h);
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_set_external_forces: bad replica");
    size_t const per = 3 * (size_t) h->n;
    if (!xyz) {
        // cleared for this replica; the array stays in use while any replica may still hold forces
        if (h->ext_on) CK(cudaMemsetAsync(h->ext.p + (size_t) replica * per, 0, per * sizeof(double), h->stream));
        if (h->replicas == 1) h->ext_on = false;
        return 0;
    }
    if (h->ext.count < per * h->replicas) {
        CK(h->ext.alloc(per * h->replicas));
        CK(cudaMemsetAsync(h->ext.p, 0, per * h->replicas * sizeof(double), h->stream));
    }
    CK(cudaMemcpyAsync(h->ext.p + (size_t) replica * per, xyz, per * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));  // the caller's array may change right away
    h->ext_on = true;
    return 0;
}

/* Slab decomposition (slab.py): beads another engine advances are GHOSTS here. */
extern "C" int bd_set_owned(bd_engine *h, const uint8_t *owned)
{
    NEED(h);
    if (h->replicas != 1) return fail("bd_set_owned: slab decomposition applies to a single system (1 replica)");
    if (!owned) return bd_set_owned_device(h, nullptr);
    if (h->d_owned.count < h->n) CK(h->d_owned.alloc(h->n));
    CK(cudaMemcpyAsync(h->d_owned.p, owned, h->n, cudaMemcpyHostToDevice, h->stream));
    int const rc = bd_set_owned_device(h, h->d_owned.p);
    CK(cudaStreamSynchronize(h->stream));  // the caller's mask may go away
    return rc;
}

extern "C" int bd_set_owned_device(bd_engine *h, const uint8_t *d_owned)
{
    NEED(h);
    if (h->replicas != 1) return fail("bd_set_owned_device: slab decomposition applies to a single system (1 replica)");
    if (h->n == 0) return 0;
    k_set_ghost<<<blocks_for(h->n, 256), 256, 0, h->stream>>>(h->pos[0].p, h->pos[1].p, h->bpos.p, h->slot_of.p, d_owned, h->n);
    h->launches++;
    CK(cudaGetLastError());
    return request_rebuild(h);  // ghosts carry no Verlet row
}

/* Device address of the running maximum of |x - x(list build)|^2 over the beads this engine has seen move
 * (an IEEE double's bits; non-negative, so integer max = floating max).  Slabs of one system all-reduce it
 * (max) after every step, so that every rank takes the rebuild decision the undecomposed system takes -- any
 * bead anywhere beyond the skin -- and they rebuild at the same step instead of stalling each other's exchange. */
extern "C" int bd_displacement_pointer(bd_engine *h, void **d_out)
{
    NEED(h);
    if (!d_out) return fail("bd_displacement_pointer: null output");
    *d_out = &h->ctl.p->max_disp2_bits;
    return 0;
}

extern "C" int bd_gather_positions_indexed(bd_engine *h, const uint32_t *d_index, int64_t count, double *d_xyz)
{
    NEED(h);
    if (count < 0 || (count > 0 && (!d_index || !d_xyz))) return fail("bd_gather_positions_indexed: bad arguments");
    if (count == 0) return 0;
    k_gather_indexed<<<blocks_for((uint32_t) count, 256), 256, 0, h->stream>>>(h->pos[h->cur].p, h->slot_of.p, d_index,
                                                                                d_xyz, (uint32_t) count);
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

extern "C" int bd_scatter_positions_indexed(bd_engine *h, const uint32_t *d_index, int64_t count, const double *d_xyz)
{
    NEED(h);
    if (count < 0 || (count > 0 && (!d_index || !d_xyz))) return fail("bd_scatter_positions_indexed: bad arguments");
    if (count == 0) return 0;
    if (upload_params(h)) return 1;
    k_scatter_indexed<<<blocks_for((uint32_t) count, 256), 256, 0, h->stream>>>(
        h->pos[h->cur].p, h->shadow[h->cur].p, h->bpos.p, h->slot_of.p, d_index, d_xyz, (uint32_t) count, h->d_params.p,
        h->ctl.p, h->staged);
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

static int set_mask_bit(std::vector<uint8_t> &mask, const uint8_t *targets, uint32_t n, int bit)
{
    for (uint32_t i = 0; i < n; i++) {
        mask[i] = (uint8_t) (mask[i] & ~(1u << bit));
        if (targets && targets[i]) mask[i] |= (uint8_t) (1u << bit);
    }
    return 0;
}

extern "C" int bd_add_pair_ff(bd_engine *h, int32_t nterms, const bd_term *table, double dcut,
                              const uint8_t *targets, int32_t *ff_out)
{
    NEED(h);
    if (!(dcut > 0)) return fail("bd_add_pair_ff: dcut must be positive");
    for (int k = 0; k < BD_MAX_PAIR_FF; k++) {
        if (h->pair[k].used) continue;
        if (store_table(h, h->pair[k], nterms, table, h->hp.ntypes * h->hp.ntypes)) return 1;
        DPair &p = h->hp.pair[k];
        p.enabled = 1;
        p.nterms = nterms;
        p.term_off = h->pair[k].term_off;
        p.has_targets = targets != nullptr;
        p.dcut = dcut;
        p.dcut2 = dcut * dcut;
        set_mask_bit(h->h_pair_mask, targets, h->n, k);
        *ff_out = k;
        return upload_words(h);
    }
    return fail("bd_add_pair_ff: too many pair forcefields");
}

extern "C" int bd_update_pair_ff(bd_engine *h, int32_t ff, const bd_term *table, double dcut)
{
    NEED(h);
    if (ff < 0 || ff >= BD_MAX_PAIR_FF || !h->pair[ff].used) return fail("bd_update_pair_ff: bad forcefield id");
    if (!(dcut > 0)) return fail("bd_update_pair_ff: dcut must be positive");
    if (table && store_table(h, h->pair[ff], h->pair[ff].nterms, table, h->pair[ff].cells)) return 1;
    h->hp.pair[ff].dcut = dcut;
    h->hp.pair[ff].dcut2 = dcut * dcut;
    h->params_dirty = true;
    return 0;
}

extern "C" int bd_add_bond_ff(bd_engine *h, int32_t nterms, const bd_term *table, const uint32_t *pairs,
                              int64_t m, int32_t *ff_out)
{
    NEED(h);
    if (m < 0) return fail("bd_add_bond_ff: negative pair count");
    for (int64_t b = 0; b < 2 * m; b++)
        if (pairs[b] >= h->n) return fail("bd_add_bond_ff: bond index out of range");
    for (int k = 0; k < BD_MAX_BOND_FF; k++) {
        if (h->bond[k].used) continue;
        if (store_table(h, h->bond[k], nterms, table, h->hp.ntypes * h->hp.ntypes)) return 1;
        DBond &b = h->hp.bond[k];
        b.enabled = 1;
        b.nterms = nterms;
        b.term_off = h->bond[k].term_off;
        h->bond_pairs[k].assign(pairs, pairs + 2 * m);
        *ff_out = k;
        return upload_topology(h);
    }
    return fail("bd_add_bond_ff: too many bond forcefields");
}

/* dynamic pair lists (glue / loop force fields of 3-sim-1kb): only the list crosses to the device; the
 * slot-space bond rows are refilled by the next list build */
extern "C" int bd_set_bond_pairs(bd_engine *h, int32_t ff, const uint32_t *pairs, int64_t m)
{
    NEED(h);
    if (ff < 0 || ff >= BD_MAX_BOND_FF || !h->bond[ff].used) return fail("bd_set_bond_pairs: bad forcefield id");
    if (m < 0 || (m > 0 && !pairs)) return fail("bd_set_bond_pairs: bad pair list");
    for (int64_t b = 0; b < 2 * m; b++)
        if (pairs[b] >= h->n) return fail("bd_set_bond_pairs: bond index out of range");
    h->bond_pairs[ff].assign(pairs, pairs + 2 * m);
    return upload_topology(h);
}

extern "C" int bd_set_bond_min_image(bd_engine *h, int32_t ff, int32_t on)
{
    NEED(h);
    if (ff < 0 || ff >= BD_MAX_BOND_FF || !h->bond[ff].used) return fail("bd_set_bond_min_image: bad forcefield id");
    h->hp.bond[ff].min_image = on != 0;
    h->params_dirty = true;
    return 0;
}

extern "C" int bd_update_bond_ff(bd_engine *h, int32_t ff, const bd_term *table)
{
    NEED(h);
    if (ff < 0 || ff >= BD_MAX_BOND_FF || !h->bond[ff].used) return fail("bd_update_bond_ff: bad forcefield id");
    return store_table(h, h->bond[ff], h->bond[ff].nterms, table, h->bond[ff].cells);
}

extern "C" int bd_add_bend_ff(bd_engine *h, const uint32_t *triples, const double *bending_energy, int64_t m,
                              int32_t *ff_out)
{
    NEED(h);
    if (m < 0 || m >= (1ll << 30)) return fail("bd_add_bend_ff: bad triple count");
    for (int64_t b = 0; b < 3 * m; b++)
        if (triples[b] >= h->n) return fail("bd_add_bend_ff: triple index out of range");
    for (int k = 0; k < BD_MAX_BEND_FF; k++) {
        if (h->bend_used[k]) continue;
        h->bend_used[k] = true;
        h->hp.bend[k].enabled = 1;
        h->bend_triples[k].assign(triples, triples + 3 * m);
        h->bend_energies[k].assign(bending_energy, bending_energy + m);
        h->params_dirty = true;
        *ff_out = k;
        return upload_topology(h);
    }
    return fail("bd_add_bend_ff: too many bending forcefields");
}

/* reach of a wall table: the largest distance at which any of its terms is non-zero, 0 if unbounded */
static double table_range(const bd_term *table, int cells, int nterms)
{
    double range = 0;
    for (int i = 0; i < cells * nterms; i++) {
        const bd_term &t = table[i];
        if (t.kind == BD_TERM_NONE || t.factor == 0.0) continue;
        double reach = 0;
        if (t.kind == BD_TERM_SOFTCORE) reach = t.length;  // zero force and energy for r >= diameter
        if (t.cutoff > 0) reach = reach > 0 ? std::min(reach, t.cutoff) : t.cutoff;
        if (!(reach > 0)) return 0;
        range = std::max(range, reach);
    }
    return range;
}

/* Write one wall's geometry for `replica` (or for every replica when replica < 0).  The device table is the
 * master copy: with the relaxation law active it moves on its own, so single entries are written, never the
 * whole table. */
static int write_wall_geometry(bd_engine *h, int ff, int replica, const double center[3], const double semiaxes[3])
{
    for (int d = 0; d < 3; d++)
        if (!(semiaxes[d] > 0)) return fail("bd: wall semiaxes must be positive");
    DWallGeom g;
    std::memset(&g, 0, sizeof g);
    for (int d = 0; d < 3; d++) {
        g.center[d] = center[d];
        g.semi[d] = semiaxes[d];
        g.quad[d] = 1.0 / (semiaxes[d] * semiaxes[d]);  // MD/forcefield/ellipsoid_surface_forcefield.hpp:72-76
    }
    if (replica < 0 && h->replicas == 1) replica = 0;
    if (replica >= 0) {
        // pageable source: the call returns once the 96 bytes are staged, the local may go away
        CK(cudaMemcpyAsync(h->wall_geom.p + (size_t) replica * BD_MAX_WALL_FF + ff, &g, sizeof g, cudaMemcpyHostToDevice,
                           h->stream));
    } else {
        std::vector<DWallGeom> all(h->replicas, g);
        CK(cudaMemcpy2DAsync(h->wall_geom.p + ff, BD_MAX_WALL_FF * sizeof(DWallGeom), all.data(), sizeof(DWallGeom),
                             sizeof(DWallGeom), h->replicas, cudaMemcpyHostToDevice, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    return 0;
}

extern "C" int bd_add_wall_ff(bd_engine *h, int32_t shape, int32_t direction, int32_t nterms,
                              const bd_term *table, const double center[3], const double semiaxes[3],
                              int32_t *ff_out)
{
    NEED(h);
    if (shape != BD_WALL_SPHERE && shape != BD_WALL_ELLIPSOID) return fail("bd_add_wall_ff: unknown shape");
    if (direction != BD_WALL_INWARD && direction != BD_WALL_OUTWARD) return fail("bd_add_wall_ff: unknown direction");
    for (int k = 0; k < BD_MAX_WALL_FF; k++) {
        if (h->wall[k].used) continue;
        if (write_wall_geometry(h, k, -1, center, semiaxes)) return 1;
        if (store_table(h, h->wall[k], nterms, table, h->hp.ntypes)) return 1;
        DWall &w = h->hp.wall[k];
        w.enabled = 1;
        w.nterms = nterms;
        w.term_off = h->wall[k].term_off;
        w.shape = shape;
        w.direction = direction;
        w.range = table_range(table, h->hp.ntypes, nterms);
        *ff_out = k;
        return 0;
    }
    return fail("bd_add_wall_ff: too many wall forcefields");
}

/* center / semiaxes may be NULL together: only the table changes (the geometry may be moving on the
 * device: bd_set_wall_relaxation) */
extern "C" int bd_update_wall_ff(bd_engine *h, int32_t ff, const bd_term *table, const double center[3],
                                 const double semiaxes[3])
{
    NEED(h);
    if (ff < 0 || ff >= BD_MAX_WALL_FF || !h->wall[ff].used) return fail("bd_update_wall_ff: bad forcefield id");
    if ((center == nullptr) != (semiaxes == nullptr)) return fail("bd_update_wall_ff: center and semiaxes go together");
    if (table && store_table(h, h->wall[ff], h->wall[ff].nterms, table, h->wall[ff].cells)) return 1;
    if (table) {
        h->hp.wall[ff].range = table_range(table, h->wall[ff].cells, h->wall[ff].nterms);
        h->params_dirty = true;
    }
    if (center) return write_wall_geometry(h, ff, -1, center, semiaxes);
    return 0;
}

extern "C" int bd_update_wall_ff_replica(bd_engine *h, int32_t ff, int32_t replica, const double center[3],
                                         const double semiaxes[3])
{
    NEED(h);
    if (ff < 0 || ff >= BD_MAX_WALL_FF || !h->wall[ff].used) return fail("bd_update_wall_ff_replica: bad forcefield id");
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_update_wall_ff_replica: bad replica");
    if (!center || !semiaxes) return fail("bd_update_wall_ff_replica: null geometry");
    return write_wall_geometry(h, ff, replica, center, semiaxes);
}

extern "C" int bd_get_wall_geometry(bd_engine *h, int32_t ff, int32_t replica, double center[3], double semiaxes[3])
{
    NEED(h);
    if (ff < 0 || ff >= BD_MAX_WALL_FF || !h->wall[ff].used) return fail("bd_get_wall_geometry: bad forcefield id");
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_get_wall_geometry: bad replica");
    DWallGeom g;
    CK(cudaMemcpyAsync(&g, h->wall_geom.p + (size_t) replica * BD_MAX_WALL_FF + ff, sizeof g, cudaMemcpyDeviceToHost,
                       h->stream));
    if (sync_and_check(h)) return 1;
    for (int d = 0; d < 3; d++) {
        if (center) center[d] = g.center[d];
        if (semiaxes) semiaxes[d] = g.semi[d];
    }
    return 0;
}

extern "C" int bd_set_wall_relaxation(bd_engine *h, const int32_t *ffs, int32_t count, const double spring[3],
                                      double mobility)
{
    NEED(h);
    if (count < 0 || count > BD_MAX_WALL_FF || (count > 0 && (!ffs || !spring)))
        return fail("bd_set_wall_relaxation: bad arguments");
    WallRelax law = {};
    for (int k = 0; k < count; k++) {
        if (ffs[k] < 0 || ffs[k] >= BD_MAX_WALL_FF || !h->wall[ffs[k]].used)
            return fail("bd_set_wall_relaxation: bad forcefield id");
        if (h->hp.wall[ffs[k]].shape != BD_WALL_ELLIPSOID) return fail("bd_set_wall_relaxation: ellipsoid walls only");
        law.mask |= 1u << ffs[k];
    }
    if (count > 0) {
        for (int d = 0; d < 3; d++) law.spring[d] = spring[d];
        law.mobility = mobility;
    }
    h->relax = law;
    return 0;
}

extern "C" int bd_add_point_ff(bd_engine *h, int32_t nterms, const bd_term *table, const double source[3],
                               const uint8_t *targets, int32_t *ff_out)
{
    NEED(h);
    for (int k = 0; k < BD_MAX_POINT_FF; k++) {
        if (h->point[k].used) continue;
        if (store_table(h, h->point[k], nterms, table, h->hp.ntypes)) return 1;
        DPoint &p = h->hp.point[k];
        p.enabled = 1;
        p.nterms = nterms;
        p.term_off = h->point[k].term_off;
        p.has_targets = targets != nullptr;
        for (int d = 0; d < 3; d++) p.source[d] = source[d];
        set_mask_bit(h->h_point_mask, targets, h->n, k);
        *ff_out = k;
        return upload_words(h);
    }
    return fail("bd_add_point_ff: too many point-source forcefields");
}

extern "C" int bd_set_ff_enabled(bd_engine *h, int32_t ff_class, int32_t ff, int32_t enabled)
{
    NEED(h);
    int32_t *flag = nullptr;
    switch (ff_class) {
    case BD_FF_PAIR: if (ff >= 0 && ff < BD_MAX_PAIR_FF && h->pair[ff].used) flag = &h->hp.pair[ff].enabled; break;
    case BD_FF_BOND: if (ff >= 0 && ff < BD_MAX_BOND_FF && h->bond[ff].used) flag = &h->hp.bond[ff].enabled; break;
    case BD_FF_BEND: if (ff >= 0 && ff < BD_MAX_BEND_FF && h->bend_used[ff]) flag = &h->hp.bend[ff].enabled; break;
    case BD_FF_WALL: if (ff >= 0 && ff < BD_MAX_WALL_FF && h->wall[ff].used) flag = &h->hp.wall[ff].enabled; break;
    case BD_FF_POINT: if (ff >= 0 && ff < BD_MAX_POINT_FF && h->point[ff].used) flag = &h->hp.point[ff].enabled; break;
    default: break;
    }
    if (!flag) return fail("bd_set_ff_enabled: bad forcefield class / id");
    *flag = enabled != 0;
    h->params_dirty = true;
    return 0;
}

// ---- the step path -------------------------------------------------------------------------------

static NlistArrays nlist_arrays(bd_engine *h)
{
    NlistArrays A;
    std::memset(&A, 0, sizeof A);  // padding included: the step graphs compare these structs byte for byte
    A.pos = h->pos[h->cur].p;
    A.pos_out = h->pos[h->cur].p;
    A.bpos = h->bpos.p;
    A.shadow = h->shadow[h->cur].p;
    A.slot_of = h->slot_of.p;
    A.cellkey = h->cellkey.p;
    A.rank = h->rank.p;
    A.stage = h->stage.p;
    A.cell_count = h->cell_count.p;
    A.cell_start = h->cell_start.p;
    A.block_sum = h->block_sum.p;
    A.nbr4 = h->nbr4.p;
    A.plen = h->plen.p;
    A.nnbr = h->nnbr.p;
    A.order = h->order.p;
    A.bondrow = h->bondrow.p;
    A.bond_start = h->bond_start.p;
    A.bond_adj = h->bond_adj.p;
    A.type = h->d_type.p;
    A.bend_start = h->bend_start.p;
    A.bend_adj = h->bend_adj.p;
    A.triples = h->triples.p;
    A.bend_energy = h->bend_energy.p;
    A.bend_ff = h->bend_ff.p;
    A.bend_pq = h->bend_pq.p;
    A.bend_e = h->bend_e.p;
    A.bend_meta = h->bend_meta.p;
    A.bend_deg = h->bend_deg;
    for (int f = 0; f < BD_MAX_BOND_FF; f++) A.bond_slot[f] = h->bond_slot[f];
    A.ntypes = (uint32_t) h->hp.ntypes;
    A.bond_deg = h->bond_deg;
    A.tile_runs = h->tile_runs.p;
    A.staged = h->staged;
    A.balance = h->balance_rows;
    A.npad = h->npad;
    A.pstride = h->pstride;
    A.ctl = h->ctl.p;
    A.params = h->d_params.p;
    A.n = h->n;
    A.replicas = h->replicas;
    A.ntot = h->ntot;
    A.cell_cap = h->cell_cap;
    A.max_bits = h->max_bits;
    A.cap = h->cap;
    return A;
}

static StepArrays step_arrays(bd_engine *h)
{
    StepArrays A;
    std::memset(&A, 0, sizeof A);  // padding included: the step graphs compare these structs byte for byte
    A.pos = h->pos[h->cur].p;
    A.pos_out = h->pos[h->cur ^ 1].p;
    A.shadow = h->shadow[h->cur].p;
    A.shadow_out = h->shadow[h->cur ^ 1].p;
    A.bpos = h->bpos.p;
    A.slot_of = h->slot_of.p;
    A.nbr4 = h->nbr4.p;
    A.tile_runs = h->tile_runs.p;
    A.staged = h->staged;
    A.pair_f32 = h->pair_f32 && !h->staged;
    A.stage_cap = h->stage_cap;
    A.plen = h->plen.p;
    A.bondrow = h->bondrow.p;
    A.bond_deg = h->bond_deg;
    A.npad = h->npad;
    A.pstride = h->pstride;
    A.bond_rows = h->d_bond_rows.p;
    for (int f = 0; f < BD_MAX_BOND_FF; f++) A.bond_slot[f] = h->bond_slot[f];
    A.bonds_compiled = h->bonds_compiled;
    A.nbond_rows = h->nbond_rows;
    A.any_bend = 0;
    for (int f = 0; f < BD_MAX_BEND_FF; f++) A.any_bend |= h->bend_used[f];
    A.order = h->order.p;
    A.bond_start = h->bond_start.p;
    A.bond_adj = h->bond_adj.p;
    A.bend_pq = h->bend_pq.p;
    A.bend_e = h->bend_e.p;
    A.bend_meta = h->bend_meta.p;
    A.bend_deg = h->bend_deg;
    A.bend_start = h->bend_start.p;
    A.bend_adj = h->bend_adj.p;
    A.triples = h->triples.p;
    A.bend_energy = h->bend_energy.p;
    A.bend_ff = h->bend_ff.p;
    A.terms = h->d_terms.p;
    A.params = h->d_params.p;
    A.wall_geom = h->wall_geom.p;
    A.ctl = h->ctl.p;
    A.wall_partial = h->wall_partial.p;
    A.inter_partial = h->inter_partial.p;
    A.force = h->force.p;
    A.wprev = h->wprev.p;
    A.ext = h->ext_on ? h->ext.p : nullptr;
    A.dt_bits = h->dt_bits.p;
    A.dt_prev = h->dt_prev.p;
    A.pair_rows = h->d_pair_rows.p;
    A.ntypes2 = (uint32_t) (h->hp.ntypes * h->hp.ntypes);
    A.stage_off = (uint32_t) table_bytes(A.ntypes2, (uint32_t) h->nbond_rows);
    A.fast_ff = h->fast_ff;
    A.k2_uniform = h->prog != PROG_GENERIC && h->k2_uniform;
    A.k2a = h->k2a;
    A.k2b = h->k2b;
    A.n = h->n;
    A.replicas = h->replicas;
    A.ntot = h->ntot;
    A.chunks = h->chunks;
    return A;
}

/* dynamic shared memory of the step kernels: compiled pair rows, compiled bond rows, then the staged shadows */
static size_t table_bytes(uint32_t ntypes2, uint32_t nbond_rows)
{
    return ((size_t) ntypes2 * sizeof(PairRow) + (size_t) nbond_rows * sizeof(BondRow) + 15) & ~(size_t) 15;
}

static size_t step_smem_bytes(const StepArrays &A)
{
    return A.stage_off + (A.staged ? STAGE_TABLE_BYTES + ((size_t) A.stage_cap + 1) * sizeof(float4) : 0);
}

constexpr int STEP_SMEM_OPT_IN = 160 * 1024;  // dynamic shared memory the step kernels may ask for

template <typename K>
static void allow_big_smem(K kernel)
{
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, STEP_SMEM_OPT_IN);
}

static void opt_in_shared_memory()
{
    // function attributes belong to the current device: once per device of this process
    static unsigned long long done = 0;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && ((done >> dev) & 1ull)) return;
    if (dev < 64) done |= 1ull << dev;
    allow_big_smem(k_step_fused<BD_BOX_OPEN, PROG_GENERIC>);
    allow_big_smem(k_step_fused<BD_BOX_OPEN, PROG_SC23_SC83>);
    allow_big_smem(k_step_fused<BD_BOX_OPEN, PROG_SC23>);
    allow_big_smem(k_step_fused<BD_BOX_PERIODIC, PROG_GENERIC>);
    allow_big_smem(k_step_fused<BD_BOX_PERIODIC, PROG_SC23_SC83>);
    allow_big_smem(k_step_fused<BD_BOX_PERIODIC, PROG_SC23>);
    allow_big_smem(k_force<BD_BOX_OPEN, PROG_GENERIC>);
    allow_big_smem(k_force<BD_BOX_OPEN, PROG_SC23_SC83>);
    allow_big_smem(k_force<BD_BOX_OPEN, PROG_SC23>);
    allow_big_smem(k_force<BD_BOX_PERIODIC, PROG_GENERIC>);
    allow_big_smem(k_force<BD_BOX_PERIODIC, PROG_SC23_SC83>);
    allow_big_smem(k_force<BD_BOX_PERIODIC, PROG_SC23>);
    allow_big_smem(k_rows_staged);
}

template <int BOX>
static void launch_step_box(int prog, dim3 grid, cudaStream_t st, const StepArrays &A, const StepScalars &S)
{
    size_t const rows_bytes = step_smem_bytes(A);
    switch (prog) {
    case PROG_SC23_SC83: k_step_fused<BOX, PROG_SC23_SC83><<<grid, STEP_BLOCK, rows_bytes, st>>>(A, S); break;
    case PROG_SC23: k_step_fused<BOX, PROG_SC23><<<grid, STEP_BLOCK, rows_bytes, st>>>(A, S); break;
    default: k_step_fused<BOX, PROG_GENERIC><<<grid, STEP_BLOCK, rows_bytes, st>>>(A, S); break;
    }
}

template <int BOX>
static void launch_force_box(int prog, dim3 grid, cudaStream_t st, const StepArrays &A, const StepScalars &S,
                             Select sel)
{
    size_t const rows_bytes = step_smem_bytes(A);
    switch (prog) {
    case PROG_SC23_SC83: k_force<BOX, PROG_SC23_SC83><<<grid, STEP_BLOCK, rows_bytes, st>>>(A, S, sel); break;
    case PROG_SC23: k_force<BOX, PROG_SC23><<<grid, STEP_BLOCK, rows_bytes, st>>>(A, S, sel); break;
    default: k_force<BOX, PROG_GENERIC><<<grid, STEP_BLOCK, rows_bytes, st>>>(A, S, sel); break;
    }
}

static void launch_step(bd_engine *h, const StepArrays &A, const StepScalars &S)
{
    dim3 const grid((h->chunks + STEP_TILES - 1) / STEP_TILES, h->replicas);
    if (h->hp.box_kind == BD_BOX_PERIODIC) launch_step_box<BD_BOX_PERIODIC>(h->prog, grid, h->stream, A, S);
    else launch_step_box<BD_BOX_OPEN>(h->prog, grid, h->stream, A, S);
}

static void launch_force(bd_engine *h, const StepArrays &A, const StepScalars &S, Select sel)
{
    dim3 const grid(h->chunks, h->replicas);
    // the specialised programs evaluate exactly one pair forcefield; other selections interpret
    bool const fast = sel.ff_class == BD_FF_ALL || (sel.ff_class == BD_FF_PAIR && sel.ff == h->fast_ff);
    int const prog = fast ? h->prog : PROG_GENERIC;
    if (h->hp.box_kind == BD_BOX_PERIODIC) launch_force_box<BD_BOX_PERIODIC>(prog, grid, h->stream, A, S, sel);
    else launch_force_box<BD_BOX_OPEN>(prog, grid, h->stream, A, S, sel);
}

static void launch_integrate(bd_engine *h, const StepArrays &A, const StepScalars &S)
{
    dim3 const grid(h->chunks, h->replicas);
    if (h->hp.box_kind == BD_BOX_PERIODIC) k_integrate<BD_BOX_PERIODIC><<<grid, STEP_BLOCK, 0, h->stream>>>(A, S);
    else k_integrate<BD_BOX_OPEN><<<grid, STEP_BLOCK, 0, h->stream>>>(A, S);
}

/* Enqueue the rebuild chain.  Every kernel returns at once unless k_rebuild_begin latched a
 * rebuild, so the host never has to know whether the list is still valid. */
static int enqueue_rebuild_chain(bd_engine *h)
{
    if (h->ntot == 0) return 0;
    if (upload_params(h)) return 1;
    NlistArrays const A = nlist_arrays(h);
    cudaStream_t const st = h->stream;
    // per-bead kernels run grid-stride: a few waves of blocks when a rebuild is due, one wave of
    // returning blocks when it is not
    unsigned const bead_blocks = std::min(blocks_for(h->ntot, NL_BLOCK), 148u * 8u);
    unsigned const row_blocks = std::min(blocks_for(h->ntot, ROWS_BLOCK), 148u * 56u);
    unsigned const tile_blocks = blocks_for(h->cell_cap, SCAN_TILE);
    k_rebuild_begin<<<1, 32, 0, st>>>(h->ctl.p, h->d_params.p, (cudaGraphConditionalHandle) 0, 0u);
    k_bbox_zero<<<std::min(bead_blocks + tile_blocks, 148u * 8u), NL_BLOCK, 0, st>>>(A);
    k_make_grid<<<1, 32, 0, st>>>(A);
    k_cell_count<<<bead_blocks, NL_BLOCK, 0, st>>>(A);
    k_scan_reduce<<<tile_blocks, NL_BLOCK, 0, st>>>(A);
    k_scan_top<<<1, 1024, 0, st>>>(A);
    k_scan_down<<<tile_blocks, NL_BLOCK, 0, st>>>(A);
    k_group<<<bead_blocks, NL_BLOCK, 0, st>>>(A);
    k_place<<<bead_blocks, NL_BLOCK, 0, st>>>(A);
    if (h->staged) {
        unsigned const tiles = h->chunks * h->replicas;
        k_rows_staged<<<std::min(tiles, 148u * 12u), ROWS_BLOCK, sizeof(RowsStagedShared), st>>>(A);
    } else {
        k_rows<<<row_blocks, ROWS_BLOCK, 0, st>>>(A);
    }
    k_balance<<<std::min(h->chunks * h->replicas, 148u * 16u), BAL_TILE, 0, st>>>(A);
    h->launches += 11;
    if (h->bond_rows_stale) {
        if (h->bond_deg || h->bend_deg) {
            k_bond_rows<<<std::min(h->chunks * h->replicas, 148u * 16u), BAL_TILE, 0, st>>>(A);
            h->launches += 1;
        }
        h->bond_rows_stale = false;
    }
    CK(cudaGetLastError());
    return 0;
}

/* Blocks until the stream is idle and turns sticky device-side conditions into errors. */
static int sync_and_check(bd_engine *h)
{
    CK(cudaStreamSynchronize(h->stream));
    uint32_t overflow = 0;
    CK(cudaMemcpy(&overflow, &h->ctl.p->overflow, sizeof overflow, cudaMemcpyDeviceToHost));
    if (overflow) {
        char msg[200];
        std::snprintf(msg, sizeof msg,
                      "bd: %u neighbour rows exceeded list_capacity=%u; raise it with bd_set_option", overflow, h->cap);
        return fail(msg);
    }
    if (h->staged) {
        uint32_t spilled = 0;
        CK(cudaMemcpy(&spilled, &h->ctl.p->stage_overflow, sizeof spilled, cudaMemcpyDeviceToHost));
        if (spilled && h->list_verified)
            return fail("bd: a tile's neighbourhood outgrew the staging tables during a run; set option staged=0");
    }
    return 0;
}

/* Make the list current, make sure no row overflowed and keep 2x headroom over the longest row,
 * so that a run of steps without host synchronisation does not outgrow the rows. */
static int prepare_list(bd_engine *h)
{
    for (int attempt = 0; attempt < 12; attempt++) {
        if (enqueue_rebuild_chain(h)) return 1;
        CK(cudaStreamSynchronize(h->stream));
        DControl c;
        CK(cudaMemcpy(&c, h->ctl.p, sizeof c, cudaMemcpyDeviceToHost));
        if (h->staged && c.stage_overflow) {
            // some tile's neighbourhood does not fit the staging tables (a very sparse or badly sorted system):
            // this engine keeps rows of slots and gathers from global memory
            h->staged = false;
            CK(cudaMemset(&h->ctl.p->stage_overflow, 0, sizeof(uint32_t)));
            if (request_rebuild(h)) return 1;
            continue;
        }
        if (h->staged && h->stage_auto) {
            // room for the largest tile plus what it may gain before the next host visit; tiles beyond the
            // ceiling keep working through the run table (WALK_SPILLED)
            uint32_t const want_stage = std::min<uint32_t>(STAGE_CAP_CEILING, ((c.max_stage + c.max_stage / 8 + 63u) & ~63u));
            h->stage_cap = std::max(h->stage_cap, want_stage);
        }
        uint32_t const limit = std::max<uint32_t>(h->n, 16u);
        uint32_t want = h->cap;
        if (c.overflow) {
            if (h->cap >= h->n) return fail("bd: neighbour rows overflow although they hold every bead");
            want = std::min<uint32_t>(h->cap * 4, limit);
        } else if (2 * c.max_row > h->cap && h->cap < limit) {
            want = std::min<uint32_t>((2 * c.max_row + 15u) & ~15u, limit);
        }
        want = (want + ROW_PAD - 1) & ~(ROW_PAD - 1);
        if (want == h->cap) {
            h->list_verified = true;
            return 0;
        }
        if ((double) want * h->ntot * sizeof(uint32_t) > 64e9)
            return fail("bd: neighbour rows would exceed 64 GB (list radius too large for this density)");
        h->cap = want;
        if (alloc_rows(h)) return 1;
        CK(cudaMemset(&h->ctl.p->overflow, 0, sizeof(uint32_t)));
        if (request_rebuild(h)) return 1;
        h->capacity_rebuilds++;  // not a rebuild the reference's criterion asked for
    }
    return fail("bd: neighbour rows keep overflowing (positions degenerate?)");
}

// ---- one BD step as a CUDA graph with a conditional rebuild -----------------------------------------
//
// Launched kernel by kernel, a step is k_rebuild_begin + ten chain kernels that return at once nine times out
// of ten + the step kernel (+ the wall relaxation): thirteen launches, ~40 us of launch gaps that dominate a
// single 62 600-bead system (15 us of work) and cost the 64-replica ensemble 4 %.  As a graph the chain sits
// in a conditional IF node whose handle k_rebuild_begin sets on the device, so a step that keeps its list
// executes two (three) kernels, and the host issues one cudaGraphLaunch.  The graph of a ping-pong side is
// rebuilt whenever anything it was built from differs (compared byte for byte: pointers, sizes, kernels,
// launch shapes), which makes it safe against every setter of this file; only the step kernel's scalars
// (step number, ...) are patched per launch.

template <int BOX>
static const void *step_kernel_of(int prog)
{
    switch (prog) {
    case PROG_SC23_SC83: return (const void *) k_step_fused<BOX, PROG_SC23_SC83>;
    case PROG_SC23: return (const void *) k_step_fused<BOX, PROG_SC23>;
    default: return (const void *) k_step_fused<BOX, PROG_GENERIC>;
    }
}

static void destroy_step_graph(bd_engine::StepGraph &g)
{
    if (g.exec) cudaGraphExecDestroy(g.exec);
    if (g.graph) cudaGraphDestroy(g.graph);
    g.exec = nullptr;
    g.graph = nullptr;
    g.step_node = nullptr;
    g.signature.clear();
}

template <typename T>
static void sign(std::vector<unsigned char> &sig, const T &v)
{
    const unsigned char *p = reinterpret_cast<const unsigned char *>(&v);
    sig.insert(sig.end(), p, p + sizeof(T));
}

static int add_kernel(cudaGraph_t graph, cudaGraphNode_t *node, cudaGraphNode_t after, const void *func, dim3 grid,
                      dim3 block, size_t smem, void **args)
{
    cudaKernelNodeParams kp = {};
    kp.func = const_cast<void *>(func);
    kp.gridDim = grid;
    kp.blockDim = block;
    kp.sharedMemBytes = (unsigned) smem;
    kp.kernelParams = args;
    CK(cudaGraphAddKernelNode(node, graph, after ? &after : nullptr, after ? 1 : 0, &kp));
    return 0;
}

/* make graphs[h->cur] current for (A, S-independent launch data); dt is the wall relaxation's timestep */
static int ensure_step_graph(bd_engine *h, const StepArrays &A, const StepScalars &S, double dt)
{
    bd_engine::StepGraph &g = h->graphs[h->cur];
    NlistArrays NA = nlist_arrays(h);
    const void *step_fn = h->hp.box_kind == BD_BOX_PERIODIC ? step_kernel_of<BD_BOX_PERIODIC>(h->prog)
                                                            : step_kernel_of<BD_BOX_OPEN>(h->prog);
    size_t const step_smem = step_smem_bytes(A);
    dim3 const step_grid((h->chunks + STEP_TILES - 1) / STEP_TILES, h->replicas);
    std::vector<unsigned char> sig;
    sign(sig, NA);
    sign(sig, A);
    sign(sig, step_fn);
    sign(sig, step_smem);
    sign(sig, step_grid);
    sign(sig, h->relax);
    sign(sig, dt);
    sign(sig, h->cell_cap);
    if (g.exec && sig == g.signature) return 0;
    destroy_step_graph(g);

    CK(cudaGraphCreate(&g.graph, 0));
    cudaGraphConditionalHandle gate;
    CK(cudaGraphConditionalHandleCreate(&gate, g.graph, 0, cudaGraphCondAssignDefault));
    DControl *ctl = h->ctl.p;
    const DParams *params = h->d_params.p;
    uint32_t use_gate = 1;
    cudaGraphNode_t begin = nullptr, cond = nullptr, relax = nullptr, prev = nullptr;
    {
        void *args[] = {&ctl, &params, &gate, &use_gate};
        if (add_kernel(g.graph, &begin, nullptr, (const void *) k_rebuild_begin, 1, 32, 0, args)) return 1;
    }
    cudaGraphNodeParams cp = {};
    cp.type = cudaGraphNodeTypeConditional;
    cp.conditional.handle = gate;
    cp.conditional.type = cudaGraphCondTypeIf;
    cp.conditional.size = 1;
    CK(cudaGraphAddNode(&cond, g.graph, &begin, 1, &cp));
    cudaGraph_t body = cp.conditional.phGraph_out[0];
    {
        unsigned const bead_blocks = std::min(blocks_for(h->ntot, NL_BLOCK), 148u * 8u);
        unsigned const row_blocks = std::min(blocks_for(h->ntot, ROWS_BLOCK), 148u * 56u);
        unsigned const tile_blocks = blocks_for(h->cell_cap, SCAN_TILE);
        unsigned const tiles = h->chunks * h->replicas;
        void *args[] = {&NA};
        struct { const void *fn; unsigned grid, block; size_t smem; } chain[] = {
            {(const void *) k_bbox_zero, std::min(bead_blocks + tile_blocks, 148u * 8u), NL_BLOCK, 0},
            {(const void *) k_make_grid, 1, 32, 0},
            {(const void *) k_cell_count, bead_blocks, NL_BLOCK, 0},
            {(const void *) k_scan_reduce, tile_blocks, NL_BLOCK, 0},
            {(const void *) k_scan_top, 1, 1024, 0},
            {(const void *) k_scan_down, tile_blocks, NL_BLOCK, 0},
            {(const void *) k_group, bead_blocks, NL_BLOCK, 0},
            {(const void *) k_place, bead_blocks, NL_BLOCK, 0},
            {h->staged ? (const void *) k_rows_staged : (const void *) k_rows, h->staged ? std::min(tiles, 148u * 12u) : row_blocks,
             ROWS_BLOCK, h->staged ? sizeof(RowsStagedShared) : 0},
            {(const void *) k_balance, std::min(tiles, 148u * 16u), BAL_TILE, 0},
        };
        for (auto &k : chain) {
            cudaGraphNode_t node = nullptr;
            if (add_kernel(body, &node, prev, k.fn, k.grid, k.block, k.smem, args)) return 1;
            prev = node;
        }
    }
    {
        StepArrays a = A;
        StepScalars s0 = S;
        void *args[] = {&a, &s0};
        if (add_kernel(g.graph, &g.step_node, cond, step_fn, step_grid, STEP_BLOCK, step_smem, args)) return 1;
        g.step_params = {};
        g.step_params.func = const_cast<void *>(step_fn);
        g.step_params.gridDim = step_grid;
        g.step_params.blockDim = dim3(STEP_BLOCK);
        g.step_params.sharedMemBytes = (unsigned) step_smem;
    }
    if (h->relax.mask) {
        const double *partial = h->wall_partial.p;
        double *out = h->wall_out.p;
        DWallGeom *geom = h->wall_geom.p;
        WallRelax law = h->relax;
        uint32_t chunks = h->chunks;
        double step_dt = dt;
        void *args[] = {&partial, &out, &geom, &law, &chunks, &step_dt};
        if (add_kernel(g.graph, &relax, g.step_node, (const void *) k_wall_relax, h->replicas, RELAX_WIDTH * RELAX_LANES, 0, args))
            return 1;
    }
    CK(cudaGraphInstantiate(&g.exec, g.graph, 0));
    g.signature.swap(sig);
    return 0;
}

/* one step through the graph of the current ping-pong side */
static int launch_step_graph(bd_engine *h, const StepArrays &A, const StepScalars &S, double dt)
{
    if (upload_params(h)) return 1;
    if (ensure_step_graph(h, A, S, dt)) return 1;
    bd_engine::StepGraph &g = h->graphs[h->cur];
    StepArrays a = A;
    StepScalars s = S;
    void *args[] = {&a, &s};
    cudaKernelNodeParams kp = g.step_params;
    kp.kernelParams = args;
    CK(cudaGraphExecKernelNodeSetParams(g.exec, g.step_node, &kp));
    CK(cudaGraphLaunch(g.exec, h->stream));
    h->launches += h->relax.mask ? 3 : 2;  // + 10 per rebuild inside the conditional node (counted on the device)
    h->graph_steps++;
    return 0;
}

extern "C" int bd_step(bd_engine *h, const bd_step_config *cfg)
{
    NEED(h);
    if (!cfg) return fail("bd_step: null config");
    if (cfg->steps < 0 || cfg->first_step < 1) return fail("bd_step: bad step range");
    if (!(cfg->timestep > 0) || cfg->temperature < 0 || cfg->spacestep < 0) return fail("bd_step: bad parameters");
    if (h->ntot == 0 || cfg->steps == 0) return 0;
    bool const adaptive = cfg->spacestep != 0;
    bool const replay = cfg->replay_normals != nullptr;
    size_t const per_step = 3 * (size_t) h->ntot;
    cudaStream_t const st = h->stream;

    if (replay && h->wprev.count == 0) {
        CK(h->wprev.alloc(per_step));
        CK(cudaMemsetAsync(h->wprev.p, 0, per_step * sizeof(double), st));
    }
    if (cfg->first_step == 1) {
        // brownian_simulator constructor: previous noise starts at zero (brownian_simulator.hpp:37-40)
        if (h->wprev.count) CK(cudaMemsetAsync(h->wprev.p, 0, per_step * sizeof(double), st));
        h->last_dt = cfg->timestep;
        h->last_kT = cfg->temperature;
    } else if (replay && !h->wprev_valid) {
        return fail("bd_step: a replayed run must start at first_step = 1");
    }
    h->wprev_valid = replay;
    if (adaptive) {
        std::vector<unsigned long long> init(h->replicas);
        double const max_dt = cfg->timestep;
        unsigned long long bits;
        std::memcpy(&bits, &max_dt, sizeof bits);
        std::fill(init.begin(), init.end(), bits);
        CK(cudaMemcpyAsync(h->dt_bits.p, init.data(), init.size() * sizeof(bits), cudaMemcpyHostToDevice, st));
        if (cfg->first_step == 1 || !h->stats_adaptive_last) {
            // no adaptive history: the previous step (if any) used the fixed timestep
            std::vector<double> prev(h->replicas, cfg->first_step == 1 ? 0.0 : h->last_dt);
            CK(cudaMemcpyAsync(h->dt_prev.p, prev.data(), prev.size() * sizeof(double), cudaMemcpyHostToDevice, st));
        }
    }

    if (!h->list_verified && prepare_list(h)) return 1;

    // replayed normals are staged through a bounded device buffer
    size_t const max_stage_steps = std::max<size_t>(1, ((size_t) 256 << 20) / (per_step * sizeof(double)));
    size_t staged_from = 0, staged_count = 0;

    StepScalars S;
    S.kT = cfg->temperature;
    S.spacestep = cfg->spacestep;
    S.seed = cfg->seed;
    S.adaptive = adaptive;
    S.replay = nullptr;
    S.replica_base = h->replica_base;
    S.replica_stride = h->replica_stride;
    Select const all = {BD_FF_ALL, -1};

    for (int64_t k = 0; k < cfg->steps; k++) {
        if (replay && (size_t) k >= staged_from + staged_count) {
            staged_from = (size_t) k;
            staged_count = std::min<size_t>(max_stage_steps, (size_t) (cfg->steps - k));
            if (h->replay.count < staged_count * per_step) {
                CK(cudaStreamSynchronize(st));
                CK(h->replay.alloc(staged_count * per_step));
            }
            CK(cudaStreamSynchronize(st));  // the previous chunk is still being read
            CK(cudaMemcpyAsync(h->replay.p, cfg->replay_normals + staged_from * per_step,
                               staged_count * per_step * sizeof(double), cudaMemcpyHostToDevice, st));
        }
        bool const graphed = h->use_graph && !adaptive && !replay && !h->time_kernel && !h->bond_rows_stale;
        if (!graphed && enqueue_rebuild_chain(h)) return 1;
        // step_arrays reads what upload_params derives (pair program, uniform k2, compiled bond rows): bring the
        // parameters up to date first on the graph path too (tests/test_gpu_graph.py caught a stale first step
        // after bd_set_ff_enabled)
        if (graphed && upload_params(h)) return 1;
        StepArrays const A = step_arrays(h);
        S.step = cfg->first_step + k;
        S.dt = cfg->timestep;
        S.dt_prev = h->last_dt;
        S.kT_prev = h->last_kT;
        S.want_stats = k == cfg->steps - 1 || h->relax.mask != 0;
        if (replay) S.replay = h->replay.p + ((size_t) k - staged_from) * per_step;
        if (graphed) {
            if (launch_step_graph(h, A, S, cfg->timestep)) return 1;
            h->wall_stats_fresh = h->relax.mask != 0;
            h->cur ^= 1;
            h->last_dt = cfg->timestep;
            h->last_kT = cfg->temperature;
            h->steps_done++;
            continue;
        }
        if (!adaptive) {
            cudaEvent_t e0 = nullptr, e1 = nullptr;
            if (h->time_kernel) {
                CK(cudaEventCreate(&e0));
                CK(cudaEventCreate(&e1));
                CK(cudaEventRecord(e0, st));
            }
            launch_step(h, A, S);
            if (h->time_kernel) {
                CK(cudaEventRecord(e1, st));
                h->kernel_events.emplace_back(e0, e1);
            }
            h->launches += 1;
        } else {
            launch_force(h, A, S, all);
            launch_integrate(h, A, S);
            k_dt_rotate<<<blocks_for(h->replicas, 128), 128, 0, st>>>(h->dt_bits.p, h->dt_prev.p, h->dt_last.p,
                                                                      cfg->timestep, h->replicas);
            h->launches += 3;
        }
        h->wall_stats_fresh = false;
        if (h->relax.mask) {
            // the callback's wall update of every replica, on the device (bd_set_wall_relaxation)
            k_wall_relax<<<h->replicas, RELAX_WIDTH * RELAX_LANES, 0, st>>>(h->wall_partial.p, h->wall_out.p, h->wall_geom.p,
                                                                           h->relax, h->chunks, cfg->timestep);
            h->launches += 1;
            h->wall_stats_fresh = true;
        }
        CK(cudaGetLastError());
        h->cur ^= 1;
        h->last_dt = cfg->timestep;
        h->last_kT = cfg->temperature;
        h->steps_done++;
    }
    h->stats_adaptive_last = adaptive;
    if (replay) return sync_and_check(h);  // the caller's noise buffer must outlive the copies
    return 0;
}

extern "C" int bd_get_last_timestep(bd_engine *h, int32_t replica, double *out)
{
    NEED(h);
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_get_last_timestep: bad replica");
    if (!h->stats_adaptive_last) {
        *out = h->last_dt;
        return 0;
    }
    CK(cudaMemcpyAsync(out, h->dt_last.p + replica, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    return sync_and_check(h);
}

static int check_select(bd_engine *h, int32_t ff_class, int32_t ff)
{
    bool ok = false;
    switch (ff_class) {
    case BD_FF_ALL: ok = true; break;
    case BD_FF_PAIR: ok = ff >= 0 && ff < BD_MAX_PAIR_FF && h->pair[ff].used; break;
    case BD_FF_BOND: ok = ff >= 0 && ff < BD_MAX_BOND_FF && h->bond[ff].used; break;
    case BD_FF_BEND: ok = ff >= 0 && ff < BD_MAX_BEND_FF && h->bend_used[ff]; break;
    case BD_FF_WALL: ok = ff >= 0 && ff < BD_MAX_WALL_FF && h->wall[ff].used; break;
    case BD_FF_POINT: ok = ff >= 0 && ff < BD_MAX_POINT_FF && h->point[ff].used; break;
    default: break;
    }
    return ok ? 0 : fail("bd: bad forcefield class / id");
}

extern "C" int bd_compute_forces(bd_engine *h, int32_t replica, int32_t ff_class, int32_t ff, double *out_xyz)
{
    NEED(h);
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_compute_forces: bad replica");
    if (check_select(h, ff_class, ff)) return 1;
    if (h->ntot == 0) return 0;
    if (prepare_list(h)) return 1;
    StepArrays const A = step_arrays(h);
    StepScalars S;
    std::memset(&S, 0, sizeof S);
    Select const sel = {ff_class, ff};
    launch_force(h, A, S, sel);
    h->wall_stats_fresh = false;
    k_gather_forces<<<blocks_for(h->n, 256), 256, 0, h->stream>>>(h->force.p, h->slot_of.p, h->xyz_stage.p,
                                                                  (uint32_t) replica * h->n, h->n);
    h->launches += 2;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_xyz, h->xyz_stage.p, 3 * (size_t) h->n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    return sync_and_check(h);
}

extern "C" int bd_compute_energy(bd_engine *h, int32_t replica, int32_t ff_class, int32_t ff, double *out)
{
    NEED(h);
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_compute_energy: bad replica");
    if (check_select(h, ff_class, ff)) return 1;
    *out = 0;
    if (h->ntot == 0) return 0;
    if (prepare_list(h)) return 1;
    StepArrays const A = step_arrays(h);
    Select const sel = {ff_class, ff};
    k_energy<<<dim3(h->chunks, h->replicas), STEP_BLOCK, 0, h->stream>>>(A, sel, h->energy_partial.p);
    k_reduce_partials<<<h->replicas, 32, 0, h->stream>>>(h->energy_partial.p, h->energy_out.p, h->chunks, 1);
    h->launches += 2;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out, h->energy_out.p + replica, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    return sync_and_check(h);
}

/* wall_out <- per-replica sums of the per-chunk statistics of the latest force evaluation */
static int reduce_wall_stats(bd_engine *h)
{
    if (h->wall_stats_fresh) return 0;
    WallRelax none = {};
    k_wall_relax<<<h->replicas, RELAX_WIDTH * RELAX_LANES, 0, h->stream>>>(h->wall_partial.p, h->wall_out.p, h->wall_geom.p,
                                                                          none, h->chunks, 0.0);
    h->launches += 1;
    CK(cudaGetLastError());
    h->wall_stats_fresh = true;
    return 0;
}

extern "C" int bd_get_wall_stats(bd_engine *h, int32_t ff, int32_t replica, double out[4])
{
    NEED(h);
    if (ff < 0 || ff >= BD_MAX_WALL_FF || !h->wall[ff].used) return fail("bd_get_wall_stats: bad forcefield id");
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_get_wall_stats: bad replica");
    for (int c = 0; c < 4; c++) out[c] = 0;
    if (h->ntot == 0) return 0;
    if (reduce_wall_stats(h)) return 1;
    CK(cudaMemcpyAsync(out, h->wall_out.p + (size_t) replica * RELAX_WIDTH + (size_t) ff * WALL_STATS,
                       WALL_STATS * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    return sync_and_check(h);
}

/* the same for every replica at once: out[n_replicas][4] */
extern "C" int bd_get_wall_stats_all(bd_engine *h, int32_t ff, double *out)
{
    NEED(h);
    if (ff < 0 || ff >= BD_MAX_WALL_FF || !h->wall[ff].used) return fail("bd_get_wall_stats_all: bad forcefield id");
    if (!out) return fail("bd_get_wall_stats_all: null output");
    for (size_t c = 0; c < 4 * (size_t) h->replicas; c++) out[c] = 0;
    if (h->ntot == 0) return 0;
    if (reduce_wall_stats(h)) return 1;
    CK(cudaMemcpy2DAsync(out, WALL_STATS * sizeof(double), h->wall_out.p + (size_t) ff * WALL_STATS,
                         RELAX_WIDTH * sizeof(double), WALL_STATS * sizeof(double), h->replicas, cudaMemcpyDeviceToHost,
                         h->stream));
    return sync_and_check(h);
}

extern "C" int bd_get_neighbor_list(bd_engine *h, int32_t replica, uint32_t *pairs_out, int64_t cap,
                                    int64_t *count_out)
{
    NEED(h);
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_get_neighbor_list: bad replica");
    *count_out = 0;
    if (h->ntot == 0) return 0;
    if (prepare_list(h)) return 1;
    DevBuf<uint32_t> d_pairs;
    CK(d_pairs.alloc(2 * (size_t) std::max<int64_t>(cap, 1)));
    CK(cudaMemsetAsync(h->pair_count.p, 0, sizeof(unsigned long long), h->stream));
    k_export_pairs<<<blocks_for(h->n, NL_BLOCK), NL_BLOCK, 0, h->stream>>>(
        nlist_arrays(h), (uint32_t) replica, d_pairs.p, (unsigned long long) std::max<int64_t>(cap, 0), h->pair_count.p);
    h->launches += 1;
    unsigned long long count = 0;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(&count, h->pair_count.p, sizeof count, cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e == cudaSuccess && cap > 0)
        e = cudaMemcpy(pairs_out, d_pairs.p, 2 * (size_t) std::min<unsigned long long>(count, (unsigned long long) cap) * sizeof(uint32_t),
                       cudaMemcpyDeviceToHost);
    d_pairs.release();
    if (e != cudaSuccess) return fail(std::string("bd_get_neighbor_list: ") + cudaGetErrorString(e));
    *count_out = (int64_t) count;
    return 0;
}

extern "C" int bd_get_counters(bd_engine *h, bd_counters *out)
{
    NEED(h);
    std::memset(out, 0, sizeof *out);
    if (sync_and_check(h)) return 1;
    DControl c;
    CK(cudaMemcpy(&c, h->ctl.p, sizeof c, cudaMemcpyDeviceToHost));
    std::vector<unsigned long long> part(h->inter_partial.count);
    if (!part.empty())
        CK(cudaMemcpy(part.data(), h->inter_partial.p, part.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    unsigned long long inter = 0;
    for (unsigned long long v : part) inter += v;
    out->steps_done = h->steps_done;
    out->list_rebuilds = (int64_t) c.rebuilds - h->capacity_rebuilds;
    out->pair_interactions = (int64_t) (inter / 2);  // every pair is counted from both of its beads
    out->listed_pairs = (int64_t) (c.listed_entries / 2);
    out->overflow_beads = c.overflow;
    return 0;
}

// ---- engine-only controls ------------------------------------------------------------------------

extern "C" int bd_set_stream(bd_engine *h, void *cuda_stream)
{
    NEED(h);
    CK(cudaStreamSynchronize(h->stream));
    h->stream = cuda_stream ? (cudaStream_t) cuda_stream : h->own_stream;
    return 0;
}

extern "C" int bd_sync(bd_engine *h)
{
    NEED(h);
    return sync_and_check(h);
}

extern "C" int bd_set_option(bd_engine *h, const char *name, double value)
{
    NEED(h);
    std::string const key = name ? name : "";
    if (key == "verlet_factor") {
        if (!(value >= 1.0)) return fail("bd_set_option: verlet_factor must be >= 1");
        h->hp.verlet_factor = value;
        h->params_dirty = true;
        return request_rebuild(h);
    }
    if (key == "specialize") {
        h->specialize = value != 0;
        h->params_dirty = true;
        return 0;
    }
    if (key == "staged") {
        h->staged = value != 0;
        CK(cudaStreamSynchronize(h->stream));
        CK(cudaMemset(&h->ctl.p->stage_overflow, 0, sizeof(uint32_t)));
        return request_rebuild(h);
    }
    if (key == "pair_precision") {
        if (value != 32 && value != 64) return fail("bd_set_option: pair_precision is 32 or 64");
        h->pair_f32 = value == 32;
        return 0;
    }
    if (key == "balance_rows") {
        h->balance_rows = value != 0;
        return request_rebuild(h);
    }
    if (key == "stage_capacity") {  // 0 = follow the largest tile
        if (!(value >= 0 && value <= STAGE_CAP_CEILING)) return fail("bd_set_option: stage_capacity out of range");
        h->stage_auto = value == 0;
        h->stage_cap = value == 0 ? 512u : (uint32_t) value;
        return 0;
    }
    if (key == "time_kernel") {
        h->time_kernel = value != 0;
        return 0;
    }
    if (key == "step_graph") {
        h->use_graph = value != 0;
        return 0;
    }
    if (key == "replica_id_base" || key == "replica_id_stride") {
        if (!(value >= 0 && value < 4294967296.0)) return fail("bd_set_option: replica id out of range");
        (key == "replica_id_base" ? h->replica_base : h->replica_stride) = (uint32_t) value;
        return 0;
    }
    if (key == "list_capacity") {
        if (!(value >= 1 && value <= 4096)) return fail("bd_set_option: list_capacity out of range");
        CK(cudaStreamSynchronize(h->stream));
        h->cap = ((uint32_t) value + ROW_PAD - 1) & ~(ROW_PAD - 1);
        if (alloc_rows(h)) return 1;
        CK(cudaMemset(&h->ctl.p->overflow, 0, sizeof(uint32_t)));
        return request_rebuild(h);
    }
    return fail("bd_set_option: unknown option '" + key + "'");
}

extern "C" int bd_get_option(bd_engine *h, const char *name, double *value)
{
    NEED(h);
    std::string const key = name ? name : "";
    if (key == "verlet_factor") *value = h->hp.verlet_factor;
    else if (key == "list_capacity") *value = h->cap;
    else if (key == "block_size") *value = STEP_BLOCK;
    else if (key == "replica_id_base") *value = h->replica_base;
    else if (key == "replica_id_stride") *value = h->replica_stride;
    else if (key == "specialize") *value = h->specialize;
    else if (key == "staged") *value = h->staged;
    else if (key == "step_graph") *value = h->use_graph;
    else if (key == "balance_rows") *value = h->balance_rows;
    else if (key == "pair_precision") *value = h->pair_f32 ? 32 : 64;
    else if (key == "stage_capacity") *value = h->stage_cap;
    else if (key == "max_stage") {
        CK(cudaStreamSynchronize(h->stream));
        uint32_t v = 0;
        CK(cudaMemcpy(&v, &h->ctl.p->max_stage, sizeof v, cudaMemcpyDeviceToHost));
        *value = v;
    }
    else if (key == "pair_program") {
        if (upload_params(h)) return 1;
        *value = h->prog;
    }
    else if (key == "kernel_ms_avg") {
        // mean duration of the fused step kernel over the launches recorded since the last query
        CK(cudaStreamSynchronize(h->stream));
        double total = 0;
        for (auto &ev : h->kernel_events) {
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, ev.first, ev.second));
            total += ms;
            cudaEventDestroy(ev.first);
            cudaEventDestroy(ev.second);
        }
        *value = h->kernel_events.empty() ? 0.0 : total / (double) h->kernel_events.size();
        h->kernel_events.clear();
    }
    else return fail("bd_get_option: unknown option '" + key + "'");
    return 0;
}

extern "C" int bd_get_launch_count(bd_engine *h, int64_t *out)
{
    NEED(h);
    *out = h->launches;
    if (h->graph_steps) {
        // the ten chain kernels inside the step graphs' conditional nodes ran once per rebuild (device count)
        CK(cudaStreamSynchronize(h->stream));
        unsigned long long in_graph = 0;
        CK(cudaMemcpy(&in_graph, &h->ctl.p->graph_rebuilds, sizeof in_graph, cudaMemcpyDeviceToHost));
        *out += 10 * (int64_t) in_graph;
    }
    return 0;
}

// ---- statistics of the production noise -------------------------------------------------------------
//
// The reference draws double-precision ziggurat normals from one sequential sfc64 stream
// (MD/simulation/detail/brownian_simulator.hpp:61-70); the engine draws philox_normals (bd_device.cuh)
// keyed by (seed, step, replica, bead).  The streams cannot be compared value by value, so the generator is
// pinned by its statistics: this entry point draws the normals of `beads` beads for `steps` steps exactly as
// the step kernel does and returns their moments, tail counts and correlations.

constexpr int NOISE_STATS = 20;

__global__ void __launch_bounds__(256) k_noise_stats(unsigned long long seed, long long first_step, uint32_t steps,
                                                     uint32_t beads, uint32_t replica, double *out)
{
    double acc[NOISE_STATS];
    for (int c = 0; c < NOISE_STATS; c++) acc[c] = 0.0;
    unsigned long long const total = (unsigned long long) beads * steps;
    for (unsigned long long t = (unsigned long long) blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (unsigned long long) gridDim.x * blockDim.x) {
        uint32_t const bead = (uint32_t) (t % beads);
        unsigned long long const step = (unsigned long long) first_step + t / beads;
        double z[3], zn[3];
        philox_normals(seed, step, replica, bead, z);
        philox_normals(seed, step + 1, replica, bead, zn);
        for (int d = 0; d < 3; d++) {
            double const v = z[d], v2 = v * v, a = fabs(v);
            acc[0] += 1.0;
            acc[1] += v;
            acc[2] += v2;
            acc[3] += v2 * v;
            acc[4] += v2 * v2;
            acc[5] += v2 * v2 * v2;
            acc[6] = fmax(acc[6], a);
            for (int k = 1; k <= 7; k++) acc[6 + k] += a > (double) k ? 1.0 : 0.0;  // acc[7..13]: |z| > 1..7
            acc[14] += v * zn[d];                                                   // lag-1 in time, same bead
        }
        acc[15] += z[0] * z[1];  // within one Box-Muller pair
        acc[16] += z[0] * z[2];  // across the two pairs of one Philox call
        acc[17] += z[1] * z[2];
        acc[18] += z[0] * z[0] * z[1] * z[1];  // radius coupling inside a pair: E = 1 for independent normals
        acc[19] += 1.0;
    }
    for (int c = 0; c < NOISE_STATS; c++) {
        double v = acc[c];
        for (int o = 16; o > 0; o >>= 1) {
            double const w = __shfl_xor_sync(0xffffffffu, v, o);
            v = c == 6 ? fmax(v, w) : v + w;
        }
        if ((threadIdx.x & 31) == 0) {
            if (c == 6) atomicMax((unsigned long long *) &out[c], (unsigned long long) __double_as_longlong(v));
            else atomicAdd(&out[c], v);
        }
    }
}

extern "C" int bd_noise_statistics(uint64_t seed, int64_t first_step, int64_t steps, int64_t beads, int32_t replica,
                                   int32_t device, double *out, int32_t out_len)
{
    if (!out || out_len < NOISE_STATS) return fail("bd_noise_statistics: out must hold 20 doubles");
    if (steps < 1 || beads < 1 || first_step < 1 || beads >= (1ll << 32) || steps >= (1ll << 32))
        return fail("bd_noise_statistics: bad range");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail("bd_noise_statistics: no CUDA device");
    CK(cudaSetDevice(device));
    double *d = nullptr;
    CK(cudaMalloc((void **) &d, NOISE_STATS * sizeof(double)));
    cudaError_t e = cudaMemset(d, 0, NOISE_STATS * sizeof(double));
    if (e == cudaSuccess) {
        k_noise_stats<<<148 * 8, 256>>>(seed, first_step, (uint32_t) steps, (uint32_t) beads, (uint32_t) replica, d);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(out, d, NOISE_STATS * sizeof(double), cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) return fail(std::string("bd_noise_statistics: ") + cudaGetErrorString(e));
    return 0;
}

int bd_internal_fail(const char *msg)
{
    return fail(msg);
}

// ---- contact map on the device -------------------------------------------------------------------
//
// Replaces contact_map::update / accumulate / clear of the interphase driver
// (5-sim-genome/src/simulation_interphase/contact_map.cc:25-91), which downloads all positions every
// contactmap_update_interval steps, runs md::neighbor_searcher<open_box> with the default hash modulus
// on the host and adds the pairs into an Eigen sparse matrix.  Here the pairs within the contact
// distance are found by a second engine used as the searcher (list radius = contact distance, fp64
// test, the reference's pair set), turned into 64-bit keys (i << 32 | j) and merged into a sorted
// (key, count) table by one radix sort + one reduce-by-key (CUB).  Positions never leave the device;
// only bd_contact_accumulate copies the (i, j, count) triplets out.

__global__ void k_contact_keys(const uint32_t *pairs, unsigned long long count, unsigned long long *keys, uint32_t *ones)
{
    unsigned long long const k = (unsigned long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    uint32_t const a = pairs[2 * k], b = pairs[2 * k + 1];
    uint32_t const i = a < b ? a : b, j = a < b ? b : a;
    keys[k] = ((unsigned long long) i << 32) | j;
    ones[k] = 1u;
}

extern "C" int bd_contact_clear(bd_engine *h, int32_t replica)
{
    NEED(h);
    if (replica != 0 && (uint32_t) replica >= h->replicas) return fail("bd_contact_clear: bad replica");
    h->contact_size = 0;
    return 0;
}

/* Pairs of `replica`'s CURRENT configuration within `distance`, left in h->contact_pairs on the device
 * (unordered, i < j).  A second engine acts as the searcher: list radius = distance, the reference's fp64
 * test.  use_box = 0 searches in an open box whatever the simulation's unit cell is (contact_map.cc:64),
 * 1 in the simulation's own box (the searcher glue_simulator.cpp:42 keeps). */
static int search_current(bd_engine *h, int32_t replica, double distance, int use_box, unsigned long long *found_out)
{
    *found_out = 0;
    if (h->n < 2) return 0;
    cudaStream_t const st = h->stream;
    bd_engine *q = h->contact_search;
    int const kind = use_box ? h->hp.box_kind : BD_BOX_OPEN;
    double const unit[3] = {1, 1, 1};
    const double *period = kind == BD_BOX_PERIODIC ? h->hp.period : unit;
    bd_term none;
    std::memset(&none, 0, sizeof none);
    if (!q) {
        if (bd_create(h->n, 1, h->device, &q)) return 1;
        h->contact_search = q;
        int32_t ff = -1;
        // a list radius of exactly the distance: the rows hold precisely the wanted pairs
        if (bd_set_box(q, kind, period) || bd_set_option(q, "verlet_factor", 1.0)
            || bd_add_pair_ff(q, 1, &none, distance, nullptr, &ff))
            return 1;
        h->contact_distance = distance;
    } else {
        if (q->hp.box_kind != kind || (kind == BD_BOX_PERIODIC && std::memcmp(q->hp.period, period, 3 * sizeof(double)) != 0))
            if (bd_set_box(q, kind, period)) return 1;
        if (distance != h->contact_distance) {
            // the engine keeps a list whose radius still covers the new cutoff; the searcher wants rows at
            // exactly the new distance
            if (bd_update_pair_ff(q, 0, &none, distance) || request_rebuild(q)) return 1;
            h->contact_distance = distance;
        }
    }
    if (bd_set_stream(q, st)) return 1;
    // positions of the replica in original order, device to device
    k_gather_positions<<<blocks_for(h->n, 256), 256, 0, st>>>(h->pos[h->cur].p, h->slot_of.p, h->xyz_stage.p,
                                                              (uint32_t) replica * h->n, h->n);
    h->launches++;
    CK(cudaGetLastError());
    if (bd_set_positions_device(q, h->xyz_stage.p)) return 1;
    if (prepare_list(q)) return 1;

    DControl c;
    CK(cudaMemcpy(&c, q->ctl.p, sizeof c, cudaMemcpyDeviceToHost));
    size_t const cap = (size_t) (c.listed_entries / 2) + 1;  // the rows are a superset by the shadow bound
    if (h->contact_pairs.count < 2 * cap) CK(h->contact_pairs.alloc(2 * cap + cap / 4));
    CK(cudaMemsetAsync(q->pair_count.p, 0, sizeof(unsigned long long), st));
    k_export_pairs<<<blocks_for(q->n, NL_BLOCK), NL_BLOCK, 0, st>>>(nlist_arrays(q), 0u, h->contact_pairs.p,
                                                                     (unsigned long long) cap, q->pair_count.p);
    h->launches++;
    CK(cudaGetLastError());
    unsigned long long found = 0;
    CK(cudaMemcpyAsync(&found, q->pair_count.p, sizeof found, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (found > cap) return fail("bd: pair export overflow in the searcher");
    *found_out = found;
    return 0;
}

extern "C" int bd_contact_update(bd_engine *h, int32_t replica, double distance)
{
    NEED(h);
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_contact_update: bad replica");
    if (!(distance > 0)) return fail("bd_contact_update: the contact distance must be positive");
    cudaStream_t const st = h->stream;
    unsigned long long found = 0;
    if (search_current(h, replica, distance, 0, &found)) return 1;
    if (found == 0) return 0;

    // merge: [accumulated keys + counts | new keys + ones] -> sort by key -> sum counts of equal keys
    int const cur = h->contact_cur, nxt = cur ^ 1;
    size_t const total = h->contact_size + (size_t) found;
    for (DevBuf<unsigned long long> *b : {&h->contact_keys[cur], &h->contact_keys[nxt], &h->contact_new})
        if (b->count < total) {
            DevBuf<unsigned long long> grown;
            CK(grown.alloc(total + total / 2));
            if (b == &h->contact_keys[cur] && h->contact_size)
                CK(cudaMemcpyAsync(grown.p, b->p, h->contact_size * sizeof(unsigned long long), cudaMemcpyDeviceToDevice, st));
            CK(cudaStreamSynchronize(st));
            b->release();
            *b = grown;
            grown.p = nullptr;
            grown.count = 0;
        }
    for (DevBuf<uint32_t> *b : {&h->contact_counts[cur], &h->contact_counts[nxt], &h->contact_ones})
        if (b->count < total) {
            DevBuf<uint32_t> grown;
            CK(grown.alloc(total + total / 2));
            if (b == &h->contact_counts[cur] && h->contact_size)
                CK(cudaMemcpyAsync(grown.p, b->p, h->contact_size * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
            CK(cudaStreamSynchronize(st));
            b->release();
            *b = grown;
            grown.p = nullptr;
            grown.count = 0;
        }
    if (h->contact_runs.count < 1) CK(h->contact_runs.alloc(1));
    k_contact_keys<<<(unsigned) ((found + 255) / 256), 256, 0, st>>>(h->contact_pairs.p, found,
                                                                     h->contact_keys[cur].p + h->contact_size,
                                                                     h->contact_counts[cur].p + h->contact_size);
    h->launches++;
    CK(cudaGetLastError());
    size_t need_sort = 0, need_reduce = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, need_sort, h->contact_keys[cur].p, h->contact_new.p, h->contact_counts[cur].p,
                                       h->contact_ones.p, (int) total, 0, 64, st));
    CK(cub::DeviceReduce::ReduceByKey(nullptr, need_reduce, h->contact_new.p, h->contact_keys[nxt].p, h->contact_ones.p,
                                      h->contact_counts[nxt].p, h->contact_runs.p, cuda::std::plus<>(), (int) total, st));
    size_t const need = std::max(need_sort, need_reduce);
    if (h->contact_temp.count < need) CK(h->contact_temp.alloc(need + need / 4));
    size_t bytes = h->contact_temp.count;
    CK(cub::DeviceRadixSort::SortPairs(h->contact_temp.p, bytes, h->contact_keys[cur].p, h->contact_new.p,
                                       h->contact_counts[cur].p, h->contact_ones.p, (int) total, 0, 64, st));
    bytes = h->contact_temp.count;
    CK(cub::DeviceReduce::ReduceByKey(h->contact_temp.p, bytes, h->contact_new.p, h->contact_keys[nxt].p, h->contact_ones.p,
                                      h->contact_counts[nxt].p, h->contact_runs.p, cuda::std::plus<>(), (int) total, st));
    h->launches += 2;
    unsigned long long runs = 0;
    CK(cudaMemcpyAsync(&runs, h->contact_runs.p, sizeof runs, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    h->contact_cur = nxt;
    h->contact_size = (size_t) runs;
    return 0;
}

extern "C" int bd_contact_accumulate(bd_engine *h, int32_t replica, uint32_t *triplets_out, int64_t cap, int64_t *count_out)
{
    NEED(h);
    (void) replica;
    if (!count_out || cap < 0) return fail("bd_contact_accumulate: bad arguments");
    *count_out = (int64_t) h->contact_size;
    size_t const m = std::min<size_t>(h->contact_size, (size_t) cap);
    if (m == 0) return 0;
    std::vector<unsigned long long> keys(m);
    std::vector<uint32_t> counts(m);
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaMemcpy(keys.data(), h->contact_keys[h->contact_cur].p, m * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(counts.data(), h->contact_counts[h->contact_cur].p, m * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    for (size_t k = 0; k < m; k++) {  // row-major (i, j) order: the order of the reference's sparse matrix walk
        triplets_out[3 * k] = (uint32_t) (keys[k] >> 32);
        triplets_out[3 * k + 1] = (uint32_t) keys[k];
        triplets_out[3 * k + 2] = counts[k];
    }
    return 0;
}

// ---- single-precision snapshots ------------------------------------------------------------------
//
// The step either side of the hot path on sampling steps.  The drivers' stores convert the fp64
// positions to float on the host after a blocking download: 5-sim-genome quantises to 16 fraction bits
// (simulation_common/simulation_store.cc:253-276 with quantize() at :403-407), 4-sim-ab and 3-sim-1kb
// hand the doubles to an HDF5 float dataset, i.e. a plain conversion (4-sim-ab/box/src/simulation/
// simulation_store.cc:57-66, 3-sim-1kb/src/simulation/store.cpp).  Here the conversion runs on the device in
// original bead order, the float buffer (half the bytes) crosses PCIe on a second stream into pinned
// memory, and the engine keeps stepping meanwhile.

__global__ void k_snapshot(const bead4 *pos, const uint32_t *slot_of, float *out, uint32_t first, uint32_t count,
                           int fraction_bits)
{
    uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const bead4 *p = pos + slot_of[first + i];
    float v[3] = {__double2float_rn(p->x), __double2float_rn(p->y), __double2float_rn(p->z)};
    if (fraction_bits >= 0) {
        // quantize(): nearbyint(float(val) * scale) / scale, scale = 1 << bits, all in float
        float const scale = (float) (1u << fraction_bits);
#pragma unroll
        for (int d = 0; d < 3; d++) v[d] = __fdiv_rn(rintf(__fmul_rn(v[d], scale)), scale);
    }
    out[3 * (size_t) i] = v[0];
    out[3 * (size_t) i + 1] = v[1];
    out[3 * (size_t) i + 2] = v[2];
}

extern "C" int bd_snapshot_begin(bd_engine *h, int32_t replica, int32_t fraction_bits)
{
    NEED(h);
    if (replica >= (int32_t) h->replicas) return fail("bd_snapshot_begin: bad replica");
    if (fraction_bits > 30) return fail("bd_snapshot_begin: at most 30 fraction bits");
    if (h->snap_pending) return fail("bd_snapshot_begin: the previous snapshot has not been collected (bd_snapshot_wait)");
    uint32_t const first = replica < 0 ? 0u : (uint32_t) replica * h->n;
    uint32_t const count = replica < 0 ? h->ntot : h->n;
    h->snap_count = count;
    if (count == 0) {
        h->snap_pending = true;
        return 0;
    }
    if (!h->snap_stream) {
        CK(cudaStreamCreateWithFlags(&h->snap_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&h->snap_ready, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&h->snap_done, cudaEventDisableTiming));
    }
    size_t const floats = 3 * (size_t) count;
    if (h->snap_dev.count < floats) {
        CK(cudaStreamSynchronize(h->snap_stream));
        CK(h->snap_dev.alloc(floats));
    }
    if (h->snap_host_count < floats) {
        if (h->snap_host) CK(cudaFreeHost(h->snap_host));
        h->snap_host = nullptr;
        h->snap_host_count = 0;
        CK(cudaHostAlloc((void **) &h->snap_host, floats * sizeof(float), cudaHostAllocDefault));
        h->snap_host_count = floats;
    }
    k_snapshot<<<blocks_for(count, 256), 256, 0, h->stream>>>(h->pos[h->cur].p, h->slot_of.p, h->snap_dev.p, first, count,
                                                             fraction_bits);
    h->launches++;
    CK(cudaGetLastError());
    CK(cudaEventRecord(h->snap_ready, h->stream));
    CK(cudaStreamWaitEvent(h->snap_stream, h->snap_ready, 0));
    CK(cudaMemcpyAsync(h->snap_host, h->snap_dev.p, floats * sizeof(float), cudaMemcpyDeviceToHost, h->snap_stream));
    CK(cudaEventRecord(h->snap_done, h->snap_stream));
    h->snap_pending = true;
    return 0;
}

extern "C" int bd_snapshot_wait(bd_engine *h, const float **xyz_out, int64_t *count_out)
{
    NEED(h);
    if (!h->snap_pending) return fail("bd_snapshot_wait: no snapshot in flight (bd_snapshot_begin)");
    if (h->snap_count) CK(cudaEventSynchronize(h->snap_done));
    h->snap_pending = false;
    if (xyz_out) *xyz_out = h->snap_host;
    if (count_out) *count_out = (int64_t) h->snap_count;
    return 0;
}

// ---- pair search and pair distances on the resident positions -----------------------------------
//
// What 3-sim-1kb's kinetic models ask of the positions every update (glues/glue_simulator.cpp:35-80): the
// pairs within the binding distance (search) and the current distance of every glued pair (unbond).  Both
// are answered from the device-resident positions; only pairs and distances cross PCIe.

__global__ void k_unpack_keys(const unsigned long long *keys, unsigned long long count, uint32_t *pairs)
{
    unsigned long long const k = (unsigned long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    pairs[2 * k] = (uint32_t) (keys[k] >> 32);
    pairs[2 * k + 1] = (uint32_t) keys[k];
}

__global__ void k_pair_distances(const bead4 *pos, const uint32_t *slot_of, uint32_t base, const uint32_t *pairs, uint32_t m,
                                 const DParams *params, int min_image, double *out)
{
    uint32_t const k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    bead4 const a = pos[slot_of[base + pairs[2 * k]]], b = pos[slot_of[base + pairs[2 * k + 1]]];
    d3 r;
    if (min_image) {
        r = displacement(*params, a.x, a.y, a.z, b.x, b.y, b.z);
    } else {
        r.x = __dsub_rn(a.x, b.x);
        r.y = __dsub_rn(a.y, b.y);
        r.z = __dsub_rn(a.z, b.z);
    }
    out[k] = sqrt(sqnorm_exact(r.x, r.y, r.z));  // md::vector::norm (MD/basic_types/point.hpp:120-134)
}

extern "C" int bd_search_pairs(bd_engine *h, int32_t replica, double distance, int32_t use_box, uint32_t *pairs_out,
                               int64_t cap, int64_t *count_out)
{
    NEED(h);
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_search_pairs: bad replica");
    if (!(distance > 0) || cap < 0 || !count_out) return fail("bd_search_pairs: bad arguments");
    cudaStream_t const st = h->stream;
    unsigned long long found = 0;
    if (search_current(h, replica, distance, use_box != 0, &found)) return 1;
    *count_out = (int64_t) found;
    size_t const m = std::min<size_t>((size_t) found, (size_t) cap);
    if (found == 0 || m == 0) return 0;
    // sorted by (i, j): the device emits the pairs in no particular order
    DevBuf<unsigned long long> keys, sorted;
    DevBuf<uint32_t> ones;  // k_contact_keys also writes a count column
    CK(keys.alloc(found));
    CK(sorted.alloc(found));
    CK(ones.alloc(found));
    k_contact_keys<<<(unsigned) ((found + 255) / 256), 256, 0, st>>>(h->contact_pairs.p, found, keys.p, ones.p);
    size_t need = 0;
    CK(cub::DeviceRadixSort::SortKeys(nullptr, need, keys.p, sorted.p, (int) found, 0, 64, st));
    if (h->contact_temp.count < need) CK(h->contact_temp.alloc(need + need / 4));
    size_t bytes = h->contact_temp.count;
    CK(cub::DeviceRadixSort::SortKeys(h->contact_temp.p, bytes, keys.p, sorted.p, (int) found, 0, 64, st));
    k_unpack_keys<<<(unsigned) ((found + 255) / 256), 256, 0, st>>>(sorted.p, found, h->contact_pairs.p);
    h->launches += 3;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(pairs_out, h->contact_pairs.p, 2 * m * sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    keys.release();
    sorted.release();
    ones.release();
    if (e != cudaSuccess) return fail(std::string("bd_search_pairs: ") + cudaGetErrorString(e));
    return 0;
}

extern "C" int bd_pair_distances(bd_engine *h, int32_t replica, const uint32_t *pairs, int64_t m, int32_t min_image,
                                 double *out)
{
    NEED(h);
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_pair_distances: bad replica");
    if (m < 0 || (m > 0 && (!pairs || !out))) return fail("bd_pair_distances: bad arguments");
    for (int64_t k = 0; k < 2 * m; k++)
        if (pairs[k] >= h->n) return fail("bd_pair_distances: bead index out of range");
    if (m == 0) return 0;
    if (upload_params(h)) return 1;
    DevBuf<uint32_t> d_pairs;
    DevBuf<double> d_out;
    CK(d_pairs.alloc(2 * (size_t) m));
    CK(d_out.alloc((size_t) m));
    cudaError_t e = cudaMemcpyAsync(d_pairs.p, pairs, 2 * (size_t) m * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream);
    k_pair_distances<<<blocks_for((size_t) m, 256), 256, 0, h->stream>>>(h->pos[h->cur].p, h->slot_of.p, (uint32_t) replica * h->n,
                                                                         d_pairs.p, (uint32_t) m, h->d_params.p, min_image != 0,
                                                                         d_out.p);
    h->launches++;
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out.p, (size_t) m * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    d_pairs.release();
    d_out.release();
    if (e != cudaSuccess) return fail(std::string("bd_pair_distances: ") + cudaGetErrorString(e));
    return 0;
}

// ---- kinetic glue update on the device (second half) ----------------------------------
//
// Replaces glue_simulator::update of the 1kb chromatin driver (3-sim-1kb/src/simulation/glues/
// glue_simulator.cpp:35-80), which at every update downloads all positions, searches them on the host and
// walks the pairs through one sequential generator:
//   unbond: a glued pair comes apart if its box distance exceeds max_distance or with probability p_unbind;
//   bond:   every pair within max_distance that is not glued passes with probability p_bind; the passing pairs
//           fill the free capacity (max_glues - glued), a uniformly random subset of them if they are too many
//           (reservoir sampling, glues/reservoir_sampler.hpp:57-93).
// Here the glue list is a bonded pair list of the engine (bd_set_bond_pairs); the update runs on the resident
// positions: distances and Bernoulli draws per glued pair, the candidate search of bd_search_pairs, a binary
// search against the sorted survivors, Bernoulli draws per free candidate, and -- when the accepted ones
// exceed the capacity -- a radix sort by a random 64-bit tag whose first `capacity` entries are a uniform
// subset.  Only the new list (<= max_glues pairs) goes to the host, which rebuilds the bond adjacency.
// The draws are counter-based (Philox keyed by seed, update index and the pair), not the reference's
// sequential stream in search order: the update is the same stochastic process, not the same sample path --
// tests/test_gpu_glue.py holds its rates and the glue-count distribution against the reference algorithm.

__device__ inline void pair_uniforms(unsigned long long seed, unsigned long long update, uint32_t i, uint32_t j, uint32_t stream,
                                     double &u, unsigned long long &tag)
{
    uint32_t c[4] = {i, j, (uint32_t) update, ((uint32_t) (update >> 32) << 8) | stream};
    philox4x32_10(c, (uint32_t) seed, (uint32_t) (seed >> 32));
    u = ((double) c[0] * 4294967296.0 + (double) c[1] + 0.5) * 5.421010862427522e-20;  // 2^-64: uniform in (0, 1)
    tag = ((unsigned long long) c[2] << 32) | c[3];
}

__global__ void k_glue_unbond(const bead4 *pos, const uint32_t *slot_of, uint32_t base, const uint32_t *pairs, uint32_t m,
                              const DParams *params, int use_box, double max_distance, double p_unbind,
                              unsigned long long seed, unsigned long long update, unsigned long long *keys, uint8_t *keep)
{
    uint32_t const k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    uint32_t const a = pairs[2 * k], b = pairs[2 * k + 1];
    uint32_t const i = a < b ? a : b, j = a < b ? b : a;
    bead4 const p = pos[slot_of[base + i]], q = pos[slot_of[base + j]];
    d3 r = {__dsub_rn(p.x, q.x), __dsub_rn(p.y, q.y), __dsub_rn(p.z, q.z)};
    if (use_box) r = displacement(*params, p.x, p.y, p.z, q.x, q.y, q.z);
    double const dist = sqrt(sqnorm_exact(r.x, r.y, r.z));
    double u;
    unsigned long long tag;
    pair_uniforms(seed, update, i, j, 0u, u, tag);
    keys[k] = ((unsigned long long) i << 32) | j;
    keep[k] = !(dist > max_distance || u < p_unbind);  // glue_simulator.cpp:57
}

__global__ void k_glue_candidates(const uint32_t *cand, unsigned long long found, const unsigned long long *glued,
                                  unsigned long long nglued, double p_bind, unsigned long long seed, unsigned long long update,
                                  unsigned long long *keys, unsigned long long *tags, uint8_t *accept)
{
    unsigned long long const k = (unsigned long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= found) return;
    uint32_t const a = cand[2 * k], b = cand[2 * k + 1];
    uint32_t const i = a < b ? a : b, j = a < b ? b : a;
    unsigned long long const key = ((unsigned long long) i << 32) | j;
    unsigned long long lo = 0, hi = nglued;  // is the pair glued already? (glue_simulator.cpp:74)
    while (lo < hi) {
        unsigned long long const mid = (lo + hi) >> 1;
        if (glued[mid] < key) lo = mid + 1; else hi = mid;
    }
    bool const already = lo < nglued && glued[lo] == key;
    double u;
    unsigned long long tag;
    pair_uniforms(seed, update, i, j, 1u, u, tag);
    keys[k] = key;
    tags[k] = tag;
    accept[k] = !already && u < p_bind;
}

extern "C" int bd_glue_update(bd_engine *h, int32_t bond_ff, int32_t replica, double max_distance, double p_bind,
                              double p_unbind, int64_t max_glues, uint64_t seed, int64_t update_index, int32_t use_box,
                              int64_t *unbound_out, int64_t *bound_out, int64_t *count_out)
{
    NEED(h);
    if (bond_ff < 0 || bond_ff >= BD_MAX_BOND_FF || !h->bond[bond_ff].used) return fail("bd_glue_update: bad forcefield id");
    if (replica < 0 || (uint32_t) replica >= h->replicas) return fail("bd_glue_update: bad replica");
    if (h->replicas != 1) return fail("bd_glue_update: the pair lists of an ensemble are shared; one replica per engine");
    if (!(max_distance > 0) || !(p_bind >= 0 && p_bind <= 1) || !(p_unbind >= 0 && p_unbind <= 1) || max_glues < 0)
        return fail("bd_glue_update: bad parameters");
    std::vector<uint32_t> &list = h->bond_pairs[bond_ff];
    size_t const m = list.size() / 2;
    if (max_glues == 0) {
        // glue_simulator::update returns before unbond when there is no capacity (glue_simulator.cpp:37-39):
        // the list, whatever it holds, is left alone
        if (unbound_out) *unbound_out = 0;
        if (bound_out) *bound_out = 0;
        if (count_out) *count_out = (int64_t) m;
        return 0;
    }
    if (upload_params(h)) return 1;
    cudaStream_t const st = h->stream;

    // ---- unbond ----------------------------------------------------------------------------------
    // scratch of this call: freed on every way out, the early error returns of CK included
    ScopedBuf<uint32_t> d_pairs;
    ScopedBuf<unsigned long long> keys, kept, sorted, d_count;
    ScopedBuf<uint8_t> flags;
    CK(d_count.alloc(1));
    unsigned long long nkept = 0;
    size_t need = 0;
    if (m) {
        CK(d_pairs.alloc(2 * m));
        CK(keys.alloc(m));
        CK(kept.alloc(m));
        CK(sorted.alloc(m));
        CK(flags.alloc(m));
        CK(cudaMemcpyAsync(d_pairs.p, list.data(), 2 * m * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
        k_glue_unbond<<<blocks_for(m, 256), 256, 0, st>>>(h->pos[h->cur].p, h->slot_of.p, (uint32_t) replica * h->n, d_pairs.p,
                                                         (uint32_t) m, h->d_params.p, use_box != 0, max_distance, p_unbind, seed,
                                                         (unsigned long long) update_index, keys.p, flags.p);
        CK(cub::DeviceSelect::Flagged(nullptr, need, keys.p, flags.p, kept.p, d_count.p, (int) m, st));
        if (h->contact_temp.count < need) CK(h->contact_temp.alloc(need + need / 4));
        size_t bytes = h->contact_temp.count;
        CK(cub::DeviceSelect::Flagged(h->contact_temp.p, bytes, keys.p, flags.p, kept.p, d_count.p, (int) m, st));
        CK(cudaMemcpyAsync(&nkept, d_count.p, sizeof nkept, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (nkept) {
            CK(cub::DeviceRadixSort::SortKeys(nullptr, need, kept.p, sorted.p, (int) nkept, 0, 64, st));
            if (h->contact_temp.count < need) CK(h->contact_temp.alloc(need + need / 4));
            bytes = h->contact_temp.count;
            CK(cub::DeviceRadixSort::SortKeys(h->contact_temp.p, bytes, kept.p, sorted.p, (int) nkept, 0, 64, st));
        }
        h->launches += 3;
    }
    if (unbound_out) *unbound_out = (int64_t) (m - nkept);

    // ---- bond ------------------------------------------------------------------------------------
    unsigned long long found = 0, naccepted = 0;
    size_t const capacity = (size_t) max_glues > nkept ? (size_t) max_glues - (size_t) nkept : 0;
    ScopedBuf<unsigned long long> ckeys, ctags, akeys, atags, skeys, stags;
    ScopedBuf<uint8_t> accept;
    if (capacity && p_bind > 0) {
        if (search_current(h, replica, max_distance, use_box != 0, &found)) return 1;
    }
    if (found) {
        CK(ckeys.alloc(found));
        CK(ctags.alloc(found));
        CK(akeys.alloc(found));
        CK(atags.alloc(found));
        CK(accept.alloc(found));
        k_glue_candidates<<<(unsigned) ((found + 255) / 256), 256, 0, st>>>(h->contact_pairs.p, found, sorted.p, nkept, p_bind, seed,
                                                                          (unsigned long long) update_index, ckeys.p, ctags.p,
                                                                          accept.p);
        size_t bytes = 0;
        CK(cub::DeviceSelect::Flagged(nullptr, need, ckeys.p, accept.p, akeys.p, d_count.p, (int) found, st));
        if (h->contact_temp.count < need) CK(h->contact_temp.alloc(need + need / 4));
        bytes = h->contact_temp.count;
        CK(cub::DeviceSelect::Flagged(h->contact_temp.p, bytes, ckeys.p, accept.p, akeys.p, d_count.p, (int) found, st));
        bytes = h->contact_temp.count;
        CK(cub::DeviceSelect::Flagged(h->contact_temp.p, bytes, ctags.p, accept.p, atags.p, d_count.p, (int) found, st));
        CK(cudaMemcpyAsync(&naccepted, d_count.p, sizeof naccepted, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        h->launches += 3;
    }
    size_t const nbound = std::min<size_t>((size_t) naccepted, capacity);
    const unsigned long long *chosen = akeys.p;
    if (naccepted > capacity) {
        // too many: the `capacity` entries with the smallest random tags are a uniformly random subset
        CK(skeys.alloc(naccepted));
        CK(stags.alloc(naccepted));
        CK(cub::DeviceRadixSort::SortPairs(nullptr, need, atags.p, stags.p, akeys.p, skeys.p, (int) naccepted, 0, 64, st));
        if (h->contact_temp.count < need) CK(h->contact_temp.alloc(need + need / 4));
        size_t bytes = h->contact_temp.count;
        CK(cub::DeviceRadixSort::SortPairs(h->contact_temp.p, bytes, atags.p, stags.p, akeys.p, skeys.p, (int) naccepted, 0, 64, st));
        chosen = skeys.p;
        h->launches += 1;
    }
    if (bound_out) *bound_out = (int64_t) nbound;

    // ---- the new list: survivors (sorted) then the newly bound pairs -----------------------------------
    std::vector<unsigned long long> out(nkept + nbound);
    if (nkept) CK(cudaMemcpyAsync(out.data(), sorted.p, nkept * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    if (nbound) CK(cudaMemcpyAsync(out.data() + nkept, chosen, nbound * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    list.resize(2 * out.size());
    for (size_t k = 0; k < out.size(); k++) {
        list[2 * k] = (uint32_t) (out[k] >> 32);
        list[2 * k + 1] = (uint32_t) out[k];
    }
    if (count_out) *count_out = (int64_t) out.size();
    return upload_topology(h);
}

extern "C" int bd_get_bond_pairs(bd_engine *h, int32_t bond_ff, uint32_t *pairs_out, int64_t cap, int64_t *count_out)
{
    NEED(h);
    if (bond_ff < 0 || bond_ff >= BD_MAX_BOND_FF || !h->bond[bond_ff].used) return fail("bd_get_bond_pairs: bad forcefield id");
    const std::vector<uint32_t> &list = h->bond_pairs[bond_ff];
    int64_t const m = (int64_t) (list.size() / 2);
    if (count_out) *count_out = m;
    int64_t const take = std::min<int64_t>(m, std::max<int64_t>(cap, 0));
    if (take > 0 && pairs_out) std::memcpy(pairs_out, list.data(), (size_t) take * 2 * sizeof(uint32_t));
    return 0;
}
