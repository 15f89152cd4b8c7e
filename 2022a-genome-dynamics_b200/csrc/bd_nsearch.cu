/*
 * bd_nsearch.cu -- stateless neighbour-search entry points (validation and analysis callers).
 *
 *   bd_nsearch_buckets  md::neighbor_searcher<Box>::set_points -> search_grid::locate_bucket
 *                       (MD/misc/nsearch_detail/search_grid.hpp:80-86,142-145,231-252,
 *                        MD/misc/linear_hash.hpp:21-43, MD/misc/nsearch_detail/math.hpp:27-86):
 *                       the reference's own bucket index of every point, bit for bit.
 *   bd_nsearch_pairs    md::neighbor_searcher<Box>::search (MD/misc/neighbor_searcher.hpp:54-100):
 *                       all unordered pairs within dcut, found with the engine's cell list (list
 *                       radius = dcut) and the reference's fp64 distance test.
 */
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/bd_cuda.h"
#include "bd_device.cuh"

using namespace bd;

extern int bd_internal_fail(const char *msg);

struct BucketGrid {
    int box_kind;
    double freq;        // open: 1 / spacing
    uint32_t modulus;   // open: hash modulus
    double step[3];     // periodic: bin width
    double inv_step[3];
    double span[3], inv_span[3];
    uint32_t count[3];
};

__global__ void k_locate_bucket(BucketGrid g, const double *xyz, uint32_t *out, size_t n)
{
    size_t const i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double const p[3] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    if (g.box_kind == BD_BOX_OPEN) {
        // search_grid.hpp:231-244: uint32(2^20 + freq * x), then linear_hash mod M
        uint32_t c[3];
        for (int d = 0; d < 3; d++) c[d] = (uint32_t) __dadd_rn(1048576.0, __dmul_rn(g.freq, p[d]));
        unsigned long long sum = 0;
        sum += 3929498747ull * c[0];
        sum += 1008281837ull * c[1];
        sum += 1832832077ull * c[2];
        out[i] = (uint32_t) (sum % g.modulus);
    } else {
        // math.hpp:78-86 locate_bin per axis, search_grid.hpp:142-145 x + nx*(y + ny*z)
        uint32_t b[3];
        for (int d = 0; d < 3; d++) {
            double const image = floor_mod_exact(p[d], g.span[d], g.inv_span[d]);
            uint32_t const idx = (uint32_t) (int32_t) __dmul_rn(image, g.inv_step[d]);
            b[d] = idx < g.count[d] ? idx : g.count[d] - 1;
        }
        out[i] = b[0] + g.count[0] * (b[1] + g.count[1] * b[2]);
    }
}

extern "C" int bd_nsearch_buckets(int32_t box_kind, const double period[3], int64_t particle_count,
                                  double spacing, const double *xyz, int64_t n, uint32_t *bucket_out,
                                  int32_t device)
{
    if (n < 0 || !(spacing > 0)) return bd_internal_fail("bd_nsearch_buckets: bad arguments");
    if (cudaSetDevice(device) != cudaSuccess) return bd_internal_fail("bd_nsearch_buckets: no such CUDA device");
    BucketGrid g;
    std::memset(&g, 0, sizeof g);
    g.box_kind = box_kind;
    if (box_kind == BD_BOX_OPEN) {
        g.freq = 1 / spacing;
        g.modulus = (uint32_t) ((uint64_t) particle_count * 2 / 11) | 1u;  // search_grid.hpp:247-252
    } else if (box_kind == BD_BOX_PERIODIC) {
        for (int d = 0; d < 3; d++) {
            // math.hpp:68-76 define_bins(span, min_step)
            uint32_t const c = (uint32_t) (int32_t) (period[d] / spacing);
            g.count[d] = c > 1 ? c : 1;
            g.step[d] = period[d] / (double) g.count[d];
            g.inv_step[d] = 1 / g.step[d];
            g.span[d] = g.step[d] * (double) g.count[d];
            g.inv_span[d] = 1 / g.span[d];
        }
    } else {
        return bd_internal_fail("bd_nsearch_buckets: unknown box kind");
    }
    if (n == 0) return 0;
    double *d_xyz = nullptr;
    uint32_t *d_out = nullptr;
    cudaError_t e = cudaMalloc((void **) &d_xyz, 3 * (size_t) n * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc((void **) &d_out, (size_t) n * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemcpy(d_xyz, xyz, 3 * (size_t) n * sizeof(double), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        k_locate_bucket<<<(unsigned) ((n + 255) / 256), 256>>>(g, d_xyz, d_out, (size_t) n);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(bucket_out, d_out, (size_t) n * sizeof(uint32_t), cudaMemcpyDeviceToHost);
    cudaFree(d_xyz);
    cudaFree(d_out);
    if (e != cudaSuccess) return bd_internal_fail((std::string("bd_nsearch_buckets: ") + cudaGetErrorString(e)).c_str());
    return 0;
}

extern "C" int bd_nsearch_pairs(int32_t box_kind, const double period[3], int64_t particle_count, double dcut,
                                const double *xyz, int64_t n, uint32_t *pairs_out, int64_t cap,
                                int64_t *count_out, int32_t device)
{
    (void) particle_count;  // only shapes the reference's hash table, not the result
    *count_out = 0;
    if (n < 0 || !(dcut > 0)) return bd_internal_fail("bd_nsearch_pairs: bad arguments");
    if (n == 0) return 0;
    bd_engine *h = nullptr;
    if (bd_create(n, 1, device, &h)) return 1;
    int rc = bd_set_box(h, box_kind, period);
    // a list radius of exactly dcut: the rows then hold precisely the pairs with r^2 <= dcut^2
    if (!rc) rc = bd_set_option(h, "verlet_factor", 1.0);
    bd_term none;
    std::memset(&none, 0, sizeof none);
    int32_t ff = -1;
    if (!rc) rc = bd_add_pair_ff(h, 1, &none, dcut, nullptr, &ff);
    if (!rc) rc = bd_set_positions(h, 0, xyz);
    if (!rc) rc = bd_get_neighbor_list(h, 0, pairs_out, cap, count_out);
    bd_destroy(h);
    return rc;
}

/* The same pair set in the ORDER md::neighbor_searcher<Box>::search emits it (MD/misc/neighbor_searcher.hpp:54-100):
 * buckets in index order; for each, its sorted directed neighbours (open box: neighbour >= centre,
 * search_grid.hpp:279-283; periodic: adjacent <= centre, :124-127); members of the neighbour bucket in the
 * outer loop, members of the centre bucket in the inner loop, both in insertion (= index) order.  So a pair
 * sorts by (centre bucket, neighbour bucket, index of the bead in the neighbour bucket, index of the bead in
 * the centre bucket); within one bucket the larger index is the outer member.  3-sim-1kb's glue simulator
 * feeds a sequential RNG in this order (glues/glue_simulator.cpp:66-80), so a relinked driver takes the same
 * stochastic decisions only if the order is the reference's.  Pairs come out as (min, max). */
extern "C" int bd_nsearch_pairs_ordered(int32_t box_kind, const double period[3], int64_t particle_count, double dcut,
                                        const double *xyz, int64_t n, uint32_t *pairs_out, int64_t cap,
                                        int64_t *count_out, int32_t device)
{
    if (bd_nsearch_pairs(box_kind, period, particle_count, dcut, xyz, n, pairs_out, cap, count_out, device)) return 1;
    int64_t const m = *count_out;
    if (m > cap || m == 0) return 0;  // the caller retries with a larger buffer
    std::vector<uint32_t> bucket((size_t) n);
    if (bd_nsearch_buckets(box_kind, period, particle_count, dcut, xyz, n, bucket.data(), device)) return 1;
    struct Key { uint32_t centre, other, outer, inner; };
    std::vector<Key> keys((size_t) m);
    for (int64_t k = 0; k < m; k++) {
        uint32_t const p = pairs_out[2 * k], q = pairs_out[2 * k + 1];
        uint32_t const lo = p < q ? p : q, hi = p < q ? q : p;
        uint32_t const bl = bucket[lo], bh = bucket[hi];
        Key key;
        if (bl == bh) {
            key = {bl, bl, hi, lo};
        } else {
            bool const lo_is_centre = box_kind == BD_BOX_OPEN ? bl < bh : bl > bh;
            key = lo_is_centre ? Key{bl, bh, hi, lo} : Key{bh, bl, lo, hi};
        }
        keys[(size_t) k] = key;
    }
    std::sort(keys.begin(), keys.end(), [](const Key &a, const Key &b) {
        if (a.centre != b.centre) return a.centre < b.centre;
        if (a.other != b.other) return a.other < b.other;
        if (a.outer != b.outer) return a.outer < b.outer;
        return a.inner < b.inner;
    });
    for (int64_t k = 0; k < m; k++) {
        Key const &key = keys[(size_t) k];
        pairs_out[2 * k] = key.outer < key.inner ? key.outer : key.inner;
        pairs_out[2 * k + 1] = key.outer < key.inner ? key.inner : key.outer;
    }
    return 0;
}

// ---- md::neighbor_searcher<Box>::query, batched -------------------------------------------------------
//
// The reference answers one point at a time from its bucket grid (MD/misc/neighbor_searcher.hpp:65-70,104-131:
// every stored point with shortest_displacement(point, query).squared_norm() < dcut^2, strict).  Its one caller,
// the hetero-RDF tool (4-sim-ab/box/src/rdf_analysis_hetero/distance_histogram.cc:52), asks for every particle in
// turn; here the whole batch is one call: a warp per query scans the points 32 at a time (coalesced) and a
// ballot keeps the hits of a query in ascending index order.  Two passes (count, fill) around one scan.

struct QueryBox { int32_t periodic; double period[3], inv_period[3]; };

template <bool FILL>
__global__ void k_query(QueryBox box, double dcut2, const double *xyz, uint32_t n, const double *queries, uint32_t m,
                        uint32_t *count, const unsigned long long *start, uint32_t *out, unsigned long long cap)
{
    uint32_t const q = (blockIdx.x * blockDim.x + threadIdx.x) / 32, lane = threadIdx.x & 31;
    if (q >= m) return;
    double const qx = queries[3 * (size_t) q], qy = queries[3 * (size_t) q + 1], qz = queries[3 * (size_t) q + 2];
    unsigned long long at = FILL ? start[q] : 0;
    uint32_t found = 0;
    for (uint32_t base = 0; base < n; base += 32) {
        uint32_t const i = base + lane;
        bool hit = false;
        if (i < n) {
            double dx = __dsub_rn(xyz[3 * (size_t) i], qx), dy = __dsub_rn(xyz[3 * (size_t) i + 1], qy),
                   dz = __dsub_rn(xyz[3 * (size_t) i + 2], qz);
            if (box.periodic) {
                dx = round_mod_exact(dx, box.period[0], box.inv_period[0]);
                dy = round_mod_exact(dy, box.period[1], box.inv_period[1]);
                dz = round_mod_exact(dz, box.period[2], box.inv_period[2]);
            }
            hit = sqnorm_exact(dx, dy, dz) < dcut2;
        }
        uint32_t const mask = __ballot_sync(0xffffffffu, hit);
        if (FILL && hit) {
            unsigned long long const slot = at + found + __popc(mask & ((1u << lane) - 1u));
            if (slot < cap) out[slot] = i;
        }
        found += __popc(mask);
    }
    if (!FILL && lane == 0) count[q] = found;
}

extern "C" int bd_nsearch_query(int32_t box_kind, const double period[3], double dcut, const double *xyz, int64_t n,
                                const double *queries, int64_t m, int64_t *start_out, uint32_t *index_out, int64_t cap,
                                int32_t device)
{
    if (n < 0 || m < 0 || !(dcut > 0) || !start_out) return bd_internal_fail("bd_nsearch_query: bad arguments");
    if (box_kind != BD_BOX_OPEN && box_kind != BD_BOX_PERIODIC) return bd_internal_fail("bd_nsearch_query: unknown box kind");
    if (n >= (1ll << 32) || m >= (1ll << 27)) return bd_internal_fail("bd_nsearch_query: too many points");
    for (int64_t q = 0; q <= m; q++) start_out[q] = 0;
    if (m == 0 || n == 0) return 0;
    if (cudaSetDevice(device) != cudaSuccess) return bd_internal_fail("bd_nsearch_query: no such CUDA device");
    QueryBox box;
    std::memset(&box, 0, sizeof box);
    box.periodic = box_kind == BD_BOX_PERIODIC;
    for (int d = 0; d < 3; d++) {
        box.period[d] = box.periodic ? period[d] : 1.0;
        box.inv_period[d] = 1 / box.period[d];
    }
    double *d_xyz = nullptr, *d_q = nullptr;
    uint32_t *d_count = nullptr, *d_out = nullptr;
    unsigned long long *d_start = nullptr;
    std::vector<uint32_t> counts((size_t) m);
    std::vector<unsigned long long> starts((size_t) m + 1, 0ull);
    unsigned const blocks = (unsigned) ((m * 32 + 255) / 256);
    cudaError_t e = cudaMalloc((void **) &d_xyz, 3 * (size_t) n * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc((void **) &d_q, 3 * (size_t) m * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc((void **) &d_count, (size_t) m * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc((void **) &d_start, ((size_t) m + 1) * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemcpy(d_xyz, xyz, 3 * (size_t) n * sizeof(double), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_q, queries, 3 * (size_t) m * sizeof(double), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        k_query<false><<<blocks, 256>>>(box, dcut * dcut, d_xyz, (uint32_t) n, d_q, (uint32_t) m, d_count, nullptr, nullptr, 0);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(counts.data(), d_count, (size_t) m * sizeof(uint32_t), cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) {
        for (int64_t q = 0; q < m; q++) starts[(size_t) q + 1] = starts[(size_t) q] + counts[(size_t) q];
        for (int64_t q = 0; q <= m; q++) start_out[q] = (int64_t) starts[(size_t) q];
        unsigned long long const total = starts[(size_t) m];
        unsigned long long const take = std::min<unsigned long long>(total, (unsigned long long) std::max<int64_t>(cap, 0));
        if (take && index_out) {
            e = cudaMalloc((void **) &d_out, (size_t) take * sizeof(uint32_t));
            if (e == cudaSuccess) e = cudaMemcpy(d_start, starts.data(), ((size_t) m + 1) * sizeof(unsigned long long), cudaMemcpyHostToDevice);
            if (e == cudaSuccess) {
                k_query<true><<<blocks, 256>>>(box, dcut * dcut, d_xyz, (uint32_t) n, d_q, (uint32_t) m, d_count, d_start, d_out, take);
                e = cudaGetLastError();
            }
            if (e == cudaSuccess) e = cudaMemcpy(index_out, d_out, (size_t) take * sizeof(uint32_t), cudaMemcpyDeviceToHost);
        }
    }
    cudaFree(d_xyz);
    cudaFree(d_q);
    cudaFree(d_count);
    cudaFree(d_start);
    cudaFree(d_out);
    if (e != cudaSuccess) return bd_internal_fail((std::string("bd_nsearch_query: ") + cudaGetErrorString(e)).c_str());
    return 0;
}
