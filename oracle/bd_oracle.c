/*
 * bd_oracle.c -- TEST INFRASTRUCTURE ONLY (see bd_oracle.h).
 *
 * CPU restatement of the reference's Brownian-dynamics hot path in plain C.  Every function
 * cites the reference file:line it follows.  Arithmetic is fp64 and keeps the reference's
 * operation order; build with -ffp-contract=off so no multiply-add is fused (the reference
 * binaries are built for baseline x86-64 without FMA: 5-sim-genome/Makefile:7-12).
 *
 * Path shorthand:  MD/ = submodules/github.com/snsinfu/micromd/include/md/
 */
#include "bd_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * small helpers
 * ---------------------------------------------------------------------------------------- */

static char g_error[512];

static int fail(const char *msg)
{
    snprintf(g_error, sizeof g_error, "%s", msg);
    return 1;
}

const char *orc_last_error(void)
{
    return g_error;
}

/* md::vector / md::point (MD/basic_types/point.hpp:22-26,97-189) */
typedef struct {
    double x, y, z;
} v3;

static v3 v3_make(double x, double y, double z)
{
    v3 v = {x, y, z};
    return v;
}
static v3 v3_add(v3 a, v3 b) { return v3_make(a.x + b.x, a.y + b.y, a.z + b.z); }
static v3 v3_sub(v3 a, v3 b) { return v3_make(a.x - b.x, a.y - b.y, a.z - b.z); }
static v3 v3_scale(double s, v3 a) { return v3_make(a.x * s, a.y * s, a.z * s); }
static v3 v3_hadamard(v3 a, v3 b) { return v3_make(a.x * b.x, a.y * b.y, a.z * b.z); }
/* point.hpp:120-127: x*x + y*y + z*z, left to right */
static double v3_dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static double v3_sqnorm(v3 a) { return v3_dot(a, a); }
static double v3_norm(v3 a) { return sqrt(v3_sqnorm(a)); }
/* point.hpp:113-116,161-164: division multiplies by the reciprocal */
static v3 v3_normalize(v3 a) { return v3_scale(1 / v3_norm(a), a); }
/* point.hpp:166-169: axis * (dot(axis) / axis.squared_norm()) */
static v3 v3_project(v3 a, v3 axis) { return v3_scale(v3_dot(a, axis) / v3_sqnorm(axis), axis); }

/* ------------------------------------------------------------------------------------------
 * RNG: sfc64 + 128-layer ziggurat normal
 * ---------------------------------------------------------------------------------------- */

/* MD/basic_types/sfc.hpp:70-76 (state init), :331-346 (step), :392 (sfc64 = <u64,24,11,3,12,18>) */
typedef struct {
    uint64_t a, b, c, counter;
} sfc64;

static uint64_t sfc64_next(sfc64 *s)
{
    uint64_t const out = s->a + s->b + s->counter++;
    s->a = s->b ^ (s->b >> 11);
    s->b = s->c + (s->c << 3);
    s->c = ((s->c << 24) | (s->c >> 40)) + out;
    return out;
}

static void sfc64_seed(sfc64 *s, uint64_t seed)
{
    s->a = s->b = s->c = seed;
    s->counter = 1;
    for (int i = 0; i < 12; i++) {
        sfc64_next(s); /* seed64(): discard(initial_round_1 = 12), sfc.hpp:316-320 */
    }
}

static const double zig_edges[0x81] = {
#include "ziggurat_edges.inc"
};

/* libstdc++ std::generate_canonical<double,53>(sfc64): one 64-bit draw, converted to double
 * (round to nearest), divided by 2^64, clamped below 1.  The reference reaches it through
 * std::uniform_real_distribution in the two rare ziggurat branches (ziggurat.hpp:296,314). */
static double canonical(sfc64 *rng)
{
    double r = (double) sfc64_next(rng) / 18446744073709551616.0;
    if (r >= 1.0) {
        r = nextafter(1.0, 0.0);
    }
    return r;
}

static double uniform_real(sfc64 *rng, double a, double b)
{
    return (b - a) * canonical(rng) + a; /* libstdc++ uniform_real_distribution::operator() */
}

/* ziggurat.hpp:262-290 sample(); :292-306 sample_from_tail(); :308-318 check_accept() */
static double zig_normal(sfc64 *rng)
{
    for (;;) {
        /* bit_count = log2(max - min) truncated = 63 for a 64-bit engine (ziggurat.hpp:40-43,267):
           generate_bits<63> masks the top bit off, canonicalize<63,double> keeps 53 bits */
        uint64_t const bits = sfc64_next(rng) & 0x7FFFFFFFFFFFFFFFull;
        double const uniform = (1.0 / 9007199254740992.0) * (double) (bits >> 10);
        unsigned const layer = (unsigned) (bits & 0x7F);
        double const sign = (bits & 0x80) ? 1.0 : -1.0;
        double const lower_edge = zig_edges[layer];
        double const upper_edge = zig_edges[layer + 1];
        double const x = uniform * lower_edge;

        if (x < upper_edge) {
            return sign * x;
        }
        if (layer == 0) {
            double const tail_edge = zig_edges[1];
            double tx, ty;
            do {
                tx = -log(uniform_real(rng, 0.0, 1.0)) / tail_edge;
                ty = -log(uniform_real(rng, 0.0, 1.0));
            } while (2 * ty < tx * tx);
            return sign * (tail_edge + tx);
        }
        {
            double const lo = exp(-0.5 * lower_edge * lower_edge);
            double const hi = exp(-0.5 * upper_edge * upper_edge);
            if (uniform_real(rng, lo, hi) < exp(-0.5 * x * x)) {
                return sign * x;
            }
        }
    }
}

/* ziggurat.hpp:207-211: mean + stddev * sample with mean 0, stddev 1 */
static double normal_draw(sfc64 *rng)
{
    return 0.0 + 1.0 * zig_normal(rng);
}

int orc_sfc64(uint64_t seed, int64_t count, uint64_t *out)
{
    sfc64 rng;
    sfc64_seed(&rng, seed);
    for (int64_t i = 0; i < count; i++) {
        out[i] = sfc64_next(&rng);
    }
    return 0;
}

int orc_normals(uint64_t seed, int64_t count, double *out)
{
    sfc64 rng;
    sfc64_seed(&rng, seed);
    for (int64_t i = 0; i < count; i++) {
        out[i] = normal_draw(&rng);
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * potentials
 * ---------------------------------------------------------------------------------------- */

/* MD/misc/math.hpp:18-33: power by repeated squaring, same multiplication order */
static double power_n(int n, double x)
{
    if (n == 0) {
        return 1;
    }
    double pw = x;
    for (int k = n - 1; k > 0; k /= 2) {
        if (k % 2) {
            pw *= x;
        }
        x *= x;
    }
    return pw;
}

/* MD/misc/math.hpp:37-44 */
static double power_sqrt_n(int n, double x)
{
    if (n % 2 == 0) {
        return power_n(n / 2, x);
    }
    return power_n(n, sqrt(x));
}

static int term_is_cut(const bd_term *t, v3 r)
{
    /* MD/potential/cutoff_potential.hpp:42-46: r.squared_norm() >= cutoff^2 */
    return t->cutoff != 0 && v3_sqnorm(r) >= t->cutoff * t->cutoff;
}

static double term_energy(const bd_term *t, v3 r)
{
    if (term_is_cut(t, r)) {
        return 0;
    }
    switch (t->kind) {
    case BD_TERM_SOFTCORE: {
        /* MD/potential/softcore_potential.hpp:35-47 */
        double const k2 = 1 / (t->length * t->length);
        double const u2 = k2 * v3_sqnorm(r);
        double const v = power_sqrt_n(t->p, u2);
        double const g = 1 - v;
        if (g < 0) {
            return 0;
        }
        return t->energy * power_n(t->q, g);
    }
    case BD_TERM_SOFTWELL: {
        /* MD/potential/softwell_potential.hpp:35-42 */
        double const k2 = 1 / (t->length * t->length);
        double const u2 = k2 * v3_sqnorm(r);
        double const up = power_sqrt_n(t->p, u2);
        double const v1 = 1 + up;
        return -t->energy / v1;
    }
    case BD_TERM_HARMONIC:
        /* MD/potential/harmonic_potential.hpp:24-27 */
        return 0.5 * t->energy * v3_sqnorm(r);
    case BD_TERM_SPRING: {
        /* MD/potential/spring_potential.hpp:26-30 */
        double const u = v3_norm(r) - t->length;
        return 0.5 * t->energy * u * u;
    }
    case BD_TERM_SEMISPRING: {
        /* MD/potential/semispring_potential.hpp:28-35 */
        double const u = v3_norm(r) - t->length;
        if (u <= 0) {
            return 0;
        }
        return 0.5 * t->energy * u * u;
    }
    default:
        return 0; /* constant_potential{0} */
    }
}

static v3 term_force(const bd_term *t, v3 r)
{
    v3 const zero = {0, 0, 0};
    if (term_is_cut(t, r)) {
        return zero;
    }
    switch (t->kind) {
    case BD_TERM_SOFTCORE: {
        /* MD/potential/softcore_potential.hpp:49-61 */
        double const k2 = 1 / (t->length * t->length);
        double const u2 = k2 * v3_sqnorm(r);
        double const v = power_sqrt_n(t->p - 2, u2);
        double const g = 1 - v * u2;
        if (g < 0) {
            return zero;
        }
        return v3_scale((double) (t->p * t->q) * t->energy * k2 * power_n(t->q - 1, g) * v, r);
    }
    case BD_TERM_SOFTWELL: {
        /* MD/potential/softwell_potential.hpp:44-52 */
        double const k2 = 1 / (t->length * t->length);
        double const u2 = k2 * v3_sqnorm(r);
        double const up_2 = power_sqrt_n(t->p - 2, u2);
        double const up = up_2 * u2;
        double const v1 = 1 + up;
        return v3_scale(((double) (-t->p) * t->energy * k2) * up_2 / (v1 * v1), r);
    }
    case BD_TERM_HARMONIC:
        /* MD/potential/harmonic_potential.hpp:29-32 */
        return v3_scale(-t->energy, r);
    case BD_TERM_SPRING: {
        /* MD/potential/spring_potential.hpp:32-39 */
        double const r1 = v3_norm(r);
        if (r1 == 0) {
            return zero;
        }
        return v3_scale(t->energy * t->length / r1 - t->energy, r);
    }
    case BD_TERM_SEMISPRING: {
        /* MD/potential/semispring_potential.hpp:37-44 */
        double const r1 = v3_norm(r);
        if (r1 <= t->length) {
            return zero;
        }
        return v3_scale(t->energy * t->length / r1 - t->energy, r);
    }
    default:
        return zero;
    }
}

/* sum_potential / scaled_potential: factor_0*p_0 + factor_1*p_1 + ...  evaluated left to right
 * (MD/potential/sum_potential.hpp:32-40, scaled_potential.hpp:32-40) */
static double pot_energy(const bd_term *terms, int nterms, v3 r)
{
    double e = terms[0].factor * term_energy(&terms[0], r);
    for (int k = 1; k < nterms; k++) {
        e = e + terms[k].factor * term_energy(&terms[k], r);
    }
    return e;
}

static v3 pot_force(const bd_term *terms, int nterms, v3 r)
{
    v3 f = v3_scale(terms[0].factor, term_force(&terms[0], r));
    for (int k = 1; k < nterms; k++) {
        f = v3_add(f, v3_scale(terms[k].factor, term_force(&terms[k], r)));
    }
    return f;
}

int orc_eval_potential(const bd_term *terms, int32_t nterms, const double r[3],
                       double *energy_out, double force_out[3])
{
    v3 const rv = {r[0], r[1], r[2]};
    v3 const f = pot_force(terms, nterms, rv);
    *energy_out = pot_energy(terms, nterms, rv);
    force_out[0] = f.x;
    force_out[1] = f.y;
    force_out[2] = f.z;
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * boxes
 * ---------------------------------------------------------------------------------------- */

typedef struct {
    int kind;              /* bd_box_kind */
    double period[3];      /* periodic_box::{x,y,z}_period */
    size_t particle_count; /* open_box::particle_count hint */
} box_t;

static box_t box_default(int kind)
{
    /* MD/misc/box.hpp:20 (particle_count = 1000), :35-41 (periods = 1) */
    box_t b;
    b.kind = kind;
    b.period[0] = b.period[1] = b.period[2] = 1;
    b.particle_count = 1000;
    return b;
}

/* MD/misc/nsearch_detail/math.hpp:34-38 -- multiply by the reciprocal, not divide */
static double round_mod(double x, double m)
{
    return x - nearbyint(x * (1 / m)) * m;
}

/* MD/misc/nsearch_detail/math.hpp:27-31 */
static double floor_mod(double x, double m)
{
    return x - floor(x * (1 / m)) * m;
}

/* MD/misc/box.hpp:25-28 (open), :45-51 (periodic) */
static v3 shortest_displacement(const box_t *box, v3 p1, v3 p2)
{
    if (box->kind == BD_BOX_PERIODIC) {
        return v3_make(round_mod(p1.x - p2.x, box->period[0]),
                       round_mod(p1.y - p2.y, box->period[1]),
                       round_mod(p1.z - p2.z, box->period[2]));
    }
    return v3_sub(p1, p2);
}

/* ------------------------------------------------------------------------------------------
 * neighbour searcher
 * ---------------------------------------------------------------------------------------- */

/* MD/misc/linear_hash.hpp:21-43 */
static uint32_t linear_hash_general(const uint32_t coeff[3], uint32_t modulus, uint32_t x,
                                    uint32_t y, uint32_t z)
{
    uint64_t sum = 0;
    sum += (uint64_t) coeff[0] * x;
    sum += (uint64_t) coeff[1] * y;
    sum += (uint64_t) coeff[2] * z;
    return (uint32_t) (sum % modulus);
}

static uint32_t linear_hash(uint32_t modulus, uint32_t x, uint32_t y, uint32_t z)
{
    static const uint32_t coeff[3] = {3929498747u, 1008281837u, 1832832077u};
    return linear_hash_general(coeff, modulus, x, y, z);
}

uint32_t orc_linear_hash(const uint32_t coeff[3], uint32_t modulus, uint32_t x, uint32_t y,
                         uint32_t z)
{
    return linear_hash_general(coeff, modulus, x, y, z);
}

/* MD/misc/nsearch_detail/math.hpp:40-48: static_cast<uint32_t>(static_cast<int32_t>(x)) */
static uint32_t trunc_uint(double x)
{
    return (uint32_t) (int32_t) x;
}

typedef struct {
    double step;
    uint32_t count;
} bin_layout;

/* MD/misc/nsearch_detail/math.hpp:68-76 */
static bin_layout define_bins(double span, double min_step)
{
    bin_layout bins;
    uint32_t const c = trunc_uint(span / min_step);
    bins.count = c > 1 ? c : 1;
    bins.step = span / (double) bins.count;
    return bins;
}

/* MD/misc/nsearch_detail/math.hpp:78-86 */
static uint32_t locate_bin(const bin_layout *bins, double coord)
{
    double const span = bins->step * (double) bins->count;
    double const image = floor_mod(coord, span);
    uint32_t const index = trunc_uint(image * (1 / bins->step));
    return index < bins->count ? index : bins->count - 1;
}

/* MD/misc/nsearch_detail/math.hpp:56-64 */
static uint32_t add_mod(uint32_t x, uint32_t y, uint32_t m)
{
    uint32_t sum = x + y;
    if (sum >= m) {
        sum -= m;
    }
    return sum;
}

static int cmp_u32(const void *a, const void *b)
{
    uint32_t const x = *(const uint32_t *) a, y = *(const uint32_t *) b;
    return (x > y) - (x < y);
}

/* search_grid.hpp:25-30 */
static size_t sort_unique(uint32_t *v, size_t n)
{
    if (n == 0) {
        return 0;
    }
    qsort(v, n, sizeof *v, cmp_u32);
    size_t m = 1;
    for (size_t i = 1; i < n; i++) {
        if (v[i] != v[m - 1]) {
            v[m++] = v[i];
        }
    }
    return m;
}

/* md::neighbor_searcher<Box> (MD/misc/neighbor_searcher.hpp:23-131) on top of
 * search_grid<open_box> (search_grid.hpp:218-295) or search_grid<periodic_box>
 * (search_grid.hpp:67-160).  Buckets are stored CSR instead of vector-of-vectors; the
 * iteration order is the reference's. */
typedef struct {
    box_t box;      /* box_ : the box distances are measured in */
    double dcut;    /* dcut_ == grid spacing */
    /* open: hash grid */
    uint32_t modulus;
    /* periodic: bin grid */
    bin_layout bins[3];
    size_t nbuckets;
    size_t *dir_start;  /* [nbuckets+1] */
    uint32_t *dir_nb;   /* directed_neighbors, sorted unique per bucket */
    /* members, filled by set_points */
    size_t npoints;
    size_t *mem_start;  /* [nbuckets+1] */
    uint32_t *mem_index;
    v3 *mem_point;
} searcher_t;

static void searcher_free(searcher_t *s)
{
    free(s->dir_start);
    free(s->dir_nb);
    free(s->mem_start);
    free(s->mem_index);
    free(s->mem_point);
    memset(s, 0, sizeof *s);
}

static size_t searcher_locate_bucket(const searcher_t *s, v3 pt)
{
    if (s->box.kind == BD_BOX_OPEN) {
        /* search_grid.hpp:231-244 */
        double const offset = (double) (1L << 20);
        double const freq = 1 / s->dcut;
        uint32_t const x = (uint32_t) (offset + freq * pt.x);
        uint32_t const y = (uint32_t) (offset + freq * pt.y);
        uint32_t const z = (uint32_t) (offset + freq * pt.z);
        return linear_hash(s->modulus, x, y, z);
    }
    /* search_grid.hpp:80-86,142-145 */
    uint32_t const x = locate_bin(&s->bins[0], pt.x);
    uint32_t const y = locate_bin(&s->bins[1], pt.y);
    uint32_t const z = locate_bin(&s->bins[2], pt.z);
    return x + (size_t) s->bins[0].count * (y + (size_t) s->bins[1].count * z);
}

static void searcher_init(searcher_t *s, box_t box, double dcut)
{
    memset(s, 0, sizeof *s);
    s->box = box;
    s->dcut = dcut;

    if (box.kind == BD_BOX_OPEN) {
        /* search_grid.hpp:247-252: modulus = uint32(particle_count * 2 / 11) | 1 */
        s->modulus = (uint32_t) (box.particle_count * 2 / 11);
        s->modulus |= 1;
        s->nbuckets = s->modulus;

        /* search_grid.hpp:254-275: the 27 hash deltas */
        uint32_t deltas[27];
        uint32_t const cd[3] = {s->modulus - 1, s->modulus, s->modulus + 1};
        size_t nd = 0;
        for (int a = 0; a < 3; a++) {
            for (int b = 0; b < 3; b++) {
                for (int c = 0; c < 3; c++) {
                    deltas[nd++] = linear_hash(s->modulus, cd[a], cd[b], cd[c]);
                }
            }
        }
        nd = sort_unique(deltas, nd);

        /* search_grid.hpp:277-292: neighbor >= center is "directed" */
        s->dir_start = malloc((s->nbuckets + 1) * sizeof *s->dir_start);
        s->dir_nb = malloc(s->nbuckets * 27 * sizeof *s->dir_nb);
        size_t fill = 0;
        for (uint32_t center = 0; center < s->modulus; center++) {
            s->dir_start[center] = fill;
            size_t k = 0;
            uint32_t *row = s->dir_nb + fill;
            for (size_t d = 0; d < nd; d++) {
                uint32_t const nb = (center + deltas[d]) % s->modulus;
                if (nb >= center) {
                    row[k++] = nb;
                }
            }
            fill += sort_unique(row, k);
        }
        s->dir_start[s->nbuckets] = fill;
    } else {
        /* search_grid.hpp:148-160 basic_binner<periodic_box> */
        for (int d = 0; d < 3; d++) {
            s->bins[d] = define_bins(box.period[d], dcut);
        }
        uint32_t const nx = s->bins[0].count, ny = s->bins[1].count, nz = s->bins[2].count;
        s->nbuckets = (size_t) nx * ny * nz;

        /* search_grid.hpp:97-140 init_buckets / init_bucket_adjacents: adj <= self is "directed" */
        s->dir_start = malloc((s->nbuckets + 1) * sizeof *s->dir_start);
        s->dir_nb = malloc(s->nbuckets * 27 * sizeof *s->dir_nb);
        size_t fill = 0;
        /* buckets are addressed x + nx*(y + ny*z); iterate in index order */
        for (uint32_t z = 0; z < nz; z++) {
            for (uint32_t y = 0; y < ny; y++) {
                for (uint32_t x = 0; x < nx; x++) {
                    size_t const self = x + (size_t) nx * (y + (size_t) ny * z);
                    uint32_t const dxs[3] = {0, 1, nx - 1};
                    uint32_t const dys[3] = {0, 1, ny - 1};
                    uint32_t const dzs[3] = {0, 1, nz - 1};
                    s->dir_start[self] = fill;
                    uint32_t *row = s->dir_nb + fill;
                    size_t k = 0;
                    for (int c = 0; c < 3; c++) {
                        for (int b = 0; b < 3; b++) {
                            for (int a = 0; a < 3; a++) {
                                uint32_t const ax = add_mod(x, dxs[a], nx);
                                uint32_t const ay = add_mod(y, dys[b], ny);
                                uint32_t const az = add_mod(z, dzs[c], nz);
                                size_t const adj = ax + (size_t) nx * (ay + (size_t) ny * az);
                                if (adj <= self) {
                                    row[k++] = (uint32_t) adj;
                                }
                            }
                        }
                    }
                    fill += sort_unique(row, k);
                }
            }
        }
        s->dir_start[s->nbuckets] = fill;
    }
    s->mem_start = calloc(s->nbuckets + 1, sizeof *s->mem_start);
}

/* neighbor_searcher.hpp:36-48 set_points: members are appended in index order */
static void searcher_set_points(searcher_t *s, const v3 *points, size_t n)
{
    free(s->mem_index);
    free(s->mem_point);
    s->npoints = n;
    s->mem_index = malloc((n ? n : 1) * sizeof *s->mem_index);
    s->mem_point = malloc((n ? n : 1) * sizeof *s->mem_point);
    size_t *bucket_of = malloc((n ? n : 1) * sizeof *bucket_of);
    memset(s->mem_start, 0, (s->nbuckets + 1) * sizeof *s->mem_start);
    for (size_t i = 0; i < n; i++) {
        bucket_of[i] = searcher_locate_bucket(s, points[i]);
        s->mem_start[bucket_of[i] + 1]++;
    }
    for (size_t b = 0; b < s->nbuckets; b++) {
        s->mem_start[b + 1] += s->mem_start[b];
    }
    size_t *cursor = malloc((s->nbuckets ? s->nbuckets : 1) * sizeof *cursor);
    memcpy(cursor, s->mem_start, s->nbuckets * sizeof *cursor);
    for (size_t i = 0; i < n; i++) {
        size_t const at = cursor[bucket_of[i]]++;
        s->mem_index[at] = (uint32_t) i;
        s->mem_point[at] = points[i];
    }
    free(cursor);
    free(bucket_of);
}

typedef struct {
    uint32_t *data; /* [cap][2] */
    size_t count, cap;
    int growable;
} pair_sink;

static void sink_push(pair_sink *out, uint32_t i, uint32_t j)
{
    if (out->count >= out->cap) {
        if (!out->growable) {
            out->count++;
            return;
        }
        out->cap = out->cap ? out->cap * 2 : 1024;
        out->data = realloc(out->data, out->cap * 2 * sizeof *out->data);
    }
    out->data[2 * out->count] = i;
    out->data[2 * out->count + 1] = j;
    out->count++;
}

/* neighbor_searcher.hpp:54-61 search + :73-100 search_among.
 * `map` (may be NULL) is neighbor_list's index_mapper (neighbor_list.hpp:199-227). */
static void searcher_search(const searcher_t *s, const uint32_t *map, pair_sink *out)
{
    double const dcut2 = s->dcut * s->dcut;
    for (size_t a = 0; a < s->nbuckets; a++) {
        for (size_t d = s->dir_start[a]; d < s->dir_start[a + 1]; d++) {
            size_t const b = s->dir_nb[d];
            for (size_t mj = s->mem_start[b]; mj < s->mem_start[b + 1]; mj++) {
                for (size_t mi = s->mem_start[a]; mi < s->mem_start[a + 1]; mi++) {
                    uint32_t const ii = s->mem_index[mi], jj = s->mem_index[mj];
                    if (ii == jj) {
                        break; /* sic: break, not continue (neighbor_searcher.hpp:83-86) */
                    }
                    v3 const r = shortest_displacement(&s->box, s->mem_point[mi], s->mem_point[mj]);
                    if (v3_sqnorm(r) > dcut2) {
                        continue;
                    }
                    uint32_t const lo = ii < jj ? ii : jj, hi = ii < jj ? jj : ii;
                    if (map) {
                        sink_push(out, map[lo], map[hi]);
                    } else {
                        sink_push(out, lo, hi);
                    }
                }
            }
        }
    }
}

int orc_nsearch_buckets(int32_t box_kind, const double period[3], int64_t particle_count,
                        double spacing, const double *xyz, int64_t n, uint32_t *bucket_out,
                        int32_t device)
{
    (void) device;
    box_t box = box_default(box_kind);
    if (box_kind == BD_BOX_PERIODIC) {
        memcpy(box.period, period, sizeof box.period);
    } else {
        box.particle_count = (size_t) particle_count;
    }
    searcher_t s;
    searcher_init(&s, box, spacing);
    for (int64_t i = 0; i < n; i++) {
        bucket_out[i] = (uint32_t) searcher_locate_bucket(
            &s, v3_make(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
    }
    searcher_free(&s);
    return 0;
}

int orc_nsearch_pairs(int32_t box_kind, const double period[3], int64_t particle_count,
                      double dcut, const double *xyz, int64_t n, uint32_t *pairs_out, int64_t cap,
                      int64_t *count_out, int32_t device)
{
    (void) device;
    box_t box = box_default(box_kind);
    if (box_kind == BD_BOX_PERIODIC) {
        memcpy(box.period, period, sizeof box.period);
    } else {
        box.particle_count = (size_t) particle_count;
    }
    searcher_t s;
    searcher_init(&s, box, dcut);
    searcher_set_points(&s, (const v3 *) xyz, (size_t) n);
    pair_sink sink = {pairs_out, 0, (size_t) cap, 0};
    searcher_search(&s, NULL, &sink);
    *count_out = (int64_t) sink.count;
    searcher_free(&s);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Verlet neighbour list  (MD/forcefield/detail/neighbor_list.hpp:63-239)
 * ---------------------------------------------------------------------------------------- */

typedef struct {
    int box_kind;
    box_t prev_box;
    double prev_verlet_radius;  /* = 1 */
    double prev_dcut;           /* = 1 */
    searcher_t searcher;
    int searcher_ready;
    v3 *prev_points;
    size_t prev_count;
    pair_sink pairs;
    uint32_t *targets;
    size_t ntargets;
    int64_t rebuilds;
} nlist_t;

static void nlist_init(nlist_t *nl, int box_kind)
{
    memset(nl, 0, sizeof *nl);
    nl->box_kind = box_kind;
    nl->prev_box = box_default(box_kind);
    nl->prev_verlet_radius = 1;
    nl->prev_dcut = 1;
    nl->pairs.growable = 1;
    /* neighbor_list.hpp:72-75: searcher_{prev_box_, prev_verlet_radius_}; built lazily here,
       which is equivalent because a searcher is a pure function of (box, radius). */
}

static void nlist_free(nlist_t *nl)
{
    if (nl->searcher_ready) {
        searcher_free(&nl->searcher);
    }
    free(nl->prev_points);
    free(nl->pairs.data);
    free(nl->targets);
    memset(nl, 0, sizeof *nl);
}

/* neighbor_list.hpp:31-35 */
static int approx_scalar(double x, double y)
{
    return fabs(x - y) < 1e-6 * fabs(y);
}

/* neighbor_list.hpp:37-49 */
static int approx_box(const box_t *a, const box_t *b)
{
    if (a->kind == BD_BOX_OPEN) {
        return a->particle_count == b->particle_count;
    }
    return approx_scalar(a->period[0], b->period[0]) && approx_scalar(a->period[1], b->period[1])
        && approx_scalar(a->period[2], b->period[2]);
}

/* neighbor_list.hpp:101-159 */
static int nlist_check_consistency(const nlist_t *nl, const v3 *points, size_t n, double dcut,
                                   const box_t *box)
{
    if (nl->prev_count == 0) {
        return 0;
    }
    if (!approx_box(box, &nl->prev_box) || !approx_scalar(dcut, nl->prev_dcut)) {
        return 0;
    }
    if (nl->ntargets == 0) {
        if (n != nl->prev_count) {
            return 0;
        }
    } else if (nl->targets[nl->ntargets - 1] >= n) {
        return 0;
    }
    double const threshold = (nl->prev_verlet_radius - dcut) / 2;
    if (threshold <= 0) {
        return 0;
    }
    for (size_t i = 0; i < nl->prev_count; i++) {
        v3 const cur = nl->ntargets ? points[nl->targets[i]] : points[i];
        v3 const disp = shortest_displacement(box, cur, nl->prev_points[i]);
        if (v3_sqnorm(disp) > threshold * threshold) {
            return 0;
        }
    }
    return 1;
}

/* neighbor_list.hpp:162-230 */
static void nlist_rebuild(nlist_t *nl, const v3 *points, size_t n, double dcut, box_t box)
{
    size_t const m = nl->ntargets ? nl->ntargets : n;
    nl->prev_points = realloc(nl->prev_points, (m ? m : 1) * sizeof *nl->prev_points);
    for (size_t i = 0; i < m; i++) {
        nl->prev_points[i] = nl->ntargets ? points[nl->targets[i]] : points[i];
    }
    nl->prev_count = m;

    /* neighbor_list_heuristics.hpp:87-92 set_box_hints<open_box> */
    if (box.kind == BD_BOX_OPEN) {
        box.particle_count = m;
    }

    double const verlet_radius = 1.5 * dcut;
    int const box_changed = !approx_box(&box, &nl->prev_box);
    int const verlet_changed = !approx_scalar(verlet_radius, nl->prev_verlet_radius);
    if (box_changed || verlet_changed || !nl->searcher_ready) {
        if (nl->searcher_ready) {
            searcher_free(&nl->searcher);
        }
        if (box_changed || verlet_changed) {
            searcher_init(&nl->searcher, box, verlet_radius);
        } else {
            /* the default-constructed searcher_{prev_box_, prev_verlet_radius_} is kept */
            searcher_init(&nl->searcher, nl->prev_box, nl->prev_verlet_radius);
        }
        nl->searcher_ready = 1;
    }

    nl->prev_box = box;
    nl->prev_verlet_radius = verlet_radius;
    nl->prev_dcut = dcut;

    nl->pairs.count = 0;
    searcher_set_points(&nl->searcher, nl->prev_points, m);
    searcher_search(&nl->searcher, nl->ntargets ? nl->targets : NULL, &nl->pairs);
    nl->rebuilds++;
}

/* neighbor_list.hpp:80-85 */
static void nlist_update(nlist_t *nl, const v3 *points, size_t n, double dcut, box_t box)
{
    if (!nlist_check_consistency(nl, points, n, dcut, &box)) {
        nlist_rebuild(nl, points, n, dcut, box);
    }
}

/* ------------------------------------------------------------------------------------------
 * the system and its force fields
 * ---------------------------------------------------------------------------------------- */

typedef struct {
    int used, enabled, nterms;
    bd_term *table; /* [T*T*nterms] */
    double dcut;
    nlist_t nlist;
} pair_ff;

typedef struct {
    int used, enabled, nterms;
    int min_image; /* glue_forcefield.cpp:24,40: displacement through the periodic box */
    bd_term *table; /* [T*T*nterms] */
    uint32_t *pairs;
    size_t m;
} bond_ff;

typedef struct {
    int used, enabled;
    uint32_t *triples;
    double *energy;
    size_t m;
} bend_ff;

typedef struct {
    int used, enabled, shape, direction, nterms;
    bd_term *table; /* [T*nterms] */
    v3 center;
    double semi[3];
    double stats[4]; /* reaction_force, axial_reaction.{x,y,z} */
} wall_ff;

typedef struct {
    int used, enabled, nterms;
    bd_term *table; /* [T*nterms] */
    v3 source;
    uint32_t *targets;
    size_t ntargets;
} point_ff;

struct orc_system {
    size_t n;
    int ntypes;
    uint8_t *type;
    double mobility[BD_MAX_TYPES];
    v3 *pos;
    box_t box;
    pair_ff pair[BD_MAX_PAIR_FF];
    bond_ff bond[BD_MAX_BOND_FF];
    bend_ff bend[BD_MAX_BEND_FF];
    wall_ff wall[BD_MAX_WALL_FF];
    point_ff point[BD_MAX_POINT_FF];
    /* brownian_simulator state (MD/simulation/detail/brownian_simulator.hpp:79-84) */
    v3 *forces, *weiners;
    v3 *external; /* host-evaluated user forcefields (set_external_forces), or NULL */
    sfc64 rng;
    double last_dt;
    bd_counters counters;
};

int orc_create(int64_t n_beads, int32_t n_replicas, int32_t device, orc_system **out)
{
    (void) device;
    if (n_replicas != 1) {
        return fail("oracle: n_replicas must be 1");
    }
    if (n_beads < 0) {
        return fail("oracle: negative bead count");
    }
    orc_system *s = calloc(1, sizeof *s);
    s->n = (size_t) n_beads;
    s->ntypes = 1;
    s->type = calloc(s->n ? s->n : 1, 1);
    s->mobility[0] = 1; /* mobility_attribute default, MD/system.hpp:41-44 */
    s->pos = calloc(s->n ? s->n : 1, sizeof *s->pos);
    s->forces = calloc(s->n ? s->n : 1, sizeof *s->forces);
    s->weiners = calloc(s->n ? s->n : 1, sizeof *s->weiners);
    s->box = box_default(BD_BOX_OPEN);
    *out = s;
    return 0;
}

int orc_destroy(orc_system *s)
{
    if (!s) {
        return 0;
    }
    for (int k = 0; k < BD_MAX_PAIR_FF; k++) {
        free(s->pair[k].table);
        if (s->pair[k].used) {
            nlist_free(&s->pair[k].nlist);
        }
    }
    for (int k = 0; k < BD_MAX_BOND_FF; k++) {
        free(s->bond[k].table);
        free(s->bond[k].pairs);
    }
    for (int k = 0; k < BD_MAX_BEND_FF; k++) {
        free(s->bend[k].triples);
        free(s->bend[k].energy);
    }
    for (int k = 0; k < BD_MAX_WALL_FF; k++) {
        free(s->wall[k].table);
    }
    for (int k = 0; k < BD_MAX_POINT_FF; k++) {
        free(s->point[k].table);
        free(s->point[k].targets);
    }
    free(s->type);
    free(s->pos);
    free(s->forces);
    free(s->weiners);
    free(s->external);
    free(s);
    return 0;
}

int orc_set_box(orc_system *s, int32_t box_kind, const double period[3])
{
    s->box = box_default(box_kind);
    if (box_kind == BD_BOX_PERIODIC) {
        memcpy(s->box.period, period, sizeof s->box.period);
    }
    for (int k = 0; k < BD_MAX_PAIR_FF; k++) {
        if (s->pair[k].used && s->pair[k].nlist.box_kind != box_kind) {
            return fail("oracle: set the box kind before adding pair forcefields");
        }
    }
    return 0;
}

int orc_set_types(orc_system *s, const uint8_t *type, int32_t ntypes, const double *mobility)
{
    if (ntypes < 1 || ntypes > BD_MAX_TYPES) {
        return fail("oracle: ntypes out of range");
    }
    for (size_t i = 0; i < s->n; i++) {
        if (type[i] >= ntypes) {
            return fail("oracle: type id out of range");
        }
    }
    s->ntypes = ntypes;
    memcpy(s->type, type, s->n);
    memcpy(s->mobility, mobility, (size_t) ntypes * sizeof(double));
    return 0;
}

int orc_set_positions(orc_system *s, int32_t replica, const double *xyz)
{
    (void) replica;
    memcpy(s->pos, xyz, s->n * sizeof *s->pos);
    return 0;
}

int orc_get_positions(orc_system *s, int32_t replica, double *xyz)
{
    (void) replica;
    memcpy(xyz, s->pos, s->n * sizeof *s->pos);
    return 0;
}

static bd_term *copy_terms(const bd_term *table, size_t count)
{
    bd_term *t = malloc(count * sizeof *t);
    memcpy(t, table, count * sizeof *t);
    return t;
}

static int mask_to_indices(const uint8_t *mask, size_t n, uint32_t **out, size_t *count)
{
    *out = NULL;
    *count = 0;
    if (!mask) {
        return 0;
    }
    size_t m = 0;
    for (size_t i = 0; i < n; i++) {
        m += mask[i] != 0;
    }
    *out = malloc((m ? m : 1) * sizeof **out);
    for (size_t i = 0; i < n; i++) {
        if (mask[i]) {
            (*out)[(*count)++] = (uint32_t) i;
        }
    }
    return 0;
}

int orc_add_pair_ff(orc_system *s, int32_t nterms, const bd_term *table, double dcut,
                    const uint8_t *targets, int32_t *ff_out)
{
    if (nterms < 1 || nterms > BD_MAX_TERMS) {
        return fail("oracle: nterms out of range");
    }
    for (int k = 0; k < BD_MAX_PAIR_FF; k++) {
        if (!s->pair[k].used) {
            pair_ff *ff = &s->pair[k];
            ff->used = ff->enabled = 1;
            ff->nterms = nterms;
            ff->table = copy_terms(table, (size_t) s->ntypes * s->ntypes * nterms);
            ff->dcut = dcut;
            nlist_init(&ff->nlist, s->box.kind);
            mask_to_indices(targets, s->n, &ff->nlist.targets, &ff->nlist.ntargets);
            *ff_out = k;
            return 0;
        }
    }
    return fail("oracle: too many pair forcefields");
}

int orc_update_pair_ff(orc_system *s, int32_t ff, const bd_term *table, double dcut)
{
    if (ff < 0 || ff >= BD_MAX_PAIR_FF || !s->pair[ff].used) {
        return fail("oracle: bad pair forcefield id");
    }
    if (table) {
        memcpy(s->pair[ff].table, table,
               (size_t) s->ntypes * s->ntypes * s->pair[ff].nterms * sizeof *table);
    }
    s->pair[ff].dcut = dcut;
    return 0;
}

int orc_add_bond_ff(orc_system *s, int32_t nterms, const bd_term *table, const uint32_t *pairs,
                    int64_t m, int32_t *ff_out)
{
    if (nterms < 1 || nterms > BD_MAX_TERMS) {
        return fail("oracle: nterms out of range");
    }
    for (int64_t b = 0; b < 2 * m; b++) {
        if (pairs[b] >= s->n) {
            return fail("oracle: bond index out of range");
        }
    }
    for (int k = 0; k < BD_MAX_BOND_FF; k++) {
        if (!s->bond[k].used) {
            bond_ff *ff = &s->bond[k];
            ff->used = ff->enabled = 1;
            ff->nterms = nterms;
            ff->table = copy_terms(table, (size_t) s->ntypes * s->ntypes * nterms);
            ff->m = (size_t) m;
            ff->pairs = malloc((m ? m : 1) * 2 * sizeof *ff->pairs);
            memcpy(ff->pairs, pairs, (size_t) m * 2 * sizeof *ff->pairs);
            *ff_out = k;
            return 0;
        }
    }
    return fail("oracle: too many bond forcefields");
}

int orc_set_bond_pairs(orc_system *s, int32_t ff, const uint32_t *pairs, int64_t m)
{
    if (ff < 0 || ff >= BD_MAX_BOND_FF || !s->bond[ff].used || m < 0) {
        return fail("oracle: bad bond forcefield id or pair count");
    }
    for (int64_t b = 0; b < 2 * m; b++) {
        if (pairs[b] >= s->n) {
            return fail("oracle: bond index out of range");
        }
    }
    free(s->bond[ff].pairs);
    s->bond[ff].m = (size_t) m;
    s->bond[ff].pairs = malloc((m ? m : 1) * 2 * sizeof *s->bond[ff].pairs);
    memcpy(s->bond[ff].pairs, pairs, (size_t) m * 2 * sizeof *s->bond[ff].pairs);
    return 0;
}

int orc_set_bond_min_image(orc_system *s, int32_t ff, int32_t on)
{
    if (ff < 0 || ff >= BD_MAX_BOND_FF || !s->bond[ff].used) {
        return fail("oracle: bad bond forcefield id");
    }
    s->bond[ff].min_image = on != 0;
    return 0;
}

int orc_update_bond_ff(orc_system *s, int32_t ff, const bd_term *table)
{
    if (ff < 0 || ff >= BD_MAX_BOND_FF || !s->bond[ff].used) {
        return fail("oracle: bad bond forcefield id");
    }
    memcpy(s->bond[ff].table, table,
           (size_t) s->ntypes * s->ntypes * s->bond[ff].nterms * sizeof *table);
    return 0;
}

int orc_add_bend_ff(orc_system *s, const uint32_t *triples, const double *bending_energy,
                    int64_t m, int32_t *ff_out)
{
    for (int64_t b = 0; b < 3 * m; b++) {
        if (triples[b] >= s->n) {
            return fail("oracle: triple index out of range");
        }
    }
    for (int k = 0; k < BD_MAX_BEND_FF; k++) {
        if (!s->bend[k].used) {
            bend_ff *ff = &s->bend[k];
            ff->used = ff->enabled = 1;
            ff->m = (size_t) m;
            ff->triples = malloc((m ? m : 1) * 3 * sizeof *ff->triples);
            ff->energy = malloc((m ? m : 1) * sizeof *ff->energy);
            memcpy(ff->triples, triples, (size_t) m * 3 * sizeof *ff->triples);
            memcpy(ff->energy, bending_energy, (size_t) m * sizeof *ff->energy);
            *ff_out = k;
            return 0;
        }
    }
    return fail("oracle: too many bending forcefields");
}

int orc_add_wall_ff(orc_system *s, int32_t shape, int32_t direction, int32_t nterms,
                    const bd_term *table, const double center[3], const double semiaxes[3],
                    int32_t *ff_out)
{
    if (nterms < 1 || nterms > BD_MAX_TERMS) {
        return fail("oracle: nterms out of range");
    }
    for (int k = 0; k < BD_MAX_WALL_FF; k++) {
        if (!s->wall[k].used) {
            wall_ff *ff = &s->wall[k];
            ff->used = ff->enabled = 1;
            ff->shape = shape;
            ff->direction = direction;
            ff->nterms = nterms;
            ff->table = copy_terms(table, (size_t) s->ntypes * nterms);
            ff->center = v3_make(center[0], center[1], center[2]);
            memcpy(ff->semi, semiaxes, sizeof ff->semi);
            *ff_out = k;
            return 0;
        }
    }
    return fail("oracle: too many wall forcefields");
}

int orc_update_wall_ff(orc_system *s, int32_t ff, const bd_term *table, const double center[3],
                       const double semiaxes[3])
{
    if (ff < 0 || ff >= BD_MAX_WALL_FF || !s->wall[ff].used) {
        return fail("oracle: bad wall forcefield id");
    }
    if (table) {
        memcpy(s->wall[ff].table, table, (size_t) s->ntypes * s->wall[ff].nterms * sizeof *table);
    }
    s->wall[ff].center = v3_make(center[0], center[1], center[2]);
    memcpy(s->wall[ff].semi, semiaxes, sizeof s->wall[ff].semi);
    return 0;
}

int orc_get_wall_stats(orc_system *s, int32_t ff, int32_t replica, double out[4])
{
    (void) replica;
    if (ff < 0 || ff >= BD_MAX_WALL_FF || !s->wall[ff].used) {
        return fail("oracle: bad wall forcefield id");
    }
    memcpy(out, s->wall[ff].stats, sizeof s->wall[ff].stats);
    return 0;
}

int orc_add_point_ff(orc_system *s, int32_t nterms, const bd_term *table, const double source[3],
                     const uint8_t *targets, int32_t *ff_out)
{
    if (nterms < 1 || nterms > BD_MAX_TERMS) {
        return fail("oracle: nterms out of range");
    }
    for (int k = 0; k < BD_MAX_POINT_FF; k++) {
        if (!s->point[k].used) {
            point_ff *ff = &s->point[k];
            ff->used = ff->enabled = 1;
            ff->nterms = nterms;
            ff->table = copy_terms(table, (size_t) s->ntypes * nterms);
            ff->source = v3_make(source[0], source[1], source[2]);
            mask_to_indices(targets, s->n, &ff->targets, &ff->ntargets);
            *ff_out = k;
            return 0;
        }
    }
    return fail("oracle: too many point-source forcefields");
}

int orc_set_ff_enabled(orc_system *s, int32_t ff_class, int32_t ff, int32_t enabled)
{
    switch (ff_class) {
    case BD_FF_PAIR:
        if (ff >= 0 && ff < BD_MAX_PAIR_FF && s->pair[ff].used) {
            s->pair[ff].enabled = enabled;
            return 0;
        }
        break;
    case BD_FF_BOND:
        if (ff >= 0 && ff < BD_MAX_BOND_FF && s->bond[ff].used) {
            s->bond[ff].enabled = enabled;
            return 0;
        }
        break;
    case BD_FF_BEND:
        if (ff >= 0 && ff < BD_MAX_BEND_FF && s->bend[ff].used) {
            s->bend[ff].enabled = enabled;
            return 0;
        }
        break;
    case BD_FF_WALL:
        if (ff >= 0 && ff < BD_MAX_WALL_FF && s->wall[ff].used) {
            s->wall[ff].enabled = enabled;
            return 0;
        }
        break;
    case BD_FF_POINT:
        if (ff >= 0 && ff < BD_MAX_POINT_FF && s->point[ff].used) {
            s->point[ff].enabled = enabled;
            return 0;
        }
        break;
    default:
        break;
    }
    return fail("oracle: bad forcefield class / id");
}

/* ---- per-forcefield kernels ------------------------------------------------------------- */

static const bd_term *pair_row(const orc_system *s, const bd_term *table, int nterms, size_t i,
                               size_t j)
{
    return table + ((size_t) s->type[i] * s->ntypes + s->type[j]) * nterms;
}

/* MD/forcefield/neighbor_pairwise_forcefield.hpp:48-84 */
static void pair_ff_eval(orc_system *s, pair_ff *ff, v3 *forces, double *energy)
{
    nlist_update(&ff->nlist, s->pos, s->n, ff->dcut, s->box);
    s->counters.list_rebuilds = ff->nlist.rebuilds;
    for (size_t k = 0; k < ff->nlist.pairs.count; k++) {
        size_t const i = ff->nlist.pairs.data[2 * k], j = ff->nlist.pairs.data[2 * k + 1];
        const bd_term *row = pair_row(s, ff->table, ff->nterms, i, j);
        v3 const r = shortest_displacement(&s->box, s->pos[i], s->pos[j]);
        if (forces) {
            v3 const f = pot_force(row, ff->nterms, r);
            forces[i] = v3_add(forces[i], f);
            forces[j] = v3_sub(forces[j], f);
        }
        if (energy) {
            *energy += pot_energy(row, ff->nterms, r);
        }
    }
}

/* MD/forcefield/bonded_pairwise_forcefield.hpp:56-93 -- no minimum image -- or, with min_image,
 * 3-sim-1kb/src/simulation/forces/glue_forcefield.cpp:17-45 */
static void bond_ff_eval(orc_system *s, bond_ff *ff, v3 *forces, double *energy)
{
    for (size_t k = 0; k < ff->m; k++) {
        size_t const i = ff->pairs[2 * k], j = ff->pairs[2 * k + 1];
        const bd_term *row = pair_row(s, ff->table, ff->nterms, i, j);
        v3 const r = ff->min_image ? shortest_displacement(&s->box, s->pos[i], s->pos[j])
                                   : v3_sub(s->pos[i], s->pos[j]);
        if (forces) {
            v3 const f = pot_force(row, ff->nterms, r);
            forces[i] = v3_add(forces[i], f);
            forces[j] = v3_sub(forces[j], f);
        }
        if (energy) {
            *energy += pot_energy(row, ff->nterms, r);
        }
    }
}

/* MD/forcefield/bonded_triplewise_forcefield.hpp:56-103 +
 * MD/potential/cosine_bending_potential.hpp:27-60 */
static void bend_ff_eval(orc_system *s, bend_ff *ff, v3 *forces, double *energy)
{
    for (size_t t = 0; t < ff->m; t++) {
        size_t const i = ff->triples[3 * t], j = ff->triples[3 * t + 1], k = ff->triples[3 * t + 2];
        double const e = ff->energy[t];
        v3 const rij = v3_sub(s->pos[i], s->pos[j]);
        v3 const rjk = v3_sub(s->pos[j], s->pos[k]);
        double const dij_sq = v3_sqnorm(rij);
        double const djk_sq = v3_sqnorm(rjk);
        if (energy) {
            if (dij_sq * djk_sq != 0) {
                double const dot_ijk = v3_dot(rij, rjk);
                double const dij_djk = sqrt(dij_sq * djk_sq);
                double const cos_ijk = dot_ijk / dij_djk;
                *energy += e - e * cos_ijk;
            } else {
                *energy += 0;
            }
        }
        if (forces) {
            double const dij_djk = sqrt(dij_sq * djk_sq);
            if (dij_djk == 0) {
                continue;
            }
            double const dot_ijk = v3_dot(rij, rjk);
            double const e_div_dd = e / dij_djk;
            v3 const fij = v3_scale(e_div_dd, v3_sub(rjk, v3_scale(dot_ijk / dij_sq, rij)));
            v3 const fjk = v3_scale(e_div_dd, v3_sub(rij, v3_scale(dot_ijk / djk_sq, rjk)));
            forces[i] = v3_add(forces[i], fij);
            forces[j] = v3_add(forces[j], v3_sub(fjk, fij));
            forces[k] = v3_add(forces[k], v3_make(-fjk.x, -fjk.y, -fjk.z));
        }
    }
}

/* MD/forcefield/sphere_surface_forcefield.hpp:73-143.  An inward forcefield acts on beads
 * inside the sphere only (the other side is constant_potential{0}, :145-153), and vice versa. */
static void sphere_ff_eval(orc_system *s, wall_ff *ff, v3 *forces, double *energy)
{
    double const radius = ff->semi[0];
    double const radius2 = radius * radius;
    if (forces) {
        memset(ff->stats, 0, sizeof ff->stats);
    }
    for (size_t i = 0; i < s->n; i++) {
        v3 const r = v3_sub(s->pos[i], ff->center);
        double const r2 = v3_sqnorm(r);
        double const r1 = sqrt(r2);
        if (r2 == 0) {
            continue;
        }
        double const scale = radius / r1;
        v3 const sv = v3_sub(r, v3_scale(scale, r));
        int const inside = r2 < radius2;
        int const active = inside ? ff->direction == BD_WALL_INWARD : ff->direction == BD_WALL_OUTWARD;
        const bd_term *row = ff->table + (size_t) s->type[i] * ff->nterms;
        if (energy) {
            *energy += active ? pot_energy(row, ff->nterms, sv) : 0.0;
        }
        if (forces) {
            v3 const force = active ? pot_force(row, ff->nterms, sv) : v3_make(0, 0, 0);
            v3 const aniso = v3_scale(scale, v3_sub(v3_project(force, r), force));
            forces[i] = v3_add(forces[i], v3_add(force, aniso));
            ff->stats[0] -= v3_dot(force, r) / v3_norm(r);
        }
    }
}

/* MD/forcefield/ellipsoid_surface_forcefield.hpp:54-99 (evaluate_point), :137-211 */
static void ellipsoid_ff_eval(orc_system *s, wall_ff *ff, v3 *forces, double *energy)
{
    v3 const quadform = {1 / (ff->semi[0] * ff->semi[0]), 1 / (ff->semi[1] * ff->semi[1]),
                         1 / (ff->semi[2] * ff->semi[2])};
    if (forces) {
        memset(ff->stats, 0, sizeof ff->stats);
    }
    for (size_t i = 0; i < s->n; i++) {
        v3 const radial = v3_sub(s->pos[i], ff->center);
        v3 const dual = v3_hadamard(quadform, radial);
        if (v3_sqnorm(dual) == 0) {
            continue; /* ev.undefined */
        }
        double const implicit = v3_dot(dual, radial) - 1;
        double const scale = implicit / (2 * v3_sqnorm(dual));
        v3 const delta = v3_scale(scale, dual);
        v3 const strain = v3_scale(scale, quadform);
        int const inside = implicit < 0;
        int const active = inside ? ff->direction == BD_WALL_INWARD : ff->direction == BD_WALL_OUTWARD;
        const bd_term *row = ff->table + (size_t) s->type[i] * ff->nterms;
        if (energy) {
            *energy += active ? pot_energy(row, ff->nterms, delta) : 0.0;
        }
        if (forces) {
            v3 const basic = active ? pot_force(row, ff->nterms, delta) : v3_make(0, 0, 0);
            v3 const iso = v3_project(basic, delta);
            v3 const aniso = v3_sub(basic, iso);
            v3 const strain_force = v3_hadamard(v3_sub(aniso, iso), strain);
            v3 const force = v3_add(iso, strain_force);
            forces[i] = v3_add(forces[i], force);
            v3 const normal = v3_scale(inside ? -1.0 : 1.0, v3_normalize(delta));
            ff->stats[0] -= v3_dot(force, normal);
            ff->stats[1] -= (normal.x > 0 ? force.x : -force.x);
            ff->stats[2] -= (normal.y > 0 ? force.y : -force.y);
            ff->stats[3] -= (normal.z > 0 ? force.z : -force.z);
        }
    }
}

/* MD/forcefield/point_source_forcefield.hpp:53-103 */
static void point_ff_eval(orc_system *s, point_ff *ff, v3 *forces, double *energy)
{
    size_t const m = ff->ntargets ? ff->ntargets : s->n;
    for (size_t t = 0; t < m; t++) {
        size_t const i = ff->ntargets ? ff->targets[t] : t;
        const bd_term *row = ff->table + (size_t) s->type[i] * ff->nterms;
        v3 const r = v3_sub(s->pos[i], ff->source);
        if (energy) {
            *energy += pot_energy(row, ff->nterms, r);
        }
        if (forces) {
            forces[i] = v3_add(forces[i], pot_force(row, ff->nterms, r));
        }
    }
}

/* sum_forcefield (MD/system/detail/sum_forcefield.hpp:48-64): each component adds into the
 * same array, in registration order.  The canonical order here is pair, bond, bend, wall,
 * point; the reference harness registers in the same order. */
static int eval_all(orc_system *s, int ff_class, int ff, v3 *forces, double *energy)
{
    int hit = 0;
    for (int k = 0; k < BD_MAX_PAIR_FF; k++) {
        pair_ff *p = &s->pair[k];
        int const want = ff_class == BD_FF_ALL ? p->enabled : (ff_class == BD_FF_PAIR && ff == k);
        if (p->used && want) {
            double e = 0;
            pair_ff_eval(s, p, forces, energy ? &e : NULL);
            if (energy) {
                *energy += e;
            }
            hit = 1;
        }
    }
    for (int k = 0; k < BD_MAX_BOND_FF; k++) {
        bond_ff *p = &s->bond[k];
        int const want = ff_class == BD_FF_ALL ? p->enabled : (ff_class == BD_FF_BOND && ff == k);
        if (p->used && want) {
            double e = 0;
            bond_ff_eval(s, p, forces, energy ? &e : NULL);
            if (energy) {
                *energy += e;
            }
            hit = 1;
        }
    }
    for (int k = 0; k < BD_MAX_BEND_FF; k++) {
        bend_ff *p = &s->bend[k];
        int const want = ff_class == BD_FF_ALL ? p->enabled : (ff_class == BD_FF_BEND && ff == k);
        if (p->used && want) {
            double e = 0;
            bend_ff_eval(s, p, forces, energy ? &e : NULL);
            if (energy) {
                *energy += e;
            }
            hit = 1;
        }
    }
    for (int k = 0; k < BD_MAX_WALL_FF; k++) {
        wall_ff *p = &s->wall[k];
        int const want = ff_class == BD_FF_ALL ? p->enabled : (ff_class == BD_FF_WALL && ff == k);
        if (p->used && want) {
            double e = 0;
            if (p->shape == BD_WALL_SPHERE) {
                sphere_ff_eval(s, p, forces, energy ? &e : NULL);
            } else {
                ellipsoid_ff_eval(s, p, forces, energy ? &e : NULL);
            }
            if (energy) {
                *energy += e;
            }
            hit = 1;
        }
    }
    for (int k = 0; k < BD_MAX_POINT_FF; k++) {
        point_ff *p = &s->point[k];
        int const want = ff_class == BD_FF_ALL ? p->enabled : (ff_class == BD_FF_POINT && ff == k);
        if (p->used && want) {
            double e = 0;
            point_ff_eval(s, p, forces, energy ? &e : NULL);
            if (energy) {
                *energy += e;
            }
            hit = 1;
        }
    }
    if (ff_class != BD_FF_ALL && !hit) {
        return fail("oracle: bad forcefield class / id");
    }
    /* a user forcefield is one more member of the sum forcefield */
    if (ff_class == BD_FF_ALL && forces && s->external) {
        for (size_t i = 0; i < s->n; i++) {
            forces[i] = v3_add(forces[i], s->external[i]);
        }
    }
    return 0;
}

int orc_set_external_forces(orc_system *s, int32_t replica, const double *xyz)
{
    (void) replica;
    if (!xyz) {
        free(s->external);
        s->external = NULL;
        return 0;
    }
    if (!s->external) {
        s->external = malloc((s->n ? s->n : 1) * sizeof(v3));
    }
    memcpy(s->external, xyz, s->n * sizeof(v3));
    return 0;
}

/* MD/system.hpp:264-272: zero-fill, then the sum forcefield adds */
int orc_compute_forces(orc_system *s, int32_t replica, int32_t ff_class, int32_t ff,
                       double *out_xyz)
{
    (void) replica;
    v3 *forces = (v3 *) out_xyz;
    for (size_t i = 0; i < s->n; i++) {
        forces[i] = v3_make(0, 0, 0);
    }
    return eval_all(s, ff_class, ff, forces, NULL);
}

/* MD/system.hpp:249-252 */
int orc_compute_energy(orc_system *s, int32_t replica, int32_t ff_class, int32_t ff, double *out)
{
    (void) replica;
    double e = 0;
    int const rc = eval_all(s, ff_class, ff, NULL, &e);
    *out = e;
    return rc;
}

/* MD/simulation/detail/brownian_timestepper.hpp:72-91 */
static double solve_brownian_timestep(double distance, double mobility, double force2,
                                      double temperature)
{
    double const epsilon = 1e-6;
    double const random_walk_factor = 2.55;
    double const a = mobility * mobility * force2;
    double const b = random_walk_factor * mobility * temperature;
    double const c = distance * distance;
    if (a * c < epsilon * b * b) {
        return c / b;
    }
    return (-b + sqrt(b * b + 4 * a * c)) / (2 * a);
}

/* MD/simulation/detail/brownian_timestepper.hpp:39-55 (fixed), :109-136 (adaptive) */
static double determine_timestep(const orc_system *s, double max_timestep, double spacestep,
                                 double temperature)
{
    if (spacestep == 0) {
        return max_timestep;
    }
    double timestep = max_timestep;
    for (size_t i = 0; i < s->n; i++) {
        double dt = max_timestep;
        double const mobility = s->mobility[s->type[i]];
        double const force2 = v3_sqnorm(s->forces[i]);
        if (mobility != 0 || force2 != 0) {
            dt = solve_brownian_timestep(spacestep, mobility, force2, temperature);
        }
        timestep = timestep < dt ? timestep : dt;
    }
    return timestep;
}

/* MD/simulation/brownian_dynamics.hpp:49-73 + MD/simulation/detail/brownian_simulator.hpp:27-76 */
int orc_step(orc_system *s, const bd_step_config *cfg)
{
    if (cfg->first_step == 1) {
        /* brownian_simulator ctor: seed the engine, zero the previous noise */
        sfc64_seed(&s->rng, cfg->seed);
        for (size_t i = 0; i < s->n; i++) {
            s->weiners[i] = v3_make(0, 0, 0);
        }
    }
    for (int64_t step = 0; step < cfg->steps; step++) {
        for (size_t i = 0; i < s->n; i++) {
            s->forces[i] = v3_make(0, 0, 0);
        }
        int const rc = eval_all(s, BD_FF_ALL, -1, s->forces, NULL);
        if (rc) {
            return rc;
        }
        double const timestep = determine_timestep(s, cfg->timestep, cfg->spacestep, cfg->temperature);
        s->last_dt = timestep;
        const double *z = cfg->replay_normals ? cfg->replay_normals + (size_t) step * s->n * 3 : NULL;
        for (size_t i = 0; i < s->n; i++) {
            double const mu_dt = timestep * s->mobility[s->type[i]];
            double const sigma = sqrt(2 * cfg->temperature * mu_dt);
            v3 weiner;
            if (z) {
                weiner = v3_make(sigma * z[3 * i], sigma * z[3 * i + 1], sigma * z[3 * i + 2]);
            } else {
                weiner.x = sigma * normal_draw(&s->rng);
                weiner.y = sigma * normal_draw(&s->rng);
                weiner.z = sigma * normal_draw(&s->rng);
            }
            s->pos[i] = v3_add(s->pos[i], v3_scale(mu_dt, s->forces[i]));
            s->pos[i] = v3_add(s->pos[i], v3_scale(0.5, v3_add(weiner, s->weiners[i])));
            s->weiners[i] = weiner;
        }
        s->counters.steps_done++;
    }
    return 0;
}

int orc_get_last_timestep(orc_system *s, int32_t replica, double *out)
{
    (void) replica;
    *out = s->last_dt;
    return 0;
}

int orc_get_neighbor_list(orc_system *s, int32_t replica, uint32_t *pairs_out, int64_t cap,
                          int64_t *count_out)
{
    (void) replica;
    if (!s->pair[0].used) {
        return fail("oracle: no pair forcefield");
    }
    nlist_update(&s->pair[0].nlist, s->pos, s->n, s->pair[0].dcut, s->box);
    size_t const count = s->pair[0].nlist.pairs.count;
    size_t const ncopy = count < (size_t) cap ? count : (size_t) cap;
    memcpy(pairs_out, s->pair[0].nlist.pairs.data, ncopy * 2 * sizeof *pairs_out);
    *count_out = (int64_t) count;
    return 0;
}

int orc_get_counters(orc_system *s, bd_counters *out)
{
    *out = s->counters;
    if (s->pair[0].used) {
        out->list_rebuilds = s->pair[0].nlist.rebuilds;
        out->listed_pairs = (int64_t) s->pair[0].nlist.pairs.count;
    }
    return 0;
}

/* simulation_store.cc:403-407: float const scale = 1 << bits; nearbyint(float(val) * scale) / scale */
int orc_quantize_positions(const double *xyz, int64_t count, int32_t fraction_bits, float *out)
{
    for (int64_t k = 0; k < 3 * count; k++) {
        float const v = (float) xyz[k];
        if (fraction_bits < 0) {
            out[k] = v;
        } else {
            float const scale = (float) (1 << fraction_bits);
            out[k] = nearbyintf(v * scale) / scale;
        }
    }
    return 0;
}

/* refinement.py:11-21 + run.py:27-40.  The not-a-knot cubic spline from its B-spline-free definition:
 * a dense solve of the (m x m) system in the second derivatives by Gaussian elimination with partial
 * pivoting -- deliberately a different algorithm from the engine's closed-form elimination. */
int orc_refine_spline(const double *coarse_xyz, const int64_t *coarse_ranges, int64_t n_chains, int32_t factor,
                      const int64_t *fine_ranges, const int64_t *attach, int64_t n_attach, int64_t n_coarse,
                      int64_t n_fine, double *fine_xyz_out)
{
    (void) n_coarse;
    (void) n_fine;
    for (int64_t c = 0; c < n_chains; c++) {
        int64_t const beg = coarse_ranges[2 * c], m = coarse_ranges[2 * c + 1] - beg;
        int64_t const fbeg = fine_ranges[2 * c], fcount = fine_ranges[2 * c + 1] - fbeg;
        if (m < 4) {
            return fail("oracle: a coarse chain needs at least 4 points");
        }
        double const h = 1.0 / (double) m;
        double *A = malloc((size_t) m * m * sizeof *A);
        double *M = malloc((size_t) m * 3 * sizeof *M);
        for (int d = 0; d < 3; d++) {
            memset(A, 0, (size_t) m * m * sizeof *A);
            double *rhs = M + (size_t) d * m;
            /* row 0 and row m-1: continuity of the third derivative at u_1 and u_{m-2} */
            A[0] = 1; A[1] = -2; A[2] = 1; rhs[0] = 0;
            A[(m - 1) * m + m - 3] = 1; A[(m - 1) * m + m - 2] = -2; A[(m - 1) * m + m - 1] = 1; rhs[m - 1] = 0;
            for (int64_t i = 1; i < m - 1; i++) {
                A[i * m + i - 1] = 1; A[i * m + i] = 4; A[i * m + i + 1] = 1;
                const double *y = coarse_xyz + 3 * (beg + i) + d;
                rhs[i] = 6.0 * (y[-3] - 2.0 * y[0] + y[3]) / (h * h);
            }
            for (int64_t col = 0; col < m; col++) {
                int64_t piv = col;
                for (int64_t r = col + 1; r < m; r++) {
                    if (fabs(A[r * m + col]) > fabs(A[piv * m + col])) piv = r;
                }
                if (piv != col) {
                    for (int64_t k = 0; k < m; k++) { double t = A[col * m + k]; A[col * m + k] = A[piv * m + k]; A[piv * m + k] = t; }
                    double t = rhs[col]; rhs[col] = rhs[piv]; rhs[piv] = t;
                }
                for (int64_t r = col + 1; r < m; r++) {
                    double const f = A[r * m + col] / A[col * m + col];
                    if (f == 0.0) continue;
                    for (int64_t k = col; k < m; k++) A[r * m + k] -= f * A[col * m + k];
                    rhs[r] -= f * rhs[col];
                }
            }
            for (int64_t r = m - 1; r >= 0; r--) {
                double acc = rhs[r];
                for (int64_t k = r + 1; k < m; k++) acc -= A[r * m + k] * rhs[k];
                rhs[r] = acc / A[r * m + r];
            }
        }
        int64_t const n = m * factor;
        for (int64_t f = 0; f < fcount && f < n; f++) {
            double const u = ((double) f + 0.5) / (double) n;
            int64_t i = (int64_t) floor((u - 0.5 * h) / h);
            i = i < 0 ? 0 : (i > m - 2 ? m - 2 : i);
            double const t = u - ((double) i + 0.5) * h;
            for (int d = 0; d < 3; d++) {
                double const y0 = coarse_xyz[3 * (beg + i) + d], y1 = coarse_xyz[3 * (beg + i + 1) + d];
                double const m0 = M[(size_t) d * m + i], m1 = M[(size_t) d * m + i + 1];
                double const b = (y1 - y0) / h - h * (2.0 * m0 + m1) / 6.0;
                fine_xyz_out[3 * (fbeg + f) + d] = y0 + b * t + 0.5 * m0 * t * t + (m1 - m0) / (6.0 * h) * t * t * t;
            }
        }
        free(A);
        free(M);
    }
    for (int64_t a = 0; a < n_attach; a++) {
        for (int d = 0; d < 3; d++) {
            fine_xyz_out[3 * attach[2 * a + 1] + d] = fine_xyz_out[3 * attach[2 * a] + d];
        }
    }
    return 0;
}
