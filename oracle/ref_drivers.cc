/*
 * ref_drivers.cc -- TEST INFRASTRUCTURE ONLY: the timed CPU baseline as the paper's drivers run it.
 *
 * oracle/ref_harness.cc puts micromd behind the generic C entry points: its pair potentials are evaluated
 * through a run-time switch over flattened terms.  The drivers do not pay that: their lambdas return
 * expression templates (a * softcore<2,3> + b * softcore<8,3>) which the compiler inlines into the pair
 * loop.  This file registers the force fields of the 100kb interphase model the way
 * 5-sim-genome/src/simulation_interphase/simulation_driver_forcefield.cc:19-235 does -- the same factory
 * calls with the same kind of lambdas over a per-particle (a, b) attribute -- on the UNMODIFIED micromd, and
 * runs md::simulate_brownian_dynamics with the driver's per-step wall feedback
 * (simulation_driver_interphase.cc:70-80) as the callback.  Nothing here is used for parity, only for
 * timing (bench.py cpu_baseline / --impl reference).  Built by oracle/Makefile into oracle/_ref/ twice:
 * -O2 (the paper's optimisation level) and -O3 -march=x86-64-v3.
 */
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include <md.hpp>

namespace
{
    struct ab_factors { md::scalar a = 0, b = 0; };
    constexpr md::attribute_key<ab_factors, struct ab_factors_tag> ab_attribute = {};

    struct c4_params {
        double a_diameter, b_diameter, a_repulsion, b_repulsion;
        double a_bond_spring, a_bond_length, b_bond_spring, b_bond_length;
        double a_loop_spring, b_loop_spring;
        double nuc_bond_spring, nuc_bond_length;
        double wall_packing_spring, wall_a_factor, wall_b_factor;
        double wall_mobility, wall_spring;
        double semiaxes[3];
        double bead_scale, bond_scale;
    };

    struct c4_system {
        md::system system;
        c4_params p;
        md::vector semiaxes;
        std::function<md::vector()> packing_reaction;
    };
}

extern "C" {

/* types: per-bead (a, b) factors; chains: [nchains][2] = (start, end); nuc: [n_nuc][2] = (nor, side bead) */
void *drv_c4_create(int64_t n, const double *ab, const int64_t *chains, int64_t nchains, const int64_t *nuc,
                    int64_t n_nuc, const double *xyz, const double *params, int64_t nparams)
{
    if (nparams != (int64_t) (sizeof(c4_params) / sizeof(double))) return nullptr;
    auto *h = new c4_system;
    std::memcpy(&h->p, params, sizeof h->p);
    c4_params const &p = h->p;
    md::system &system = h->system;
    system.require(ab_attribute);
    for (int64_t i = 0; i < n; i++) {
        auto part = system.add_particle();
        part.mobility = 1;
        part.position = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
        part.view(ab_attribute) = {ab[2 * i], ab[2 * i + 1]};
    }
    h->semiaxes = {p.semiaxes[0], p.semiaxes[1], p.semiaxes[2]};
    md::system *sys = &system;
    c4_system *ctx = h;

    // general A/B repulsion: the potential is an expression template, inlined into micromd's pair loop
    system.add_forcefield(
        md::make_neighbor_pairwise_forcefield([=](md::index i, md::index j) {
            auto const data = sys->view(ab_attribute);
            auto const a = 0.5 * (data[i].a + data[j].a);
            auto const b = 0.5 * (data[i].b + data[j].b);
            md::softcore_potential<2, 3> const pa{ctx->p.a_repulsion, ctx->p.a_diameter * ctx->p.bead_scale};
            md::softcore_potential<8, 3> const pb{ctx->p.b_repulsion, ctx->p.b_diameter * ctx->p.bead_scale};
            return a * pa + b * pb;
        }).set_neighbor_distance([=] { return std::max(ctx->p.a_diameter, ctx->p.b_diameter) * ctx->p.bead_scale; }));

    auto chain_bonds = system.add_forcefield(md::make_bonded_pairwise_forcefield([=](md::index i, md::index j) {
        auto const data = sys->view(ab_attribute);
        auto const a = 0.5 * (data[i].a + data[j].a);
        auto const b = 0.5 * (data[i].b + data[j].b);
        auto const K = a * ctx->p.a_bond_spring + b * ctx->p.b_bond_spring;
        auto const l = a * ctx->p.a_bond_length + b * ctx->p.b_bond_length;
        auto const s = ctx->p.bond_scale;
        return md::semispring_potential{K * (1 / (s * s)), l * s};
    }));
    auto loop_bonds = system.add_forcefield(md::make_bonded_pairwise_forcefield([=](md::index i, md::index j) {
        auto const data = sys->view(ab_attribute);
        auto const a = 0.5 * (data[i].a + data[j].a);
        auto const b = 0.5 * (data[i].b + data[j].b);
        auto const s = ctx->p.bond_scale;
        return md::harmonic_potential{(a * ctx->p.a_loop_spring + b * ctx->p.b_loop_spring) * (1 / (s * s))};
    }));
    for (int64_t c = 0; c < nchains; c++) {
        chain_bonds->add_bonded_range((md::index) chains[2 * c], (md::index) chains[2 * c + 1]);
        for (int64_t i = chains[2 * c]; i + 2 < chains[2 * c + 1]; i++) loop_bonds->add_bonded_pair((md::index) i, (md::index) i + 2);
    }
    auto nuc_bonds = system.add_forcefield(md::make_bonded_pairwise_forcefield([=](md::index, md::index) {
        auto const s = ctx->p.bond_scale;
        return md::semispring_potential{ctx->p.nuc_bond_spring * (1 / (s * s)), ctx->p.nuc_bond_length * s};
    }));
    for (int64_t k = 0; k < n_nuc; k++) nuc_bonds->add_bonded_pair((md::index) nuc[2 * k], (md::index) nuc[2 * k + 1]);

    auto ellipsoid = [=] {
        return md::ellipsoid{.semiaxis_x = ctx->semiaxes.x, .semiaxis_y = ctx->semiaxes.y, .semiaxis_z = ctx->semiaxes.z};
    };
    auto inward = system.add_forcefield(md::make_ellipsoid_inward_forcefield([=](md::index i) {
        auto const data = sys->view(ab_attribute);
        auto const a = 0.5 * (data[i].a + ctx->p.wall_a_factor);
        auto const b = 0.5 * (data[i].b + ctx->p.wall_b_factor);
        md::softcore_potential<2, 3> const pa{ctx->p.a_repulsion, ctx->p.a_diameter / 2 * ctx->p.bead_scale};
        md::softcore_potential<8, 3> const pb{ctx->p.b_repulsion, ctx->p.b_diameter / 2 * ctx->p.bead_scale};
        return a * pa + b * pb;
    }).set_ellipsoid(ellipsoid));
    auto outward = system.add_forcefield(
        md::make_ellipsoid_outward_forcefield(md::harmonic_potential{ctx->p.wall_packing_spring}).set_ellipsoid(ellipsoid));
    h->packing_reaction = [=] { return inward->stats.axial_reaction + outward->stats.axial_reaction; };
    return h;
}

/* md::simulate_brownian_dynamics for `steps` steps; feedback != 0 runs the driver's wall update after every step */
int drv_c4_run(void *handle, int64_t steps, double temperature, double timestep, uint64_t seed, int32_t feedback)
{
    auto *h = static_cast<c4_system *>(handle);
    md::brownian_dynamics_config config;
    config.temperature = temperature;
    config.timestep = timestep;
    config.steps = steps;
    config.seed = seed;
    if (feedback) {
        config.callback = [=](md::step) {
            md::vector net;
            net += h->packing_reaction();
            net -= md::vector{h->p.wall_spring, h->p.wall_spring, h->p.wall_spring}.hadamard(h->semiaxes);
            h->semiaxes += timestep * h->p.wall_mobility * net;
        };
    }
    md::simulate_brownian_dynamics(h->system, config);
    return 0;
}

int drv_c4_get(void *handle, double *xyz, double semiaxes[3])
{
    auto *h = static_cast<c4_system *>(handle);
    auto const pos = h->system.view_positions();
    for (md::index i = 0; i < pos.size(); i++) {
        xyz[3 * i] = pos[i].x; xyz[3 * i + 1] = pos[i].y; xyz[3 * i + 2] = pos[i].z;
    }
    semiaxes[0] = h->semiaxes.x; semiaxes[1] = h->semiaxes.y; semiaxes[2] = h->semiaxes.z;
    return 0;
}

void drv_c4_destroy(void *handle)
{
    delete static_cast<c4_system *>(handle);
}

}  // extern "C"
