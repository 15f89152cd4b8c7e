// ref_harness.cc -- TEST INFRASTRUCTURE ONLY.
//
// Thin extern "C" wrapper that drives the UNMODIFIED reference (micromd, header-only) through the
// same entry points as the engine (include/bd_api_decl.inc) under the ref_ prefix.  It is
// compiled by oracle/Makefile against the headers where they lie under /root/reference and the
// output goes to oracle/_ref/libmicromd_ref.so (git-ignored; travels to the GPU box as a binary).
// No reference source is copied: everything below only CALLS md:: classes.
//
// The only non-reference arithmetic is `table_potential`, which plays the role of the drivers'
// potential lambdas: it looks the bead classes up in a table and returns
//     factor_0 * pot_0 + factor_1 * pot_1 + ...
// built from the reference's own potential classes and evaluated left to right, exactly like
// md::sum_potential / md::scaled_potential (checked by ref_selfcheck_expression_templates()).

#include <array>
#include <cmath>
#include <cstdint>
#include <functional>
#include <map>
#include <cstring>
#include <memory>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include <md.hpp>

#include "../include/bd_types.h"

namespace
{
    thread_local std::string g_error;

    int fail(std::string const& msg)
    {
        g_error = msg;
        return 1;
    }

    // ---- one flattened term, evaluated by the reference's potential classes -------------

    template<int P, int Q>
    md::vector softcore_force(bd_term const& t, md::vector r)
    {
        md::softcore_potential<P, Q> pot;
        pot.energy = t.energy;
        pot.diameter = t.length;
        return pot.evaluate_force(r);
    }

    template<int P, int Q>
    md::scalar softcore_energy(bd_term const& t, md::vector r)
    {
        md::softcore_potential<P, Q> pot;
        pot.energy = t.energy;
        pot.diameter = t.length;
        return pot.evaluate_energy(r);
    }

    template<int P>
    md::vector softwell_force(bd_term const& t, md::vector r)
    {
        md::softwell_potential<P> pot;
        pot.energy = t.energy;
        pot.decay_distance = t.length;
        return pot.evaluate_force(r);
    }

    template<int P>
    md::scalar softwell_energy(bd_term const& t, md::vector r)
    {
        md::softwell_potential<P> pot;
        pot.energy = t.energy;
        pot.decay_distance = t.length;
        return pot.evaluate_energy(r);
    }

#define FOR_Q(P, FN)                                                                               \
    switch (t.q) {                                                                                 \
    case 1: return FN<P, 1>(t, r);                                                                 \
    case 2: return FN<P, 2>(t, r);                                                                 \
    case 3: return FN<P, 3>(t, r);                                                                 \
    case 4: return FN<P, 4>(t, r);                                                                 \
    case 5: return FN<P, 5>(t, r);                                                                 \
    default: throw std::runtime_error("ref: unsupported softcore Q");                              \
    }

#define FOR_PQ(FN)                                                                                 \
    switch (t.p) {                                                                                 \
    case 2: FOR_Q(2, FN)                                                                           \
    case 3: FOR_Q(3, FN)                                                                           \
    case 4: FOR_Q(4, FN)                                                                           \
    case 5: FOR_Q(5, FN)                                                                           \
    case 6: FOR_Q(6, FN)                                                                           \
    case 8: FOR_Q(8, FN)                                                                           \
    case 12: FOR_Q(12, FN)                                                                         \
    default: throw std::runtime_error("ref: unsupported softcore P");                              \
    }

    md::vector bare_force(bd_term const& t, md::vector r)
    {
        switch (t.kind) {
        case BD_TERM_SOFTCORE:
            FOR_PQ(softcore_force)
        case BD_TERM_SOFTWELL:
            switch (t.p) {
            case 2: return softwell_force<2>(t, r);
            case 4: return softwell_force<4>(t, r);
            case 5: return softwell_force<5>(t, r);
            case 6: return softwell_force<6>(t, r);
            case 8: return softwell_force<8>(t, r);
            default: throw std::runtime_error("ref: unsupported softwell P");
            }
        case BD_TERM_HARMONIC:
            return md::harmonic_potential{t.energy}.evaluate_force(r);
        case BD_TERM_SPRING:
            return md::spring_potential{t.energy, t.length}.evaluate_force(r);
        case BD_TERM_SEMISPRING:
            return md::semispring_potential{t.energy, t.length}.evaluate_force(r);
        default:
            return md::constant_potential{0}.evaluate_force(r);
        }
    }

    md::scalar bare_energy(bd_term const& t, md::vector r)
    {
        switch (t.kind) {
        case BD_TERM_SOFTCORE:
            FOR_PQ(softcore_energy)
        case BD_TERM_SOFTWELL:
            switch (t.p) {
            case 2: return softwell_energy<2>(t, r);
            case 4: return softwell_energy<4>(t, r);
            case 5: return softwell_energy<5>(t, r);
            case 6: return softwell_energy<6>(t, r);
            case 8: return softwell_energy<8>(t, r);
            default: throw std::runtime_error("ref: unsupported softwell P");
            }
        case BD_TERM_HARMONIC:
            return md::harmonic_potential{t.energy}.evaluate_energy(r);
        case BD_TERM_SPRING:
            return md::spring_potential{t.energy, t.length}.evaluate_energy(r);
        case BD_TERM_SEMISPRING:
            return md::semispring_potential{t.energy, t.length}.evaluate_energy(r);
        default:
            return 0;
        }
    }

    // A bare potential object the reference's combinators accept.
    struct term_potential
    {
        bd_term t;

        md::scalar evaluate_energy(md::vector r) const
        {
            return bare_energy(t, r);
        }

        md::vector evaluate_force(md::vector r) const
        {
            return bare_force(t, r);
        }
    };

    // The drivers' `a * pot_a + b * pot_b` with run-time arity.
    struct table_potential
    {
        bd_term const* terms;
        int nterms;

        md::scalar term_energy(int k, md::vector r) const
        {
            term_potential const base{terms[k]};
            if (terms[k].cutoff != 0) {
                return (terms[k].factor * md::apply_cutoff(base, terms[k].cutoff)).evaluate_energy(r);
            }
            return md::operator*(terms[k].factor, base).evaluate_energy(r);
        }

        md::vector term_force(int k, md::vector r) const
        {
            term_potential const base{terms[k]};
            if (terms[k].cutoff != 0) {
                return (terms[k].factor * md::apply_cutoff(base, terms[k].cutoff)).evaluate_force(r);
            }
            return md::operator*(terms[k].factor, base).evaluate_force(r);
        }

        md::scalar evaluate_energy(md::vector r) const
        {
            md::scalar sum = term_energy(0, r);
            for (int k = 1; k < nterms; k++) {
                sum = sum + term_energy(k, r);
            }
            return sum;
        }

        md::vector evaluate_force(md::vector r) const
        {
            md::vector sum = term_force(0, r);
            for (int k = 1; k < nterms; k++) {
                sum = sum + term_force(k, r);
            }
            return sum;
        }
    };

    struct pair_slot
    {
        bool used = false;
        bool enabled = false;
        int nterms = 0;
        std::vector<bd_term> table;
        md::scalar dcut = 0;
        std::vector<md::index> targets;
        std::shared_ptr<md::forcefield> ff;
    };

    struct bond_slot
    {
        bool used = false;
        bool enabled = false;
        bool min_image = false;
        int nterms = 0;
        std::vector<bd_term> table;
        std::vector<std::uint32_t> pairs;
        std::shared_ptr<md::forcefield> ff;
    };

    struct bend_slot
    {
        bool used = false;
        bool enabled = false;
        std::shared_ptr<md::forcefield> ff;
    };

    struct wall_slot
    {
        bool used = false;
        bool enabled = false;
        int shape = 0;
        int nterms = 0;
        std::vector<bd_term> table;
        md::point center;
        md::vector semi;
        std::shared_ptr<md::forcefield> ff;
        std::function<void(double*)> read_stats;
    };

    struct point_slot
    {
        bool used = false;
        bool enabled = false;
        int nterms = 0;
        std::vector<bd_term> table;
        std::shared_ptr<md::forcefield> ff;
    };
}

// a driver-defined forcefield: adds a force array the caller evaluated (set_external_forces)
struct external_forcefield : md::forcefield
{
    std::vector<md::vector> forces;
    md::scalar compute_energy(md::system const&) override { return 0; }
    void compute_force(md::system const&, md::array_view<md::vector> out) override
    {
        for (std::size_t i = 0; i < forces.size(); i++) out[i] += forces[i];
    }
};

struct ref_system
{
    md::system system;
    std::shared_ptr<external_forcefield> external;
    std::vector<std::uint8_t> type;
    int ntypes = 1;
    int box_kind = BD_BOX_OPEN;
    md::periodic_box pbox;
    pair_slot pair[BD_MAX_PAIR_FF];
    bond_slot bond[BD_MAX_BOND_FF];
    bend_slot bend[BD_MAX_BEND_FF];
    wall_slot wall[BD_MAX_WALL_FF];
    point_slot point[BD_MAX_POINT_FF];

    // persistent integrator so that a run can be advanced in several ref_step() calls
    std::unique_ptr<md::detail::brownian_timestepper> timestepper;
    std::unique_ptr<md::detail::brownian_simulator> simulator;
    std::vector<md::vector> scratch_forces;
    md::scalar last_dt = 0;
    std::int64_t steps_done = 0;

    // sum_forcefield order is registration order; re-register in the canonical order
    // pair, bond, bend, wall, point whenever the set changes
    std::vector<std::shared_ptr<md::forcefield>> registered;

    void reregister()
    {
        for (auto& ff : registered) {
            system.remove_forcefield(ff);
        }
        registered.clear();
        auto add = [&](bool on, std::shared_ptr<md::forcefield> const& ff) {
            if (on && ff) {
                system.add_forcefield(ff);
                registered.push_back(ff);
            }
        };
        for (auto& s : pair) add(s.used && s.enabled, s.ff);
        for (auto& s : bond) add(s.used && s.enabled, s.ff);
        for (auto& s : bend) add(s.used && s.enabled, s.ff);
        for (auto& s : wall) add(s.used && s.enabled, s.ff);
        for (auto& s : point) add(s.used && s.enabled, s.ff);
        add(external != nullptr, external);
    }
};

// A pair-list force field with the displacement taken through the periodic box: the shape of
// 3-sim-1kb's glue_forcefield (src/simulation/forces/glue_forcefield.cpp:17-45), on micromd's
// md::periodic_box::shortest_displacement and the same potential evaluation as the bonded force field.
template<typename PotFun>
struct boxed_pair_forcefield : md::forcefield
{
    ref_system* owner;
    int slot;
    PotFun potfun;

    boxed_pair_forcefield(ref_system* s, int k, PotFun f) : owner{s}, slot{k}, potfun{f} {}

    md::vector displacement(md::point a, md::point b) const
    {
        if (owner->box_kind == BD_BOX_PERIODIC) {
            return owner->pbox.shortest_displacement(a, b);
        }
        return md::open_box{}.shortest_displacement(a, b);
    }

    md::scalar compute_energy(md::system const& system) override
    {
        auto const positions = system.view_positions();
        auto const& pairs = owner->bond[slot].pairs;
        md::scalar total = 0;
        for (std::size_t b = 0; b + 1 < pairs.size(); b += 2) {
            md::index const i = pairs[b], j = pairs[b + 1];
            total += potfun(i, j).evaluate_energy(displacement(positions[i], positions[j]));
        }
        return total;
    }

    void compute_force(md::system const& system, md::array_view<md::vector> forces) override
    {
        auto const positions = system.view_positions();
        auto const& pairs = owner->bond[slot].pairs;
        for (std::size_t b = 0; b + 1 < pairs.size(); b += 2) {
            md::index const i = pairs[b], j = pairs[b + 1];
            auto const f = potfun(i, j).evaluate_force(displacement(positions[i], positions[j]));
            forces[i] += f;
            forces[j] -= f;
        }
    }
};

// (re)build the micromd force field object of bond slot k from its stored pairs
static void build_bond_ff(ref_system* s, int k)
{
    bond_slot& slot = s->bond[k];
    auto potfun = [s, k](md::index i, md::index j) {
        bond_slot const& sl = s->bond[k];
        return table_potential{
            sl.table.data() + (std::size_t(s->type[i]) * s->ntypes + s->type[j]) * sl.nterms,
            sl.nterms
        };
    };
    if (slot.min_image) {
        slot.ff = std::make_shared<boxed_pair_forcefield<decltype(potfun)>>(s, k, potfun);
    } else {
        auto ff = md::make_bonded_pairwise_forcefield(potfun);
        for (std::size_t b = 0; b + 1 < slot.pairs.size(); b += 2) {
            ff.add_bonded_pair(slot.pairs[b], slot.pairs[b + 1]);
        }
        slot.ff = std::make_shared<decltype(ff)>(ff);
    }
    s->reregister();
}

#define REF_TRY try {
#define REF_CATCH                                                                                  \
    }                                                                                              \
    catch (std::exception const& e) { return fail(e.what()); }

extern "C" {

const char* ref_last_error(void)
{
    return g_error.c_str();
}

int ref_create(std::int64_t n_beads, std::int32_t n_replicas, std::int32_t, ref_system** out)
{
    REF_TRY
    if (n_replicas != 1) {
        return fail("ref: n_replicas must be 1");
    }
    auto s = new ref_system;
    for (std::int64_t i = 0; i < n_beads; i++) {
        s->system.add_particle();
    }
    s->type.assign(std::size_t(n_beads), 0);
    *out = s;
    return 0;
    REF_CATCH
}

int ref_destroy(ref_system* s)
{
    delete s;
    return 0;
}

int ref_set_box(ref_system* s, std::int32_t box_kind, const double period[3])
{
    s->box_kind = box_kind;
    if (box_kind == BD_BOX_PERIODIC) {
        s->pbox.x_period = period[0];
        s->pbox.y_period = period[1];
        s->pbox.z_period = period[2];
    }
    return 0;
}

int ref_set_types(ref_system* s, const std::uint8_t* type, std::int32_t ntypes, const double* mobility)
{
    s->ntypes = ntypes;
    auto mob = s->system.view_mobilities();
    for (std::size_t i = 0; i < s->type.size(); i++) {
        s->type[i] = type[i];
        mob[i] = mobility[type[i]];
    }
    return 0;
}

int ref_set_positions(ref_system* s, std::int32_t, const double* xyz)
{
    auto pos = s->system.view_positions();
    for (std::size_t i = 0; i < pos.size(); i++) {
        pos[i] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    }
    return 0;
}

int ref_get_positions(ref_system* s, std::int32_t, double* xyz)
{
    auto pos = s->system.view_positions();
    for (std::size_t i = 0; i < pos.size(); i++) {
        xyz[3 * i] = pos[i].x;
        xyz[3 * i + 1] = pos[i].y;
        xyz[3 * i + 2] = pos[i].z;
    }
    return 0;
}

int ref_add_pair_ff(ref_system* s, std::int32_t nterms, const bd_term* table, double dcut,
                    const std::uint8_t* targets, std::int32_t* ff_out)
{
    REF_TRY
    for (int k = 0; k < BD_MAX_PAIR_FF; k++) {
        if (s->pair[k].used) {
            continue;
        }
        pair_slot& slot = s->pair[k];
        slot.used = slot.enabled = true;
        slot.nterms = nterms;
        slot.table.assign(table, table + std::size_t(s->ntypes) * s->ntypes * nterms);
        slot.dcut = dcut;
        if (targets) {
            for (std::size_t i = 0; i < s->type.size(); i++) {
                if (targets[i]) {
                    slot.targets.push_back(i);
                }
            }
        }
        auto potfun = [s, k](md::index i, md::index j) {
            pair_slot const& sl = s->pair[k];
            return table_potential{
                sl.table.data() + (std::size_t(s->type[i]) * s->ntypes + s->type[j]) * sl.nterms,
                sl.nterms
            };
        };
        auto ndist = [s, k] { return s->pair[k].dcut; };
        if (s->box_kind == BD_BOX_PERIODIC) {
            auto ff = md::make_neighbor_pairwise_forcefield<md::periodic_box>(potfun);
            ff.set_unit_cell([s] { return s->pbox; });
            ff.set_neighbor_distance(ndist);
            if (!slot.targets.empty()) {
                ff.set_neighbor_targets(slot.targets);
            }
            slot.ff = std::make_shared<decltype(ff)>(ff);
        } else {
            auto ff = md::make_neighbor_pairwise_forcefield(potfun);
            ff.set_neighbor_distance(ndist);
            if (!slot.targets.empty()) {
                ff.set_neighbor_targets(slot.targets);
            }
            slot.ff = std::make_shared<decltype(ff)>(ff);
        }
        s->reregister();
        *ff_out = k;
        return 0;
    }
    return fail("ref: too many pair forcefields");
    REF_CATCH
}

int ref_update_pair_ff(ref_system* s, std::int32_t ff, const bd_term* table, double dcut)
{
    pair_slot& slot = s->pair[ff];
    if (table) {
        slot.table.assign(table, table + slot.table.size());
    }
    slot.dcut = dcut;
    return 0;
}

int ref_add_bond_ff(ref_system* s, std::int32_t nterms, const bd_term* table,
                    const std::uint32_t* pairs, std::int64_t m, std::int32_t* ff_out)
{
    REF_TRY
    for (int k = 0; k < BD_MAX_BOND_FF; k++) {
        if (s->bond[k].used) {
            continue;
        }
        bond_slot& slot = s->bond[k];
        slot.used = slot.enabled = true;
        slot.nterms = nterms;
        slot.table.assign(table, table + std::size_t(s->ntypes) * s->ntypes * nterms);
        slot.pairs.assign(pairs, pairs + 2 * m);
        build_bond_ff(s, k);
        *ff_out = k;
        return 0;
    }
    return fail("ref: too many bond forcefields");
    REF_CATCH
}

int ref_set_bond_pairs(ref_system* s, std::int32_t ff, const std::uint32_t* pairs, std::int64_t m)
{
    REF_TRY
    if (ff < 0 || ff >= BD_MAX_BOND_FF || !s->bond[ff].used || m < 0) {
        return fail("ref: bad bond forcefield id or pair count");
    }
    s->bond[ff].pairs.assign(pairs, pairs + 2 * m);
    build_bond_ff(s, ff);
    return 0;
    REF_CATCH
}

int ref_set_bond_min_image(ref_system* s, std::int32_t ff, std::int32_t on)
{
    REF_TRY
    if (ff < 0 || ff >= BD_MAX_BOND_FF || !s->bond[ff].used) {
        return fail("ref: bad bond forcefield id");
    }
    s->bond[ff].min_image = on != 0;
    build_bond_ff(s, ff);
    return 0;
    REF_CATCH
}

int ref_update_bond_ff(ref_system* s, std::int32_t ff, const bd_term* table)
{
    bond_slot& slot = s->bond[ff];
    slot.table.assign(table, table + slot.table.size());
    return 0;
}

int ref_add_bend_ff(ref_system* s, const std::uint32_t* triples, const double* bending_energy,
                    std::int64_t m, std::int32_t* ff_out)
{
    REF_TRY
    for (int k = 0; k < BD_MAX_BEND_FF; k++) {
        if (s->bend[k].used) {
            continue;
        }
        bend_slot& slot = s->bend[k];
        slot.used = slot.enabled = true;
        auto energy_of = std::make_shared<std::map<std::array<md::index, 3>, double>>();
        for (std::int64_t t = 0; t < m; t++) {
            (*energy_of)[{triples[3 * t], triples[3 * t + 1], triples[3 * t + 2]}] = bending_energy[t];
        }
        auto ff = md::make_bonded_triplewise_forcefield(
            [energy_of](md::index i, md::index j, md::index k) {
                return md::cosine_bending_potential{energy_of->at({i, j, k})};
            });
        for (std::int64_t t = 0; t < m; t++) {
            ff.add_bonded_triple(triples[3 * t], triples[3 * t + 1], triples[3 * t + 2]);
        }
        slot.ff = std::make_shared<decltype(ff)>(ff);
        s->reregister();
        *ff_out = k;
        return 0;
    }
    return fail("ref: too many bending forcefields");
    REF_CATCH
}

int ref_add_wall_ff(ref_system* s, std::int32_t shape, std::int32_t direction, std::int32_t nterms,
                    const bd_term* table, const double center[3], const double semiaxes[3],
                    std::int32_t* ff_out)
{
    REF_TRY
    for (int k = 0; k < BD_MAX_WALL_FF; k++) {
        if (s->wall[k].used) {
            continue;
        }
        wall_slot& slot = s->wall[k];
        slot.used = slot.enabled = true;
        slot.shape = shape;
        slot.nterms = nterms;
        slot.table.assign(table, table + std::size_t(s->ntypes) * nterms);
        slot.center = {center[0], center[1], center[2]};
        slot.semi = {semiaxes[0], semiaxes[1], semiaxes[2]};

        auto potfun = [s, k](md::index i) {
            wall_slot const& sl = s->wall[k];
            return table_potential{sl.table.data() + std::size_t(s->type[i]) * sl.nterms, sl.nterms};
        };
        auto get_sphere = [s, k] {
            md::sphere sp;
            sp.center = s->wall[k].center;
            sp.radius = s->wall[k].semi.x;
            return sp;
        };
        auto get_ellipsoid = [s, k] {
            md::ellipsoid el;
            el.center = s->wall[k].center;
            el.semiaxis_x = s->wall[k].semi.x;
            el.semiaxis_y = s->wall[k].semi.y;
            el.semiaxis_z = s->wall[k].semi.z;
            return el;
        };

        auto keep_sphere = [&](auto ff) {
            ff.set_sphere(std::function<md::sphere()>(get_sphere));
            auto ptr = std::make_shared<decltype(ff)>(ff);
            slot.ff = ptr;
            slot.read_stats = [ptr](double* out) {
                out[0] = ptr->stats.reaction_force;
                out[1] = out[2] = out[3] = 0;
            };
        };
        auto keep_ellipsoid = [&](auto ff) {
            ff.set_ellipsoid(std::function<md::ellipsoid()>(get_ellipsoid));
            auto ptr = std::make_shared<decltype(ff)>(ff);
            slot.ff = ptr;
            slot.read_stats = [ptr](double* out) {
                out[0] = ptr->stats.reaction_force;
                out[1] = ptr->stats.axial_reaction.x;
                out[2] = ptr->stats.axial_reaction.y;
                out[3] = ptr->stats.axial_reaction.z;
            };
        };

        if (shape == BD_WALL_SPHERE && direction == BD_WALL_INWARD) {
            keep_sphere(md::make_sphere_inward_forcefield(potfun));
        } else if (shape == BD_WALL_SPHERE) {
            keep_sphere(md::make_sphere_outward_forcefield(potfun));
        } else if (direction == BD_WALL_INWARD) {
            keep_ellipsoid(md::make_ellipsoid_inward_forcefield(potfun));
        } else {
            keep_ellipsoid(md::make_ellipsoid_outward_forcefield(potfun));
        }
        s->reregister();
        *ff_out = k;
        return 0;
    }
    return fail("ref: too many wall forcefields");
    REF_CATCH
}

int ref_update_wall_ff(ref_system* s, std::int32_t ff, const bd_term* table, const double center[3],
                       const double semiaxes[3])
{
    wall_slot& slot = s->wall[ff];
    if (table) {
        slot.table.assign(table, table + slot.table.size());
    }
    slot.center = {center[0], center[1], center[2]};
    slot.semi = {semiaxes[0], semiaxes[1], semiaxes[2]};
    return 0;
}

int ref_get_wall_stats(ref_system* s, std::int32_t ff, std::int32_t, double out[4])
{
    s->wall[ff].read_stats(out);
    return 0;
}

int ref_add_point_ff(ref_system* s, std::int32_t nterms, const bd_term* table,
                     const double source[3], const std::uint8_t* targets, std::int32_t* ff_out)
{
    REF_TRY
    for (int k = 0; k < BD_MAX_POINT_FF; k++) {
        if (s->point[k].used) {
            continue;
        }
        point_slot& slot = s->point[k];
        slot.used = slot.enabled = true;
        slot.nterms = nterms;
        slot.table.assign(table, table + std::size_t(s->ntypes) * nterms);
        auto ff = md::make_point_source_forcefield([s, k](md::index i) {
            point_slot const& sl = s->point[k];
            return table_potential{sl.table.data() + std::size_t(s->type[i]) * sl.nterms, sl.nterms};
        });
        ff.set_point_source({source[0], source[1], source[2]});
        if (targets) {
            std::vector<md::index> idx;
            for (std::size_t i = 0; i < s->type.size(); i++) {
                if (targets[i]) {
                    idx.push_back(i);
                }
            }
            ff.set_point_source_targets(idx);
        }
        slot.ff = std::make_shared<decltype(ff)>(ff);
        s->reregister();
        *ff_out = k;
        return 0;
    }
    return fail("ref: too many point-source forcefields");
    REF_CATCH
}

int ref_set_ff_enabled(ref_system* s, std::int32_t ff_class, std::int32_t ff, std::int32_t enabled)
{
    switch (ff_class) {
    case BD_FF_PAIR: s->pair[ff].enabled = enabled; break;
    case BD_FF_BOND: s->bond[ff].enabled = enabled; break;
    case BD_FF_BEND: s->bend[ff].enabled = enabled; break;
    case BD_FF_WALL: s->wall[ff].enabled = enabled; break;
    case BD_FF_POINT: s->point[ff].enabled = enabled; break;
    default: return fail("ref: bad forcefield class");
    }
    s->reregister();
    return 0;
}

static md::forcefield* pick(ref_system* s, std::int32_t ff_class, std::int32_t ff)
{
    switch (ff_class) {
    case BD_FF_PAIR: return s->pair[ff].ff.get();
    case BD_FF_BOND: return s->bond[ff].ff.get();
    case BD_FF_BEND: return s->bend[ff].ff.get();
    case BD_FF_WALL: return s->wall[ff].ff.get();
    case BD_FF_POINT: return s->point[ff].ff.get();
    default: return nullptr;
    }
}

int ref_set_external_forces(ref_system* s, std::int32_t, double const* xyz)
{
    REF_TRY
    if (!xyz) {
        s->external.reset();
    } else {
        if (!s->external) s->external = std::make_shared<external_forcefield>();
        std::size_t const n = s->system.particle_count();
        s->external->forces.resize(n);
        for (std::size_t i = 0; i < n; i++) s->external->forces[i] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    }
    s->reregister();
    return 0;
    REF_CATCH
}

int ref_compute_forces(ref_system* s, std::int32_t, std::int32_t ff_class, std::int32_t ff,
                       double* out_xyz)
{
    REF_TRY
    std::size_t const n = s->system.particle_count();
    s->scratch_forces.assign(n, md::vector{});
    if (ff_class == BD_FF_ALL) {
        s->system.compute_force(s->scratch_forces);
    } else {
        md::forcefield* one = pick(s, ff_class, ff);
        if (!one) {
            return fail("ref: bad forcefield class / id");
        }
        one->compute_force(s->system, s->scratch_forces);
    }
    for (std::size_t i = 0; i < n; i++) {
        out_xyz[3 * i] = s->scratch_forces[i].x;
        out_xyz[3 * i + 1] = s->scratch_forces[i].y;
        out_xyz[3 * i + 2] = s->scratch_forces[i].z;
    }
    return 0;
    REF_CATCH
}

int ref_compute_energy(ref_system* s, std::int32_t, std::int32_t ff_class, std::int32_t ff, double* out)
{
    REF_TRY
    if (ff_class == BD_FF_ALL) {
        *out = s->system.compute_potential_energy();
    } else {
        md::forcefield* one = pick(s, ff_class, ff);
        if (!one) {
            return fail("ref: bad forcefield class / id");
        }
        *out = one->compute_energy(s->system);
    }
    return 0;
    REF_CATCH
}

int ref_step(ref_system* s, const bd_step_config* cfg)
{
    REF_TRY
    if (cfg->replay_normals) {
        return fail("ref: the reference draws its own noise; use ref_normals() to obtain it");
    }
    if (cfg->first_step == 1 || !s->simulator) {
        // md::simulate_brownian_dynamics (simulation/brownian_dynamics.hpp:49-64)
        if (cfg->spacestep != 0) {
            s->timestepper = std::make_unique<md::detail::adaptive_brownian_timestepper>(
                cfg->timestep, cfg->spacestep);
        } else {
            s->timestepper = std::make_unique<md::detail::monotonic_brownian_timestepper>(cfg->timestep);
        }
        s->simulator = std::make_unique<md::detail::brownian_simulator>(
            s->system, *s->timestepper, cfg->temperature, cfg->seed);
    }
    for (std::int64_t k = 0; k < cfg->steps; k++) {
        s->simulator->simulate_step();
        s->steps_done++;
    }
    return 0;
    REF_CATCH
}

// The public entry point itself, for cross-checking ref_step().
int ref_simulate(ref_system* s, double temperature, double timestep, double spacestep,
                 std::int64_t steps, std::uint64_t seed)
{
    REF_TRY
    md::brownian_dynamics_config config;
    config.temperature = temperature;
    config.timestep = timestep;
    config.spacestep = spacestep;
    config.steps = steps;
    config.seed = seed;
    md::simulate_brownian_dynamics(s->system, config);
    return 0;
    REF_CATCH
}

int ref_get_last_timestep(ref_system* s, std::int32_t, double* out)
{
    *out = s->last_dt;
    return 0;
}

int ref_get_neighbor_list(ref_system* s, std::int32_t, std::uint32_t* pairs_out, std::int64_t cap,
                          std::int64_t* count_out)
{
    REF_TRY
    // The forcefield's own list is private; a fresh md::neighbor_list updated with the current
    // positions holds exactly what a rebuild at these positions produces.
    pair_slot const& slot = s->pair[0];
    std::int64_t count = 0;
    auto emit = [&](auto const& list) {
        for (auto const& pair : list) {
            if (count < cap) {
                pairs_out[2 * count] = std::uint32_t(pair.first);
                pairs_out[2 * count + 1] = std::uint32_t(pair.second);
            }
            count++;
        }
    };
    if (s->box_kind == BD_BOX_PERIODIC) {
        md::neighbor_list<md::periodic_box> list;
        if (!slot.targets.empty()) {
            list.set_targets(slot.targets);
        }
        list.update(s->system.view_positions(), slot.dcut, s->pbox);
        emit(list);
    } else {
        md::neighbor_list<md::open_box> list;
        if (!slot.targets.empty()) {
            list.set_targets(slot.targets);
        }
        list.update(s->system.view_positions(), slot.dcut, md::open_box{});
        emit(list);
    }
    *count_out = count;
    return 0;
    REF_CATCH
}

int ref_get_counters(ref_system* s, bd_counters* out)
{
    std::memset(out, 0, sizeof *out);
    out->steps_done = s->steps_done;
    return 0;
}

int ref_nsearch_buckets(std::int32_t box_kind, const double period[3], std::int64_t particle_count,
                        double spacing, const double* xyz, std::int64_t n, std::uint32_t* bucket_out,
                        std::int32_t)
{
    REF_TRY
    if (box_kind == BD_BOX_PERIODIC) {
        md::periodic_box box;
        box.x_period = period[0];
        box.y_period = period[1];
        box.z_period = period[2];
        md::nsearch_detail::search_grid<md::periodic_box> grid{box, spacing};
        for (std::int64_t i = 0; i < n; i++) {
            bucket_out[i] = std::uint32_t(grid.locate_bucket({xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]}));
        }
    } else {
        md::open_box box;
        box.particle_count = md::index(particle_count);
        md::nsearch_detail::search_grid<md::open_box> grid{box, spacing};
        for (std::int64_t i = 0; i < n; i++) {
            bucket_out[i] = std::uint32_t(grid.locate_bucket({xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]}));
        }
    }
    return 0;
    REF_CATCH
}

int ref_nsearch_pairs(std::int32_t box_kind, const double period[3], std::int64_t particle_count,
                      double dcut, const double* xyz, std::int64_t n, std::uint32_t* pairs_out,
                      std::int64_t cap, std::int64_t* count_out, std::int32_t)
{
    REF_TRY
    std::vector<md::point> points;
    points.reserve(std::size_t(n));
    for (std::int64_t i = 0; i < n; i++) {
        points.push_back({xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]});
    }
    std::vector<std::pair<md::index, md::index>> pairs;
    if (box_kind == BD_BOX_PERIODIC) {
        md::periodic_box box;
        box.x_period = period[0];
        box.y_period = period[1];
        box.z_period = period[2];
        md::neighbor_searcher<md::periodic_box> searcher{box, dcut};
        searcher.set_points(points);
        searcher.search(std::back_inserter(pairs));
    } else {
        md::open_box box;
        box.particle_count = md::index(particle_count);
        md::neighbor_searcher<md::open_box> searcher{box, dcut};
        searcher.set_points(points);
        searcher.search(std::back_inserter(pairs));
    }
    for (std::size_t k = 0; k < pairs.size() && std::int64_t(k) < cap; k++) {
        pairs_out[2 * k] = std::uint32_t(pairs[k].first);
        pairs_out[2 * k + 1] = std::uint32_t(pairs[k].second);
    }
    *count_out = std::int64_t(pairs.size());
    return 0;
    REF_CATCH
}

int ref_normals(std::uint64_t seed, std::int64_t count, double* out)
{
    md::random_engine random;
    random.seed(seed);
    md::normal_distribution<md::scalar> normal;
    for (std::int64_t i = 0; i < count; i++) {
        out[i] = normal(random);
    }
    return 0;
}

int ref_sfc64(std::uint64_t seed, std::int64_t count, std::uint64_t* out)
{
    md::random_engine random;
    random.seed(seed);
    for (std::int64_t i = 0; i < count; i++) {
        out[i] = random();
    }
    return 0;
}

int ref_eval_potential(const bd_term* terms, std::int32_t nterms, const double r[3],
                       double* energy_out, double force_out[3])
{
    REF_TRY
    table_potential const pot{terms, nterms};
    md::vector const rv{r[0], r[1], r[2]};
    md::vector const f = pot.evaluate_force(rv);
    *energy_out = pot.evaluate_energy(rv);
    force_out[0] = f.x;
    force_out[1] = f.y;
    force_out[2] = f.z;
    return 0;
    REF_CATCH
}

// table_potential vs the drivers' compile-time expression templates
// (5-sim-genome/src/simulation_interphase/simulation_driver_forcefield.cc:30-45 shape):
// returns the largest absolute difference over a sweep of separations; expected exactly 0.
double ref_selfcheck_expression_templates(void)
{
    double worst = 0;
    std::mt19937 random;
    std::uniform_real_distribution<double> coord{-0.4, 0.4};
    for (int trial = 0; trial < 2000; trial++) {
        double const a = 0.5 * (trial % 3), b = 1 - 0.25 * (trial % 5);
        md::softcore_potential<2, 3> pa;
        pa.energy = 2.0;
        pa.diameter = 0.30;
        md::softcore_potential<8, 3> pb;
        pb.energy = 2.5;
        pb.diameter = 0.24;
        auto const expr = a * pa + b * pb;

        bd_term terms[2] = {};
        terms[0] = {BD_TERM_SOFTCORE, 2, 3, 0, a, 2.0, 0.30, 0};
        terms[1] = {BD_TERM_SOFTCORE, 8, 3, 0, b, 2.5, 0.24, 0};
        table_potential const tab{terms, 2};

        md::vector const r{coord(random), coord(random), coord(random)};
        md::vector const df = expr.evaluate_force(r) - tab.evaluate_force(r);
        worst = std::max(worst, std::fabs(df.x) + std::fabs(df.y) + std::fabs(df.z));
        worst = std::max(worst, std::fabs(expr.evaluate_energy(r) - tab.evaluate_energy(r)));
    }
    return worst;
}

} // extern "C"
