// A miniature of the paper's simulation drivers, written against <md.hpp> ONLY.
//
// The same source is compiled twice by tests/test_md_shim.py:
//   * against the reference's micromd headers (where /root/reference exists) -> expected output,
//     committed as tests/golden/shim_driver_expected.txt
//   * against 2022a-genome-dynamics_b200/md/md.hpp, linked with libbd_b200.so (GPU) or, for the
//     host-logic test without a GPU, with bd_ -> orc_ forwarders and the CPU oracle
//     (-DSHIM_DRIVER_HOST_QUERIES: the engine-only extensions are not there)
// and the printed numbers must agree.  Temperature is zero, so the trajectories are deterministic
// whatever noise generator is behind simulate_brownian_dynamics.
//
// The setup mirrors, in small, what the drivers register (citations are to /root/reference):
//   interphase: 5-sim-genome/src/simulation_interphase/simulation_driver_forcefield.cc:19-236,
//               simulation_driver_interphase.cc:9-88 (per-step callback moving scales and walls)
//   blend:      4-sim-ab/box/src/simulation/simulation_driver.cc, 4-sim-ab/sphere/src/simulation_driver.cc:95-233
//   spindle:    5-sim-genome/src/simulation_spindle/simulation_driver.cc:90-175 (point sources, bending)
//   chromatin:  3-sim-1kb/src/simulation/simulation.cpp:100-175 + its own glue / loop md::forcefield subclasses
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <functional>
#include <iterator>
#include <memory>
#include <random>
#include <string>
#include <utility>
#include <vector>

#include <md.hpp>

namespace
{
    struct particle_data
    {
        md::scalar a_factor = 0;
        md::scalar b_factor = 0;
    };
    md::attribute_key<particle_data> particle_data_attribute;

    struct monomer_data
    {
        md::scalar bending_energy = 0;
    };
    md::attribute_key<monomer_data> monomer_data_attribute;

    void print_points(char const* tag, md::array_view<md::point const> pts, md::index stride)
    {
        for (md::index i = 0; i < pts.size(); i += stride)
            std::printf("%s %zu %.17g %.17g %.17g\n", tag, i, pts[i].x, pts[i].y, pts[i].z);
    }

    void print_vectors(char const* tag, std::vector<md::vector> const& v, md::index stride)
    {
        for (md::index i = 0; i < v.size(); i += stride)
            std::printf("%s %zu %.17g %.17g %.17g\n", tag, i, v[i].x, v[i].y, v[i].z);
    }

    // random walk chains inside a ball
    void place_chains(md::system& system, std::vector<std::pair<md::index, md::index>> const& chains,
                      double bond, double radius, std::uint64_t seed)
    {
        std::mt19937_64 random(seed);
        std::normal_distribution<double> normal;
        std::uniform_real_distribution<double> coord(-0.5 * radius, 0.5 * radius);
        auto positions = system.view_positions();
        for (auto const& chain : chains) {
            md::point pos{coord(random), coord(random), coord(random)};
            for (md::index i = chain.first; i < chain.second; i++) {
                positions[i] = pos;
                for (;;) {
                    md::vector const step = bond * md::normalize({normal(random), normal(random), normal(random)});
                    md::point const next = pos + step;
                    if (next.distance({0, 0, 0}) < radius) {
                        pos = next;
                        break;
                    }
                }
            }
        }
    }

    // ---- 5-sim-genome interphase in small --------------------------------------------------------

    struct interphase
    {
        md::system system;
        std::vector<std::pair<md::index, md::index>> chains;
        double bead_scale = 0.5, bond_scale = 0.6;
        md::vector wall_semiaxes{2.4, 2.0, 1.7};
        std::function<md::vector()> packing_reaction;

        double const a_rep = 2.5, b_rep = 2.0, a_diam = 0.30, b_diam = 0.24;
        double const a_spring = 100, b_spring = 80, a_len = 0.25, b_len = 0.2;
        double const a_spring2 = 5, b_spring2 = 3;
        double const wall_a = 0, wall_b = 1, wall_spring = 50;
        double const dt = 2e-5;

        void setup()
        {
            system.require(particle_data_attribute);
            std::mt19937_64 random(5);
            md::index const lengths[] = {150, 110, 90};
            for (md::index len : lengths) {
                md::index const start = system.particle_count();
                int kind = 0;
                for (md::index i = 0; i < len; i++) {
                    if (random() % 8 == 0) kind = int(random() % 3);
                    auto part = system.add_particle();
                    part.mobility = kind == 2 ? 0.8 : 1.0;
                    part.view(particle_data_attribute) =
                        kind == 0 ? particle_data{1, 0} : kind == 1 ? particle_data{0, 1} : particle_data{0.5, 0.5};
                }
                chains.push_back({start, system.particle_count()});
            }
            // nucleolar side beads on a few chain beads
            std::vector<std::pair<md::index, md::index>> nucleolus_bonds;
            for (md::index nor : {md::index(20), md::index(21), md::index(200)}) {
                for (int k = 0; k < 2; k++) {
                    auto part = system.add_particle();
                    part.mobility = 0.5;
                    part.view(particle_data_attribute) = particle_data{2, 2};
                    nucleolus_bonds.push_back({nor, part.index});
                }
            }
            place_chains(system, chains, 0.2, 1.4, 17);
            {
                auto positions = system.view_positions();
                for (auto const& bond : nucleolus_bonds)
                    positions[bond.second] = positions[bond.first] + md::vector{0.05, -0.04, 0.03 * double(bond.second % 3)};
            }

            auto const max_diameter = std::max(a_diam, b_diam);
            system.add_forcefield(
                md::make_neighbor_pairwise_forcefield(
                    [this](md::index i, md::index j) {
                        auto const data = system.view(particle_data_attribute);
                        auto const a = 0.5 * (data[i].a_factor + data[j].a_factor);
                        auto const b = 0.5 * (data[i].b_factor + data[j].b_factor);
                        md::softcore_potential<2, 3> const a_potential{.energy = a_rep, .diameter = a_diam * bead_scale};
                        md::softcore_potential<8, 3> const b_potential{.energy = b_rep, .diameter = b_diam * bead_scale};
                        return a * a_potential + b * b_potential;
                    })
                .set_neighbor_distance([this, max_diameter] { return max_diameter * bead_scale; }));

            auto chrom_bonds = system.add_forcefield(md::make_bonded_pairwise_forcefield(
                [this](md::index i, md::index j) {
                    auto const data = system.view(particle_data_attribute);
                    auto const a = 0.5 * (data[i].a_factor + data[j].a_factor);
                    auto const b = 0.5 * (data[i].b_factor + data[j].b_factor);
                    auto const K = a * a_spring + b * b_spring;
                    auto const l = a * a_len + b * b_len;
                    auto const s = bond_scale;
                    return md::semispring_potential{.spring_constant = K * (1 / (s * s)), .equilibrium_distance = l * s};
                }));
            for (auto const& chain : chains) chrom_bonds->add_bonded_range(chain.first, chain.second);

            auto loop_bonds = system.add_forcefield(md::make_bonded_pairwise_forcefield(
                [this](md::index i, md::index j) {
                    auto const data = system.view(particle_data_attribute);
                    auto const a = 0.5 * (data[i].a_factor + data[j].a_factor);
                    auto const b = 0.5 * (data[i].b_factor + data[j].b_factor);
                    auto const s = bond_scale;
                    return md::harmonic_potential{.spring_constant = (a * a_spring2 + b * b_spring2) * (1 / (s * s))};
                }));
            for (auto const& chain : chains)
                for (md::index i = chain.first; i + 2 < chain.second; i++) loop_bonds->add_bonded_pair(i, i + 2);

            auto nucleo = system.add_forcefield(md::make_bonded_pairwise_forcefield(
                [this](md::index, md::index) {
                    auto const s = bond_scale;
                    return md::semispring_potential{.spring_constant = 30 * (1 / (s * s)), .equilibrium_distance = 0.1 * s};
                }));
            for (auto const& bond : nucleolus_bonds) nucleo->add_bonded_pair(bond.first, bond.second);

            // droplet-forming attraction among the nucleolar beads only
            std::vector<md::index> nucleolar;
            for (auto const& bond : nucleolus_bonds) nucleolar.push_back(bond.second);
            system.add_forcefield(
                md::make_neighbor_pairwise_forcefield(
                    md::apply_cutoff(md::softwell_potential<6>{.energy = 0.3, .decay_distance = 0.1}, 0.25))
                .set_neighbor_distance(0.25)
                .set_neighbor_targets(nucleolar));

            auto set_ellipsoid = [this] {
                return md::ellipsoid{.semiaxis_x = wall_semiaxes.x, .semiaxis_y = wall_semiaxes.y, .semiaxis_z = wall_semiaxes.z};
            };
            auto inward = system.add_forcefield(
                md::make_ellipsoid_inward_forcefield(
                    [this](md::index i) {
                        auto const data = system.view(particle_data_attribute);
                        auto const a = 0.5 * (data[i].a_factor + wall_a);
                        auto const b = 0.5 * (data[i].b_factor + wall_b);
                        md::softcore_potential<2, 3> const a_potential{.energy = a_rep, .diameter = a_diam / 2 * bead_scale};
                        md::softcore_potential<8, 3> const b_potential{.energy = b_rep, .diameter = b_diam / 2 * bead_scale};
                        return a * a_potential + b * b_potential;
                    })
                .set_ellipsoid(set_ellipsoid));
            auto outward = system.add_forcefield(
                md::make_ellipsoid_outward_forcefield(md::harmonic_potential{.spring_constant = wall_spring})
                .set_ellipsoid(set_ellipsoid));
            packing_reaction = [=] { return inward->stats.axial_reaction + outward->stats.axial_reaction; };
        }

        void run()
        {
            // push a few beads through the envelope so that both wall force fields act
            {
                auto positions = system.view_positions();
                positions[5] = {2.6, 0.1, 0.0};
                positions[170] = {0.0, -2.15, 0.2};
                positions[300] = {0.1, 0.2, 1.69};
            }
            std::vector<md::vector> forces(system.particle_count());
            system.compute_force(forces);
            print_vectors("interphase.force", forces, 7);
            auto const reaction = packing_reaction();
            std::printf("interphase.reaction %.17g %.17g %.17g\n", reaction.x, reaction.y, reaction.z);
            std::printf("interphase.energy %.17g\n", system.compute_energy());

            auto callback = [this](md::step step) {
                double const time = double(step) * dt;
                bead_scale = 1 - (1 - 0.5) * std::exp(-time / 2e-3);
                bond_scale = 1 - (1 - 0.6) * std::exp(-time / 1e-3);
                md::vector net = packing_reaction();
                net -= md::vector{3, 3, 3}.hadamard(wall_semiaxes);
                wall_semiaxes += dt * 40.0 * net;
                if (step % 40 == 0) std::printf("interphase.mean_energy %lld %.17g\n", (long long) step,
                                                system.compute_energy() / double(system.particle_count()));
            };
            callback(0);
            md::simulate_brownian_dynamics(system, {.temperature = 0, .timestep = dt, .steps = 120, .seed = 3, .callback = callback});
            print_points("interphase.pos", system.view_positions(), 5);
            std::printf("interphase.semiaxes %.17g %.17g %.17g\n", wall_semiaxes.x, wall_semiaxes.y, wall_semiaxes.z);
            std::printf("interphase.scales %.17g %.17g\n", bead_scale, bond_scale);
        }
    };

    // ---- 4-sim-ab in small: periodic blend, then the same beads inside sphere walls ----------------

    struct blend
    {
        md::system system;
        std::vector<std::pair<md::index, md::index>> chains;
        double const box = 2.2;

        void setup(bool periodic)
        {
            system.require(particle_data_attribute);
            for (int c = 0; c < 12; c++) {
                md::index const start = system.particle_count();
                for (int i = 0; i < 20; i++) {
                    auto part = system.add_particle();
                    part.view(particle_data_attribute) = c % 2 ? particle_data{0, 1} : particle_data{1, 0};
                }
                chains.push_back({start, system.particle_count()});
            }
            place_chains(system, chains, 0.18, 1.0, 23);

            md::softcore_potential<2> const a_potential{.energy = 2, .diameter = 0.30};
            md::softcore_potential<8> const b_potential{.energy = 2, .diameter = 0.24};
            auto potential = [=, this](md::index i, md::index j) {
                auto const data = system.view(particle_data_attribute);
                auto const a = (data[i].a_factor + data[j].a_factor) / 2;
                auto const b = (data[i].b_factor + data[j].b_factor) / 2;
                return a * a_potential + b * b_potential;
            };
            if (periodic) {
                system.add_forcefield(
                    md::make_neighbor_pairwise_forcefield<md::periodic_box>(potential)
                    .set_unit_cell({.x_period = box, .y_period = box, .z_period = box})
                    .set_neighbor_distance(0.30));
            } else {
                system.add_forcefield(md::make_neighbor_pairwise_forcefield(potential).set_neighbor_distance(0.30));
            }
            auto bonds = system.add_forcefield(md::make_bonded_pairwise_forcefield(md::harmonic_potential{.spring_constant = 70}));
            for (auto const& chain : chains) bonds->add_bonded_range(chain.first, chain.second);

            if (!periodic) {
                md::sphere const outer{.radius = 1.15};
                md::sphere const inner{.radius = 0.35};
                md::softcore_potential<2> const wa{.energy = 2, .diameter = 0.15};
                md::softcore_potential<8> const wb{.energy = 2, .diameter = 0.12};
                auto wall = [=, this](md::index i) {
                    auto const data = system.view(particle_data_attribute);
                    auto const a = (data[i].a_factor + 0) / 2;
                    auto const b = (data[i].b_factor + 1) / 2;
                    return 1.5 * (a * wa + b * wb);
                };
                outer_in = system.add_forcefield(md::make_sphere_inward_forcefield(wall).set_sphere(outer));
                system.add_forcefield(md::make_sphere_outward_forcefield(md::harmonic_potential{.spring_constant = 10}).set_sphere(outer));
                system.add_forcefield(md::make_sphere_inward_forcefield(md::harmonic_potential{.spring_constant = 10}).set_sphere(inner));
                system.add_forcefield(md::make_sphere_outward_forcefield(wall).set_sphere(inner));
            }
        }

        std::shared_ptr<md::forcefield> outer_in;

        void run(char const* tag)
        {
            char name[64];
            if (outer_in) {
                // beads within reach of the outer and the inner wall
                auto positions = system.view_positions();
                positions[3] = {1.1, 0.02, -0.01};
                positions[47] = {-0.02, 0.4, 0.03};
                positions[101] = {0.0, 0.0, 1.2};
                positions[150] = {0.1, -0.2, 0.15};
            }
            std::vector<md::vector> forces(system.particle_count());
            system.compute_force(forces);
            std::snprintf(name, sizeof name, "%s.force", tag);
            print_vectors(name, forces, 9);
            std::printf("%s.energy %.17g\n", tag, system.compute_potential_energy());
            if (outer_in) {
                // one force field on its own, through the md::forcefield interface
                std::vector<md::vector> own(system.particle_count());
                outer_in->compute_force(system, own);
                double sum = 0;
                for (auto const& f : own) sum += f.norm();
                std::printf("%s.outer_wall_force_sum %.17g\n", tag, sum);
                std::printf("%s.outer_wall_energy %.17g\n", tag, outer_in->compute_energy(system));
            }
            md::step samples = 0;
            md::simulate_brownian_dynamics(system, {.temperature = 0, .timestep = 1e-4, .steps = 60, .callback = [&](md::step step) {
                if (step % 20 == 0) samples++;
            }});
            // a run without callback is one engine call
            md::simulate_brownian_dynamics(system, {.temperature = 0, .timestep = 1e-4, .steps = 40});
            std::snprintf(name, sizeof name, "%s.pos", tag);
            print_points(name, system.view_positions(), 6);
            std::printf("%s.samples %lld\n", tag, (long long) samples);
        }
    };

    // ---- 5-sim-genome spindle phase / 3-sim-1kb pieces: point sources, bending, phase changes -----

    void spindle()
    {
        md::system system;
        system.require(monomer_data_attribute);
        std::vector<std::pair<md::index, md::index>> chains;
        for (int c = 0; c < 6; c++) {
            md::index const start = system.particle_count();
            for (int i = 0; i < 16 + 3 * c; i++) {
                auto part = system.add_particle();
                part.mobility = 1;
                part.view(monomer_data_attribute) = {.bending_energy = i < 8 ? 0.7 : 0.2};
            }
            chains.push_back({start, system.particle_count()});
        }
        place_chains(system, chains, 0.2, 1.2, 31);

        system.add_forcefield(
            md::make_neighbor_pairwise_forcefield(md::softcore_potential<2, 3>{.energy = 5, .diameter = 0.2})
            .set_neighbor_distance(0.2));
        auto bonds = system.add_forcefield(
            md::make_bonded_pairwise_forcefield(md::semispring_potential{.spring_constant = 500, .equilibrium_distance = 0.2}));
        auto bends = system.add_forcefield(md::make_bonded_triplewise_forcefield(
            [](md::system const& sys, md::index, md::index j, md::index) {
                return md::cosine_bending_potential{sys.view(monomer_data_attribute)[j].bending_energy};
            }));
        for (auto const& chain : chains) {
            bonds->add_bonded_range(chain.first, chain.second);
            bends->add_bonded_range(chain.first, chain.second);
        }
        std::vector<md::index> centromeres;
        for (auto const& chain : chains) centromeres.push_back((chain.first + chain.second) / 2);

        std::shared_ptr<md::forcefield> spindle_ff = std::make_shared<decltype(md::make_point_source_forcefield(md::harmonic_potential{}))>(
            md::make_point_source_forcefield(md::harmonic_potential{.spring_constant = 2})
            .set_point_source({0.3, 0, 0})
            .set_point_source_targets(centromeres));
        std::shared_ptr<md::forcefield> packing_ff = std::make_shared<decltype(md::make_point_source_forcefield(md::semispring_potential{}))>(
            md::make_point_source_forcefield(md::semispring_potential{.spring_constant = 4, .equilibrium_distance = 0.6})
            .set_point_source({0.3, 0, 0}));

        system.add_forcefield(spindle_ff);
        md::simulate_brownian_dynamics(system, {.temperature = 0, .timestep = 1e-4, .steps = 50});
        std::printf("spindle.energy1 %.17g\n", system.compute_potential_energy());
        system.remove_forcefield(spindle_ff);
        system.add_forcefield(packing_ff);
        md::simulate_brownian_dynamics(system, {.temperature = 0, .timestep = 1e-4, .spacestep = 0.002, .steps = 50});
        std::printf("spindle.energy2 %.17g\n", system.compute_potential_energy());
        print_points("spindle.pos", system.view_positions(), 4);

        // the contact map's use of md::neighbor_searcher (5-sim-genome/src/simulation_interphase/contact_map.cc:64-66)
        md::neighbor_searcher<md::open_box> searcher{{}, 0.25};
        searcher.set_points(system.view_positions());
        std::vector<std::pair<md::index, md::index>> pairs;
        searcher.search(std::back_inserter(pairs));
        std::uint64_t checksum = 0;
        for (auto const& p : pairs) {
            auto const lo = std::min(p.first, p.second), hi = std::max(p.first, p.second);
            checksum += (lo * 1000003u + hi) * 2654435761u % 4294967291u;
        }
        std::printf("spindle.contacts %zu %llu\n", pairs.size(), (unsigned long long) checksum);
    }
}

// ---- 3-sim-1kb in small: driver-defined force fields over dynamic pair lists ------------------------
// (3-sim-1kb/src/simulation/forces/glue_forcefield.cpp:3-43, loop_forcefield.cpp:8-50,
//  simulation.cpp:100-175: periodic softcore repulsion + attraction, spring bonds, per-monomer bending)
namespace
{
    struct pair_list
    {
        std::vector<std::pair<md::index, md::index>> pairs;
    };

    class glue_forcefield : public md::forcefield
    {
    public:
        glue_forcefield(std::shared_ptr<pair_list> list, md::periodic_box const& box, md::scalar energy, md::scalar cutoff)
            : _list{list}, _box{box}, _potential{.energy = -energy, .diameter = cutoff}
        {
        }

        md::scalar compute_energy(md::system const& system) override
        {
            auto const positions = system.view_positions();
            md::scalar total = 0;
            for (auto const& pair : _list->pairs)
                total += _potential.evaluate_energy(_box.shortest_displacement(positions[pair.first], positions[pair.second]));
            return total;
        }

        void compute_force(md::system const& system, md::array_view<md::vector> forces) override
        {
            auto const positions = system.view_positions();
            for (auto const& pair : _list->pairs) {
                auto const f = _potential.evaluate_force(_box.shortest_displacement(positions[pair.first], positions[pair.second]));
                forces[pair.first] += f;
                forces[pair.second] -= f;
            }
        }

    private:
        std::shared_ptr<pair_list> _list;
        md::periodic_box _box;
        md::softcore_potential<8, 3> _potential;
    };

    class loop_forcefield : public md::forcefield
    {
    public:
        loop_forcefield(std::shared_ptr<pair_list> list, md::scalar stiffness, md::scalar length)
            : _list{list}, _potential{stiffness, length}  // positional, as 3-sim-1kb/src/simulation/forces/loop_forcefield.cpp:13
        {
        }

        md::scalar compute_energy(md::system const& system) override
        {
            auto const positions = system.view_positions();
            md::scalar total = 0;
            for (auto const& pair : _list->pairs) total += _potential.evaluate_energy(positions[pair.first] - positions[pair.second]);
            return total;
        }

        void compute_force(md::system const& system, md::array_view<md::vector> forces) override
        {
            auto const positions = system.view_positions();
            for (auto const& pair : _list->pairs) {
                auto const f = _potential.evaluate_force(positions[pair.first] - positions[pair.second]);
                forces[pair.first] += f;
                forces[pair.second] -= f;
            }
        }

    private:
        std::shared_ptr<pair_list> _list;
        md::spring_potential _potential;
    };

    // device_lists (B200 header only): the same model with the two pair-list force fields evaluated by
    // the engine -- a bonded force field whose list is replaced while the simulation runs
    // (b200_set_bonded_pairs), the glue one through the periodic box (b200_use_shortest_displacement)
    void chromatin(bool device_lists = false)
    {
        char const* const tag = device_lists ? "chromatin_dev" : "chromatin";
        auto const name = [&](char const* what) { return std::string(tag) + what; };
        md::system system;
        system.require(monomer_data_attribute);
        double const box_size = 2.4;
        std::vector<std::pair<md::index, md::index>> chains;
        for (int c = 0; c < 4; c++) {
            md::index const start = system.particle_count();
            for (int i = 0; i < 60; i++) {
                auto part = system.add_particle();
                part.mobility = 1;
                part.view(monomer_data_attribute) = {.bending_energy = (i / 20) % 2 ? 2.0 : 1.0};
            }
            chains.push_back({start, system.particle_count()});
        }
        place_chains(system, chains, 0.28, 1.1, 41);

        md::softcore_potential<2, 3> const repulsive{.energy = 3, .diameter = 0.30};
        md::softcore_potential<8, 3> const attractive{.energy = 0.3 * -1, .diameter = 0.45};
        md::periodic_box const box{.x_period = box_size, .y_period = box_size, .z_period = box_size};
        system.add_forcefield(
            md::make_neighbor_pairwise_forcefield<md::periodic_box>(repulsive + attractive)
            .set_unit_cell(box)
            .set_neighbor_distance(0.45));
        auto bonds = system.add_forcefield(
            md::make_bonded_pairwise_forcefield(md::spring_potential{100, 0.3}));  // positional aggregate initialisation
        auto bends = system.add_forcefield(md::make_bonded_triplewise_forcefield(
            [](md::system const& sys, md::index, md::index i, md::index) {
                return md::cosine_bending_potential{sys.view(monomer_data_attribute)[i].bending_energy};
            }));
        for (auto const& chain : chains) {
            bonds->add_bonded_range(chain.first, chain.second);
            bends->add_bonded_range(chain.first, chain.second);
        }
        auto loops = std::make_shared<pair_list>();
        auto glues = std::make_shared<pair_list>();
        loops->pairs = {{5, 40}, {70, 110}, {130, 170}};
        glues->pairs = {{10, 14}, {65, 71}, {200, 204}};
        std::function<void()> lists_changed = [] {};
#ifdef MD_B200_MD_HPP
        if (device_lists) {
            auto loop_ff = system.add_forcefield(
                md::make_bonded_pairwise_forcefield(md::spring_potential{.spring_constant = 50, .equilibrium_distance = 0.3}));
            auto glue_ff = system.add_forcefield(
                md::make_bonded_pairwise_forcefield(md::softcore_potential<8, 3>{.energy = -2.0, .diameter = 0.45}));
            glue_ff->b200_use_shortest_displacement(true);
            lists_changed = [=] {
                loop_ff->b200_set_bonded_pairs(loops->pairs);
                glue_ff->b200_set_bonded_pairs(glues->pairs);
            };
            lists_changed();
        }
#endif
        if (!device_lists) {
            system.add_forcefield(loop_forcefield{loops, 50, 0.3});
            system.add_forcefield(glue_forcefield{glues, box, 2.0, 0.45});
        }

        std::vector<md::vector> forces(system.particle_count());
        system.compute_force(forces);
        print_vectors(name(".force").c_str(), forces, 6);
        std::printf("%s.energy %.17g\n", tag, system.compute_potential_energy());
        // the kinetic part of the driver: the pair lists change while the simulation runs
        md::simulate_brownian_dynamics(system, {.temperature = 0, .timestep = 1e-5, .steps = 60, .callback = [&](md::step step) {
            if (step == 20) loops->pairs.push_back({185, 230});
            if (step == 40) glues->pairs.erase(glues->pairs.begin());
            if (step == 20 || step == 40) lists_changed();
        }});
        md::simulate_brownian_dynamics(system, {.temperature = 0, .timestep = 1e-5, .steps = 20});
        print_points(name(".pos").c_str(), system.view_positions(), 5);
        std::printf("%s.energy_end %.17g\n", tag, system.compute_potential_energy());

        // what the kinetic glue model asks of the positions at every update (glues/glue_simulator.cpp:35-80):
        // the pairs within the binding distance and the current length of every glued pair
        std::vector<std::pair<md::index, md::index>> near;
        std::vector<md::scalar> lengths;
        bool asked_device = false;
#if defined(MD_B200_ENGINE) && !defined(SHIM_DRIVER_HOST_QUERIES)
        if (device_lists) {
            near = md::b200::search_pairs(system, 0.4);
            lengths = md::b200::pair_distances(system, glues->pairs, true);
            asked_device = true;
        }
#endif
        if (!asked_device) {
            md::neighbor_searcher<md::periodic_box> searcher{box, 0.4};
            searcher.set_points(system.view_positions());
            searcher.search(std::back_inserter(near));
            std::sort(near.begin(), near.end());
            auto const positions = system.view_positions();
            for (auto const& pair : glues->pairs)
                lengths.push_back(box.shortest_displacement(positions[pair.first], positions[pair.second]).norm());
        }
        double checksum = 0;
        for (auto const& pair : near) checksum += double(pair.first) * 1009 + double(pair.second);
        std::printf("%s.near %zu %.17g\n", tag, near.size(), checksum);
        std::printf("%s.gluelen", tag);
        for (md::scalar l : lengths) std::printf(" %.17g", l);
        std::printf("\n");
        // neighbor_searcher::query as the hetero-RDF tool uses it (4-sim-ab/box/src/rdf_analysis_hetero/distance_histogram.cc:52)
        {
            md::neighbor_searcher<md::periodic_box> searcher{box, 0.4};
            searcher.set_points(system.view_positions());
            std::vector<md::index> around;
            searcher.query(system.view_positions()[7], std::back_inserter(around));
            std::sort(around.begin(), around.end());
            double sum = 0;
            for (md::index i : around) sum += double(i) * 31 + 1;
            std::printf("%s.query %zu %.17g\n", tag, around.size(), sum);
#if defined(MD_B200_ENGINE) && !defined(SHIM_DRIVER_HOST_QUERIES)
            // the batched device form answers every particle at once and must agree with the one-point query
            auto const all = searcher.b200_query_all(system.view_positions());
            if (all.size() != system.particle_count() || all[7] != around) {
                std::printf("%s.query_all MISMATCH\n", tag);
                std::abort();
            }
#endif
        }
    }
}

int main()
{
    {
        interphase sim;
        sim.setup();
        sim.run();
    }
    {
        blend sim;
        sim.setup(true);
        sim.run("blend_box");
    }
    {
        blend sim;
        sim.setup(false);
        sim.run("blend_sphere");
    }
    spindle();
    chromatin();
#ifdef MD_B200_MD_HPP
    chromatin(true);
#endif
    return 0;
}
