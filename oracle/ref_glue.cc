/*
 * ref_glue.cc -- TEST INFRASTRUCTURE: the reference's kinetic glue model behind plain C entry points.
 *
 * oracle/Makefile compiles this file TOGETHER WITH the unmodified
 *     /root/reference/3-sim-1kb/src/simulation/glues/glue_simulator.cpp
 * (which needs only <md.hpp>, its two header neighbours and the standard library) into
 * oracle/_ref/libglue_ref.so.  tests/test_gpu_glue.py drives the class on the positions the engine holds and
 * compares the engine's bd_glue_update with IT: glue-count trajectories, pairs lost and gained per update,
 * the uniformity of the subset chosen when more pairs pass than fit.  The generator is the driver's own
 * (std::mt19937_64, simulation.hpp), owned by the handle and seeded by the caller.  Nothing in the product
 * links or loads this library.
 */
#include <cstdint>
#include <random>
#include <vector>

#include <md.hpp>

#include "glue_simulator.hpp"  // -I 3-sim-1kb/src/simulation/glues

namespace {
struct handle {
    glue_simulator glues;
    glue_simulator::random_engine random;
    std::vector<md::point> positions;
};
}  // namespace

extern "C" {

void *ref_glue_create(uint64_t max_glues, double max_distance, double binding_rate, double unbinding_rate,
                      const double period[3], uint64_t seed)
{
    glue_simulator::config_type config;
    config.max_glues = max_glues;
    config.max_distance = max_distance;
    config.binding_rate = binding_rate;
    config.unbinding_rate = unbinding_rate;
    config.box.x_period = period[0];
    config.box.y_period = period[1];
    config.box.z_period = period[2];
    return new handle{glue_simulator{config}, glue_simulator::random_engine{seed}, {}};
}

void ref_glue_destroy(void *h) { delete static_cast<handle *>(h); }

/* glue_simulator::update on n points; returns the number of glued pairs afterwards */
int64_t ref_glue_update(void *hv, double timestep, const double *xyz, int64_t n)
{
    handle *h = static_cast<handle *>(hv);
    h->positions.resize((std::size_t) n);
    for (int64_t i = 0; i < n; i++) h->positions[(std::size_t) i] = md::point{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    h->glues.update(timestep, h->positions, h->random);
    return (int64_t) h->glues.size();
}

/* the glued pairs in the set's iteration order; returns the total */
int64_t ref_glue_pairs(void *hv, uint32_t *pairs_out, int64_t cap)
{
    handle *h = static_cast<handle *>(hv);
    int64_t k = 0;
    for (auto const &pair : h->glues) {
        if (k < cap) {
            pairs_out[2 * k] = pair.i;
            pairs_out[2 * k + 1] = pair.j;
        }
        k++;
    }
    return k;
}

}  // extern "C"
