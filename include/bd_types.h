/*
 * bd_types.h -- plain-old-data types shared by the C-ABI (bd_cuda.h), the CPU
 * oracle (oracle/bd_oracle.c) and the reference harness (oracle/ref_harness.cc).
 *
 * A micromd potential expression such as
 *     a * softcore_potential<2,3>{eps_a, sigma_a} + b * softcore_potential<8,3>{eps_b, sigma_b}
 * (reference: 5-sim-genome/src/simulation_interphase/simulation_driver_forcefield.cc:30-45,
 *  micromd include/md/potential/sum_potential.hpp:20-52, scaled_potential.hpp:20-64)
 * is flattened into an ordered list of `bd_term`s.  The value of the expression is
 *     sum_k  factor_k * term_k(r)          (left to right)
 * which is exactly how sum_potential / scaled_potential evaluate it.
 */
#ifndef BD_TYPES_H
#define BD_TYPES_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BD_MAX_TERMS 4      /* terms per flattened potential                       */
#define BD_MAX_TYPES 16     /* bead classes (rows of the type table)               */
#define BD_MAX_PAIR_FF 2    /* neighbor_pairwise forcefields per system            */
#define BD_MAX_BOND_FF 8    /* bonded_pairwise forcefields per system              */
#define BD_MAX_BEND_FF 2    /* bonded_triplewise (cosine bending) forcefields      */
#define BD_MAX_WALL_FF 6    /* sphere / ellipsoid surface forcefields              */
#define BD_MAX_POINT_FF 4   /* point_source forcefields                            */

/* md::open_box / md::periodic_box  (micromd include/md/misc/box.hpp:18-52) */
enum bd_box_kind {
    BD_BOX_OPEN = 0,
    BD_BOX_PERIODIC = 1
};

/* potential kinds (micromd include/md/potential/, one header each) */
enum bd_term_kind {
    BD_TERM_NONE = 0,       /* constant_potential{0}: no force, no energy          */
    BD_TERM_SOFTCORE = 1,   /* softcore_potential<P,Q>{energy, diameter=length}    */
    BD_TERM_SOFTWELL = 2,   /* softwell_potential<P>{energy, decay_distance=length}*/
    BD_TERM_HARMONIC = 3,   /* harmonic_potential{spring_constant=energy}          */
    BD_TERM_SPRING = 4,     /* spring_potential{K=energy, b=length}                */
    BD_TERM_SEMISPRING = 5  /* semispring_potential{K=energy, b=length}            */
};

typedef struct bd_term {
    int32_t kind;       /* bd_term_kind                                            */
    int32_t p;          /* softcore P / softwell P                                 */
    int32_t q;          /* softcore Q                                              */
    int32_t reserved;
    double factor;      /* scaled_potential factor (1 for a bare potential)        */
    double energy;      /* epsilon, or spring constant K                           */
    double length;      /* diameter sigma, decay distance, or equilibrium length b */
    double cutoff;      /* apply_cutoff distance; 0 = no cutoff wrapper            */
} bd_term;

/* forcefield classes, used to address a single forcefield in queries */
enum bd_ff_class {
    BD_FF_ALL = 0,
    BD_FF_PAIR = 1,     /* md::make_neighbor_pairwise_forcefield                   */
    BD_FF_BOND = 2,     /* md::make_bonded_pairwise_forcefield                     */
    BD_FF_BEND = 3,     /* md::make_bonded_triplewise_forcefield + cosine bending  */
    BD_FF_WALL = 4,     /* md::make_{sphere,ellipsoid}_{inward,outward}_forcefield */
    BD_FF_POINT = 5     /* md::make_point_source_forcefield                        */
};

enum bd_wall_shape { BD_WALL_SPHERE = 0, BD_WALL_ELLIPSOID = 1 };
enum bd_wall_direction { BD_WALL_INWARD = 0, BD_WALL_OUTWARD = 1 };

/* md::brownian_dynamics_config (micromd include/md/simulation/brownian_dynamics.hpp:24-45)
 * plus the controls the C boundary needs. */
typedef struct bd_step_config {
    double temperature;     /* kT                                                  */
    double timestep;        /* dt (maximum dt when spacestep != 0)                 */
    double spacestep;       /* 0 = fixed dt; otherwise adaptive dt                 */
    int64_t steps;          /* number of steps to run in this call                 */
    uint64_t seed;          /* config.seed                                         */
    int64_t first_step;     /* 1-based index of the first step of this call within
                               the simulate_brownian_dynamics run; step 1 starts
                               with a zero previous-noise vector                   */
    /* validation mode: standard normals drawn by the reference RNG on the host,
       layout [steps][replicas][beads][3]; NULL = device Philox noise            */
    const double *replay_normals;
} bd_step_config;

/* counters the engine keeps on the device and returns on request */
typedef struct bd_counters {
    int64_t steps_done;         /* BD steps executed since creation                 */
    int64_t list_rebuilds;      /* Verlet list rebuilds                             */
    int64_t pair_interactions;  /* unordered pairs with r^2 <= dcut^2, summed over
                                   steps and replicas                               */
    int64_t listed_pairs;       /* unordered pairs in the current Verlet list       */
    int64_t overflow_beads;     /* beads whose list row overflowed at last rebuild  */
    int64_t reserved[3];
} bd_counters;

#ifdef __cplusplus
}
#endif

#endif /* BD_TYPES_H */
