/*
 * bd_cuda.h -- C ABI of the B200 Brownian-dynamics engine (libbd_b200.so).
 *
 * Plain pointers and sizes only; no C++ or torch types cross this boundary.  All pointer
 * arguments are HOST pointers unless a function says otherwise; the engine owns every device
 * buffer.  One host thread drives one engine handle (the reference is single-threaded and its
 * forcefield objects are stateful: MD/forcefield/detail/neighbor_list.hpp:232-239).
 *
 * The entry points and the reference interface each one replaces are documented in
 * bd_api_decl.inc; INTEGRATION.md shows the binding the reference side adds.
 */
#ifndef BD_CUDA_H
#define BD_CUDA_H

#include "bd_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bd_engine bd_engine;

#define BD_FN(name) bd_##name
#define BD_HANDLE bd_engine
#include "bd_api_decl.inc"
#undef BD_FN
#undef BD_HANDLE

/* ---- engine-only controls (no reference counterpart) ------------------------------------ */

/* Run the engine's kernels on an existing CUDA stream (cudaStream_t passed as void*), e.g. the
 * stream a benchmark records its CUDA events on.  NULL restores the engine's own stream. */
int bd_set_stream(bd_engine *h, void *cuda_stream);
/* Block until all work queued by this engine has finished. */
int bd_sync(bd_engine *h);

/* Tunables.  Names: "verlet_factor" (list radius / dcut, reference value 1.5,
 * MD/forcefield/detail/neighbor_list.hpp:182), "list_capacity" (ELL row length),
 * "time_kernel" (0/1: bracket every fused step kernel with CUDA events), "replica_id_base" /
 * "replica_id_stride" (noise key of local replica r = base + r * stride, for sharded ensembles), and read-only
 * "block_size", "kernel_ms_avg" (mean event-timed duration of the step kernel since last read). */
int bd_set_option(bd_engine *h, const char *name, double value);
int bd_get_option(bd_engine *h, const char *name, double *value);

/* md::neighbor_searcher<Box>::search with the reference's EMISSION ORDER (MD/misc/neighbor_searcher.hpp:54-100:
 * bucket by bucket, directed neighbours ascending, neighbour-bucket member outer, centre-bucket member inner),
 * pairs as (min, max): the pair set of bd_nsearch_pairs found on the device, ordered on the host by the
 * device-computed bucket ids.  Callers that feed a sequential generator in search order (3-sim-1kb's
 * glue_simulator.cpp:66-80) then take the reference's decisions.  Same arguments as bd_nsearch_pairs. */
int bd_nsearch_pairs_ordered(int32_t box_kind, const double period[3], int64_t particle_count, double dcut,
                             const double *xyz, int64_t n, uint32_t *pairs_out, int64_t cap, int64_t *count_out,
                             int32_t device);

/* md::neighbor_searcher<Box>::query (MD/misc/neighbor_searcher.hpp:65-70,104-131), batched: for query q the indices i
 * of the points with shortest_displacement(x_i, query_q).squared_norm() < dcut^2 (strict, fp64), ascending, at
 * index_out[start_out[q] .. start_out[q + 1]); start_out has m + 1 entries and is complete even when the total
 * exceeds cap (call with cap = 0 to size index_out).  The reference's one caller asks for every particle of a
 * frame in turn (4-sim-ab/box/src/rdf_analysis_hetero/distance_histogram.cc:52); here that is one call. */
int bd_nsearch_query(int32_t box_kind, const double period[3], double dcut, const double *xyz, int64_t n,
                     const double *queries, int64_t m, int64_t *start_out, uint32_t *index_out, int64_t cap, int32_t device);

/* ---- per-replica walls (the real 100kb interphase model in an ensemble) ---------------------------
 * The interphase driver moves the wall semiaxes of its system after every step from that system's own
 * wall reaction (5-sim-genome/src/simulation_interphase/simulation_driver_interphase.cc:70-80:
 * net = inward.stats.axial_reaction + outward.stats.axial_reaction - spring o semiaxes;
 * semiaxes += timestep * wall_mobility * net), so in an ensemble every replica has its own geometry.
 * bd_update_wall_ff (bd_api_decl.inc) sets the geometry of EVERY replica (its center/semiaxes may be NULL
 * together: table-only update); the calls below address one replica, read all replicas' statistics in one
 * copy, or hand the relaxation law to the engine so that it runs after every step on the device with no
 * host round trip (`ffs`: the walls that share the semiaxes, e.g. the inward and the outward ellipsoid;
 * count = 0 switches it off).  The host-driven and the device-resident feedback give the same trajectory
 * (tests/test_gpu_parity.py::test_wall_feedback_*). */
int bd_update_wall_ff_replica(bd_engine *h, int32_t ff, int32_t replica, const double center[3], const double semiaxes[3]);
int bd_get_wall_geometry(bd_engine *h, int32_t ff, int32_t replica, double center[3], double semiaxes[3]);
int bd_get_wall_stats_all(bd_engine *h, int32_t ff, double *out /* [n_replicas][4] */);
int bd_set_wall_relaxation(bd_engine *h, const int32_t *ffs, int32_t count, const double spring[3], double mobility);

/* Device-resident variants used by bench.py's HBM-resident timing: xyz is a DEVICE pointer to
 * [n_replicas][n_beads][3] doubles in original bead order. */
int bd_set_positions_device(bd_engine *h, const double *d_xyz);
int bd_get_positions_device(bd_engine *h, double *d_xyz);

/* ---- slab decomposition of one large system over several engines (several GPUs) ------------------
 * The reference has no counterpart (micromd is single-threaded); SURVEY.md 8e / slab.py.  Every engine
 * holds all beads; `owned[n_beads]` (HOST, 0/1, NULL = all) says which ones this engine advances.  The
 * others are ghosts: neighbours and bonded partners whose positions arrive from their owner every step.
 * The indexed gather / scatter take DEVICE pointers (original bead indices, xyz[count][3]) and are
 * enqueued on the engine's stream, so a halo exchange needs no host synchronisation. */
int bd_set_owned(bd_engine *h, const uint8_t *owned);
/* the same with a DEVICE mask, enqueued on the engine's stream (no host copy of the mask) */
int bd_set_owned_device(bd_engine *h, const uint8_t *d_owned);
int bd_gather_positions_indexed(bd_engine *h, const uint32_t *d_index, int64_t count, double *d_xyz);
/* DEVICE address of the engine's running max of |x - x(list build)|^2 (the bits of a non-negative double, 8 bytes):
 * slabs all-reduce it (integer max) after every step so that all ranks rebuild their lists at the same step */
int bd_displacement_pointer(bd_engine *h, void **d_out);
int bd_scatter_positions_indexed(bd_engine *h, const uint32_t *d_index, int64_t count, const double *d_xyz);

/* ---- contact map accumulated on the device -------------------------------------
 * Replaces contact_map::update / accumulate / clear of the interphase driver
 * (5-sim-genome/src/simulation_interphase/contact_map.cc:25-37 clear, :40-76 update, :79-91 accumulate),
 * which copies all positions to the host every contactmap_update_interval steps, searches them with
 * md::neighbor_searcher<open_box> and adds 1 per unordered pair into a sparse matrix.  The engine keeps
 * ONE map: bd_contact_update adds every pair (i < j) of `replica`'s current configuration whose OPEN-box
 * distance is <= `distance` (the reference's fp64 test; the simulation's own unit cell is ignored, as in
 * contact_map.cc:64); positions stay on the device.  bd_contact_accumulate writes up to `cap`
 * (i, j, count) triplets in (i, j) order -- the order of the reference's row-major sparse walk -- and
 * the total number of distinct pairs to *count_out (call with cap = 0 to size the buffer). */
int bd_contact_update(bd_engine *h, int32_t replica, double distance);
int bd_contact_accumulate(bd_engine *h, int32_t replica, uint32_t *triplets_out, int64_t cap, int64_t *count_out);
int bd_contact_clear(bd_engine *h, int32_t replica);

/* ---- pair search and pair distances on the resident positions -------------------
 * What 3-sim-1kb's kinetic glue model asks of the positions at every update
 * (src/simulation/glues/glue_simulator.cpp:35-80): `_searcher.search` -- all pairs within the binding
 * distance -- and the current distance of every glued pair (unbond: r_ij.norm() > max_distance).  Both are
 * answered from the device-resident positions, so an update moves pairs and distances over PCIe, not all
 * positions.  bd_search_pairs writes up to `cap` pairs (i < j) SORTED by (i, j) and the total to *count_out;
 * use_box = 1 searches in the simulation's box (md::neighbor_searcher<periodic_box>), 0 in an open box.
 * The pair set is the reference's (fp64 test r^2 <= distance^2).  bd_pair_distances writes |x_i - x_j|,
 * through the box's shortest displacement when min_image is set. */
int bd_search_pairs(bd_engine *h, int32_t replica, double distance, int32_t use_box, uint32_t *pairs_out, int64_t cap,
                    int64_t *count_out);
int bd_pair_distances(bd_engine *h, int32_t replica, const uint32_t *pairs, int64_t m, int32_t min_image, double *out);

/* ---- kinetic glue update on the device --------------------------------------------
 * Replaces glue_simulator::update of the 1kb chromatin driver (3-sim-1kb/src/simulation/glues/
 * glue_simulator.cpp:35-80) for a glue list held as the pair list of bond forcefield `bond_ff`
 * (bd_add_bond_ff / bd_set_bond_pairs + bd_set_bond_min_image): glued pairs farther apart than max_distance
 * (box distance when use_box) or hit by a Bernoulli(p_unbind) draw come apart; every pair within max_distance
 * that is not glued passes a Bernoulli(p_bind) draw, and the passing pairs fill the free capacity
 * max_glues - glued, a uniformly random subset if they are too many (the reference's reservoir sampling).
 * p_bind / p_unbind are the driver's -expm1(-rate * timestep).  Everything runs on the resident positions;
 * only the new list crosses to the host.  Draws are Philox keyed by (seed, update_index, pair): the same
 * stochastic process as the reference, not its sample path (the reference consumes one sequential generator in
 * search order).  Outputs (each may be NULL): pairs unbound, pairs bound, length of the new list. */
int bd_glue_update(bd_engine *h, int32_t bond_ff, int32_t replica, double max_distance, double p_bind, double p_unbind,
                   int64_t max_glues, uint64_t seed, int64_t update_index, int32_t use_box, int64_t *unbound_out,
                   int64_t *bound_out, int64_t *count_out);
/* the current pair list of a bond forcefield: up to `cap` pairs to pairs_out [cap][2], the total to *count_out */
int bd_get_bond_pairs(bd_engine *h, int32_t bond_ff, uint32_t *pairs_out, int64_t cap, int64_t *count_out);

/* ---- single-precision snapshots -------------------------------------------------
 * The drivers' stores write float positions: 5-sim-genome quantises to 16 fraction bits on the host,
 * quantize(val, bits) = nearbyint(float(val) * 2^bits) / 2^bits in float arithmetic
 * (5-sim-genome/src/simulation_common/simulation_store.cc:253-276, :403-407); 4-sim-ab and 3-sim-1kb
 * convert to float (4-sim-ab/box/src/simulation/simulation_store.cc:57-66).  bd_snapshot_begin converts
 * replica `replica` (-1 = every replica, [n_replicas][n_beads][3]) on the device in original bead order
 * -- fraction_bits >= 0: the quantisation above, bit for bit; < 0: round-to-nearest float -- and starts
 * the copy into pinned host memory on a second stream; the engine can keep stepping.  bd_snapshot_wait
 * blocks until the floats have arrived and returns the engine-owned buffer (valid until the next
 * bd_snapshot_begin) and the number of beads in it. */
int bd_snapshot_begin(bd_engine *h, int32_t replica, int32_t fraction_bits);
int bd_snapshot_wait(bd_engine *h, const float **xyz_out, int64_t *count_out);

/* ---- spline refinement between phases ------------------------------------------
 * Replaces 5-sim-genome/src/refine (refinement.py:11-21, run.py:27-40), a Python/scipy step between the
 * packing and the relaxation binaries: coarse chain c = rows [coarse_ranges[c][0], coarse_ranges[c][1]) of
 * coarse_xyz (m >= 4 points) is interpolated by the cubic spline of scipy.interpolate.splprep(k=3, s=0) on
 * u_i = (i + 1/2)/m, evaluated at (f + 1/2)/(m * factor), and its first fine_ranges[c][1] - fine_ranges[c][0]
 * points become rows fine_ranges[c][0].. of the output; then every attach[a] = {nor, nuc} copies bead nor to
 * bead nuc.  Rows of the output no chain covers are left as the caller passed them.  The _device variant
 * takes DEVICE position arrays (ranges stay on the host) and runs on `cuda_stream` (NULL = default). */
int bd_refine_spline(const double *coarse_xyz, const int64_t *coarse_ranges, int64_t n_chains, int32_t factor,
                     const int64_t *fine_ranges, const int64_t *attach, int64_t n_attach, int64_t n_coarse,
                     int64_t n_fine, double *fine_xyz_out, int32_t device);
int bd_refine_spline_device(const double *d_coarse, const int64_t *coarse_ranges, int64_t n_chains, int32_t factor,
                            const int64_t *fine_ranges, const int64_t *attach, int64_t n_attach, int64_t n_coarse,
                            int64_t n_fine, double *d_fine, void *cuda_stream);

/* bd_refine_spline_device for callers that cannot allocate device memory.  No counterpart in the reference
 * (its refinement is a scipy script on the host); same arguments, same result as bd_refine_spline_device.
 * Meant for a driver that refines while the card is already full -- the engines of the 100kb phase were created
 * first, or a framework's caching allocator holds all free memory, so cudaMalloc would fail -- and for one that
 * refines while other engines step on the same GPU, which every cudaMalloc / cudaFree would stall.  The spline
 * workspace (second derivatives, sweep coefficients, the chain ranges and attachments) is a static array inside
 * libbd_b200.so, BD_REFINE_NOALLOC_SCRATCH_BYTES on every GPU that loads the library, held until the process
 * ends; the call never calls cudaMalloc, cudaFree or any allocating library routine.  A call larger than the
 * limits below is refused with an error that names them (split it by chains, or use bd_refine_spline_device).
 *   - The first call on a device looks up where that device keeps the array; make it from plain host code, before
 *     beginning any stream capture, with the device of d_coarse / d_fine current.
 *   - All calls on a device share the one array.  It keeps the previous call's contents, which is harmless (a call
 *     writes every entry before reading it; nothing needs zeroing), but two calls that overlap on the GPU -- from
 *     two threads or on two streams -- race and give wrong splines.  Calls from one thread are safe: each returns
 *     after its work has finished. */
#define BD_REFINE_NOALLOC_MAX_COARSE 65536 /* coarse points per call (a 10Mb genome has about 640) */
#define BD_REFINE_NOALLOC_MAX_CHAINS 4096  /* chains per call */
#define BD_REFINE_NOALLOC_MAX_ATTACH 4096  /* {nor, nuc} attachments per call */
#define BD_REFINE_NOALLOC_SCRATCH_BYTES 3342336 /* 2 x 3 x 65536 doubles + 2 x 2 x 4096 + 2 x 4096 int64 */
int bd_refine_spline_device_noalloc(const double *d_coarse, const int64_t *coarse_ranges, int64_t n_chains, int32_t factor,
                                    const int64_t *fine_ranges, const int64_t *attach, int64_t n_attach, int64_t n_coarse,
                                    int64_t n_fine, double *d_fine, void *cuda_stream);

/* All replicas at once through HOST buffers laid out [n_replicas][n_beads][3] (original bead order);
 * with pinned buffers the copies are asynchronous on the engine's stream.  bd_get_all_positions
 * returns after the data has arrived. */
int bd_set_all_positions(bd_engine *h, const double *xyz);
int bd_get_all_positions(bd_engine *h, double *xyz);
/* the download only enqueued (pinned xyz): complete after bd_sync.  Engines that hold chunks of one replica
 * ensemble on their own streams overlap upload, step and download of different chunks this way. */
int bd_get_all_positions_async(bd_engine *h, double *xyz);

/* ---- statistics of the production noise -----------------------------------------------------------
 * The engine's step kernel draws counter-based Philox4x32-10 + single-precision Box-Muller normals keyed by
 * (seed, step, replica, bead) where the reference draws sfc64 + double ziggurat normals from one sequential
 * stream (MD/simulation/detail/brownian_simulator.hpp:61-70): same distribution, different stream.  This
 * call draws the 3 * beads * steps normals of beads [0, beads) at steps [first_step, first_step + steps)
 * exactly as the kernel does and writes 20 doubles: {count, sum z, sum z^2, sum z^3, sum z^4, sum z^6,
 * max |z|, #(|z| > 1) .. #(|z| > 7), sum z(step) z(step + 1), sum z0 z1, sum z0 z2, sum z1 z2,
 * sum z0^2 z1^2, triples}.  tests/test_gpu_noise.py holds them against the exact normal values. */
int bd_noise_statistics(uint64_t seed, int64_t first_step, int64_t steps, int64_t beads, int32_t replica,
                        int32_t device, double *out, int32_t out_len);

/* Number of kernels launched by this engine since creation (bench.py "gpu_launches").  Not a cheap getter: the
 * rebuild kernels inside the step graph's conditional node are counted on the device, so the call WAITS for the
 * engine's stream before it reads that count. */
int bd_get_launch_count(bd_engine *h, int64_t *out);

#ifdef __cplusplus
}
#endif

#endif /* BD_CUDA_H */
