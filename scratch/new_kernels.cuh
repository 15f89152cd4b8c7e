
/* MD/simulation/detail/brownian_simulator.hpp:57-75 for one bead; returns the squared distance
 * from the bead's position at the last list build (neighbor_list.hpp:138-157) */
template <int BOX>
__device__ __forceinline__ double integrate_bead(const StepArrays &A, const DParams &P, const StepScalars &S,
                                                 uint32_t s, uint32_t replica, const bead4 &me, d3 F,
                                                 double dt, double dt_prev)
{
    uint32_t const oi = w_orig(me.w);
    double const mu = P.mobility[w_type(me.w)];
    double const mu_dt = dt * mu;
    double const sigma = sqrt(2.0 * S.kT * mu_dt);
    double wx, wy, wz, px, py, pz;
    size_t const g = (size_t) replica * A.n + oi;
    if (S.replay) {
        wx = sigma * S.replay[3 * g]; wy = sigma * S.replay[3 * g + 1]; wz = sigma * S.replay[3 * g + 2];
        px = A.wprev[3 * g]; py = A.wprev[3 * g + 1]; pz = A.wprev[3 * g + 2];
        A.wprev[3 * g] = wx; A.wprev[3 * g + 1] = wy; A.wprev[3 * g + 2] = wz;
    } else {
        double z[3];
        philox_normals(S.seed, (uint64_t) S.step, replica, oi, z);
        wx = sigma * z[0]; wy = sigma * z[1]; wz = sigma * z[2];
        if (S.step > 1) {
            double const sp = sqrt(2.0 * S.kT_prev * (dt_prev * mu));
            philox_normals(S.seed, (uint64_t) (S.step - 1), replica, oi, z);
            px = sp * z[0]; py = sp * z[1]; pz = sp * z[2];
        } else {
            px = py = pz = 0.0;
        }
    }
    double x = me.x + mu_dt * F.x, y = me.y + mu_dt * F.y, z2 = me.z + mu_dt * F.z;
    x += 0.5 * (wx + px); y += 0.5 * (wy + py); z2 += 0.5 * (wz + pz);
    store_bead(A.pos_out + s, x, y, z2, me.w);

    bead4 const b0 = load_bead(A.bpos + s);
    d3 const d = displacement_t<BOX>(P, x, y, z2, b0.x, b0.y, b0.z);
    return sqnorm_exact(d.x, d.y, d.z);
}

/* fold the block's largest displacement^2 into the control block; the rebuild decision itself is
 * taken by k_rebuild_begin before the next force evaluation (neighbor_list.hpp:132-158) */
__device__ __forceinline__ void publish_displacement(const StepArrays &A, double moved2, double known)
{
    for (int o = 16; o > 0; o >>= 1) moved2 = fmax(moved2, __shfl_xor_sync(0xffffffffu, moved2, o));
    if ((threadIdx.x & 31) == 0 && moved2 > known)
        atomicMax(&A.ctl->max_disp2_bits, (unsigned long long) __double_as_longlong(moved2));
}

/* the hot kernel: forces + Brownian update + skin test, fixed timestep */
template <int BOX, int PROG>
__global__ void __launch_bounds__(STEP_BLOCK) k_step_fused(StepArrays A, StepScalars S)
{
    __shared__ StepShared sh;
    load_shared<PROG>(sh, A);
    uint32_t const replica = blockIdx.y, chunk = blockIdx.x;
    uint32_t const local = chunk * STEP_BLOCK + threadIdx.x;
    bool const live = local < A.n;
    uint32_t const s = replica * A.n + (live ? local : 0u);
    Select const all = {BD_FF_ALL, -1};
    uint32_t inter = 0;
    double const known = __longlong_as_double((long long) A.ctl->max_disp2_bits);
    bead4 const me = load_bead(A.pos + s);
    d3 const F = block_forces<BOX, PROG>(A, sh, all, live, s, replica, chunk, me, inter);
    double moved2 = 0.0;
    if (live) moved2 = integrate_bead<BOX>(A, sh.P, S, s, replica, me, F, S.dt, S.dt_prev);
    publish_displacement(A, moved2, known);
    for (int o = 16; o > 0; o >>= 1) inter += __shfl_xor_sync(0xffffffffu, inter, o);
    if ((threadIdx.x & 31) == 0) sh.inter[threadIdx.x >> 5] = inter;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < STEP_BLOCK / 32; w++) t += sh.inter[w];
        A.inter_partial[(size_t) replica * A.chunks + chunk] += t;
    }
}

/* forces only (compute_forces, and the first pass of an adaptive-timestep step) */
template <int BOX, int PROG>
__global__ void __launch_bounds__(STEP_BLOCK) k_force(StepArrays A, StepScalars S, Select sel)
{
    __shared__ StepShared sh;
    load_shared<PROG>(sh, A);
    uint32_t const replica = blockIdx.y, chunk = blockIdx.x;
    uint32_t const local = chunk * STEP_BLOCK + threadIdx.x;
    bool const live = local < A.n;
    uint32_t const s = replica * A.n + (live ? local : 0u);
    uint32_t inter = 0;
    double dt = S.dt;
    bead4 const me = load_bead(A.pos + s);
    d3 const F = block_forces<BOX, PROG>(A, sh, sel, live, s, replica, chunk, me, inter);
    if (live) {
        A.force[3 * (size_t) s] = F.x;
        A.force[3 * (size_t) s + 1] = F.y;
        A.force[3 * (size_t) s + 2] = F.z;
        if (S.adaptive) {
            // brownian_timestepper.hpp:72-91,109-136
            double const mu = sh.P.mobility[w_type(me.w)];
            double const f2 = F.x * F.x + F.y * F.y + F.z * F.z;
            if (mu != 0.0 || f2 != 0.0) {
                double const a = mu * mu * f2, b = 2.55 * mu * S.kT, c = S.spacestep * S.spacestep;
                dt = (a * c < 1e-6 * b * b) ? c / b : (-b + sqrt(b * b + 4.0 * a * c)) / (2.0 * a);
            }
            dt = fmin(dt, S.dt);
        }
    }
    if (S.adaptive) {
        for (int o = 16; o > 0; o >>= 1) dt = fmin(dt, __shfl_xor_sync(0xffffffffu, dt, o));
        if ((threadIdx.x & 31) == 0)
            atomicMin(A.dt_bits + replica, (unsigned long long) __double_as_longlong(dt));
    }
}

/* second pass of an adaptive step: update from the stored forces */
template <int BOX>
__global__ void __launch_bounds__(STEP_BLOCK) k_integrate(StepArrays A, StepScalars S)
{
    __shared__ DParams P;
    load_params(P, A.params);
    uint32_t const replica = blockIdx.y;
    uint32_t const local = blockIdx.x * STEP_BLOCK + threadIdx.x;
    double const known = __longlong_as_double((long long) A.ctl->max_disp2_bits);
    double moved2 = 0.0;
    if (local < A.n) {
        uint32_t const s = replica * A.n + local;
        bead4 const me = load_bead(A.pos + s);
        d3 const F = {A.force[3 * (size_t) s], A.force[3 * (size_t) s + 1], A.force[3 * (size_t) s + 2]};
        double const dt = __longlong_as_double((long long) A.dt_bits[replica]);
        moved2 = integrate_bead<BOX>(A, P, S, s, replica, me, F, dt, A.dt_prev[replica]);
    }
    publish_displacement(A, moved2, known);
}
