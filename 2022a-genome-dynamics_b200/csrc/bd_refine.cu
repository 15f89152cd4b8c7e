/*
 * bd_refine.cu -- the 10Mb -> 100kb hand-off between two simulation phases.
 *
 * Replaces 5-sim-genome/src/refine (refinement.py:11-21 refine_path_spline, run.py:27-40): every coarse
 * chromosome path of m points is interpolated by the cubic spline scipy.interpolate.splprep(k=3, s=0) builds
 * on u_i = (i + 1/2)/m and evaluated at (f + 1/2)/(m * factor).  With s = 0 FITPACK places the interior knots
 * at u_2 .. u_{m-3} (fpcurf.f: t(k+2..) = x(k/2+2..)), i.e. the interpolant is THE not-a-knot cubic spline,
 * and splev extrapolates the end pieces for the half bin before u_0 and after u_{m-1}.  The not-a-knot
 * spline is unique, so it is computed here from its second derivatives M_i on the uniform grid:
 *     M_0 - 2 M_1 + M_2 = 0,   M_{i-1} + 4 M_i + M_{i+1} = 6 (y_{i-1} - 2 y_i + y_{i+1}) / h^2,   mirrored at the end,
 * one thread per (chain, coordinate) for the solve and one per fine bead for the evaluation.  Agreement with
 * scipy is at rounding level (tests/test_refine.py: golden vectors made by the reference's own module).
 * Positions stay on the device when the caller passes device pointers through bd_refine_spline_device.
 */
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/bd_cuda.h"

extern int bd_internal_fail(const char *msg);

namespace {

__global__ void k_refine_solve(const double *coarse, const long long *ranges, long long n_chains, double *M, double *scratch)
{
    long long const id = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= 3 * n_chains) return;
    long long const c = id / 3;
    int const d = (int) (id % 3);
    long long const beg = ranges[2 * c], m = ranges[2 * c + 1] - beg;
    const double *y = coarse + 3 * beg + d;   // stride 3
    double *Md = M + 3 * beg + d;             // stride 3
    double *cp = scratch + 3 * beg + d;       // Thomas algorithm workspace, stride 3
    double const h = 1.0 / (double) m, ih2 = 1.0 / (h * h);
    auto rhs = [&](long long i) { return 6.0 * (y[3 * (i - 1)] - 2.0 * y[3 * i] + y[3 * (i + 1)]) * ih2; };
    // the not-a-knot conditions turn the first and last interior equations into 6 M_1 = rhs_1, 6 M_{m-2} = rhs_{m-2}
    Md[3 * 1] = rhs(1) / 6.0;
    Md[3 * (m - 2)] = rhs(m - 2) / 6.0;
    if (m > 4) {
        // unknowns M_2 .. M_{m-3}: tridiagonal (1, 4, 1) with the two known neighbours moved to the right side
        long long const lo = 2, hi = m - 3;
        double prev_c = 0.0, prev_d = 0.0;
        for (long long i = lo; i <= hi; i++) {
            double r = rhs(i);
            if (i == lo) r -= Md[3 * 1];
            if (i == hi) r -= Md[3 * (m - 2)];
            double const denom = 4.0 - (i == lo ? 0.0 : prev_c);
            prev_c = (i == hi ? 0.0 : 1.0) / denom;
            prev_d = (r - (i == lo ? 0.0 : prev_d)) / denom;
            cp[3 * i] = prev_c;
            Md[3 * i] = prev_d;
        }
        for (long long i = hi - 1; i >= lo; i--) Md[3 * i] -= cp[3 * i] * Md[3 * (i + 1)];
    }
    Md[0] = 2.0 * Md[3 * 1] - Md[3 * 2];
    Md[3 * (m - 1)] = 2.0 * Md[3 * (m - 2)] - Md[3 * (m - 3)];
}

__global__ void k_refine_eval(const double *coarse, const double *M, const long long *ranges, const long long *fine_ranges,
                              const long long *chain_of_fine_block, long long n_chains, int factor, double *fine)
{
    // blocks are assigned to chains through a prefix table: block b serves chain chain_of_fine_block[2b] from
    // fine offset chain_of_fine_block[2b+1]
    long long const c = chain_of_fine_block[2 * blockIdx.x];
    long long const f = chain_of_fine_block[2 * blockIdx.x + 1] + threadIdx.x;
    (void) n_chains;
    long long const beg = ranges[2 * c], m = ranges[2 * c + 1] - beg;
    long long const fbeg = fine_ranges[2 * c], fcount = fine_ranges[2 * c + 1] - fbeg;
    long long const n = m * factor;
    if (f >= fcount || f >= n) return;  // run.py:34: fine_positions[beg:end] = fine_chain[:(end - beg)]
    double const h = 1.0 / (double) m;
    double const u = ((double) f + 0.5) / (double) n;
    long long i = (long long) floor((u - 0.5 * h) / h);
    i = i < 0 ? 0 : (i > m - 2 ? m - 2 : i);   // the end pieces also serve the half bins outside [u_0, u_{m-1}]
    double const t = u - ((double) i + 0.5) * h;
#pragma unroll
    for (int d = 0; d < 3; d++) {
        double const y0 = coarse[3 * (beg + i) + d], y1 = coarse[3 * (beg + i + 1) + d];
        double const m0 = M[3 * (beg + i) + d], m1 = M[3 * (beg + i + 1) + d];
        double const b = (y1 - y0) / h - h * (2.0 * m0 + m1) / 6.0;
        fine[3 * (fbeg + f) + d] = y0 + t * (b + t * (0.5 * m0 + t * ((m1 - m0) / (6.0 * h))));
    }
}

__global__ void k_refine_attach(double *fine, const long long *attach, long long n_attach)
{
    long long const a = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n_attach) return;
    long long const nor = attach[2 * a], nuc = attach[2 * a + 1];  // run.py:37-38
    for (int d = 0; d < 3; d++) fine[3 * nuc + d] = fine[3 * nor + d];
}

struct Buf {
    void *p = nullptr;
    ~Buf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
};

#define RCK(call)                                                                                      \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) return bd_internal_fail((std::string("bd_refine_spline: ") + cudaGetErrorString(e_)).c_str()); \
    } while (0)

}  // namespace

extern "C" int bd_refine_spline_device(const double *d_coarse, const int64_t *coarse_ranges, int64_t n_chains, int32_t factor,
                                       const int64_t *fine_ranges, const int64_t *attach, int64_t n_attach, int64_t n_coarse,
                                       int64_t n_fine, double *d_fine, void *cuda_stream)
{
    if (n_chains < 0 || factor < 1 || n_attach < 0 || n_coarse < 0 || n_fine < 0)
        return bd_internal_fail("bd_refine_spline: bad arguments");
    cudaStream_t const st = (cudaStream_t) cuda_stream;
    constexpr int EVAL_BLOCK = 128;
    std::vector<long long> table;  // {chain, first fine offset} per evaluation block
    for (int64_t c = 0; c < n_chains; c++) {
        int64_t const beg = coarse_ranges[2 * c], end = coarse_ranges[2 * c + 1];
        int64_t const fbeg = fine_ranges[2 * c], fend = fine_ranges[2 * c + 1];
        if (beg < 0 || end > n_coarse || end - beg < 4)
            return bd_internal_fail("bd_refine_spline: a coarse chain needs at least 4 points inside the array (splprep: m > k)");
        if (fbeg < 0 || fend > n_fine || fend < fbeg) return bd_internal_fail("bd_refine_spline: bad fine range");
        if (fend - fbeg > (end - beg) * factor)
            return bd_internal_fail("bd_refine_spline: fine chain longer than factor x coarse chain");
        for (int64_t f = 0; f < fend - fbeg; f += EVAL_BLOCK) {
            table.push_back((long long) c);
            table.push_back((long long) f);
        }
    }
    for (int64_t a = 0; a < 2 * n_attach; a++)
        if (attach[a] < 0 || attach[a] >= n_fine) return bd_internal_fail("bd_refine_spline: attach index out of range");
    if (n_chains == 0) return 0;
    Buf ranges, franges, tab, att, M, scratch;
    RCK(ranges.alloc(2 * n_chains * sizeof(long long)));
    RCK(franges.alloc(2 * n_chains * sizeof(long long)));
    RCK(tab.alloc(table.size() * sizeof(long long)));
    RCK(att.alloc(2 * n_attach * sizeof(long long)));
    RCK(M.alloc(3 * (size_t) n_coarse * sizeof(double)));
    RCK(scratch.alloc(3 * (size_t) n_coarse * sizeof(double)));
    static_assert(sizeof(long long) == sizeof(int64_t), "int64 layout");
    RCK(cudaMemcpyAsync(ranges.p, coarse_ranges, 2 * n_chains * sizeof(long long), cudaMemcpyHostToDevice, st));
    RCK(cudaMemcpyAsync(franges.p, fine_ranges, 2 * n_chains * sizeof(long long), cudaMemcpyHostToDevice, st));
    if (!table.empty()) RCK(cudaMemcpyAsync(tab.p, table.data(), table.size() * sizeof(long long), cudaMemcpyHostToDevice, st));
    if (n_attach) RCK(cudaMemcpyAsync(att.p, attach, 2 * n_attach * sizeof(long long), cudaMemcpyHostToDevice, st));
    k_refine_solve<<<(unsigned) ((3 * n_chains + 63) / 64), 64, 0, st>>>(d_coarse, (const long long *) ranges.p, n_chains,
                                                                        (double *) M.p, (double *) scratch.p);
    if (!table.empty())
        k_refine_eval<<<(unsigned) (table.size() / 2), EVAL_BLOCK, 0, st>>>(d_coarse, (const double *) M.p, (const long long *) ranges.p,
                                                                          (const long long *) franges.p, (const long long *) tab.p,
                                                                          n_chains, factor, d_fine);
    if (n_attach)
        k_refine_attach<<<(unsigned) ((n_attach + 127) / 128), 128, 0, st>>>(d_fine, (const long long *) att.p, n_attach);
    RCK(cudaGetLastError());
    RCK(cudaStreamSynchronize(st));  // the temporaries above are released on return
    return 0;
}

extern "C" int bd_refine_spline(const double *coarse_xyz, const int64_t *coarse_ranges, int64_t n_chains, int32_t factor,
                                const int64_t *fine_ranges, const int64_t *attach, int64_t n_attach, int64_t n_coarse,
                                int64_t n_fine, double *fine_xyz_out, int32_t device)
{
    if (n_coarse < 0 || n_fine < 0) return bd_internal_fail("bd_refine_spline: bad arguments");
    if (cudaSetDevice(device) != cudaSuccess) return bd_internal_fail("bd_refine_spline: no such CUDA device");
    Buf coarse, fine;
    RCK(coarse.alloc(3 * (size_t) n_coarse * sizeof(double)));
    RCK(fine.alloc(3 * (size_t) n_fine * sizeof(double)));
    RCK(cudaMemcpy(coarse.p, coarse_xyz, 3 * (size_t) n_coarse * sizeof(double), cudaMemcpyHostToDevice));
    // beads no chain covers keep what the caller put there (run.py allocates with np.empty)
    RCK(cudaMemcpy(fine.p, fine_xyz_out, 3 * (size_t) n_fine * sizeof(double), cudaMemcpyHostToDevice));
    if (bd_refine_spline_device((const double *) coarse.p, coarse_ranges, n_chains, factor, fine_ranges, attach, n_attach,
                                n_coarse, n_fine, (double *) fine.p, nullptr))
        return 1;
    RCK(cudaMemcpy(fine_xyz_out, fine.p, 3 * (size_t) n_fine * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}

// ---- bd_refine_spline_device_noalloc: the same hand-off for callers that must not allocate ----------------
//
// Who needs this.  (1) A driver that runs the 10Mb -> 100kb hand-off inside a process whose device memory has
// already been handed out in full -- the 100kb ensemble engines of the next phase are created first and sized
// to fill the card, or a framework's caching allocator (torch's, for one) owns every free byte -- so a
// cudaMalloc of the spline workspace simply fails.  (2) A driver that refines one replica while other engines
// keep stepping on the same GPU: cudaMalloc and above all cudaFree serialise with all work on the device, so
// the six temporaries of bd_refine_spline_device stall every other stream twice per call.
//
// For them the workspace lives in the library image: one file-scope __device__ array that every GPU which loads
// libbd_b200.so carries from the moment this translation unit's module is loaded until the process exits
// (BD_REFINE_NOALLOC_SCRATCH_BYTES, about 3.2 MiB).  Sizes beyond the limits in bd_cuda.h are refused; nothing
// is ever allocated or freed on the device here.

namespace {

constexpr size_t NA_MAX_COARSE = BD_REFINE_NOALLOC_MAX_COARSE;
constexpr size_t NA_MAX_CHAINS = BD_REFINE_NOALLOC_MAX_CHAINS;
constexpr size_t NA_MAX_ATTACH = BD_REFINE_NOALLOC_MAX_ATTACH;
// carve-up of the array, in bytes (every piece a multiple of 256)
constexpr size_t NA_OFF_M = 0;                                                   // second derivatives [n_coarse][3]
constexpr size_t NA_OFF_SWEEP = NA_OFF_M + 3 * NA_MAX_COARSE * sizeof(double);   // Thomas sweep coefficients [n_coarse][3]
constexpr size_t NA_OFF_RANGES = NA_OFF_SWEEP + 3 * NA_MAX_COARSE * sizeof(double);
constexpr size_t NA_OFF_FRANGES = NA_OFF_RANGES + 2 * NA_MAX_CHAINS * sizeof(long long);
constexpr size_t NA_OFF_ATTACH = NA_OFF_FRANGES + 2 * NA_MAX_CHAINS * sizeof(long long);
constexpr size_t NA_BYTES = NA_OFF_ATTACH + 2 * NA_MAX_ATTACH * sizeof(long long);
static_assert(NA_BYTES == BD_REFINE_NOALLOC_SCRATCH_BYTES, "bd_cuda.h states the size of the static workspace");
static_assert(NA_OFF_SWEEP % 256 == 0 && NA_OFF_RANGES % 256 == 0 && NA_OFF_FRANGES % 256 == 0 && NA_OFF_ATTACH % 256 == 0,
              "pieces keep the array's alignment");

// In HOST code this name is only a handle for cudaGetSymbolAddress: it is not a device address, and a kernel
// handed `g_refine_workspace` itself would fault (or scribble over host memory on a system where the GPU can
// reach it).  Host code below only ever uses the per-device address looked up in noalloc_workspace().
static __device__ __align__(256) unsigned char g_refine_workspace[NA_BYTES];

constexpr int NA_MAX_DEVICES = 64;
unsigned char *g_workspace_on_device[NA_MAX_DEVICES];  // device address of the array, per CUDA device ordinal
std::mutex g_workspace_mutex;

// Device address of the static workspace on the CURRENT device; looked up once per device (each device has its
// own copy of the array at its own address).
cudaError_t noalloc_workspace(int device, unsigned char **out)
{
    std::lock_guard<std::mutex> lock(g_workspace_mutex);
    if (!g_workspace_on_device[device]) {
        void *p = nullptr;
        cudaError_t const e = cudaGetSymbolAddress(&p, g_refine_workspace);
        if (e != cudaSuccess) return e;
        g_workspace_on_device[device] = (unsigned char *) p;
    }
    *out = g_workspace_on_device[device];
    return cudaSuccess;
}

// k_refine_eval without the host-built block table (which would need its own device buffer): block (x, y) serves
// fine points [128 x, 128 x + 128) of chain y; blocks past the end of a short chain leave at once.  Same arithmetic,
// in the same order, as k_refine_eval.
__global__ void k_refine_eval_by_chain(const double *coarse, const double *M, const long long *ranges, const long long *fine_ranges,
                                       int factor, double *fine)
{
    long long const c = blockIdx.y;
    long long const f = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    long long const beg = ranges[2 * c], m = ranges[2 * c + 1] - beg;
    long long const fbeg = fine_ranges[2 * c], fcount = fine_ranges[2 * c + 1] - fbeg;
    long long const n = m * factor;
    if (f >= fcount || f >= n) return;
    double const h = 1.0 / (double) m;
    double const u = ((double) f + 0.5) / (double) n;
    long long i = (long long) floor((u - 0.5 * h) / h);
    i = i < 0 ? 0 : (i > m - 2 ? m - 2 : i);
    double const t = u - ((double) i + 0.5) * h;
#pragma unroll
    for (int d = 0; d < 3; d++) {
        double const y0 = coarse[3 * (beg + i) + d], y1 = coarse[3 * (beg + i + 1) + d];
        double const m0 = M[3 * (beg + i) + d], m1 = M[3 * (beg + i + 1) + d];
        double const b = (y1 - y0) / h - h * (2.0 * m0 + m1) / 6.0;
        fine[3 * (fbeg + f) + d] = y0 + t * (b + t * (0.5 * m0 + t * ((m1 - m0) / (6.0 * h))));
    }
}

#define NACK(call)                                                                                     \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess)                                                                         \
            return bd_internal_fail((std::string("bd_refine_spline_device_noalloc: ") + cudaGetErrorString(e_)).c_str()); \
    } while (0)

}  // namespace

/* bd_refine_spline_device without a single device allocation (see the block comment above and bd_cuda.h).
 *
 * Set-up: the first call on each device asks the runtime where that device keeps the static workspace
 * (cudaGetSymbolAddress, with that device current) and remembers the answer; make that first call from ordinary
 * host code, before any stream capture is begun.  The device that is current when the call is made must be the
 * one d_coarse and d_fine live on.
 *
 * Sharing: there is ONE workspace per device and every call on that device uses it.  Calls issued to the same
 * stream, or one after the other from one thread (the call returns after its work has finished), are safe; two
 * calls that can run on the GPU at the same time (two host threads, two streams) overwrite each other's
 * coefficients and give wrong splines -- serialise them.  The array keeps what the previous call left in it; that
 * is harmless, because a call writes every range, second derivative and sweep coefficient before it reads it, so
 * nothing in it has to start at zero and nothing is cleared. */
extern "C" int bd_refine_spline_device_noalloc(const double *d_coarse, const int64_t *coarse_ranges, int64_t n_chains, int32_t factor,
                                               const int64_t *fine_ranges, const int64_t *attach, int64_t n_attach,
                                               int64_t n_coarse, int64_t n_fine, double *d_fine, void *cuda_stream)
{
    if (n_chains < 0 || factor < 1 || n_attach < 0 || n_coarse < 0 || n_fine < 0)
        return bd_internal_fail("bd_refine_spline_device_noalloc: bad arguments");
    if ((uint64_t) n_coarse > NA_MAX_COARSE || (uint64_t) n_chains > NA_MAX_CHAINS || (uint64_t) n_attach > NA_MAX_ATTACH)
        return bd_internal_fail(("bd_refine_spline_device_noalloc: the static workspace holds at most " +
                                 std::to_string(NA_MAX_COARSE) + " coarse points, " + std::to_string(NA_MAX_CHAINS) + " chains and " +
                                 std::to_string(NA_MAX_ATTACH) + " attachments per call (asked for " + std::to_string(n_coarse) + ", " +
                                 std::to_string(n_chains) + ", " + std::to_string(n_attach) +
                                 "); split the call by chains or use bd_refine_spline_device")
                                    .c_str());
    cudaStream_t const st = (cudaStream_t) cuda_stream;
    constexpr int EVAL_BLOCK = 128;
    int64_t longest = 0;
    for (int64_t c = 0; c < n_chains; c++) {
        int64_t const beg = coarse_ranges[2 * c], end = coarse_ranges[2 * c + 1];
        int64_t const fbeg = fine_ranges[2 * c], fend = fine_ranges[2 * c + 1];
        if (beg < 0 || end > n_coarse || end - beg < 4)
            return bd_internal_fail("bd_refine_spline_device_noalloc: a coarse chain needs at least 4 points inside the array (splprep: m > k)");
        if (fbeg < 0 || fend > n_fine || fend < fbeg) return bd_internal_fail("bd_refine_spline_device_noalloc: bad fine range");
        if (fend - fbeg > (end - beg) * factor)
            return bd_internal_fail("bd_refine_spline_device_noalloc: fine chain longer than factor x coarse chain");
        longest = fend - fbeg > longest ? fend - fbeg : longest;
    }
    for (int64_t a = 0; a < 2 * n_attach; a++)
        if (attach[a] < 0 || attach[a] >= n_fine) return bd_internal_fail("bd_refine_spline_device_noalloc: attach index out of range");
    if (n_chains == 0) return 0;
    int64_t const eval_blocks = (longest + EVAL_BLOCK - 1) / EVAL_BLOCK;
    if (eval_blocks > 0x7fffffffll) return bd_internal_fail("bd_refine_spline_device_noalloc: fine chain too long");

    int device = -1;
    NACK(cudaGetDevice(&device));
    if (device < 0 || device >= NA_MAX_DEVICES) return bd_internal_fail("bd_refine_spline_device_noalloc: device ordinal out of range");
    unsigned char *ws = nullptr;
    NACK(noalloc_workspace(device, &ws));
    double *const M = (double *) (ws + NA_OFF_M);
    double *const sweep = (double *) (ws + NA_OFF_SWEEP);
    long long *const ranges = (long long *) (ws + NA_OFF_RANGES);
    long long *const franges = (long long *) (ws + NA_OFF_FRANGES);
    long long *const att = (long long *) (ws + NA_OFF_ATTACH);

    static_assert(sizeof(long long) == sizeof(int64_t), "int64 layout");
    NACK(cudaMemcpyAsync(ranges, coarse_ranges, 2 * n_chains * sizeof(long long), cudaMemcpyHostToDevice, st));
    NACK(cudaMemcpyAsync(franges, fine_ranges, 2 * n_chains * sizeof(long long), cudaMemcpyHostToDevice, st));
    if (n_attach) NACK(cudaMemcpyAsync(att, attach, 2 * n_attach * sizeof(long long), cudaMemcpyHostToDevice, st));
    k_refine_solve<<<(unsigned) ((3 * n_chains + 63) / 64), 64, 0, st>>>(d_coarse, ranges, n_chains, M, sweep);
    if (eval_blocks)
        k_refine_eval_by_chain<<<dim3((unsigned) eval_blocks, (unsigned) n_chains), EVAL_BLOCK, 0, st>>>(d_coarse, M, ranges, franges,
                                                                                                       factor, d_fine);
    if (n_attach) k_refine_attach<<<(unsigned) ((n_attach + 127) / 128), 128, 0, st>>>(d_fine, att, n_attach);
    NACK(cudaGetLastError());
    NACK(cudaStreamSynchronize(st));  // as bd_refine_spline_device: the fine positions are complete on return
    return 0;
}
