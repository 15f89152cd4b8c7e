/*
 * md.hpp -- the micromd API of the paper's simulation drivers, bound to the B200 engine.
 *
 * The reference drivers (3-sim-1kb, 4-sim-ab, 5-sim-genome) are written against micromd's
 * header-only C++ API: `#include <md.hpp>` (submodules/github.com/snsinfu/micromd/include/md.hpp:8-54).
 * This header offers the part of that API the drivers use -- the same names, argument meaning and
 * ownership rules -- and forwards the hot path (force evaluation, the Brownian update and the
 * neighbour search) to libbd_b200.so through the C ABI of include/bd_cuda.h.  A driver is pointed
 * at this directory instead of micromd's include directory and linked with -lbd_b200; its sources
 * do not change.
 *
 * Nothing here is translated from micromd: the reference evaluates `lambda(i, j)` for every listed
 * pair on the host; here the lambdas are evaluated once per *class pair* on the host (a class is a
 * distinct row of the particle attribute columns), flattened into bd_term tables and uploaded, and
 * re-evaluated before every step so that parameters which the driver's callback changes
 * (bead_scale, wall semiaxes, ...) follow exactly as they do in the reference.
 *
 * Reference interfaces mirrored (MD/ = submodules/github.com/snsinfu/micromd/include/md/):
 *   MD/basic_types.hpp:14-24, basic_types/point.hpp:22-350, basic_types/array_view.hpp
 *   MD/system.hpp:64-295, system/attribute.hpp:8-17, system/particle.hpp:14-33
 *   MD/forcefield.hpp:26-32 and the make_*_forcefield factories of MD/forcefield/
 *   MD/potential/{softcore,softwell,harmonic,spring,semispring,cosine_bending,constant,
 *                 cutoff,scaled,sum}_potential.hpp
 *   MD/misc/box.hpp:18-81, misc/math.hpp:18-44, misc/neighbor_searcher.hpp:23-131
 *   MD/simulation/brownian_dynamics.hpp:24-73
 *
 * Driver-defined md::forcefield subclasses (3-sim-1kb's glue_forcefield / loop_forcefield) keep working:
 * they are evaluated on the host on positions fetched from the device and enter the engine's force
 * sum through bd_set_external_forces -- correct, but each step then crosses PCIe twice.
 *
 * Not offered (no driver uses them): Langevin / Newtonian integrators, Lennard-Jones family,
 * plane surfaces, composite / brute-force force fields, xy_periodic_box as a unit cell.
 *
 * The header binds the bd_ entry points of libbd_b200.so and nothing else.  (The GPU-less test of its host
 * logic links generated bd_ -> orc_ forwarders instead of the engine: tests/test_md_shim.py.)
 */
#ifndef MD_B200_MD_HPP
#define MD_B200_MD_HPP

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <functional>
#include <istream>
#include <iterator>
#include <limits>
#include <map>
#include <memory>
#include <ostream>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <typeindex>
#include <utility>
#include <vector>

#include "bd_types.h"

// ---- the C ABI this header binds ------------------------------------------------------------------

#define MD_B200_HANDLE_NAME bd_engine
#define MD_B200_ENGINE 1  // the engine-only entry points of bd_cuda.h exist (extensions below)

extern "C" {
typedef struct bd_engine bd_engine;
#define BD_FN(name) bd_##name
#define BD_HANDLE bd_engine
#include "bd_api_decl.inc"
#undef BD_HANDLE
}
#define MD_B200_CALL(name) BD_FN(name)

namespace md
{
    // ---- basic types (MD/basic_types.hpp:14-24) ---------------------------------------------------

    using index = std::size_t;
    using scalar = double;
    using step = std::int64_t;

    namespace point_detail
    {
        struct coordinates
        {
            double x = 0;
            double y = 0;
            double z = 0;

            coordinates() = default;
            coordinates(double x_init, double y_init, double z_init) : x{x_init}, y{y_init}, z{z_init} {}

            double& operator[](std::size_t idx) { return idx == 0 ? x : idx == 1 ? y : z; }
            double const& operator[](std::size_t idx) const { return idx == 0 ? x : idx == 1 ? y : z; }
        };

        template<typename Char, typename Tr>
        std::basic_ostream<Char, Tr>& operator<<(std::basic_ostream<Char, Tr>& os, coordinates const& c)
        {
            Char const delim = os.widen(' ');
            return os << c.x << delim << c.y << delim << c.z;
        }

        template<typename Char, typename Tr>
        std::basic_istream<Char, Tr>& operator>>(std::basic_istream<Char, Tr>& is, coordinates& c)
        {
            return is >> c.x >> c.y >> c.z;
        }
    }

    struct vector : point_detail::coordinates
    {
        using point_detail::coordinates::coordinates;

        vector& operator+=(vector const& o) noexcept { x += o.x; y += o.y; z += o.z; return *this; }
        vector& operator-=(vector const& o) noexcept { x -= o.x; y -= o.y; z -= o.z; return *this; }
        vector& operator*=(double m) noexcept { x *= m; y *= m; z *= m; return *this; }
        vector& operator/=(double d) { x /= d; y /= d; z /= d; return *this; }

        double dot(vector const& o) const noexcept { return x * o.x + y * o.y + z * o.z; }
        vector cross(vector const& o) const noexcept
        {
            return {y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x};
        }
        vector hadamard(vector const& o) const noexcept { return {x * o.x, y * o.y, z * o.z}; }
        double squared_norm() const noexcept { return dot(*this); }
        double norm() const noexcept { return std::sqrt(squared_norm()); }
        vector normalize() const { vector v = *this; v /= norm(); return v; }
        vector project(vector const& axis) const
        {
            vector v = axis;
            v *= dot(axis) / axis.squared_norm();
            return v;
        }
    };

    inline vector operator+(vector const& v) noexcept { return v; }
    inline vector operator-(vector const& v) noexcept { return {-v.x, -v.y, -v.z}; }
    inline vector operator+(vector const& a, vector const& b) noexcept { return vector(a) += b; }
    inline vector operator-(vector const& a, vector const& b) noexcept { return vector(a) -= b; }
    inline vector operator*(vector const& v, double m) noexcept { return vector(v) *= m; }
    inline vector operator*(double m, vector const& v) noexcept { return vector(v) *= m; }
    inline vector operator/(vector const& v, double d) { return vector(v) /= d; }
    inline double dot(vector const& a, vector const& b) noexcept { return a.dot(b); }
    inline vector cross(vector const& a, vector const& b) noexcept { return a.cross(b); }
    inline vector hadamard(vector const& a, vector const& b) noexcept { return a.hadamard(b); }
    inline double norm(vector const& v) noexcept { return v.norm(); }
    inline double squared_norm(vector const& v) noexcept { return v.squared_norm(); }
    inline vector normalize(vector const& v) noexcept { return v.normalize(); }
    inline vector project(vector const& v, vector const& axis) noexcept { return v.project(axis); }

    struct point : point_detail::coordinates
    {
        using point_detail::coordinates::coordinates;
        using vector_t = struct vector;

        vector_t vector() const noexcept { return {x, y, z}; }
        point& operator+=(vector_t const& d) noexcept { x += d.x; y += d.y; z += d.z; return *this; }
        point& operator-=(vector_t const& d) noexcept { x -= d.x; y -= d.y; z -= d.z; return *this; }
        double squared_distance(point const& o) const noexcept
        {
            return vector_t{x - o.x, y - o.y, z - o.z}.squared_norm();
        }
        double distance(point const& o) const noexcept { return std::sqrt(squared_distance(o)); }
    };

    inline point operator+(point const& p, vector const& d) noexcept { return point(p) += d; }
    inline point operator-(point const& p, vector const& d) noexcept { return point(p) -= d; }
    inline vector operator-(point const& a, point const& b) noexcept { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
    inline double distance(point const& a, point const& b) noexcept { return a.distance(b); }
    inline double squared_distance(point const& a, point const& b) noexcept { return a.squared_distance(b); }

    static_assert(sizeof(point) == 3 * sizeof(double), "md::point is three contiguous doubles (bd_set_positions)");

    // span over a contiguous column (MD/basic_types/array_view.hpp)
    template<typename T>
    class array_view
    {
    public:
        using value_type = T;
        using pointer = T*;
        using reference = T&;
        using iterator = T*;
        using size_type = std::size_t;

        array_view() = default;
        array_view(T* data, std::size_t size) : data_{data}, size_{size} {}

        template<typename U, typename = typename std::enable_if<std::is_convertible<U (*)[], T (*)[]>::value>::type>
        array_view(array_view<U> const& other) : data_{other.data()}, size_{other.size()} {}

        template<typename C, typename P = decltype(std::declval<C&>().data()),
                 typename = typename std::enable_if<std::is_convertible<typename std::remove_pointer<P>::type (*)[], T (*)[]>::value>::type>
        array_view(C& container) : data_{container.data()}, size_{container.size()} {}

        bool empty() const noexcept { return size_ == 0; }
        std::size_t size() const noexcept { return size_; }
        T* data() const noexcept { return data_; }
        T& front() const { return data_[0]; }
        T& back() const { return data_[size_ - 1]; }
        T* begin() const noexcept { return data_; }
        T* end() const noexcept { return data_ + size_; }
        T& operator[](std::size_t i) const { return data_[i]; }
        T& at(std::size_t i) const
        {
            if (i >= size_) throw std::out_of_range("md::array_view: index out of range");
            return data_[i];
        }
        array_view subview(std::size_t offset, std::size_t count) const { return {data_ + offset, count}; }
        array_view<T const> as_const() const noexcept { return {data_, size_}; }

    private:
        T* data_ = nullptr;
        std::size_t size_ = 0;
    };

    // md::random_engine: sfc64 (Chris Doty-Humphrey's small fast counting generator, 64-bit variant;
    // MD/basic_types/sfc.hpp:236-392): drivers seed their own initial conditions with it.
    class sfc64
    {
    public:
        using result_type = std::uint64_t;
        static constexpr result_type min() { return 0; }
        static constexpr result_type max() { return std::numeric_limits<result_type>::max(); }

        sfc64() : sfc64(0) {}
        explicit sfc64(result_type seed) { this->seed(seed); }

        void seed(result_type s)
        {
            a_ = b_ = c_ = s;
            counter_ = 1;
            for (int i = 0; i < 12; i++) (*this)();
        }

        result_type operator()()
        {
            result_type const out = a_ + b_ + counter_++;
            a_ = b_ ^ (b_ >> 11);
            b_ = c_ + (c_ << 3);
            c_ = ((c_ << 24) | (c_ >> 40)) + out;
            return out;
        }

        void discard(unsigned long long n) { while (n--) (*this)(); }
        bool operator==(sfc64 const& o) const { return a_ == o.a_ && b_ == o.b_ && c_ == o.c_ && counter_ == o.counter_; }
        bool operator!=(sfc64 const& o) const { return !(*this == o); }

    private:
        result_type a_, b_, c_, counter_;
    };
    using random_engine = sfc64;

    // ---- misc/math.hpp:18-44 ----------------------------------------------------------------------

    template<int N, typename T>
    constexpr T power(T x)
    {
        static_assert(N >= 0, "md::power: non-negative exponent");
        if (N == 0) return T(1);
        T result = x;
        for (int k = N - 1; k > 0; k >>= 1) {
            if (k & 1) result *= x;
            x *= x;
        }
        return result;
    }

    namespace b200
    {
        // x^(n/2) from x^2
        inline double power_half(int n, double x2)
        {
            auto ipow = [](int m, double x) {
                if (m == 0) return 1.0;
                double r = x;
                for (int k = m - 1; k > 0; k >>= 1) {
                    if (k & 1) r *= x;
                    x *= x;
                }
                return r;
            };
            return (n & 1) ? ipow(n, std::sqrt(x2)) : ipow(n >> 1, x2);
        }

        inline void check(int rc, char const* what)
        {
            if (rc != 0) {
                char const* msg = MD_B200_CALL(last_error)();
                throw std::runtime_error(std::string("md (B200 engine): ") + what + ": " + (msg ? msg : "error"));
            }
        }

        using term_list = std::vector<bd_term>;

        inline bd_term make_term(int kind, int p, int q, double factor, double energy, double length, double cutoff)
        {
            bd_term t;
            std::memset(&t, 0, sizeof t);
            t.kind = kind; t.p = p; t.q = q;
            t.factor = factor; t.energy = energy; t.length = length; t.cutoff = cutoff;
            return t;
        }

        inline bool same_terms(term_list const& a, term_list const& b)
        {
            return a.size() == b.size() && (a.empty() || std::memcmp(a.data(), b.data(), a.size() * sizeof(bd_term)) == 0);
        }

        // host evaluation of a flattened term (custom md::forcefield subclasses call
        // potential.evaluate_energy / evaluate_force on the host)
        inline double term_energy(bd_term const& t, double r2)
        {
            if (t.cutoff != 0 && r2 >= t.cutoff * t.cutoff) return 0;
            double const k2 = t.length != 0 ? 1 / (t.length * t.length) : 0;
            double e = 0;
            switch (t.kind) {
            case BD_TERM_SOFTCORE: {
                double const g = 1 - power_half(t.p, k2 * r2);
                if (g >= 0) { e = t.energy; for (int i = 0; i < t.q; i++) e *= g; }
                break;
            }
            case BD_TERM_SOFTWELL: e = -t.energy / (1 + power_half(t.p, k2 * r2)); break;
            case BD_TERM_HARMONIC: e = 0.5 * t.energy * r2; break;
            case BD_TERM_SPRING: { double const u = std::sqrt(r2) - t.length; e = 0.5 * t.energy * u * u; break; }
            case BD_TERM_SEMISPRING: { double const u = std::sqrt(r2) - t.length; if (u > 0) e = 0.5 * t.energy * u * u; break; }
            default: break;
            }
            return t.factor * e;
        }

        inline double term_force_scalar(bd_term const& t, double r2)
        {
            if (t.cutoff != 0 && r2 >= t.cutoff * t.cutoff) return 0;
            double const k2 = t.length != 0 ? 1 / (t.length * t.length) : 0;
            double s = 0;
            switch (t.kind) {
            case BD_TERM_SOFTCORE: {
                double const u2 = k2 * r2, v = power_half(t.p - 2, u2), g = 1 - v * u2;
                if (g >= 0) { s = double(t.p * t.q) * t.energy * k2 * v; for (int i = 0; i < t.q - 1; i++) s *= g; }
                break;
            }
            case BD_TERM_SOFTWELL: {
                double const u2 = k2 * r2, v = power_half(t.p - 2, u2), w = 1 + v * u2;
                s = double(-t.p) * t.energy * k2 * v / (w * w);
                break;
            }
            case BD_TERM_HARMONIC: s = -t.energy; break;
            case BD_TERM_SPRING: { double const r1 = std::sqrt(r2); if (r1 != 0) s = t.energy * t.length / r1 - t.energy; break; }
            case BD_TERM_SEMISPRING: { double const r1 = std::sqrt(r2); if (r1 > t.length) s = t.energy * t.length / r1 - t.energy; break; }
            default: break;
            }
            return t.factor * s;
        }

        // Every potential of this header is described by its flattened terms.  The potentials stay plain
        // aggregates with the reference's member order (drivers initialise them positionally,
        // 3-sim-1kb/src/simulation/forces/loop_forcefield.cpp:13), so the host evaluation is injected as two
        // member functions that forward to these helpers instead of coming from a base class.
        template<typename Pot>
        md::scalar flat_energy(Pot const& pot, md::vector r)
        {
            term_list terms;
            pot.b200_flatten(terms, 1.0, 0.0);
            double const r2 = r.squared_norm();
            double e = 0;
            for (auto const& t : terms) e += term_energy(t, r2);
            return e;
        }

        template<typename Pot>
        md::vector flat_force(Pot const& pot, md::vector r)
        {
            term_list terms;
            pot.b200_flatten(terms, 1.0, 0.0);
            double const r2 = r.squared_norm();
            double s = 0;
            for (auto const& t : terms) s += term_force_scalar(t, r2);
            return s * r;
        }

#define MD_B200_POTENTIAL_MEMBERS                                                                      \
        md::scalar evaluate_energy(md::vector r) const { return b200::flat_energy(*this, r); }         \
        md::vector evaluate_force(md::vector r) const { return b200::flat_force(*this, r); }

        template<typename P, typename = void>
        struct is_potential : std::false_type {};
        template<typename P>
        struct is_potential<P, decltype(void(std::declval<P const&>().b200_flatten(std::declval<term_list&>(), 1.0, 0.0)))>
            : std::true_type {};
    }

    // ---- potentials (MD/potential/*.hpp) ----------------------------------------------------------

    struct constant_potential
    {
        MD_B200_POTENTIAL_MEMBERS
        md::scalar energy = 0;
        constant_potential() = default;
        explicit constant_potential(md::scalar e) : energy{e} {}
        void b200_flatten(b200::term_list& out, double, double) const
        {
            out.push_back(b200::make_term(BD_TERM_NONE, 0, 0, 0, 0, 0, 0));
        }
    };

    // u(r) = energy (1 - (r/diameter)^P)^Q inside the diameter (MD/potential/softcore_potential.hpp:23-62)
    template<int P, int Q = 3>
    struct softcore_potential
    {
        MD_B200_POTENTIAL_MEMBERS
        md::scalar energy = 1;
        md::scalar diameter = 1;
        void b200_flatten(b200::term_list& out, double factor, double cutoff) const
        {
            out.push_back(b200::make_term(BD_TERM_SOFTCORE, P, Q, factor, energy, diameter, cutoff));
        }
    };

    // u(r) = -energy / (1 + (r/decay_distance)^P) (MD/potential/softwell_potential.hpp:26-50)
    template<int P>
    struct softwell_potential
    {
        MD_B200_POTENTIAL_MEMBERS
        md::scalar energy = 1;
        md::scalar decay_distance = 1;
        void b200_flatten(b200::term_list& out, double factor, double cutoff) const
        {
            out.push_back(b200::make_term(BD_TERM_SOFTWELL, P, 0, factor, energy, decay_distance, cutoff));
        }
    };

    // u(r) = spring_constant r^2 / 2 (MD/potential/harmonic_potential.hpp:20-35)
    struct harmonic_potential
    {
        MD_B200_POTENTIAL_MEMBERS
        md::scalar spring_constant = 1;
        void b200_flatten(b200::term_list& out, double factor, double cutoff) const
        {
            out.push_back(b200::make_term(BD_TERM_HARMONIC, 0, 0, factor, spring_constant, 0, cutoff));
        }
    };

    // u(r) = spring_constant (r - equilibrium_distance)^2 / 2 (MD/potential/spring_potential.hpp:21-41)
    struct spring_potential
    {
        MD_B200_POTENTIAL_MEMBERS
        md::scalar spring_constant = 1;
        md::scalar equilibrium_distance = 0;
        void b200_flatten(b200::term_list& out, double factor, double cutoff) const
        {
            out.push_back(b200::make_term(BD_TERM_SPRING, 0, 0, factor, spring_constant, equilibrium_distance, cutoff));
        }
    };

    // spring that only pulls (MD/potential/semispring_potential.hpp:23-46)
    struct semispring_potential
    {
        MD_B200_POTENTIAL_MEMBERS
        md::scalar spring_constant = 1;
        md::scalar equilibrium_distance = 0;
        void b200_flatten(b200::term_list& out, double factor, double cutoff) const
        {
            out.push_back(b200::make_term(BD_TERM_SEMISPRING, 0, 0, factor, spring_constant, equilibrium_distance, cutoff));
        }
    };

    // u(i,j,k) = bending_energy (1 - cos theta) (MD/potential/cosine_bending_potential.hpp:23-62)
    struct cosine_bending_potential
    {
        md::scalar bending_energy = 1;

        md::scalar evaluate_energy(md::vector rij, md::vector rjk) const
        {
            double const dd = std::sqrt(rij.squared_norm() * rjk.squared_norm());
            return dd == 0 ? 0 : bending_energy - bending_energy * (rij.dot(rjk) / dd);
        }
    };

    template<typename Pot>
    struct scaled_potential
    {
        MD_B200_POTENTIAL_MEMBERS
        Pot potential;
        md::scalar factor;
        scaled_potential(Pot const& p, md::scalar f) : potential{p}, factor{f} {}
        void b200_flatten(b200::term_list& out, double outer, double cutoff) const
        {
            potential.b200_flatten(out, outer * factor, cutoff);
        }
    };

    template<typename P1, typename P2>
    struct sum_potential
    {
        MD_B200_POTENTIAL_MEMBERS
        P1 first;
        P2 second;
        sum_potential(P1 const& a, P2 const& b) : first{a}, second{b} {}
        void b200_flatten(b200::term_list& out, double factor, double cutoff) const
        {
            first.b200_flatten(out, factor, cutoff);
            second.b200_flatten(out, factor, cutoff);
        }
    };

    template<typename Pot>
    struct cutoff_potential
    {
        MD_B200_POTENTIAL_MEMBERS
        Pot potential;
        md::scalar cutoff_distance;
        cutoff_potential(Pot const& p, md::scalar c) : potential{p}, cutoff_distance{c} {}
        void b200_flatten(b200::term_list& out, double factor, double outer_cutoff) const
        {
            double const c = outer_cutoff != 0 ? std::min(outer_cutoff, cutoff_distance) : cutoff_distance;
            potential.b200_flatten(out, factor, c);
        }
    };

    template<typename Pot, typename = typename std::enable_if<b200::is_potential<Pot>::value>::type>
    scaled_potential<Pot> operator*(md::scalar f, Pot const& p) { return {p, f}; }
    template<typename Pot, typename = typename std::enable_if<b200::is_potential<Pot>::value>::type>
    scaled_potential<Pot> operator*(Pot const& p, md::scalar f) { return {p, f}; }
    template<typename P1, typename P2,
             typename = typename std::enable_if<b200::is_potential<P1>::value && b200::is_potential<P2>::value>::type>
    sum_potential<P1, P2> operator+(P1 const& a, P2 const& b) { return {a, b}; }
    template<typename Pot>
    cutoff_potential<Pot> apply_cutoff(Pot const& p, md::scalar cutoff) { return {p, cutoff}; }

    // ---- boxes and surfaces (MD/misc/box.hpp:18-81; sphere/ellipsoid structs of MD/forcefield/) ---

    namespace b200
    {
        inline double round_mod(double x, double m) { return x - std::nearbyint(x * (1 / m)) * m; }
    }

    struct open_box
    {
        md::index particle_count = 1000;
        md::vector shortest_displacement(md::point p1, md::point p2) const { return p1 - p2; }
    };

    struct periodic_box
    {
        md::scalar x_period = 1;
        md::scalar y_period = 1;
        md::scalar z_period = 1;
        md::vector shortest_displacement(md::point p1, md::point p2) const
        {
            return {b200::round_mod(p1.x - p2.x, x_period), b200::round_mod(p1.y - p2.y, y_period),
                    b200::round_mod(p1.z - p2.z, z_period)};
        }
    };

    struct sphere
    {
        md::point center;
        md::scalar radius = 1;
        md::scalar implicit(md::point pt) const { return (pt - center).squared_norm() / (radius * radius) - 1; }
    };

    struct ellipsoid
    {
        md::point center;
        md::scalar semiaxis_x = 1;
        md::scalar semiaxis_y = 1;
        md::scalar semiaxis_z = 1;
        md::scalar implicit(md::point pt) const
        {
            md::vector const r = pt - center;
            double const u = r.x / semiaxis_x, v = r.y / semiaxis_y, w = r.z / semiaxis_z;
            return u * u + v * v + w * w - 1;
        }
    };

    namespace b200
    {
        inline int box_kind(md::open_box const&) { return BD_BOX_OPEN; }
        inline int box_kind(md::periodic_box const&) { return BD_BOX_PERIODIC; }
        inline void box_period(md::open_box const&, double p[3]) { p[0] = p[1] = p[2] = 0; }
        inline void box_period(md::periodic_box const& b, double p[3]) { p[0] = b.x_period; p[1] = b.y_period; p[2] = b.z_period; }
        inline md::index box_count(md::open_box const& b) { return b.particle_count; }
        inline md::index box_count(md::periodic_box const&) { return 0; }
    }

    // ---- attributes (MD/system/attribute.hpp:8-17) ------------------------------------------------

    template<typename T, typename Tag = T>
    using attribute_key = T (*)(Tag*);

    template<typename T, typename Tag>
    constexpr T default_value(md::attribute_key<T, Tag> key) { return key ? key(nullptr) : T{}; }

    inline constexpr md::scalar mass_attribute(struct tag_mass_attribute*) { return 1; }
    inline constexpr md::scalar friction_attribute(struct tag_friction_attribute*) { return 0; }
    inline constexpr md::scalar mobility_attribute(struct tag_mobility_attribute*) { return 1; }
    inline md::point position_attribute(struct tag_position_attribute*) { return {}; }
    inline md::vector velocity_attribute(struct tag_velocity_attribute*) { return {}; }

    struct basic_particle_data
    {
        md::scalar mass = 1;
        md::scalar mobility = 1;
        md::scalar friction = 0;
        md::point position = {};
        md::vector velocity = {};
    };

    class system;

    struct particle_ref
    {
        md::system& system;
        md::index index;
        md::scalar& mass;
        md::scalar& friction;
        md::scalar& mobility;
        md::point& position;
        md::vector& velocity;

        particle_ref(md::system& system, md::index idx);
        template<typename T, typename Tag>
        T& view(md::attribute_key<T, Tag> key);
    };

    class particle_iterator
    {
    public:
        using value_type = md::particle_ref;
        using reference = md::particle_ref;
        using pointer = void;
        using difference_type = std::ptrdiff_t;
        using iterator_category = std::forward_iterator_tag;

        particle_iterator() = default;
        particle_iterator(md::system& system, md::index idx) : system_{&system}, index_{idx} {}
        bool operator==(particle_iterator const& o) const { return index_ == o.index_; }
        bool operator!=(particle_iterator const& o) const { return index_ != o.index_; }
        md::particle_ref operator*() const { return md::particle_ref(*system_, index_); }
        particle_iterator& operator++() { index_++; return *this; }
        particle_iterator operator++(int) { auto c = *this; index_++; return c; }

    private:
        md::system* system_ = nullptr;
        md::index index_ = 0;
    };

    // ---- force fields ------------------------------------------------------------------------------

    // MD/forcefield.hpp:26-32.  compute_force ADDS to `forces`.
    class forcefield
    {
    public:
        virtual ~forcefield() = default;
        virtual md::scalar compute_energy(md::system const& system) = 0;
        virtual void compute_force(md::system const& system, md::array_view<md::vector> forces) = 0;
    };

    namespace b200
    {
        class binding;

        // what the binding knows about the particles while it talks to a force field
        struct class_map
        {
            std::vector<std::uint8_t> cls;   // class of every particle
            std::vector<md::index> rep;      // one representative particle per class
            int ntypes() const { return int(rep.size()); }
        };

        // a force field the engine evaluates on the device
        class device_forcefield : public virtual md::forcefield
        {
        public:
            // anything that cannot be pushed as a parameter update (pair lists, target sets, term
            // counts) bumps this; the binding then re-creates the engine
            std::uint64_t b200_structure() const { return structure_; }

            virtual int b200_class() const = 0;
            virtual void b200_attach(binding& b, md::system const& sys) = 0;   // bd_add_*_ff
            virtual bool b200_refresh(binding& b, md::system const& sys) = 0;  // parameters -> engine; false = re-create
            virtual void b200_collect(binding&) {}                              // statistics <- engine

            md::scalar compute_energy(md::system const& system) override;
            void compute_force(md::system const& system, md::array_view<md::vector> forces) override;

            int b200_id = -1;

        protected:
            void touch_structure() { structure_++; }

        private:
            std::uint64_t structure_ = 1;
        };

        // type-erased attribute column
        struct column_base
        {
            virtual ~column_base() = default;
            virtual void resize(std::size_t n) = 0;
            virtual std::size_t elem_size() const = 0;
            virtual bool hashable() const = 0;
            virtual void const* bytes() const = 0;
        };

        template<typename T>
        struct column : column_base
        {
            std::vector<T> values;
            T fill;
            explicit column(T const& f) : fill{f} {}
            void resize(std::size_t n) override { values.resize(n, fill); }
            std::size_t elem_size() const override { return sizeof(T); }
            bool hashable() const override { return std::is_trivially_copyable<T>::value; }
            void const* bytes() const override { return values.data(); }
        };
    }

    // MD/system.hpp:64-295
    class system
    {
    public:
        system();
        ~system();
        system(system const&) = delete;
        system& operator=(system const&) = delete;

        md::particle_ref add_particle(md::basic_particle_data data = {});
        md::index particle_count() const { return count_; }

        struct particle_range
        {
            md::particle_iterator first, last;
            md::particle_iterator begin() const { return first; }
            md::particle_iterator end() const { return last; }
        };
        particle_range particles() { return {md::particle_iterator{*this, 0}, md::particle_iterator{*this, count_}}; }

        template<typename T, typename Tag>
        void add_attribute(md::attribute_key<T, Tag> key)
        {
            std::type_index const id{typeid(md::attribute_key<T, Tag>)};
            if (columns_.find(id) == columns_.end()) {
                auto col = std::make_unique<b200::column<T>>(md::default_value(key));
                col->resize(count_);
                columns_.emplace(id, std::move(col));
                column_order_.push_back(id);
            }
        }
        template<typename T, typename Tag>
        void require(md::attribute_key<T, Tag> key) { add_attribute(key); }

        template<typename T, typename Tag>
        md::array_view<T> view(md::attribute_key<T, Tag>)
        {
            auto& v = column_of<T, Tag>().values;
            return {v.data(), v.size()};
        }
        template<typename T, typename Tag>
        md::array_view<T const> view(md::attribute_key<T, Tag>) const
        {
            auto const& v = column_of<T, Tag>().values;
            return {v.data(), v.size()};
        }

        md::array_view<md::scalar> view_masses() noexcept { return view(md::mass_attribute); }
        md::array_view<md::scalar const> view_masses() const noexcept { return view(md::mass_attribute); }
        md::array_view<md::scalar> view_frictions() noexcept { return view(md::friction_attribute); }
        md::array_view<md::scalar const> view_frictions() const noexcept { return view(md::friction_attribute); }
        md::array_view<md::scalar> view_mobilities() noexcept { return view(md::mobility_attribute); }
        md::array_view<md::scalar const> view_mobilities() const noexcept { return view(md::mobility_attribute); }
        md::array_view<md::vector> view_velocities() noexcept { return view(md::velocity_attribute); }
        md::array_view<md::vector const> view_velocities() const noexcept { return view(md::velocity_attribute); }

        // positions live on the device while a simulation runs: a view fetches them first, and a
        // mutable view makes the host copy the master again
        md::array_view<md::point> view_positions();
        md::array_view<md::point const> view_positions() const;

        void add_forcefield(std::shared_ptr<md::forcefield> ff) { forcefields_.push_back(std::move(ff)); }

        template<typename FF, typename = typename std::enable_if<std::is_base_of<md::forcefield, FF>::value>::type>
        std::shared_ptr<FF> add_forcefield(FF const& ff)
        {
            auto ptr = std::make_shared<FF>(ff);
            add_forcefield(std::shared_ptr<md::forcefield>(ptr));
            return ptr;
        }

        void remove_forcefield(std::shared_ptr<md::forcefield> ff)
        {
            forcefields_.erase(std::remove(forcefields_.begin(), forcefields_.end(), ff), forcefields_.end());
        }

        md::scalar compute_kinetic_energy() const
        {
            auto const m = view_masses();
            auto const v = view_velocities();
            md::scalar sum = 0;
            for (md::index i = 0; i < count_; i++) sum += 0.5 * m[i] * v[i].squared_norm();
            return sum;
        }
        md::scalar compute_potential_energy();
        md::scalar compute_energy() { return compute_kinetic_energy() + compute_potential_energy(); }
        void compute_force(md::array_view<md::vector> forces);

        // -- used by the binding and by md::simulate_brownian_dynamics
        b200::binding& b200_binding() const;
        std::vector<std::shared_ptr<md::forcefield>> const& b200_forcefields() const { return forcefields_; }
        std::vector<std::type_index> const& b200_column_order() const { return column_order_; }
        b200::column_base const& b200_column(std::type_index id) const { return *columns_.at(id); }
        md::point* b200_host_positions() const
        {
            return const_cast<md::point*>(column_of<md::point, tag_position_attribute>().values.data());
        }

    private:
        template<typename T, typename Tag>
        b200::column<T>& column_of()
        {
            auto it = columns_.find(std::type_index{typeid(md::attribute_key<T, Tag>)});
            if (it == columns_.end()) throw std::logic_error("md::system: attribute was not added");
            return static_cast<b200::column<T>&>(*it->second);
        }
        template<typename T, typename Tag>
        b200::column<T> const& column_of() const
        {
            auto it = columns_.find(std::type_index{typeid(md::attribute_key<T, Tag>)});
            if (it == columns_.end()) throw std::logic_error("md::system: attribute was not added");
            return static_cast<b200::column<T> const&>(*it->second);
        }

        md::index count_ = 0;
        std::map<std::type_index, std::unique_ptr<b200::column_base>> columns_;
        std::vector<std::type_index> column_order_;
        std::vector<std::shared_ptr<md::forcefield>> forcefields_;
        mutable std::unique_ptr<b200::binding> binding_;
    };

    namespace b200
    {
        // One engine handle per md::system.  The host columns are the master copy of everything
        // except positions during a simulation.
        class binding
        {
        public:
            explicit binding(md::system const& sys) : sys_{sys} {}
            ~binding() { release(); }
            binding(binding const&) = delete;
            binding& operator=(binding const&) = delete;

            MD_B200_HANDLE_NAME* handle() const { return h_; }
            class_map const& classes() const { return classes_; }

            // make the engine describe the system as it is now: classes, force fields, parameters,
            // positions.  `full` recomputes the particle classes (start of a run, direct queries);
            // the step loop passes false and only re-evaluates the potentials.
            void sync(bool full)
            {
                auto const& ffs = sys_.b200_forcefields();
                bool rebuild = h_ == nullptr || n_ != sys_.particle_count();
                if (full || rebuild) {
                    class_map fresh = compute_classes();
                    if (fresh.cls != classes_.cls || fresh.rep != classes_.rep) rebuild = true;
                    if (rebuild) classes_ = std::move(fresh);
                }
                if (!rebuild) {
                    if (ffs.size() != attached_.size()) rebuild = true;
                    for (std::size_t k = 0; !rebuild && k < ffs.size(); k++) {
                        auto* dev = dynamic_cast<device_forcefield*>(ffs[k].get());
                        rebuild = attached_[k].ff != ffs[k].get() || (dev && dev->b200_structure() != attached_[k].structure);
                    }
                }
                if (!rebuild) {
                    // a force field whose change cannot be pushed as parameters asks for a new engine
                    for (auto const& a : attached_)
                        if (a.dev && !a.dev->b200_refresh(*this, sys_)) rebuild = true;
                }
                if (rebuild) build();
                if (!device_valid_) {
                    check(MD_B200_CALL(set_positions)(h_, 0, &sys_.b200_host_positions()->x), "set_positions");
                    device_valid_ = true;
                }
                host_forces();
            }

            bool has_host_forcefields() const
            {
                for (auto const& a : attached_)
                    if (!a.dev) return true;
                return false;
            }

            // energy of the driver-defined force fields, evaluated on the host
            md::scalar host_energy()
            {
                if (!has_host_forcefields()) return 0;
                fetch_positions();
                md::scalar e = 0;
                for (auto const& a : attached_)
                    if (!a.dev) e += a.ff->compute_energy(sys_);
                return e;
            }

            void fetch_positions()
            {
                if (!host_valid_ && h_) {
                    check(MD_B200_CALL(get_positions)(h_, 0, &sys_.b200_host_positions()->x), "get_positions");
                }
                host_valid_ = true;
            }
            void host_will_write() { fetch_positions(); device_valid_ = false; }
            void device_advanced() { host_valid_ = false; }

            void collect()
            {
                for (auto const& a : attached_)
                    if (a.dev) a.dev->b200_collect(*this);
            }

            void set_box(int kind, double const period[3])
            {
                if (box_set_ && (kind != box_kind_ || std::memcmp(period, box_period_, sizeof box_period_) != 0))
                    throw std::runtime_error("md (B200 engine): neighbour force fields of one system must share their unit cell");
                if (!box_set_ || kind != box_kind_ || std::memcmp(period, box_period_, sizeof box_period_) != 0) {
                    check(MD_B200_CALL(set_box)(h_, kind, period), "set_box");
                    box_set_ = true;
                    box_kind_ = kind;
                    std::memcpy(box_period_, period, sizeof box_period_);
                }
            }
            void update_box(int kind, double const period[3])
            {
                if (kind != box_kind_ || std::memcmp(period, box_period_, sizeof box_period_) != 0) {
                    check(MD_B200_CALL(set_box)(h_, kind, period), "set_box");
                    box_kind_ = kind;
                    std::memcpy(box_period_, period, sizeof box_period_);
                }
            }

            int device_ordinal() const { return device_; }

        private:
            // driver-defined force fields: evaluate them on the current positions and hand the sum to
            // the engine, which adds it to every whole-system force evaluation
            void host_forces()
            {
                if (!has_host_forcefields()) {
                    if (external_set_) check(MD_B200_CALL(set_external_forces)(h_, 0, nullptr), "set_external_forces");
                    external_set_ = false;
                    return;
                }
                fetch_positions();
                external_.assign(sys_.particle_count(), md::vector{});
                for (auto const& a : attached_)
                    if (!a.dev) a.ff->compute_force(sys_, md::array_view<md::vector>(external_.data(), external_.size()));
                check(MD_B200_CALL(set_external_forces)(h_, 0, &external_.data()->x), "set_external_forces");
                external_set_ = true;
            }

            struct attachment
            {
                md::forcefield* ff;
                device_forcefield* dev;
                std::uint64_t structure;
            };

            void release()
            {
                if (h_) MD_B200_CALL(destroy)(h_);
                h_ = nullptr;
            }

            // a class = a distinct row of (mobility, every trivially copyable user attribute column)
            class_map compute_classes() const
            {
                std::type_index const skip[] = {
                    typeid(md::attribute_key<md::scalar, tag_mass_attribute>),
                    typeid(md::attribute_key<md::scalar, tag_friction_attribute>),
                    typeid(md::attribute_key<md::point, tag_position_attribute>),
                    typeid(md::attribute_key<md::vector, tag_velocity_attribute>),
                };
                std::vector<column_base const*> cols;
                for (auto const& id : sys_.b200_column_order()) {
                    if (std::find(std::begin(skip), std::end(skip), id) != std::end(skip)) continue;
                    auto const& col = sys_.b200_column(id);
                    if (col.hashable()) cols.push_back(&col);
                }
                md::index const n = sys_.particle_count();
                class_map out;
                out.cls.resize(n);
                std::map<std::string, int> seen;
                std::string key;
                for (md::index i = 0; i < n; i++) {
                    key.clear();
                    for (auto const* c : cols)
                        key.append(static_cast<char const*>(c->bytes()) + i * c->elem_size(), c->elem_size());
                    auto it = seen.find(key);
                    if (it == seen.end()) {
                        if (int(seen.size()) >= BD_MAX_TYPES)
                            throw std::runtime_error("md (B200 engine): more than " + std::to_string(BD_MAX_TYPES)
                                                     + " distinct particle attribute rows");
                        it = seen.emplace(key, int(seen.size())).first;
                        out.rep.push_back(i);
                    }
                    out.cls[i] = std::uint8_t(it->second);
                }
                return out;
            }

            void build()
            {
                // the device may hold the only current copy of the positions
                if (!host_valid_ && h_) fetch_positions();
                release();
                n_ = sys_.particle_count();
                if (n_ == 0) throw std::runtime_error("md (B200 engine): the system has no particles");
                check(MD_B200_CALL(create)(std::int64_t(n_), 1, device_, &h_), "create");
                box_set_ = false;
                std::vector<double> mobility(classes_.rep.size());
                auto const mu = sys_.view_mobilities();
                for (std::size_t c = 0; c < classes_.rep.size(); c++) mobility[c] = mu[classes_.rep[c]];
                check(MD_B200_CALL(set_types)(h_, classes_.cls.data(), classes_.ntypes(), mobility.data()), "set_types");
                attached_.clear();
                for (auto const& ff : sys_.b200_forcefields()) {
                    auto* dev = dynamic_cast<device_forcefield*>(ff.get());
                    if (!dev) {
                        // a driver-defined md::forcefield subclass: evaluated by the driver on the host,
                        // its force enters the engine through set_external_forces (host_forces())
                        attached_.push_back({ff.get(), nullptr, 0});
                        continue;
                    }
                    dev->b200_id = -1;
                    dev->b200_attach(*this, sys_);
                    attached_.push_back({ff.get(), dev, dev->b200_structure()});
                }
                device_valid_ = false;
                host_valid_ = true;
            }

            md::system const& sys_;
            MD_B200_HANDLE_NAME* h_ = nullptr;
            md::index n_ = 0;
            int device_ = 0;
            class_map classes_;
            std::vector<attachment> attached_;
            std::vector<md::vector> external_;
            bool external_set_ = false;
            bool host_valid_ = true;
            bool device_valid_ = false;
            bool box_set_ = false;
            int box_kind_ = BD_BOX_OPEN;
            double box_period_[3] = {0, 0, 0};
        };

        // -- potential factories (MD/forcefield/detail/{pair,field,triple}_potfun.hpp): a potential
        //    object, lambda(indices...) or lambda(system, indices...)

        template<typename P>
        term_list flatten(P const& pot)
        {
            static_assert(is_potential<P>::value,
                          "md (B200 engine): the potential must be built from md:: potentials, operator* and operator+");
            term_list out;
            pot.b200_flatten(out, 1.0, 0.0);
            if (out.size() > BD_MAX_TERMS) throw std::runtime_error("md (B200 engine): potential has too many terms");
            return out;
        }

        template<typename P>
        term_list pair_terms(P const& p, md::system const& sys, md::index i, md::index j)
        {
            if constexpr (std::is_invocable<P const&, md::system const&, md::index, md::index>::value) return flatten(p(sys, i, j));
            else if constexpr (std::is_invocable<P const&, md::index, md::index>::value) { (void) sys; return flatten(p(i, j)); }
            else { (void) sys; (void) i; (void) j; return flatten(p); }
        }

        template<typename P>
        term_list field_terms(P const& p, md::system const& sys, md::index i)
        {
            if constexpr (std::is_invocable<P const&, md::system const&, md::index>::value) return flatten(p(sys, i));
            else if constexpr (std::is_invocable<P const&, md::index>::value) { (void) sys; return flatten(p(i)); }
            else { (void) sys; (void) i; return flatten(p); }
        }

        template<typename P>
        double triple_energy(P const& p, md::system const& sys, md::index i, md::index j, md::index k)
        {
            if constexpr (std::is_invocable<P const&, md::system const&, md::index, md::index, md::index>::value)
                return p(sys, i, j, k).bending_energy;
            else if constexpr (std::is_invocable<P const&, md::index, md::index, md::index>::value) { (void) sys; return p(i, j, k).bending_energy; }
            else { (void) sys; (void) i; (void) j; (void) k; return p.bending_energy; }
        }

        inline void pad_terms(term_list& t, std::size_t n)
        {
            while (t.size() < n) t.push_back(make_term(BD_TERM_NONE, 0, 0, 0, 0, 0, 0));
        }

        // table over class pairs: rows[(ci * T + cj)] = terms of lambda(rep[ci], rep[cj]), padded
        template<typename P>
        term_list pair_table(P const& p, md::system const& sys, class_map const& cm, int& nterms)
        {
            int const T = cm.ntypes();
            std::vector<term_list> cells(std::size_t(T) * T);
            std::size_t width = 1;
            for (int a = 0; a < T; a++)
                for (int b = 0; b < T; b++) {
                    cells[std::size_t(a) * T + b] = pair_terms(p, sys, cm.rep[a], cm.rep[b]);
                    width = std::max(width, cells[std::size_t(a) * T + b].size());
                }
            term_list table;
            for (auto& c : cells) {
                pad_terms(c, width);
                table.insert(table.end(), c.begin(), c.end());
            }
            nterms = int(width);
            return table;
        }

        template<typename P>
        term_list field_table(P const& p, md::system const& sys, class_map const& cm, int& nterms)
        {
            int const T = cm.ntypes();
            std::vector<term_list> cells(T);
            std::size_t width = 1;
            for (int a = 0; a < T; a++) {
                cells[a] = field_terms(p, sys, cm.rep[a]);
                width = std::max(width, cells[a].size());
            }
            term_list table;
            for (auto& c : cells) {
                pad_terms(c, width);
                table.insert(table.end(), c.begin(), c.end());
            }
            nterms = int(width);
            return table;
        }

        inline term_list table_row(term_list const& table, int nterms, std::size_t cell)
        {
            return term_list(table.begin() + cell * nterms, table.begin() + (cell + 1) * nterms);
        }

        // the class table must reproduce lambda(i, j) for actual particles: otherwise the potential
        // depends on something other than the attribute rows and cannot be tabulated
        inline void require_same(term_list got, term_list const& row, char const* what)
        {
            pad_terms(got, row.size());
            if (!same_terms(got, row))
                throw std::runtime_error(std::string("md (B200 engine): ") + what
                                         + " potential depends on more than the particles' attribute rows");
        }

        inline std::vector<std::uint8_t> target_mask(std::vector<md::index> const& targets, md::index n)
        {
            std::vector<std::uint8_t> mask(n, 0);
            for (auto i : targets) {
                if (i >= n) throw std::out_of_range("md (B200 engine): target index out of range");
                mask[i] = 1;
            }
            return mask;
        }
    }

    // md::make_neighbor_pairwise_forcefield<Box>(potential or lambda)
    // (MD/forcefield/neighbor_pairwise_forcefield.hpp:44-204)
    template<typename PotFun, typename Box>
    class basic_neighbor_pairwise_forcefield_impl : public b200::device_forcefield
    {
        using self = basic_neighbor_pairwise_forcefield_impl;

    public:
        explicit basic_neighbor_pairwise_forcefield_impl(PotFun p) : potfun_{std::move(p)} {}

        self& set_unit_cell(Box box) { box_ = [box] { return box; }; return *this; }
        self& set_unit_cell(std::function<Box()> cb) { box_ = std::move(cb); return *this; }
        self& set_neighbor_distance(md::scalar d) { dcut_ = [d] { return d; }; return *this; }
        self& set_neighbor_distance(std::function<md::scalar()> cb) { dcut_ = std::move(cb); return *this; }

        template<typename R>
        self& set_neighbor_targets(R const& indices)
        {
            targets_.assign(std::begin(indices), std::end(indices));
            has_targets_ = true;
            touch_structure();
            return *this;
        }

        Box unit_cell() const { return box_ ? box_() : Box{}; }
        md::scalar neighbor_distance(md::system const&) const { return dcut_ ? dcut_() : md::scalar(0); }
        auto neighbor_pairwise_potential(md::system const& system, md::index i, md::index j) const
        {
            if constexpr (std::is_invocable<PotFun const&, md::system const&, md::index, md::index>::value) return potfun_(system, i, j);
            else if constexpr (std::is_invocable<PotFun const&, md::index, md::index>::value) { (void) system; return potfun_(i, j); }
            else { (void) system; (void) i; (void) j; return potfun_; }
        }

        int b200_class() const override { return BD_FF_PAIR; }

        void b200_attach(b200::binding& b, md::system const& sys) override
        {
            Box const box = unit_cell();
            double period[3];
            b200::box_period(box, period);
            b.set_box(b200::box_kind(box), period);
            table_ = b200::pair_table(potfun_, sys, b.classes(), nterms_);
            pushed_dcut_ = neighbor_distance(sys);
            auto const& cm = b.classes();
            md::index const n = sys.particle_count();
            // spot-check the tabulation on actual particle pairs
            std::uint64_t lcg = 0x9e3779b97f4a7c15ull;
            for (int k = 0; k < 64 && n > 1; k++) {
                lcg = lcg * 6364136223846793005ull + 1442695040888963407ull;
                md::index i = md::index((lcg >> 33) % n);
                lcg = lcg * 6364136223846793005ull + 1442695040888963407ull;
                md::index j = md::index((lcg >> 33) % n);
                if (i == j) continue;
                if (i > j) std::swap(i, j);
                b200::require_same(b200::pair_terms(potfun_, sys, i, j),
                                   b200::table_row(table_, nterms_, std::size_t(cm.cls[i]) * cm.ntypes() + cm.cls[j]),
                                   "neighbor pairwise");
            }
            std::vector<std::uint8_t> mask;
            if (has_targets_) mask = b200::target_mask(targets_, n);
            b200::check(MD_B200_CALL(add_pair_ff)(b.handle(), nterms_, table_.data(), pushed_dcut_,
                                                  has_targets_ ? mask.data() : nullptr, &b200_id), "add_pair_ff");
        }

        bool b200_refresh(b200::binding& b, md::system const& sys) override
        {
            Box const box = unit_cell();
            double period[3];
            b200::box_period(box, period);
            b.update_box(b200::box_kind(box), period);
            int nterms = 0;
            b200::term_list table = b200::pair_table(potfun_, sys, b.classes(), nterms);
            double const dcut = neighbor_distance(sys);
            if (nterms != nterms_) return false;
            if (dcut != pushed_dcut_ || !b200::same_terms(table, table_)) {
                b200::check(MD_B200_CALL(update_pair_ff)(b.handle(), b200_id, table.data(), dcut), "update_pair_ff");
                table_ = std::move(table);
                pushed_dcut_ = dcut;
            }
            return true;
        }

    private:
        PotFun potfun_;
        std::function<Box()> box_;
        std::function<md::scalar()> dcut_;
        std::vector<md::index> targets_;
        bool has_targets_ = false;
        b200::term_list table_;
        int nterms_ = 0;
        double pushed_dcut_ = 0;
    };

    template<typename Box = md::open_box, typename P>
    auto make_neighbor_pairwise_forcefield(P pot)
    {
        return basic_neighbor_pairwise_forcefield_impl<P, Box>{std::move(pot)};
    }

    // md::make_bonded_pairwise_forcefield (MD/forcefield/bonded_pairwise_forcefield.hpp:36-132)
    template<typename PotFun>
    class basic_bonded_pairwise_forcefield : public b200::device_forcefield
    {
        using self = basic_bonded_pairwise_forcefield;

    public:
        explicit basic_bonded_pairwise_forcefield(PotFun p) : potfun_{std::move(p)} {}

        self& add_bonded_pair(md::index i, md::index j)
        {
            pairs_.push_back(std::uint32_t(i));
            pairs_.push_back(std::uint32_t(j));
            touch_structure();
            return *this;
        }

        self& add_bonded_range(md::index start, md::index end)
        {
            for (md::index i = start; i + 1 < end; i++) add_bonded_pair(i, i + 1);
            return *this;
        }

        // ---- extension over micromd: pair lists that change while the simulation runs --------------
        // 3-sim-1kb's loop_forcefield / glue_forcefield (src/simulation/forces/*.cpp) walk pair sets that
        // a kinetic model edits between steps.  Replacing the list does not re-create the engine: only
        // the list is uploaded (bd_set_bond_pairs).  b200_use_shortest_displacement(true) takes the pair
        // displacement through the system's periodic box, as glue_forcefield does (bd_set_bond_min_image).
        template<typename PairRange>
        self& b200_set_bonded_pairs(PairRange const& pairs)
        {
            pairs_.clear();
            for (auto const& p : pairs) {
                pairs_.push_back(std::uint32_t(std::get<0>(p)));
                pairs_.push_back(std::uint32_t(std::get<1>(p)));
            }
            pairs_dirty_ = true;
            return *this;
        }

        self& b200_use_shortest_displacement(bool on)
        {
            min_image_ = on;
            image_dirty_ = true;
            return *this;
        }

        auto bonded_pairwise_potential(md::system const& system, md::index i, md::index j) const
        {
            if constexpr (std::is_invocable<PotFun const&, md::system const&, md::index, md::index>::value) return potfun_(system, i, j);
            else if constexpr (std::is_invocable<PotFun const&, md::index, md::index>::value) { (void) system; return potfun_(i, j); }
            else { (void) system; (void) i; (void) j; return potfun_; }
        }

        int b200_class() const override { return BD_FF_BOND; }

        void b200_attach(b200::binding& b, md::system const& sys) override
        {
            table_ = b200::pair_table(potfun_, sys, b.classes(), nterms_);
            auto const& cm = b.classes();
            md::index const n = sys.particle_count();
            for (std::size_t k = 0; k + 1 < pairs_.size(); k += 2) {
                md::index const i = pairs_[k], j = pairs_[k + 1];
                if (i >= n || j >= n) throw std::out_of_range("md (B200 engine): bonded pair index out of range");
                b200::require_same(b200::pair_terms(potfun_, sys, i, j),
                                   b200::table_row(table_, nterms_, std::size_t(cm.cls[i]) * cm.ntypes() + cm.cls[j]),
                                   "bonded pairwise");
            }
            b200::check(MD_B200_CALL(add_bond_ff)(b.handle(), nterms_, table_.data(), pairs_.data(),
                                                  std::int64_t(pairs_.size() / 2), &b200_id), "add_bond_ff");
            if (min_image_) b200::check(MD_B200_CALL(set_bond_min_image)(b.handle(), b200_id, 1), "set_bond_min_image");
            pairs_dirty_ = image_dirty_ = false;
        }

        bool b200_refresh(b200::binding& b, md::system const& sys) override
        {
            int nterms = 0;
            b200::term_list table = b200::pair_table(potfun_, sys, b.classes(), nterms);
            if (nterms != nterms_) return false;
            if (!b200::same_terms(table, table_)) {
                b200::check(MD_B200_CALL(update_bond_ff)(b.handle(), b200_id, table.data()), "update_bond_ff");
                table_ = std::move(table);
            }
            if (pairs_dirty_) {
                md::index const n = sys.particle_count();
                for (std::uint32_t v : pairs_)
                    if (v >= n) throw std::out_of_range("md (B200 engine): bonded pair index out of range");
                b200::check(MD_B200_CALL(set_bond_pairs)(b.handle(), b200_id, pairs_.data(),
                                                         std::int64_t(pairs_.size() / 2)), "set_bond_pairs");
                pairs_dirty_ = false;
            }
            if (image_dirty_) {
                b200::check(MD_B200_CALL(set_bond_min_image)(b.handle(), b200_id, min_image_ ? 1 : 0), "set_bond_min_image");
                image_dirty_ = false;
            }
            return true;
        }

    private:
        PotFun potfun_;
        bool pairs_dirty_ = false, image_dirty_ = false, min_image_ = false;
        std::vector<std::uint32_t> pairs_;
        b200::term_list table_;
        int nterms_ = 0;
    };

    template<typename P>
    auto make_bonded_pairwise_forcefield(P pot) { return basic_bonded_pairwise_forcefield<P>{std::move(pot)}; }

    // md::make_bonded_triplewise_forcefield (MD/forcefield/bonded_triplewise_forcefield.hpp:36-149)
    template<typename PotFun>
    class basic_bonded_triplewise_forcefield : public b200::device_forcefield
    {
        using self = basic_bonded_triplewise_forcefield;

    public:
        explicit basic_bonded_triplewise_forcefield(PotFun p) : potfun_{std::move(p)} {}

        self& add_bonded_triple(md::index i, md::index j, md::index k)
        {
            triples_.push_back(std::uint32_t(i));
            triples_.push_back(std::uint32_t(j));
            triples_.push_back(std::uint32_t(k));
            touch_structure();
            return *this;
        }

        self& add_bonded_range(md::index start, md::index end)
        {
            for (md::index i = start; i + 2 < end; i++) add_bonded_triple(i, i + 1, i + 2);
            return *this;
        }

        int b200_class() const override { return BD_FF_BEND; }

        void b200_attach(b200::binding& b, md::system const& sys) override
        {
            energies_ = current_energies(sys);
            b200::check(MD_B200_CALL(add_bend_ff)(b.handle(), triples_.data(), energies_.data(),
                                                  std::int64_t(energies_.size()), &b200_id), "add_bend_ff");
        }

        bool b200_refresh(b200::binding& b, md::system const& sys) override
        {
            // the C ABI takes bending energies at registration: a change re-creates the engine
            (void) b;
            return current_energies(sys) == energies_;
        }

    private:
        std::vector<double> current_energies(md::system const& sys) const
        {
            std::vector<double> e(triples_.size() / 3);
            for (std::size_t t = 0; t < e.size(); t++)
                e[t] = b200::triple_energy(potfun_, sys, triples_[3 * t], triples_[3 * t + 1], triples_[3 * t + 2]);
            return e;
        }

        PotFun potfun_;
        std::vector<std::uint32_t> triples_;
        std::vector<double> energies_;
    };

    template<typename P>
    auto make_bonded_triplewise_forcefield(P pot) { return basic_bonded_triplewise_forcefield<P>{std::move(pot)}; }

    namespace b200
    {
        // shared by the four surface force fields
        template<typename Derived, typename PotFun, typename Shape, int ShapeKind, int Direction>
        class surface_forcefield : public device_forcefield
        {
        public:
            explicit surface_forcefield(PotFun p) : potfun_{std::move(p)} {}

            int b200_class() const override { return BD_FF_WALL; }

            void b200_attach(binding& b, md::system const& sys) override
            {
                table_ = field_table(potfun_, sys, b.classes(), nterms_);
                auto const& cm = b.classes();
                md::index const n = sys.particle_count();
                md::index const stride = std::max<md::index>(1, n / 257);
                for (md::index i = 0; i < n; i += stride)
                    require_same(field_terms(potfun_, sys, i), table_row(table_, nterms_, cm.cls[i]), "surface");
                geometry(center_, semi_);
                check(MD_B200_CALL(add_wall_ff)(b.handle(), ShapeKind, Direction, nterms_, table_.data(), center_, semi_,
                                                &b200_id), "add_wall_ff");
            }

            bool b200_refresh(binding& b, md::system const& sys) override
            {
                int nterms = 0;
                term_list table = field_table(potfun_, sys, b.classes(), nterms);
                if (nterms != nterms_) return false;
                double center[3], semi[3];
                geometry(center, semi);
                bool const new_table = !same_terms(table, table_);
                bool const moved = std::memcmp(center, center_, sizeof center) != 0 || std::memcmp(semi, semi_, sizeof semi) != 0;
                if (new_table || moved) {
                    check(MD_B200_CALL(update_wall_ff)(b.handle(), b200_id, new_table ? table.data() : nullptr, center, semi),
                          "update_wall_ff");
                    table_ = std::move(table);
                    std::memcpy(center_, center, sizeof center);
                    std::memcpy(semi_, semi, sizeof semi);
                }
                return true;
            }

            void b200_collect(binding& b) override
            {
                double out[4] = {0, 0, 0, 0};
                check(MD_B200_CALL(get_wall_stats)(b.handle(), b200_id, 0, out), "get_wall_stats");
                static_cast<Derived&>(*this).store_stats(out);
            }

        protected:
            virtual void geometry(double center[3], double semi[3]) const = 0;
            std::function<Shape()> shape_ = [] { return Shape{}; };

        private:
            PotFun potfun_;
            term_list table_;
            int nterms_ = 0;
            double center_[3] = {0, 0, 0}, semi_[3] = {1, 1, 1};
        };
    }

    // md::make_sphere_{inward,outward}_forcefield (MD/forcefield/sphere_surface_forcefield.hpp:62-266)
    template<typename PotFun, int Direction>
    class basic_sphere_surface_forcefield_impl
        : public b200::surface_forcefield<basic_sphere_surface_forcefield_impl<PotFun, Direction>, PotFun, md::sphere,
                                          BD_WALL_SPHERE, Direction>
    {
        using self = basic_sphere_surface_forcefield_impl;
        using base = b200::surface_forcefield<self, PotFun, md::sphere, BD_WALL_SPHERE, Direction>;

    public:
        struct statistics
        {
            md::scalar reaction_force = 0;
        };
        statistics stats;

        using base::base;
        self& set_sphere(md::sphere s) { this->shape_ = [s] { return s; }; return *this; }
        self& set_sphere(std::function<md::sphere()> cb) { this->shape_ = std::move(cb); return *this; }
        md::sphere sphere(md::system const&) const { return this->shape_(); }
        void store_stats(double const out[4]) { stats.reaction_force = out[0]; }

    protected:
        void geometry(double center[3], double semi[3]) const override
        {
            md::sphere const s = this->shape_();
            center[0] = s.center.x; center[1] = s.center.y; center[2] = s.center.z;
            semi[0] = semi[1] = semi[2] = s.radius;
        }
    };

    template<typename P>
    auto make_sphere_inward_forcefield(P pot) { return basic_sphere_surface_forcefield_impl<P, BD_WALL_INWARD>{std::move(pot)}; }
    template<typename P>
    auto make_sphere_outward_forcefield(P pot) { return basic_sphere_surface_forcefield_impl<P, BD_WALL_OUTWARD>{std::move(pot)}; }

    // md::make_ellipsoid_{inward,outward}_forcefield (MD/forcefield/ellipsoid_surface_forcefield.hpp:123-333)
    template<typename PotFun, int Direction>
    class basic_ellipsoid_surface_forcefield_impl
        : public b200::surface_forcefield<basic_ellipsoid_surface_forcefield_impl<PotFun, Direction>, PotFun, md::ellipsoid,
                                          BD_WALL_ELLIPSOID, Direction>
    {
        using self = basic_ellipsoid_surface_forcefield_impl;
        using base = b200::surface_forcefield<self, PotFun, md::ellipsoid, BD_WALL_ELLIPSOID, Direction>;

    public:
        struct statistics
        {
            md::scalar reaction_force = 0;
            md::vector axial_reaction = {};
        };
        statistics stats;

        using base::base;
        self& set_ellipsoid(md::ellipsoid e) { this->shape_ = [e] { return e; }; return *this; }
        self& set_ellipsoid(std::function<md::ellipsoid()> cb) { this->shape_ = std::move(cb); return *this; }
        md::ellipsoid ellipsoid(md::system const&) const { return this->shape_(); }
        void store_stats(double const out[4])
        {
            stats.reaction_force = out[0];
            stats.axial_reaction = {out[1], out[2], out[3]};
        }

    protected:
        void geometry(double center[3], double semi[3]) const override
        {
            md::ellipsoid const e = this->shape_();
            center[0] = e.center.x; center[1] = e.center.y; center[2] = e.center.z;
            semi[0] = e.semiaxis_x; semi[1] = e.semiaxis_y; semi[2] = e.semiaxis_z;
        }
    };

    template<typename P>
    auto make_ellipsoid_inward_forcefield(P pot) { return basic_ellipsoid_surface_forcefield_impl<P, BD_WALL_INWARD>{std::move(pot)}; }
    template<typename P>
    auto make_ellipsoid_outward_forcefield(P pot) { return basic_ellipsoid_surface_forcefield_impl<P, BD_WALL_OUTWARD>{std::move(pot)}; }

    // md::make_point_source_forcefield (MD/forcefield/point_source_forcefield.hpp:35-147)
    template<typename PotFun>
    class basic_point_source_forcefield : public b200::device_forcefield
    {
        using self = basic_point_source_forcefield;

    public:
        explicit basic_point_source_forcefield(PotFun p) : potfun_{std::move(p)} {}

        self& set_point_source(md::point pt) { source_ = pt; touch_structure(); return *this; }

        template<typename R>
        self& set_point_source_targets(R const& indices)
        {
            targets_.assign(std::begin(indices), std::end(indices));
            has_targets_ = true;
            touch_structure();
            return *this;
        }

        int b200_class() const override { return BD_FF_POINT; }

        void b200_attach(b200::binding& b, md::system const& sys) override
        {
            table_ = b200::field_table(potfun_, sys, b.classes(), nterms_);
            md::index const n = sys.particle_count();
            std::vector<std::uint8_t> mask;
            if (has_targets_) mask = b200::target_mask(targets_, n);
            double const src[3] = {source_.x, source_.y, source_.z};
            b200::check(MD_B200_CALL(add_point_ff)(b.handle(), nterms_, table_.data(), src,
                                                   has_targets_ ? mask.data() : nullptr, &b200_id), "add_point_ff");
        }

        bool b200_refresh(b200::binding& b, md::system const& sys) override
        {
            // the C ABI has no parameter update for point sources: a change re-creates the engine
            int nterms = 0;
            b200::term_list table = b200::field_table(potfun_, sys, b.classes(), nterms);
            (void) b;
            return nterms == nterms_ && b200::same_terms(table, table_);
        }

    private:
        PotFun potfun_;
        md::point source_;
        std::vector<md::index> targets_;
        bool has_targets_ = false;
        b200::term_list table_;
        int nterms_ = 0;
    };

    template<typename P>
    auto make_point_source_forcefield(P pot) { return basic_point_source_forcefield<P>{std::move(pot)}; }

    // ---- system and particle_ref members -----------------------------------------------------------

    inline system::system()
    {
        add_attribute(md::mass_attribute);
        add_attribute(md::friction_attribute);
        add_attribute(md::mobility_attribute);
        add_attribute(md::position_attribute);
        add_attribute(md::velocity_attribute);
    }

    inline system::~system() = default;

    inline md::particle_ref system::add_particle(md::basic_particle_data data)
    {
        if (binding_) binding_->fetch_positions();
        md::index const idx = count_++;
        for (auto& kv : columns_) kv.second->resize(count_);
        view_masses()[idx] = data.mass;
        view_frictions()[idx] = data.friction;
        view_mobilities()[idx] = data.mobility;
        b200_host_positions()[idx] = data.position;
        view_velocities()[idx] = data.velocity;
        if (binding_) binding_->host_will_write();
        return md::particle_ref(*this, idx);
    }

    inline b200::binding& system::b200_binding() const
    {
        if (!binding_) binding_ = std::make_unique<b200::binding>(*this);
        return *binding_;
    }

    inline md::array_view<md::point> system::view_positions()
    {
        if (binding_) binding_->host_will_write();
        return {b200_host_positions(), count_};
    }

    inline md::array_view<md::point const> system::view_positions() const
    {
        if (binding_) binding_->fetch_positions();
        return {b200_host_positions(), count_};
    }

    inline md::scalar system::compute_potential_energy()
    {
        if (forcefields_.empty()) return 0;
        auto& b = b200_binding();
        b.sync(true);
        double e = 0;
        b200::check(MD_B200_CALL(compute_energy)(b.handle(), 0, BD_FF_ALL, 0, &e), "compute_energy");
        return e + b.host_energy();
    }

    inline void system::compute_force(md::array_view<md::vector> forces)
    {
        assert(forces.size() == particle_count());
        if (forcefields_.empty()) {
            std::fill(forces.begin(), forces.end(), md::vector{});
            return;
        }
        auto& b = b200_binding();
        b.sync(true);
        static_assert(sizeof(md::vector) == 3 * sizeof(double), "md::vector is three contiguous doubles");
        b200::check(MD_B200_CALL(compute_forces)(b.handle(), 0, BD_FF_ALL, 0, &forces.data()->x), "compute_forces");
        b.collect();
    }

    inline md::particle_ref::particle_ref(md::system& system, md::index idx)
        : system{system}
        , index{idx}
        , mass{system.view_masses()[idx]}
        , friction{system.view_frictions()[idx]}
        , mobility{system.view_mobilities()[idx]}
        , position{system.view_positions()[idx]}
        , velocity{system.view_velocities()[idx]}
    {
    }

    template<typename T, typename Tag>
    T& md::particle_ref::view(md::attribute_key<T, Tag> key) { return system.view(key)[index]; }

    namespace b200
    {
        // a single force field's own compute_energy / compute_force (the force field must have been
        // added to the system: the engine evaluates it by its registered id)
        inline md::scalar device_forcefield::compute_energy(md::system const& system)
        {
            auto& b = system.b200_binding();
            b.sync(true);
            if (b200_id < 0) throw std::runtime_error("md (B200 engine): the force field is not part of the system");
            double e = 0;
            check(MD_B200_CALL(compute_energy)(b.handle(), 0, b200_class(), b200_id, &e), "compute_energy");
            return e;
        }

        inline void device_forcefield::compute_force(md::system const& system, md::array_view<md::vector> forces)
        {
            auto& b = system.b200_binding();
            b.sync(true);
            if (b200_id < 0) throw std::runtime_error("md (B200 engine): the force field is not part of the system");
            std::vector<md::vector> own(system.particle_count());
            check(MD_B200_CALL(compute_forces)(b.handle(), 0, b200_class(), b200_id, &own.data()->x), "compute_forces");
            for (md::index i = 0; i < own.size(); i++) forces[i] += own[i];
            b200_collect(b);
        }
    }

    // ---- md::simulate_brownian_dynamics (MD/simulation/brownian_dynamics.hpp:24-73) ---------------

    struct brownian_dynamics_config
    {
        md::scalar temperature = 1;
        md::scalar timestep = 1;
        md::scalar spacestep = 0;
        md::step steps = 1;
        std::uint64_t seed = 0;
        std::function<void(md::step)> callback = {};
    };

    // The callback is invoked after every step with the 1-based step number, exactly as in the
    // reference; between callbacks the potentials are re-evaluated so that whatever the callback
    // changed (scales, walls, cutoffs) acts on the next step.  Without a callback the whole run is
    // one engine call -- unless the system holds driver-defined md::forcefield subclasses, whose
    // forces are evaluated by the driver on the host before every step.  Noise: counter-based Philox keyed by (seed, step, particle) on the device
    // (the reference draws sfc64 + ziggurat normals on the host: same distribution, different stream).
    inline void simulate_brownian_dynamics(md::system& system, md::brownian_dynamics_config config)
    {
        if (config.steps <= 0) return;
        auto& b = system.b200_binding();
        b.sync(true);
        bd_step_config cfg;
        std::memset(&cfg, 0, sizeof cfg);
        cfg.temperature = config.temperature;
        cfg.timestep = config.timestep;
        cfg.spacestep = config.spacestep;
        cfg.seed = config.seed;
        cfg.replay_normals = nullptr;
        if (!config.callback && !b.has_host_forcefields()) {
            cfg.steps = config.steps;
            cfg.first_step = 1;
            b200::check(MD_B200_CALL(step)(b.handle(), &cfg), "step");
            b.device_advanced();
            b.collect();
            return;
        }
        for (md::step s = 1; s <= config.steps; s++) {
            if (s > 1) b.sync(false);
            cfg.steps = 1;
            cfg.first_step = s;
            b200::check(MD_B200_CALL(step)(b.handle(), &cfg), "step");
            b.device_advanced();
            b.collect();
            if (config.callback) config.callback(s);
        }
    }

    // ---- md::neighbor_searcher<Box> (MD/misc/neighbor_searcher.hpp:23-131) ------------------------
    //
    // Every unordered pair (i < j) within dcut, found on the device and emitted in the reference's order
    // (bucket by bucket: bd_nsearch_pairs_ordered), so that a caller which feeds a sequential generator in
    // search order -- 3-sim-1kb's glue simulator -- takes the decisions it takes with micromd.
    namespace b200
    {
        extern "C" int bd_nsearch_query(std::int32_t box_kind, const double period[3], double dcut, const double* xyz,
                                        std::int64_t n, const double* queries, std::int64_t m, std::int64_t* start_out,
                                        std::uint32_t* index_out, std::int64_t cap, std::int32_t device);
        extern "C" int bd_nsearch_pairs_ordered(std::int32_t box_kind, const double period[3], std::int64_t particle_count,
                                                double dcut, const double* xyz, std::int64_t n, std::uint32_t* pairs_out,
                                                std::int64_t cap, std::int64_t* count_out, std::int32_t device);
    }

    template<typename Box>
    class neighbor_searcher
    {
    public:
        neighbor_searcher(Box box, md::scalar dcut) : box_{box}, dcut_{dcut} {}

        void set_points(md::array_view<md::point const> points) { points_.assign(points.begin(), points.end()); }

        template<typename OutputIterator>
        void search(OutputIterator out) const
        {
            if (points_.size() < 2) return;
            double period[3];
            b200::box_period(box_, period);
            std::int64_t cap = std::int64_t(points_.size()) * 16, count = 0;
            std::vector<std::uint32_t> pairs;
            for (;;) {
                pairs.resize(std::size_t(cap) * 2);
                b200::check(b200::bd_nsearch_pairs_ordered(b200::box_kind(box_), period, std::int64_t(b200::box_count(box_)), dcut_,
                                                           &points_.data()->x, std::int64_t(points_.size()), pairs.data(), cap,
                                                           &count, 0), "nsearch_pairs_ordered");
                if (count <= cap) break;
                cap = count;
            }
            for (std::int64_t k = 0; k < count; k++)
                *out++ = std::pair<md::index, md::index>{pairs[2 * k], pairs[2 * k + 1]};
        }

        // Indices of the points closer than dcut to `point` (strict, as MD/misc/neighbor_searcher.hpp:104-131).
        // No simulation driver calls this -- only the hetero-RDF analysis tool does
        // (4-sim-ab/box/src/rdf_analysis_hetero/distance_histogram.cc:52), one point at a time, which is
        // host work by nature: a scan of the stored points in ascending index order (the reference emits
        // the same set in bucket order).
        template<typename OutputIterator>
        void query(md::point point, OutputIterator out) const
        {
            double const dcut2 = dcut_ * dcut_;
            for (md::index i = 0; i < points_.size(); i++)
                if (box_.shortest_displacement(points_[i], point).squared_norm() < dcut2) *out++ = i;
        }

        // extension over micromd: query() for many points in one device call; result[q] = indices (ascending) of
        // the stored points closer than dcut to queries[q]
        std::vector<std::vector<md::index>> b200_query_all(md::array_view<md::point const> queries) const
        {
            std::vector<std::vector<md::index>> result(queries.size());
            if (queries.size() == 0 || points_.empty()) return result;
            double period[3];
            b200::box_period(box_, period);
            std::vector<std::int64_t> start(queries.size() + 1);
            b200::check(b200::bd_nsearch_query(b200::box_kind(box_), period, dcut_, &points_.data()->x, std::int64_t(points_.size()),
                                               &queries.begin()->x, std::int64_t(queries.size()), start.data(), nullptr, 0, 0),
                        "nsearch_query");
            std::vector<std::uint32_t> index(std::size_t(std::max<std::int64_t>(start.back(), 1)));
            b200::check(b200::bd_nsearch_query(b200::box_kind(box_), period, dcut_, &points_.data()->x, std::int64_t(points_.size()),
                                               &queries.begin()->x, std::int64_t(queries.size()), start.data(), index.data(),
                                               start.back(), 0), "nsearch_query");
            for (std::size_t q = 0; q < result.size(); q++)
                result[q].assign(index.begin() + start[q], index.begin() + start[q + 1]);
            return result;
        }

    private:
        Box box_;
        md::scalar dcut_;
        std::vector<md::point> points_;
    };

#ifdef MD_B200_ENGINE
    // ---- extension over micromd: questions about the device-resident positions ----------------------
    // 3-sim-1kb's kinetic glue model (src/simulation/glues/glue_simulator.cpp:35-80) asks two things of the
    // positions at every update: the pairs within the binding distance and the current length of every
    // glued pair.  Answered here without bringing the positions to the host.
    namespace b200
    {
        extern "C" {
        int bd_search_pairs(bd_engine* h, std::int32_t replica, double distance, std::int32_t use_box, std::uint32_t* pairs_out,
                            std::int64_t cap, std::int64_t* count_out);
        int bd_pair_distances(bd_engine* h, std::int32_t replica, const std::uint32_t* pairs, std::int64_t m,
                              std::int32_t min_image, double* out);
        }

        // pairs (i < j) within `distance`, sorted; use_box: the system's unit cell (md::neighbor_searcher<periodic_box>)
        // or an open box (md::neighbor_searcher<open_box>)
        inline std::vector<std::pair<md::index, md::index>> search_pairs(md::system& system, md::scalar distance, bool use_box = true)
        {
            std::vector<std::pair<md::index, md::index>> out;
            if (system.particle_count() < 2) return out;
            auto& b = system.b200_binding();
            b.sync(true);
            std::int64_t cap = std::int64_t(system.particle_count()) * 16, count = 0;
            std::vector<std::uint32_t> pairs;
            for (;;) {
                pairs.resize(std::size_t(cap) * 2);
                check(bd_search_pairs(b.handle(), 0, distance, use_box ? 1 : 0, pairs.data(), cap, &count), "search_pairs");
                if (count <= cap) break;
                cap = count;
            }
            out.reserve(std::size_t(count));
            for (std::int64_t k = 0; k < count; k++) out.emplace_back(pairs[2 * k], pairs[2 * k + 1]);
            return out;
        }

        // |x_i - x_j| of every pair, through the box's shortest displacement when min_image is set
        template<typename PairRange>
        std::vector<md::scalar> pair_distances(md::system& system, PairRange const& pair_range, bool min_image = false)
        {
            std::vector<std::uint32_t> pairs;
            for (auto const& p : pair_range) {
                pairs.push_back(std::uint32_t(std::get<0>(p)));
                pairs.push_back(std::uint32_t(std::get<1>(p)));
            }
            std::vector<md::scalar> out(pairs.size() / 2);
            if (out.empty()) return out;
            auto& b = system.b200_binding();
            b.sync(true);
            check(bd_pair_distances(b.handle(), 0, pairs.data(), std::int64_t(out.size()), min_image ? 1 : 0, out.data()),
                  "pair_distances");
            return out;
        }
    }
#endif
}

#endif  // MD_B200_MD_HPP
