// ink_eigen_lite.h — the sliver of Eigen's Vector3f / Vector3i surface that the reference's PUBLIC headers expose
// (src/system.h:41-68: Cell, Particle, hash_func), for hosts that do not have Eigen on the include path.
// If <Eigen/Dense> is available (the reference vendors it) the real thing is used instead and this file adds nothing.
// Written from the interface, not from Eigen's sources.
#pragma once
#if defined(__has_include)
#if __has_include(<Eigen/Dense>) && !defined(INK_FORCE_EIGEN_LITE)
#include <Eigen/Dense>
#define INK_HAVE_REAL_EIGEN 1
#endif
#endif

#ifndef INK_HAVE_REAL_EIGEN
#include <cmath>
#include <cstddef>
#include <ostream>

namespace Eigen {

template <class T>
class Vec3 {
public:
    Vec3() : v_{T(0), T(0), T(0)} {}
    Vec3(T x, T y, T z) : v_{x, y, z} {}
    T& x() { return v_[0]; }
    T& y() { return v_[1]; }
    T& z() { return v_[2]; }
    const T& x() const { return v_[0]; }
    const T& y() const { return v_[1]; }
    const T& z() const { return v_[2]; }
    T& operator[](std::size_t i) { return v_[i]; }
    const T& operator[](std::size_t i) const { return v_[i]; }
    T& operator()(std::size_t i) { return v_[i]; }
    const T& operator()(std::size_t i) const { return v_[i]; }
    bool operator==(const Vec3& o) const { return v_[0] == o.v_[0] && v_[1] == o.v_[1] && v_[2] == o.v_[2]; }
    bool operator!=(const Vec3& o) const { return !(*this == o); }
    Vec3 operator+(const Vec3& o) const { return Vec3(v_[0] + o.v_[0], v_[1] + o.v_[1], v_[2] + o.v_[2]); }
    Vec3 operator-(const Vec3& o) const { return Vec3(v_[0] - o.v_[0], v_[1] - o.v_[1], v_[2] - o.v_[2]); }
    Vec3 operator*(T s) const { return Vec3(v_[0] * s, v_[1] * s, v_[2] * s); }
    friend Vec3 operator*(T s, const Vec3& a) { return a * s; }
    T squaredNorm() const { return v_[0] * v_[0] + (v_[1] * v_[1] + v_[2] * v_[2]); }
    T norm() const { return (T)std::sqrt((double)squaredNorm()); }
    const T* data() const { return v_; }
    T* data() { return v_; }
    friend std::ostream& operator<<(std::ostream& s, const Vec3& a) { return s << a.v_[0] << " " << a.v_[1] << " " << a.v_[2]; }

private:
    T v_[3];
};

using Vector3f = Vec3<float>;
using Vector3i = Vec3<int>;

}  // namespace Eigen
#endif
