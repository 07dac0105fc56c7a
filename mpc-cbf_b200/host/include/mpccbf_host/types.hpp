// Host-side value types with the reference's names, so code written against the reference's
// controller API compiles against this mirror. The reference aliases Eigen types
// (math/include/math/Types.h:8-30: Vector = Eigen::Matrix<T,Dynamic,1>, VectorDIM =
// Eigen::Matrix<T,DIM,1>); Eigen is not a dependency here, so these are minimal contiguous
// containers with the accessors the controller API uses ( v(i), v[i], size(), data() ). A
// maintainer integrating into the reference tree keeps Eigen and passes `.data()` instead — see
// INTEGRATION.md.
#pragma once
#include <array>
#include <cstddef>
#include <initializer_list>
#include <stdexcept>
#include <vector>

namespace math {

template <typename T>
class Vector {
  public:
    Vector() = default;
    explicit Vector(std::size_t n, T v = T(0)) : d_(n, v) {}
    Vector(std::initializer_list<T> il) : d_(il) {}
    std::size_t size() const { return d_.size(); }
    void resize(std::size_t n) { d_.resize(n, T(0)); }
    T& operator()(std::size_t i) { return d_[i]; }
    T operator()(std::size_t i) const { return d_[i]; }
    T& operator[](std::size_t i) { return d_[i]; }
    T operator[](std::size_t i) const { return d_[i]; }
    T* data() { return d_.data(); }
    const T* data() const { return d_.data(); }

  private:
    std::vector<T> d_;
};

template <typename T, unsigned int DIM>
class VectorDIM {
  public:
    VectorDIM() { d_.fill(T(0)); }
    VectorDIM(std::initializer_list<T> il) {
        if (il.size() != DIM) throw std::invalid_argument("VectorDIM: wrong number of initialisers");
        std::size_t i = 0;
        for (T v : il) d_[i++] = v;
    }
    static VectorDIM Zero() { return VectorDIM(); }
    static constexpr std::size_t size() { return DIM; }
    T& operator()(std::size_t i) { return d_[i]; }
    T operator()(std::size_t i) const { return d_[i]; }
    T& operator[](std::size_t i) { return d_[i]; }
    T operator[](std::size_t i) const { return d_[i]; }
    T* data() { return d_.data(); }
    const T* data() const { return d_.data(); }

  private:
    std::array<T, DIM> d_;
};

}  // namespace math

namespace model {

// model/include/model/DoubleIntegrator.h:20-24
template <typename T, unsigned int DIM>
struct State {
    using VectorDIM = math::VectorDIM<T, DIM>;
    VectorDIM pos_, vel_;
};

}  // namespace model
