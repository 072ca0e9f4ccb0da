// Shim for <boost/multi_array.hpp>: TEST INFRASTRUCTURE ONLY (oracle/_ref build).
// ccicp2d uses boost::multi_array<float,2> purely as a dense row-major 2-D
// container (kdtree.h:20; icp.cpp:51,57; kdtree.cpp:74-75,93): default ctor,
// resize(extents[n][d]), a[i][j], shape().  No arithmetic lives in Boost, so a
// minimal stand-in does not change any result.
#pragma once
#include <cstddef>
#include <vector>
namespace boost {
struct extent_pair { std::size_t n0, n1; };
struct extent_one {
  std::size_t n0;
  extent_pair operator[](std::size_t n1) const { return extent_pair{n0, n1}; }
};
struct extent_gen {
  extent_one operator[](std::size_t n0) const { return extent_one{n0}; }
};
static const extent_gen extents = extent_gen();

template <typename T, std::size_t D> class const_multi_array_ref;  // named in a typedef only

template <typename T, std::size_t D> class multi_array {
  static_assert(D == 2, "shim supports 2-D arrays only");
  std::vector<T> buf_;
  std::size_t shape_[2];
 public:
  multi_array() { shape_[0] = shape_[1] = 0; }
  void resize(const extent_pair& e) {
    shape_[0] = e.n0; shape_[1] = e.n1;
    buf_.assign(e.n0 * e.n1, T());
  }
  const std::size_t* shape() const { return shape_; }
  T* operator[](std::size_t i) { return buf_.data() + i * shape_[1]; }
  const T* operator[](std::size_t i) const { return buf_.data() + i * shape_[1]; }
};
}  // namespace boost
