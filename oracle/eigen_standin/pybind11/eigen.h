// TEST INFRASTRUCTURE -- shadows pybind11/eigen.h (which needs the real Eigen) for the stand-in of
// ../Eigen/Core: numpy array / nested sequence -> Eigen::Matrix<T, R, C> by value, which is all the
// reference binding (pointcloud_subsampling.cpp:22-26, std::vector<FMatrix3D> / <FMatrix3x4>) needs.
#pragma once
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>

#include <Eigen/Core>

namespace pybind11 {
namespace detail {

template <typename T, int R, int C>
struct type_caster<Eigen::Matrix<T, R, C>, std::enable_if_t<(C > 1)>> {
  using Mat = Eigen::Matrix<T, R, C>;
  PYBIND11_TYPE_CASTER(Mat, const_name("numpy.ndarray"));

  bool load(handle src, bool) {
    auto arr = array_t<T, array::c_style | array::forcecast>::ensure(src);
    if (!arr || arr.ndim() != 2 || arr.shape(0) != R || arr.shape(1) != C) return false;
    auto a = arr.template unchecked<2>();
    for (int i = 0; i < R; ++i)
      for (int j = 0; j < C; ++j) value(i, j) = a(i, j);
    return true;
  }
  static handle cast(const Mat& m, return_value_policy, handle) {
    array_t<T> out({R, C});
    auto a = out.template mutable_unchecked<2>();
    for (int i = 0; i < R; ++i)
      for (int j = 0; j < C; ++j) a(i, j) = m(i, j);
    return out.release();
  }
};

}  // namespace detail
}  // namespace pybind11
