// compat/tensor.hpp -- the Eigen typedef names the reference's public API is written in
// (API surface of /root/reference/src/tensor.hpp:24-28; only the float typedefs cross the boundary).
#ifndef DEMUCS_B200_COMPAT_TENSOR_HPP
#define DEMUCS_B200_COMPAT_TENSOR_HPP

#include <Eigen/Dense>
#include <complex>
#include <unsupported/Eigen/CXX11/Tensor>

namespace Eigen
{
// alias templates by rank, then the names the API spells out
template <int Rank> using TensorXf = Tensor<float, Rank>;
using Tensor1dXf = TensorXf<1>;
using Tensor2dXf = TensorXf<2>;
using Tensor3dXf = TensorXf<3>;
using Tensor4dXf = TensorXf<4>;
using Tensor3dXcf = Tensor<std::complex<float>, 3>;
} // namespace Eigen

#endif
