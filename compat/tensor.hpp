// compat/tensor.hpp -- the Eigen typedef names the reference's public API is written in
// (API surface of /root/reference/src/tensor.hpp:24-28; only the float typedefs cross the boundary).
#ifndef DEMUCS_B200_COMPAT_TENSOR_HPP
#define DEMUCS_B200_COMPAT_TENSOR_HPP

#include <Eigen/Dense>
#include <complex>
#include <unsupported/Eigen/CXX11/Tensor>

namespace Eigen
{
typedef Tensor<float, 4> Tensor4dXf;
typedef Tensor<float, 3> Tensor3dXf;
typedef Tensor<float, 2> Tensor2dXf;
typedef Tensor<float, 1> Tensor1dXf;
typedef Tensor<std::complex<float>, 3> Tensor3dXcf;
} // namespace Eigen

#endif
