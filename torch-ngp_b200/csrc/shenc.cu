// shenc.cu -- spherical-harmonics direction encoding, degree 1..4 ("next" row of the scope table).
// Contract: /root/reference/shencoder/src/shencoder.cu:27-120 (polynomial form, same constants).
// One thread per direction; writes degree^2 values at row stride `out_stride` so the caller can
// encode straight into a wider (padded) MLP input matrix.
#include "common.cuh"

namespace ngp {
namespace {

template <typename T> __device__ __forceinline__ T cvt(float v) { return (T)v; }
template <> __device__ __forceinline__ __half cvt<__half>(float v) { return __float2half_rn(v); }

template <typename T>
__global__ void __launch_bounds__(256) k_sh_encode(const float *__restrict__ inputs, T *__restrict__ outputs, uint32_t B,
                                                   uint32_t degree, uint32_t out_stride) {
    const uint32_t b = threadIdx.x + blockIdx.x * blockDim.x;
    if (b >= B) return;
    const float x = inputs[3 * (size_t)b], y = inputs[3 * (size_t)b + 1], z = inputs[3 * (size_t)b + 2];
    T *o = outputs + (size_t)b * out_stride;
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    o[0] = cvt<T>(0.28209479177387814f);
    if (degree <= 1) return;
    o[1] = cvt<T>(-0.48860251190291987f * y);
    o[2] = cvt<T>(0.48860251190291987f * z);
    o[3] = cvt<T>(-0.48860251190291987f * x);
    if (degree <= 2) return;
    o[4] = cvt<T>(1.0925484305920792f * xy);
    o[5] = cvt<T>(-1.0925484305920792f * yz);
    o[6] = cvt<T>(0.94617469575755997f * z2 - 0.31539156525251999f);
    o[7] = cvt<T>(-1.0925484305920792f * xz);
    o[8] = cvt<T>(0.54627421529603959f * x2 - 0.54627421529603959f * y2);
    if (degree <= 3) return;
    o[9] = cvt<T>(0.59004358992664352f * y * (-3.0f * x2 + y2));
    o[10] = cvt<T>(2.8906114426405538f * xy * z);
    o[11] = cvt<T>(0.45704579946446572f * y * (1.0f - 5.0f * z2));
    o[12] = cvt<T>(0.3731763325901154f * z * (5.0f * z2 - 3.0f));
    o[13] = cvt<T>(0.45704579946446572f * x * (1.0f - 5.0f * z2));
    o[14] = cvt<T>(1.4453057213202769f * z * (x2 - y2));
    o[15] = cvt<T>(0.59004358992664352f * x * (-x2 + 3.0f * y2));
}

// grad_inputs[b, :] = sum_k grad[b, k] * d o_k / d (x, y, z): the analytic derivatives of the polynomials above (what the
// reference stores as dy_dx, shencoder.cu:122-250, and contracts in kernel_sh_backward :366-383), fused into one pass.
__global__ void __launch_bounds__(256) k_sh_encode_bwd(const float *__restrict__ grad, const float *__restrict__ inputs, float *__restrict__ grad_inputs,
                                                       uint32_t B, uint32_t degree, uint32_t grad_stride) {
    const uint32_t b = threadIdx.x + blockIdx.x * blockDim.x;
    if (b >= B) return;
    const float x = inputs[3 * (size_t)b], y = inputs[3 * (size_t)b + 1], z = inputs[3 * (size_t)b + 2];
    const float *g = grad + (size_t)b * grad_stride;
    float gx = 0.f, gy = 0.f, gz = 0.f;
    if (degree > 1) {
        constexpr float c1 = 0.48860251190291987f;
        gy += g[1] * -c1;
        gz += g[2] * c1;
        gx += g[3] * -c1;
    }
    if (degree > 2) {
        constexpr float c2 = 1.0925484305920792f, a = 0.94617469575755997f, e = 0.54627421529603959f;
        gx += g[4] * (c2 * y); gy += g[4] * (c2 * x);
        gy += g[5] * (-c2 * z); gz += g[5] * (-c2 * y);
        gz += g[6] * (2.0f * a * z);
        gx += g[7] * (-c2 * z); gz += g[7] * (-c2 * x);
        gx += g[8] * (2.0f * e * x); gy += g[8] * (-2.0f * e * y);
    }
    if (degree > 3) {
        constexpr float f = 0.59004358992664352f, gg = 2.8906114426405538f, h = 0.45704579946446572f, k = 0.3731763325901154f, m = 1.4453057213202769f;
        const float x2 = x * x, y2 = y * y, z2 = z * z;
        gx += g[9] * (-6.0f * f * x * y); gy += g[9] * (f * (-3.0f * x2 + 3.0f * y2));
        gx += g[10] * (gg * y * z); gy += g[10] * (gg * x * z); gz += g[10] * (gg * x * y);
        gy += g[11] * (h * (1.0f - 5.0f * z2)); gz += g[11] * (-10.0f * h * y * z);
        gz += g[12] * (k * (15.0f * z2 - 3.0f));
        gx += g[13] * (h * (1.0f - 5.0f * z2)); gz += g[13] * (-10.0f * h * x * z);
        gx += g[14] * (2.0f * m * x * z); gy += g[14] * (-2.0f * m * y * z); gz += g[14] * (m * (x2 - y2));
        gx += g[15] * (f * (-3.0f * x2 + 3.0f * y2)); gy += g[15] * (6.0f * f * x * y);
    }
    grad_inputs[3 * (size_t)b] = gx; grad_inputs[3 * (size_t)b + 1] = gy; grad_inputs[3 * (size_t)b + 2] = gz;
}

}  // namespace
}  // namespace ngp

extern "C" int ngp_sh_encode_backward(const float *grad, const float *inputs, float *grad_inputs, uint32_t B, uint32_t degree,
                                      uint32_t grad_stride, void *stream) {
    using namespace ngp;
    NGP_CHECK_ARG(grad && inputs && grad_inputs, "sh_encode_backward: null pointer argument");
    NGP_CHECK_ARG(degree >= 1 && degree <= 4, "sh_encode_backward: degree must be in [1, 4], got %u", degree);
    NGP_CHECK_ARG(grad_stride >= degree * degree, "sh_encode_backward: grad_stride %u < degree^2", grad_stride);
    if (B == 0) return NGP_OK;
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_sh_encode_bwd", st); k_sh_encode_bwd<<<div_round_up(B, 256u), 256, 0, st>>>(grad, inputs, grad_inputs, B, degree, grad_stride); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

extern "C" int ngp_sh_encode_forward(const float *inputs, void *outputs, uint32_t B, uint32_t degree, uint32_t out_stride,
                                     int dtype, void *stream) {
    using namespace ngp;
    NGP_CHECK_ARG(inputs && outputs, "sh_encode_forward: null pointer argument");
    NGP_CHECK_ARG(degree >= 1 && degree <= 4, "sh_encode_forward: degree must be in [1, 4], got %u", degree);
    NGP_CHECK_ARG(out_stride >= degree * degree, "sh_encode_forward: out_stride %u < degree^2", out_stride);
    if (B == 0) return NGP_OK;
    cudaStream_t st = (cudaStream_t)stream;
    switch (dtype) {
        case NGP_DTYPE_F32: { NGP_PROF("k_sh_encode", st); k_sh_encode<float><<<div_round_up(B, 256u), 256, 0, st>>>(inputs, (float *)outputs, B, degree, out_stride); } break;
        case NGP_DTYPE_F16: { NGP_PROF("k_sh_encode", st); k_sh_encode<__half><<<div_round_up(B, 256u), 256, 0, st>>>(inputs, (__half *)outputs, B, degree, out_stride); } break;
        default: set_error("sh_encode_forward: outputs must be float or half (dtype code %d)", dtype); return NGP_EINVAL;
    }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}
