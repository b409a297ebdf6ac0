// shenc.cu -- spherical-harmonics direction encoding, degree 1..8 ("next" row of the scope table).
// Contract: /root/reference/shencoder/src/shencoder.cu:27-120 (polynomial form, same constants) for bands l < 4; the bands
// l = 4..7 (degree 5..8, :70-119 and their derivative table :122-355) are evaluated from the closed form those tables were
// generated from instead of 48 + 144 hand-expanded polynomials:
//     Y_l^{+m} = N_l^m Q_l^m(z) Re (x+iy)^m,   Y_l^{-m} = N_l^m Q_l^m(z) Im (x+iy)^m,   index l^2 + l +- m,
//     Q_l^m(z) = (-1)^m d^m P_l / dz^m  (Q_m^m = (-1)^m (2m-1)!!, Q_{m+1}^m = (2m+1) z Q_m^m,
//                                        (l-m) Q_l^m = (2l-1) z Q_{l-1}^m - (l+m-1) Q_{l-2}^m),
//     N_l^m = sqrt((2l+1)/(4 pi) (l-m)!/(l+m)!) * (m > 0 ? sqrt 2 : 1).
// These are the same polynomials in (x, y, z) as the reference's (also off the unit sphere: each of its entries is a
// polynomial in z times Re/Im (x+iy)^m), so values and derivatives agree to fp32 rounding of the evaluation order.
// One thread per direction; writes degree^2 values at row stride `out_stride` so the caller can
// encode straight into a wider (padded) MLP input matrix.
#include "common.cuh"

namespace ngp {
namespace {

template <typename T> __device__ __forceinline__ T cvt(float v) { return (T)v; }
template <> __device__ __forceinline__ __half cvt<__half>(float v) { return __float2half_rn(v); }

constexpr float kShNorm[8][8] = {
    {0.28209479177387814f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f},
    {0.4886025119029199f, 0.48860251190292f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f},
    {0.6307831305050401f, 0.36418281019735976f, 0.18209140509867988f, 0.f, 0.f, 0.f, 0.f, 0.f},
    {0.7463526651802308f, 0.3046971996429772f, 0.09635371475468515f, 0.03933623932844291f, 0.f, 0.f, 0.f, 0.f},
    {0.8462843753216345f, 0.2676186174229157f, 0.063078313050504f, 0.016858388283618388f, 0.005960340337611203f, 0.f, 0.f, 0.f},
    {0.9356025796273888f, 0.2415715473043717f, 0.045652731285460234f, 0.009318824751147628f, 0.002196468058075176f, 0.0006945841871324552f, 0.f, 0.f},
    {1.0171072362820548f, 0.22195099524523101f, 0.03509353369580661f, 0.005848922282634435f, 0.0010678622237644956f, 0.00022766899107568562f, 6.57223766418388e-05f, 0.f},
    {1.0925484305920792f, 0.20647224590289676f, 0.028097313806030647f, 0.003973560225074135f, 0.0005990367431114117f, 9.983945718523528e-05f, 1.958012847746254e-05f, 5.233009453691466e-06f},
};
template <int L, int M> struct sh_norm { static constexpr float value = kShNorm[L][M]; };

// bands l = 4 .. D-1: emit(k, Y_k, dY_k/dx, dY_k/dy, dY_k/dz); fully unrolled (every (l, m) is a compile-time pair)
template <int D, int M, int L, typename Emit>
__device__ __forceinline__ void sh_band_l(float z, float cm, float sm, float cp, float sp, float qmm, float q1, float q2, float d1, float d2, Emit &emit) {
    if constexpr (L < D) {
        float q, dq;
        if constexpr (L == M) { q = qmm; dq = 0.f; }
        else if constexpr (L == M + 1) { q = (float)(2 * M + 1) * z * qmm; dq = (float)(2 * M + 1) * qmm; }
        else {
            constexpr float r = 1.0f / (float)(L - M);
            q = ((float)(2 * L - 1) * z * q1 - (float)(L + M - 1) * q2) * r;
            dq = ((float)(2 * L - 1) * (q1 + z * d1) - (float)(L + M - 1) * d2) * r;
        }
        if constexpr (L >= 4) {
            constexpr float n = sh_norm<L, M>::value;
            const float nq = n * q, ndq = n * dq;
            if constexpr (M == 0) emit(L * L + L, nq, 0.f, 0.f, ndq);
            else {
                emit(L * L + L + M, nq * cm, nq * (float)M * cp, -nq * (float)M * sp, ndq * cm);
                emit(L * L + L - M, nq * sm, nq * (float)M * sp, nq * (float)M * cp, ndq * sm);
            }
        }
        sh_band_l<D, M, L + 1>(z, cm, sm, cp, sp, qmm, q, q1, dq, d1, emit);
    }
}
template <int D, int M, typename Emit>
__device__ __forceinline__ void sh_order_m(float x, float y, float z, float cm, float sm, float qmm, Emit &emit) {
    if constexpr (M < D) {
        float cp = 0.f, sp = 0.f;
        if constexpr (M > 0) {  // (x+iy)^m from (x+iy)^(m-1)
            cp = cm; sp = sm;
            cm = x * cp - y * sp; sm = x * sp + y * cp;
            qmm *= -(float)(2 * M - 1);
        }
        sh_band_l<D, M, M>(z, cm, sm, cp, sp, qmm, 0.f, 0.f, 0.f, 0.f, emit);
        sh_order_m<D, M + 1>(x, y, z, cm, sm, qmm, emit);
    }
}
template <int D, typename Emit> __device__ __forceinline__ void sh_high_bands(float x, float y, float z, Emit emit) {
    sh_order_m<D, 0>(x, y, z, 1.f, 0.f, 1.f, emit);
}

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
    if (degree <= 4) return;
    auto emit = [&](int k, float v, float, float, float) { o[k] = cvt<T>(v); };
    switch (degree) {
        case 5: sh_high_bands<5>(x, y, z, emit); break;
        case 6: sh_high_bands<6>(x, y, z, emit); break;
        case 7: sh_high_bands<7>(x, y, z, emit); break;
        default: sh_high_bands<8>(x, y, z, emit); break;
    }
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
    if (degree > 4) {
        auto emit = [&](int k, float, float dx, float dy, float dz) { const float gk = g[k]; gx += gk * dx; gy += gk * dy; gz += gk * dz; };
        switch (degree) {
            case 5: sh_high_bands<5>(x, y, z, emit); break;
            case 6: sh_high_bands<6>(x, y, z, emit); break;
            case 7: sh_high_bands<7>(x, y, z, emit); break;
            default: sh_high_bands<8>(x, y, z, emit); break;
        }
    }
    grad_inputs[3 * (size_t)b] = gx; grad_inputs[3 * (size_t)b + 1] = gy; grad_inputs[3 * (size_t)b + 2] = gz;
}

}  // namespace
}  // namespace ngp

extern "C" int ngp_sh_encode_backward(const float *grad, const float *inputs, float *grad_inputs, uint32_t B, uint32_t degree,
                                      uint32_t grad_stride, void *stream) {
    using namespace ngp;
    NGP_CHECK_ARG(grad && inputs && grad_inputs, "sh_encode_backward: null pointer argument");
    NGP_CHECK_ARG(degree >= 1 && degree <= 8, "sh_encode_backward: degree must be in [1, 8], got %u", degree);
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
    NGP_CHECK_ARG(degree >= 1 && degree <= 8, "sh_encode_forward: degree must be in [1, 8], got %u", degree);
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
