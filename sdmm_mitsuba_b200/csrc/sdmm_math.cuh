// sdmm_math.cuh -- per-component scalar device math shared by the M-step, prepare and conditioner
// kernels.  Everything here runs once per component per EM iteration (O(K) work per cell), so it
// is written for fidelity to the reference's formulas, not for throughput: IEEE sqrt/div, accurate
// sinf/cosf.  The per-(sample,component) hot loop lives in em_kernels.cu.
//
// "R:" paths are relative to /root/reference/sdmm/include/sdmm/.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace sdmmb {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kCholeskyJitter = 1e-5f;           // R: linalg/cholesky.h:32
constexpr float kInvSqrt2Pi = 0.39894228040143267793994605993438f;  // R: core/constants.h:12-14

__host__ __device__ inline float gaussian_normalization(int D) {    // R: core/constants.h:45-47
    float r = 1.f;
    for (int i = 0; i < D; ++i) r *= kInvSqrt2Pi;
    return r;
}

// R: linalg/cholesky.h:12-36.  Row-major DxD, +1e-5 on every pivot, PSD flag = all jittered
// pivots > 0.  sqrt of a negative pivot gives NaN exactly like enoki::sqrt.
template <int D>
__device__ inline bool cholesky(const float* in, float* out) {
    bool psd = true;
#pragma unroll
    for (int i = 0; i < D * D; ++i) out[i] = 0.f;
#pragma unroll
    for (int r = 0; r < D; ++r) {
#pragma unroll
        for (int c = 0; c < r; ++c) {
            float ksum = 0.f;
#pragma unroll
            for (int k = 0; k < c; ++k) ksum += out[r * D + k] * out[c * D + k];
            out[r * D + c] = (in[r * D + c] - ksum) / out[c * D + c];
        }
        float row_sum = 0.f;
#pragma unroll
        for (int k = 0; k < r; ++k) row_sum += out[r * D + k] * out[r * D + k];
        float value = in[r * D + r] - row_sum + kCholeskyJitter;
        psd = psd && (value > 0.f);
        out[r * D + r] = sqrtf(value);
    }
    return psd;
}

// R: linalg/cholesky.h:38-58 (forward substitution) and :60-78 (inverse = solve per unit vector)
template <int D>
__device__ inline void inverse_lower_tri(const float* L, float* inv) {
#pragma unroll
    for (int i = 0; i < D * D; ++i) inv[i] = 0.f;
#pragma unroll
    for (int col = 0; col < D; ++col) {
        float x[D];
#pragma unroll
        for (int r = 0; r < D; ++r) {
            float num = (r == col) ? 1.f : 0.f;
#pragma unroll
            for (int c = 0; c < r; ++c) num -= L[r * D + c] * x[c];
            x[r] = num / L[r * D + r];
        }
#pragma unroll
        for (int r = 0; r < D; ++r) inv[r * D + col] = x[r];
    }
}

// R: linalg/coordinate_system.h:21-38 (Duff et al. branch-free ONB).  from columns = (s, t, n),
// to = transpose(from); both row-major.
__device__ inline void frame_from_normal(const float n[3], float to[9], float from[9]) {
    float sign = copysignf(1.f, n[2]);
    float a = -1.f / (sign + n[2]);
    float b = n[0] * n[1] * a;
    float c0[3] = {(n[0] * n[0] * a) * sign + 1.f, b * sign, n[0] * copysignf(1.f, -n[2])};
    float c1[3] = {b, sign + n[1] * n[1] * a, -n[1]};
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        from[r * 3 + 0] = c0[r];
        from[r * 3 + 1] = c1[r];
        from[r * 3 + 2] = n[r];
    }
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) to[r * 3 + c] = from[c * 3 + r];
}

// R: spaces/spatio_directional.h:247-255 set_mean: frame from the NORMALISED direction part.
__device__ inline void frame_from_mean_dir(const float d[3], float to[9], float from[9]) {
    float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    float inv = 1.f / len;
    float n[3] = {d[0] * inv, d[1] * inv, d[2] * inv};
    frame_from_normal(n, to, from);
}

template <int N, int M, int P>
__device__ inline void matmul(const float* A, const float* B, float* C) {   // R: linalg/matrix.h:139-161
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
        for (int j = 0; j < P; ++j) {
            float s = 0.f;
#pragma unroll
            for (int k = 0; k < M; ++k) s += A[i * M + k] * B[k * P + j];
            C[i * P + j] = s;
        }
}

// C = J A J^T for 5x5 (left to right, as `jacobian * cov * transpose(jacobian)`, R: opt/em.h:475-476)
__device__ inline void congruence5(const float* J, const float* A, float* C) {
    float tmp[25];
    matmul<5, 5, 5>(J, A, tmp);
#pragma unroll
    for (int i = 0; i < 5; ++i)
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            float s = 0.f;
#pragma unroll
            for (int k = 0; k < 5; ++k) s += tmp[i * 5 + k] * J[j * 5 + k];
            C[i * 5 + j] = s;
        }
}

// R: spaces/spatio_directional.h:173-245 from_jacobian (exp map at the CURRENT frame + its 6x5
// Jacobian) composed with :145-169 to_center_jacobian of the NEW frame -> the 5x5 transport
// Jacobian of opt/em.h:470-474.  Only the 2x2 directional block is non-trivial:
//   J = [ I3 0 ; 0 to_new[0:2,:] * from_old * dj ]   with dj the 3x2 exp-map derivative.
// Outputs the new embedded mean (6) and J (5x5 row-major).
__device__ inline void recentre_jacobian(const float mean_old[6], const float from_old[9],
                                         const float t[5], float mean_new[6], float to_new[9],
                                         float from_new[9], float J[25]) {
    const float cos_angle_min = -0.98f;
    float tx = t[3], ty = t[4];
    float length_sqr = tx * tx + ty * ty;
    float length = sqrtf(length_sqr);
    float sin_angle, cos_angle;
    sincosf(length, &sin_angle, &cos_angle);
    bool degenerate = (sin_angle < 1e-2f) || (cos_angle < cos_angle_min);
    float sinc_angle = degenerate ? 1.f : sin_angle / length;
    float cmsol = degenerate ? 0.f : (cos_angle - sinc_angle) / length_sqr;
    float dj[6];
    dj[0] = sinc_angle + tx * tx * cmsol;
    dj[3] = sinc_angle + ty * ty * cmsol;
    dj[1] = tx * ty * cmsol;
    dj[2] = dj[1];
    dj[4] = degenerate ? 0.f : -tx * sinc_angle;
    dj[5] = degenerate ? 0.f : -ty * sinc_angle;
    float fdj[6];
    matmul<3, 3, 2>(from_old, dj, fdj);
    float local[3] = {tx * sinc_angle, ty * sinc_angle, cos_angle};
#pragma unroll
    for (int d = 0; d < 3; ++d) mean_new[d] = t[d] + mean_old[d];
#pragma unroll
    for (int r = 0; r < 3; ++r)
        mean_new[3 + r] = from_old[r * 3 + 0] * local[0] + from_old[r * 3 + 1] * local[1] +
                          from_old[r * 3 + 2] * local[2];
    frame_from_mean_dir(mean_new + 3, to_new, from_new);
#pragma unroll
    for (int i = 0; i < 25; ++i) J[i] = 0.f;
    J[0] = J[6] = J[12] = 1.f;
    // rows 3,4 x cols 3,4: (to_new rows 0,1) (3) . fdj (3x2); position block of from_jacobian is I3
    // and the direction rows of to_center_jacobian have zero position columns, so the full 5x6 * 6x5
    // product reduces to this (the zero terms the reference adds do not change float results).
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            float s = 0.f;
#pragma unroll
            for (int k = 0; k < 3; ++k) s += to_new[r * 3 + k] * fdj[k * 2 + c];
            J[(3 + r) * 5 + (3 + c)] = s;
        }
}

}  // namespace sdmmb
