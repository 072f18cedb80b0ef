// fmpnp_device.cuh -- float64 device math of the feature-metric PnP refinement (sm_100a).
//
// Compiled with -fmad=false: every fused multiply-add in this library is an explicit fma(), so
// the rounding of the projection (which decides integer pixel indices) is fully specified.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace fmpnp {

constexpr double kEpsF32 = 1.1920928955078125e-07;  // torch.finfo(torch.float32).eps

// ---------------------------------------------------------------------------------------------
// Projection (featurePnP/model.py:303-308):  P = R X + t ;  p = K P ;  px = int32(round(p.xy / p.z)) - 1
// Rounding order follows what torch.mm produces for the reference's operand layouts: R X is a
// plain multiply / add chain (k = 0,1,2), K P is the fma chain fma(K2, Z, fma(K1, Y, K0 * X)).
// round = half-to-even (rint); double -> int32 as x86 cvttsd2si (out of range / NaN -> INT32_MIN);
// the "- 1" wraps in int32.
// ---------------------------------------------------------------------------------------------
struct Proj {
    double X, Y, Z;
    int32_t px, py;
    bool supported;
};

__device__ __forceinline__ int32_t cast_i32_x86(double r) {
    // cvt.rzi.s32.f64 saturates: everything at or below -2^31 already gives INT32_MIN; what is left to redirect are the
    // values at or above 2^31 (would saturate to INT32_MAX) and NaN (would give 0) -- one comparison that both fail
    return r < 2147483648.0 ? __double2int_rz(r) : INT32_MIN;
}

// P = R X + t alone (the same multiply / add chain as in project_point)
__device__ __forceinline__ void transform_point(const double* __restrict__ R, const double* __restrict__ t, double x, double y,
                                                double z, double& X, double& Y, double& Z) {
    X = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(R[0], x), __dmul_rn(R[1], y)), __dmul_rn(R[2], z)), t[0]);
    Y = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(R[3], x), __dmul_rn(R[4], y)), __dmul_rn(R[5], z)), t[1]);
    Z = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(R[6], x), __dmul_rn(R[7], y)), __dmul_rn(R[8], z)), t[2]);
}

__device__ __forceinline__ Proj project_point(const double* __restrict__ R, const double* __restrict__ t,
                                               const double* __restrict__ K, double x, double y, double z,
                                               int32_t im_w, int32_t im_h) {
    Proj o;
    o.X = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(R[0], x), __dmul_rn(R[1], y)), __dmul_rn(R[2], z)), t[0]);
    o.Y = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(R[3], x), __dmul_rn(R[4], y)), __dmul_rn(R[5], z)), t[1]);
    o.Z = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(R[6], x), __dmul_rn(R[7], y)), __dmul_rn(R[8], z)), t[2]);
    const double q0 = fma(K[2], o.Z, fma(K[1], o.Y, __dmul_rn(K[0], o.X)));
    const double q1 = fma(K[5], o.Z, fma(K[4], o.Y, __dmul_rn(K[3], o.X)));
    const double q2 = fma(K[8], o.Z, fma(K[7], o.Y, __dmul_rn(K[6], o.X)));
    const double u = q0 / q2;  // IEEE-754 correctly rounded division
    const double v = q1 / q2;
    o.px = (int32_t)((uint32_t)cast_i32_x86(rint(u)) - 1u);
    o.py = (int32_t)((uint32_t)cast_i32_x86(rint(v)) - 1u);
    // points_within_image, pad = 0 (model.py:99-117)
    o.supported = (o.px >= 0) && (o.py >= 0) && (o.px < im_w) && (o.py < im_h);
    return o;
}

// indexing_ (model.py:74-97): row = floor(double(y) * Hf / height), col = floor(double(x) * Wf / width).
// For the non-negative integers of a supported point the float64 expression equals integer
// floor division exactly (the quotient is either an integer or at least 1/height away from one).
__device__ __forceinline__ int32_t cell_index(int32_t px, int32_t py, int32_t Hf, int32_t Wf, int32_t im_w,
                                              int32_t im_h) {
    // (called for supported points only: 0 <= px < im_w, 0 <= py < im_h.)  64-bit integer division is a ~90-instruction
    // dependent chain on the GPU; the products fit 32 bits for every realistic size, and then a 32-bit division
    // gives the same integers.  The test is uniform across a query.
    if ((int64_t)im_h * Hf <= 0x7fffffffll && (int64_t)im_w * Wf <= 0x7fffffffll) {
        const uint32_t row = ((uint32_t)py * (uint32_t)Hf) / (uint32_t)im_h;
        const uint32_t col = ((uint32_t)px * (uint32_t)Wf) / (uint32_t)im_w;
        return (int32_t)(row * (uint32_t)Wf + col);
    }
    const int32_t row = (int32_t)(((int64_t)py * Hf) / im_h);
    const int32_t col = (int32_t)(((int64_t)px * Wf) / im_w);
    return row * Wf + col;
}

// The usual case, image = 2^k x feature map along an axis: x * f / (f << k) == x >> k for every non-negative x, a shift
// instead of a ~20-instruction division.  cell_shift returns k, or -1 when the axis needs the division; the callers
// compute it once per evaluation (it is uniform across a query).
__device__ __forceinline__ int32_t cell_shift(int32_t im, int32_t f) {
    const uint32_t k = (uint32_t)im / (uint32_t)f;
    if (k == 0u || k * (uint32_t)f != (uint32_t)im || (k & (k - 1u)) != 0u) return -1;
    return 31 - __clz((int)k);
}
__device__ __forceinline__ int32_t cell_index_shifted(int32_t px, int32_t py, int32_t Hf, int32_t Wf, int32_t im_w, int32_t im_h,
                                                      int32_t sx, int32_t sy) {
    if (sx >= 0 && sy >= 0) return (int32_t)(((uint32_t)py >> sy) * (uint32_t)Wf + ((uint32_t)px >> sx));
    return cell_index(px, py, Hf, Wf, im_w, im_h);
}

// ---------------------------------------------------------------------------------------------
// Robust losses (featurePnP/helpers/utils.py:15-78): x = 0.5 * ||e||^2  ->  (rho, w = rho')
// ---------------------------------------------------------------------------------------------
// pow(b, e) for the exponents the common losses produce, as correctly rounded products / divisions / square roots
// instead of a pow() call (~200 instructions): Geman-McClure (alpha = -2) needs b^-1 and b^-2 per point.  The reference
// calls torch.pow(tensor, tensor) (utils.py:67-68; SLEEF, 1 ulp); these forms are within 1 ulp of the same value, like
// CUDA's pow() they replace (costs are compared at 1e-9 relative, masks only move at exact ties).
__device__ __forceinline__ double pow_scalar(double b, double e) {
    if (e == -1.0) return 1.0 / b;
    if (e == -2.0) return 1.0 / __dmul_rn(b, b);
    if (e == 2.0) return __dmul_rn(b, b);
    if (e == 3.0) return __dmul_rn(__dmul_rn(b, b), b);
    if (e == 0.5) return sqrt(b);
    if (e == -0.5) return 1.0 / sqrt(b);
    return pow(b, e);
}

__device__ __forceinline__ void robust_loss(int loss_id, double alpha, double x, double& rho, double& w) {
    switch (loss_id) {
        case 0:  // squared_loss :15-17
            rho = x;
            w = 1.0;
            return;
        case 1: {  // huber_loss :19-29 (rho = 1)
            const bool m = x <= 1.0;
            const double sx = sqrt(x);
            const double isx = fmax(kEpsF32, 1.0 / sx);
            rho = m ? x : __dadd_rn(__dmul_rn(2.0, sx), -1.0);
            w = m ? 1.0 : isx;
            return;
        }
        case 2: alpha = 0.0; break;   // cauchy_loss :31-34
        case 3: alpha = -2.0; break;  // geman_mcclure_loss :36-38
        default: break;               // barron_loss(alpha)
    }
    if (alpha == 0.0) {
        rho = __dmul_rn(2.0, log1p(fmin(__dmul_rn(0.5, x), 33e37)));
        w = 2.0 / __dadd_rn(x, 2.0);
    } else if (alpha == 2.0) {
        rho = x;
        w = 1.0;
    } else {
        const double beta_safe = fmax(kEpsF32, fabs(__dadd_rn(alpha, -2.0)));
        const double alpha_safe = (alpha >= 0.0 ? 1.0 : -1.0) * fmax(kEpsF32, fabs(alpha));
        const double base = __dadd_rn(x / beta_safe, 1.0);
        rho = __dmul_rn(__dmul_rn(2.0, beta_safe / alpha_safe), __dadd_rn(pow_scalar(base, __dmul_rn(0.5, alpha)), -1.0));
        w = pow_scalar(base, __dadd_rn(__dmul_rn(0.5, alpha), -1.0));
    }
}

// ---------------------------------------------------------------------------------------------
// Per-point contraction of the reduced quantities with the 2x6 projection Jacobian.
//   J_n = G_n A_n with G_n = [Gx_n | Gy_n] (C x 2) and A_n = J_px_p J_p_T (2 x 6)  (model.py:369-394)
//   =>  J^T J = A^T S A  with S = G^T G = [[Sxx,Sxy],[Sxy,Syy]],   J^T e = A^T b  with b = G^T e.
// acc[0..20] += w * upper triangle of A^T S A (row-major i <= j), acc[21..26] += w * A^T b.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void contract_point(double fx, double fy, double X, double Y, double Z, double w,
                                               double Sxx, double Sxy, double Syy, double bx, double by,
                                               double* __restrict__ acc) {
    // J_px_p (model.py:377-382): [[fx, 0, -fx X / Z], [0, fy, -fy Y / Z]] / Z.  One division (1 / Z) and
    // multiplications instead of the reference's six divisions: a float64 division is a ~30-instruction dependent
    // chain, and this function runs for every supported point of every evaluation.  The entries differ from the
    // quotients by at most 2 ulp -- the same size as the re-association of J = G A that this path has anyway; no
    // integer (pixel index, mask) depends on them except through the pose, like on every other rounding of H, g.
    const double rZ = 1.0 / Z;
    const double j00 = __dmul_rn(fx, rZ);
    const double j02 = __dmul_rn(__dmul_rn(__dmul_rn(-fx, X), rZ), rZ);
    const double j11 = __dmul_rn(fy, rZ);
    const double j12 = __dmul_rn(__dmul_rn(__dmul_rn(-fy, Y), rZ), rZ);
    // A = J_px_p [I | -skew(P)],  -skew(P) = [[0, Z, -Y], [-Z, 0, X], [Y, -X, 0]]
    double ax[6], ay[6];
    ax[0] = j00;  ax[1] = 0.0;  ax[2] = j02;
    ax[3] = __dmul_rn(j02, Y);
    ax[4] = fma(j00, Z, -__dmul_rn(j02, X));
    ax[5] = -__dmul_rn(j00, Y);
    ay[0] = 0.0;  ay[1] = j11;  ay[2] = j12;
    ay[3] = fma(j12, Y, -__dmul_rn(j11, Z));
    ay[4] = -__dmul_rn(j12, X);
    ay[5] = __dmul_rn(j11, X);
    const double wxx = __dmul_rn(w, Sxx), wxy = __dmul_rn(w, Sxy), wyy = __dmul_rn(w, Syy);
    double u[6], v[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        u[i] = fma(wxx, ax[i], __dmul_rn(wxy, ay[i]));
        v[i] = fma(wxy, ax[i], __dmul_rn(wyy, ay[i]));
    }
    int k = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = i; j < 6; ++j) {
            acc[k] = fma(ax[i], u[j], fma(ay[i], v[j], acc[k]));
            ++k;
        }
    const double wbx = __dmul_rn(w, bx), wby = __dmul_rn(w, by);
#pragma unroll
    for (int i = 0; i < 6; ++i) acc[21 + i] = fma(ax[i], wbx, fma(ay[i], wby, acc[21 + i]));
}

// ---------------------------------------------------------------------------------------------
// so3exp_map (helpers/utils.py:209-221) and the pose update R' = dr R, t' = dr t + dt
// (model.py:416-426). Single thread.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void so3exp(const double* w, double* dr) {
    const double theta = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(w[0], w[0]), __dmul_rn(w[1], w[1])), __dmul_rn(w[2], w[2])));
    const double a = w[0] / theta, b = w[1] / theta, c = w[2] / theta;
    const double W[9] = {0.0, -c, b, c, 0.0, -a, -b, a, 0.0};
    double WW[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            WW[3 * i + j] = __dadd_rn(__dadd_rn(__dmul_rn(W[3 * i], W[j]), __dmul_rn(W[3 * i + 1], W[3 + j])),
                                      __dmul_rn(W[3 * i + 2], W[6 + j]));
    const double s = sin(theta), c1 = __dadd_rn(1.0, -cos(theta));
#pragma unroll
    for (int i = 0; i < 9; ++i) {
        double res = __dadd_rn(__dmul_rn(W[i], s), __dmul_rn(WW[i], c1));
        if (theta < 1e-12) res = 0.0;  // torch.where(theta < 1e-12, 0, res): also removes the 0/0 NaNs
        dr[i] = __dadd_rn((i % 4) == 0 ? 1.0 : 0.0, res);
    }
}

__device__ __forceinline__ void compose_pose(const double* dr, const double* dt, const double* R, const double* t,
                                             double* Rn, double* tn) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j)
            Rn[3 * i + j] = __dadd_rn(__dadd_rn(__dmul_rn(dr[3 * i], R[j]), __dmul_rn(dr[3 * i + 1], R[3 + j])),
                                      __dmul_rn(dr[3 * i + 2], R[6 + j]));
        tn[i] = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(dr[3 * i], t[0]), __dmul_rn(dr[3 * i + 1], t[1])),
                                    __dmul_rn(dr[3 * i + 2], t[2])),
                          dt[i]);
    }
}

// ---------------------------------------------------------------------------------------------
// Warp-level 6x6 solve of the damped normal equations (optimizer_step, model.py:37-72).
// A (6x6, row-major, shared memory) and b (6, shared memory) are overwritten; x -> b.
// Executed by ONE full warp.  Returns false if the Cholesky factorisation broke down
// (non-positive / NaN pivot).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool warp_cholesky_solve(double* A, double* b, int lane) {
    bool ok = true;
    for (int j = 0; j < 6; ++j) {
        const double d = A[7 * j];
        if (!(d > 0.0)) ok = false;  // uniform: every lane reads the same value
        const double ljj = sqrt(d);
        __syncwarp();
        if (lane == 0) A[7 * j] = ljj;
        if (lane > j && lane < 6) A[6 * lane + j] = A[6 * lane + j] / ljj;
        __syncwarp();
        for (int e = lane; e < 36; e += 32) {
            const int i = e / 6, k = e % 6;
            if (k > j && i >= k) A[e] = fma(-A[6 * i + j], A[6 * k + j], A[e]);
        }
        __syncwarp();
    }
    if (!ok) return false;
    // L y = b (forward), then L^T x = y (backward); lanes 0..5 own one unknown each
    for (int j = 0; j < 6; ++j) {
        if (lane == 0) b[j] = b[j] / A[7 * j];
        __syncwarp();
        if (lane > j && lane < 6) b[lane] = fma(-A[6 * lane + j], b[j], b[lane]);
        __syncwarp();
    }
    for (int j = 5; j >= 0; --j) {
        if (lane == 0) b[j] = b[j] / A[7 * j];
        __syncwarp();
        if (lane < j) b[lane] = fma(-A[6 * j + lane], b[j], b[lane]);
        __syncwarp();
    }
    return true;
}

// LU with partial pivoting in LAPACK getrf / getrs order (first largest |entry| of the column,
// multipliers formed with the reciprocal of the pivot).  A zero pivot is carried through and
// produces inf / NaN exactly like torch.lu_solve on a singular factorisation.
__device__ __forceinline__ void warp_lu_solve(double* A, double* b, int lane) {
    for (int j = 0; j < 6; ++j) {
        int p = j;
        double best = fabs(A[7 * j]);
        for (int i = j + 1; i < 6; ++i) {  // uniform across lanes (same shared values)
            const double v = fabs(A[6 * i + j]);
            if (v > best) {
                best = v;
                p = i;
            }
        }
        __syncwarp();
        if (p != j) {
            if (lane < 6) {
                const double tmp = A[6 * j + lane];
                A[6 * j + lane] = A[6 * p + lane];
                A[6 * p + lane] = tmp;
            } else if (lane == 6) {
                const double tb = b[j];
                b[j] = b[p];
                b[p] = tb;
            }
        }
        __syncwarp();
        const double piv = A[7 * j];
        if (piv != 0.0) {
            const double rp = 1.0 / piv;
            if (lane > j && lane < 6) A[6 * lane + j] = __dmul_rn(A[6 * lane + j], rp);
            __syncwarp();
            for (int e = lane; e < 36; e += 32) {
                const int i = e / 6, k = e % 6;
                if (i > j && k > j) A[e] = __dadd_rn(A[e], -__dmul_rn(A[6 * i + j], A[6 * j + k]));
            }
        }
        __syncwarp();
    }
    if (lane == 0) {  // getrs: unit-lower forward, then upper backward substitution (30 flops, serial)
        for (int i = 1; i < 6; ++i) {
            double s = b[i];
            for (int k = 0; k < i; ++k) s = __dadd_rn(s, -__dmul_rn(A[6 * i + k], b[k]));
            b[i] = s;
        }
        for (int i = 5; i >= 0; --i) {
            double s = b[i];
            for (int k = i + 1; k < 6; ++k) s = __dadd_rn(s, -__dmul_rn(A[6 * i + k], b[k]));
            b[i] = s / A[7 * i];
        }
    }
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// PTX wrappers: cp.async (SASS LDGSTS) staging for sm_100a.  The TMA bulk-copy (cp.async.bulk + mbarrier)
// variants of the streaming loop measured slower for this access pattern (one 0.5-4 KB row per copy, a
// different row every time): see DESIGN.md section 4.1.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// 16-byte asynchronous copy global -> shared (SASS LDGSTS.E.BYPASS.128), cached in L2 only
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async16_s(uint32_t smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gmem_src) : "memory");
}
// L2 eviction policies (createpolicy): the rows a gather streams are read once -- evict_first keeps them from displacing
// what IS read again, the per-point records and coordinates (evict_last), which the light evaluations of a launch with
// record reuse read from L2 twenty times over.
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_normal() {
    uint64_t p;
    asm("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
    return p;
}
// 16-byte asynchronous copies with an L2 policy
__device__ __forceinline__ void cp_async16_pol(uint32_t smem_dst, const void* gmem_src, uint64_t pol) {
    asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(smem_dst), "l"(gmem_src), "l"(pol) : "memory");
}
template <int OFF>
__device__ __forceinline__ void cp_async16_pol_off(uint32_t smem_dst, const void* gmem_src, uint64_t pol) {
    asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0+%2], [%1+%2], 16, %3;" ::"r"(smem_dst), "l"(gmem_src), "n"(OFF), "l"(pol)
                 : "memory");
}
// 8-byte asynchronous copy (rows of three doubles are only 8-byte aligned); .ca is the only cache operator for this size
__device__ __forceinline__ void cp_async8_s(uint32_t smem_dst, const void* gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_dst), "l"(gmem_src) : "memory");
}
// same, both addresses displaced by the compile-time OFF (one address computation serves a whole row)
template <int OFF>
__device__ __forceinline__ void cp_async16_off(uint32_t smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0+%2], [%1+%2], 16;" ::"r"(smem_dst), "l"(gmem_src), "n"(OFF) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
// base + idx * pitch as ONE instruction (IMAD.WIDE.U32); written in PTX so that the compiler does not share the
// product between rows and then pay two adds per row for the different bases
__device__ __forceinline__ const char* row_address(const char* base, uint32_t idx, uint32_t pitch) {
    unsigned long long r;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(idx), "r"(pitch), "l"((unsigned long long)base));
    return reinterpret_cast<const char*>(r);
}
__device__ __forceinline__ void sts64(uint32_t addr, double v) {
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ double2 lds128d(uint32_t addr) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ double lds64d(uint32_t addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// waits until at most `n` of this thread's commit groups are still pending (n is uniform)
__device__ __forceinline__ void cp_async_wait_pending(int n) {
    switch (n) {
        case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
        case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
        case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
        case 3: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
        case 4: asm volatile("cp.async.wait_group 4;" ::: "memory"); break;
        case 5: asm volatile("cp.async.wait_group 5;" ::: "memory"); break;
        case 6: asm volatile("cp.async.wait_group 6;" ::: "memory"); break;
        default: asm volatile("cp.async.wait_group 7;" ::: "memory"); break;
    }
}

__device__ __forceinline__ double shfl_xor_f64(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }

}  // namespace fmpnp
