// fmpnp_adapter.cu -- the steps either side of the refinement loop, kept on the device:
//   * sobel_pack_kernel: reference `sobel_filter` (featurePnP/helpers/utils.py:81-104, i.e.
//     kornia.filters.spatial_gradient; spec of the operator in helpers/sobel_pytorch.py:9-59) fused
//     with the CHW -> HWC float32 layout pack: one pass over the query hypercolumn produces the
//     three maps the refinement kernel gathers from (F, Gx, Gy);
//   * gather_chw_rows_kernel: the reference-descriptor lookup of the adapter
//     (s2dhm/pose_prediction/optimize_feature_pnp.py:56): feature_ref[n] = hypercolumn[:, i_n, j_n];
//   * gauss_blur_chw_kernel / bilinear_resize_chw_kernel: the optional per-level blur and resize of
//     multilevel_optimization (featurePnP/model.py:196-200).
// All arithmetic is float64 on float32 / float64 inputs, rounded once to the float32 storage.
#include <cstdint>

#include "fmpnp_kernels.cuh"

namespace fmpnp {

// ---------------------------------------------------------------------------------------------
// Sobel + pack.  Block = 32 channels x (kTileY x 32) pixels.  The CHW tile (with a 1-pixel halo) is
// read coalesced along x into shared memory; each thread then owns one channel and walks the
// pixels, so the HWC stores are coalesced along c.
//   Gx = [[-1,0,1],[-2,0,2],[-1,0,1]] * F,  Gy = its transpose  (cross-correlation, as conv2d),
//   border: replicate (kornia) or zero (helpers/sobel_pytorch.py), optional 1/8 normalisation.
// ---------------------------------------------------------------------------------------------
constexpr int kTileY = 8;
constexpr int kTileX = 32;
constexpr int kChStride = (kTileY + 2) * (kTileX + 2) + 1;  // odd: conflict-free across channels

template <typename T>
__global__ void __launch_bounds__(256) sobel_pack_kernel(const T* __restrict__ src, float* __restrict__ F,
                                                         float* __restrict__ Gx, float* __restrict__ Gy, int C, int H,
                                                         int W, int zero_pad, int normalized) {
    extern __shared__ __align__(16) unsigned char tile_raw[];  // 32 x kChStride elements of T
    T* tile = reinterpret_cast<T*>(tile_raw);
    const int x0 = blockIdx.x * kTileX, y0 = blockIdx.y * kTileY, c0 = blockIdx.z * 32;
    const int tid = threadIdx.x;
    const int per_ch = (kTileY + 2) * (kTileX + 2);
    for (int e = tid; e < 32 * per_ch; e += 256) {
        const int ch = e / per_ch, rem = e - ch * per_ch;
        const int ry = rem / (kTileX + 2), rx = rem - ry * (kTileX + 2);
        int y = y0 + ry - 1, x = x0 + rx - 1;
        const int c = c0 + ch;
        T v = (T)0;
        const bool inside = y >= 0 && y < H && x >= 0 && x < W;
        if (c < C && (inside || !zero_pad)) {
            y = min(max(y, 0), H - 1);
            x = min(max(x, 0), W - 1);
            v = src[((size_t)c * H + y) * W + x];
        }
        tile[ch * kChStride + rem] = v;
    }
    __syncthreads();
    const int ch = tid & 31, c = c0 + ch;
    if (c >= C) return;
    const double scale = normalized ? 0.125 : 1.0;
    for (int p = tid >> 5; p < kTileY * kTileX; p += 8) {
        const int py = p / kTileX, px = p - py * kTileX;
        const int y = y0 + py, x = x0 + px;
        if (y >= H || x >= W) continue;
        auto at = [&](int dy, int dx) -> double {
            return (double)tile[ch * kChStride + (py + 1 + dy) * (kTileX + 2) + (px + 1 + dx)];
        };
        const double a = at(-1, -1), b = at(-1, 0), cc = at(-1, 1), d = at(0, -1), f = at(0, 1), g = at(1, -1),
                     h = at(1, 0), i = at(1, 1);
        // float64 sums of float32 values: differences of neighbours are exact unless their exponents
        // are > 29 bits apart, so the result does not depend on the summation order in practice
        const double gx = (cc - a) + 2.0 * (f - d) + (i - g);
        const double gy = (g - a) + 2.0 * (h - b) + (i - cc);
        const size_t o = ((size_t)y * W + x) * C + c;
        F[o] = (float)at(0, 0);
        Gx[o] = (float)(gx * scale);
        Gy[o] = (float)(gy * scale);
    }
}

cudaError_t launch_sobel_pack(const void* src, int src_is_f64, float* F, float* Gx, float* Gy, int C, int H, int W,
                              int zero_pad, int normalized, cudaStream_t stream) {
    dim3 grid((W + kTileX - 1) / kTileX, (H + kTileY - 1) / kTileY, (C + 31) / 32);
    if (src_is_f64) {
        const size_t dyn = (size_t)32 * kChStride * sizeof(double);
        cudaError_t e = cudaFuncSetAttribute(sobel_pack_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
        if (e != cudaSuccess) return e;
        sobel_pack_kernel<double><<<grid, 256, dyn, stream>>>(static_cast<const double*>(src), F, Gx, Gy, C, H, W,
                                                             zero_pad, normalized);
    } else {
        const size_t dyn = (size_t)32 * kChStride * sizeof(float);
        sobel_pack_kernel<float><<<grid, 256, dyn, stream>>>(static_cast<const float*>(src), F, Gx, Gy, C, H, W,
                                                            zero_pad, normalized);
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Descriptor gather: dst[n][c] = src[c][rows[n]][cols[n]]  (optimize_feature_pnp.py:51-56).
// One warp per point; lanes stride the channels, so the N x C float32 output is written coalesced.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void gather_chw_rows_kernel(const T* __restrict__ src, const int32_t* __restrict__ rows,
                                       const int32_t* __restrict__ cols, float* __restrict__ dst, int C, int H, int W,
                                       int N) {
    const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (n >= N) return;
    int i = rows[n], j = cols[n];
    if (i < 0) i += H;  // torch indexing: negative indices wrap once
    if (j < 0) j += W;
    const size_t pix = (size_t)i * W + j, hw = (size_t)H * W;
    for (int c = lane; c < C; c += 32) dst[(size_t)n * C + c] = (float)src[(size_t)c * hw + pix];
}

cudaError_t launch_gather_chw_rows(const void* src, int src_is_f64, const int32_t* rows, const int32_t* cols, float* dst,
                                   int C, int H, int W, int N, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    const int warps = 8;
    dim3 grid((N + warps - 1) / warps);
    if (src_is_f64)
        gather_chw_rows_kernel<double><<<grid, warps * 32, 0, stream>>>(static_cast<const double*>(src), rows, cols, dst,
                                                                         C, H, W, N);
    else
        gather_chw_rows_kernel<float><<<grid, warps * 32, 0, stream>>>(static_cast<const float*>(src), rows, cols, dst,
                                                                        C, H, W, N);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Pyramid level preparation (featurePnP/model.py:194-200), CHW in -> CHW float64 out, on the channel
// slice [c_start, c_end):
//   resize: nn.functional.interpolate(..., size = (Ho, Wo), mode = 'bilinear') (align_corners = False):
//           src = max(0, (dst + 0.5) * in / out - 0.5), the two neighbours clamped to the border;
//   blur:   depthwise conv2d with a k x k kernel (cross-correlation), zero padding `pad` -- the reference
//           fixes pad = 1 whatever k is, so the output is (H + 2 - k + 1) x (W + 2 - k + 1).
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void bilinear_resize_chw_kernel(const T* __restrict__ src, double* __restrict__ dst, int Cs, int H, int W,
                                           int Ho, int Wo) {
    const size_t total = (size_t)Cs * Ho * Wo;
    const double sy = (double)H / (double)Ho, sx = (double)W / (double)Wo;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int xo = (int)(e % Wo), yo = (int)((e / Wo) % Ho), c = (int)(e / ((size_t)Wo * Ho));
        // torch's CPU kernel contracts scale * (dst + 0.5) - 0.5 into one fma (measured: 1 ulp agreement
        // with this form, 6 ulp without)
        double fy = fma(sy, __dadd_rn((double)yo, 0.5), -0.5);
        double fx = fma(sx, __dadd_rn((double)xo, 0.5), -0.5);
        if (fy < 0.0) fy = 0.0;
        if (fx < 0.0) fx = 0.0;
        const int y0 = (int)fy, x0 = (int)fx;
        const int y1 = y0 + (y0 < H - 1 ? 1 : 0), x1 = x0 + (x0 < W - 1 ? 1 : 0);
        const double ly1 = __dadd_rn(fy, -(double)y0), lx1 = __dadd_rn(fx, -(double)x0);
        const double ly0 = __dadd_rn(1.0, -ly1), lx0 = __dadd_rn(1.0, -lx1);
        const T* p = src + (size_t)c * H * W;
        const double v00 = (double)p[(size_t)y0 * W + x0], v01 = (double)p[(size_t)y0 * W + x1];
        const double v10 = (double)p[(size_t)y1 * W + x0], v11 = (double)p[(size_t)y1 * W + x1];
        // torch (UpSampleKernel.cpp): w0y * (w0x * v00 + w1x * v01) + w1y * (w0x * v10 + w1x * v11)
        const double top = __dadd_rn(__dmul_rn(lx0, v00), __dmul_rn(lx1, v01));
        const double bot = __dadd_rn(__dmul_rn(lx0, v10), __dmul_rn(lx1, v11));
        dst[e] = __dadd_rn(__dmul_rn(ly0, top), __dmul_rn(ly1, bot));
    }
}

template <typename T>
__global__ void blur_chw_kernel(const T* __restrict__ src, const double* __restrict__ kern, double* __restrict__ dst,
                                int Cs, int H, int W, int k, int pad, int Ho, int Wo) {
    const size_t total = (size_t)Cs * Ho * Wo;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int xo = (int)(e % Wo), yo = (int)((e / Wo) % Ho), c = (int)(e / ((size_t)Wo * Ho));
        const T* p = src + (size_t)c * H * W;
        double s = 0.0;
        for (int a = 0; a < k; ++a) {
            const int y = yo - pad + a;
            if (y < 0 || y >= H) continue;
            for (int b = 0; b < k; ++b) {
                const int x = xo - pad + b;
                if (x < 0 || x >= W) continue;
                s = fma(kern[a * k + b], (double)p[(size_t)y * W + x], s);
            }
        }
        dst[e] = s;
    }
}

cudaError_t launch_bilinear_resize_chw(const void* src, int src_is_f64, double* dst, int Cs, int H, int W, int Ho, int Wo,
                                       cudaStream_t stream) {
    const size_t total = (size_t)Cs * Ho * Wo;
    const int grid = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    if (total == 0) return cudaSuccess;
    if (src_is_f64)
        bilinear_resize_chw_kernel<double><<<grid, 256, 0, stream>>>(static_cast<const double*>(src), dst, Cs, H, W, Ho, Wo);
    else
        bilinear_resize_chw_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(src), dst, Cs, H, W, Ho, Wo);
    return cudaGetLastError();
}

cudaError_t launch_blur_chw(const void* src, int src_is_f64, const double* kern, double* dst, int Cs, int H, int W, int k,
                            int pad, cudaStream_t stream) {
    const int Ho = H + 2 * pad - k + 1, Wo = W + 2 * pad - k + 1;
    if (Ho <= 0 || Wo <= 0) return cudaErrorInvalidValue;
    const size_t total = (size_t)Cs * Ho * Wo;
    const int grid = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    if (src_is_f64)
        blur_chw_kernel<double><<<grid, 256, 0, stream>>>(static_cast<const double*>(src), kern, dst, Cs, H, W, k, pad, Ho, Wo);
    else
        blur_chw_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(src), kern, dst, Cs, H, W, k, pad, Ho, Wo);
    return cudaGetLastError();
}

}  // namespace fmpnp
