// Guiding network: DenseGrid multiresolution encoding + 4-matrix ReLU MLP (in -> 64 -> 64 -> 64 -> 48).
//
// Replaces tcnn::GuidingNetwork (util/network.h:21-196) = GridEncoding "DenseGrid"
// (ext/tcnn/include/tiny-cuda-nn/encodings/grid.h:171-206,221-393,960-1016) -> FullyFusedMLP
// (ext/tcnn/src/fully_fused_mlp.cu:499-558). Same function class and parameter count (C1: 61 536 grid + 13 312
// MLP parameters; C2: 1 056 + 12 288); fp16 operands like tcnn, but fp32 accumulation (TMEM) instead of tcnn's fp16
// wmma accumulators, fp32 interpolation sums, and an fp32 EMA copy.
//
// Parameter slab (fp32 master, same order as GuidingNetwork::set_params, util/network.h:93-111):
//   [ W0 (64 x K0) | W1 (64 x 64) | W2 (64 x 64) | W3 (48 x 64) | grid level 0 | ... | grid level L-1 ]
// with W row-major [out][in] and grid points x 4 features. Two fp16 working copies (training weights, EMA
// "inference" weights) are kept in the layout the tensor cores consume directly:
//   W_l as UMMA K-major no-swizzle canonical tiles: half offset(o, k) = (k / 8) * (N_l * 8) + o * 8 + (k % 8)
//   grid: unchanged order, 4 halfs (8 bytes) per grid point -> one 64-bit load per corner.
#pragma once
#include <cuda_fp16.h>
#include "geom.cuh"

namespace wost
{
    constexpr int kHidden = 64;      // n_neurons
    constexpr int kOutPad = 48;      // N_DIM_PADDED_OUTPUT (parameters.h:25,36)
    constexpr int kGridFeat = 4;     // n_features_per_level
    constexpr int kMaxLevels = 16;   // K0 = 4 * levels <= 64
    constexpr int kTileRows = 128;   // rows per CTA tile == UMMA M

    struct GuideNetView
    {
        int dim;
        int n_levels;
        int k0;                       // encoded width = 4 * n_levels rounded up to 16 (16 / 32 / 48 / 64)
        float scale[kMaxLevels];      // grid_scale(level), grid.h:193-198
        uint32_t res[kMaxLevels];     // grid_resolution(scale), grid.h:200-202
        uint32_t off[kMaxLevels + 1]; // first grid point of each level (params_in_level rounded up to 8, grid.h:982-1012)
        const __half *grid;           // fp16 grid features
        const __half *mlp;            // fp16 MLP weights, canonical tiles, W0|W1|W2|W3
        float lo[3], hi[3];           // scene AABB inflated by 0.5% of its diagonal (normalizeSpatialCoord, train.h:148-155)
        int n_mlp_params, n_grid_params;
    };
    WOST_HOSTDEV int mlp_layer_in(const GuideNetView &g, int l) { return l == 0 ? g.k0 : kHidden; }
    WOST_HOSTDEV int mlp_layer_out(int l) { return l == 3 ? kOutPad : kHidden; }
    WOST_HOSTDEV int mlp_layer_offset(int k0, int l) // in elements
    {
        return l == 0 ? 0 : (kHidden * k0 + (l - 1) * kHidden * kHidden);
    }
    // canonical K-major tile: element (row r of R rows, column k) -> half offset
    WOST_HOSTDEV int tile_off(int R, int r, int k) { return (k >> 3) * (R * 8) + r * 8 + (k & 7); }

    // normalizeSpatialCoord (train.h:148-155)
    template <int D>
    WOST_HOSTDEV void normalize_coord(const GuideNetView &g, const Vec<D> &x, float *xn)
    {
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            const float c = (g.lo[k] + g.hi[k]) * 0.5f;
            xn[k] = 0.5f + dm_div(x[k] - c, g.hi[k] - g.lo[k]);
        }
    }
    template <int D>
    WOST_HOSTDEV bool aabb_contains(const float *lo, const float *hi, const Vec<D> &x) // Eigen AlignedBox::contains (inclusive)
    {
        bool in = true;
#pragma unroll
        for (int k = 0; k < D; ++k)
            in = in && (lo[k] <= x[k]) && (x[k] <= hi[k]);
        return in;
    }

    // One level of kernel_grid (grid.h:221-350, Linear interpolation, Dense indexing): corner indices and weights.
    template <int D>
    struct GridCell
    {
        uint32_t idx[1 << D];
        float w[1 << D];
    };
    template <int D>
    WOST_HOSTDEV void grid_cell(const GuideNetView &g, int level, const float *xn, GridCell<D> &c)
    {
        const float scale = g.scale[level];
        const uint32_t res = g.res[level], size = g.off[level + 1] - g.off[level];
        float fr[D];
        uint32_t pg[D];
#pragma unroll
        for (int d = 0; d < D; ++d)
        {
            const float p = xn[d] * scale + 0.5f; // pos_fract, common_device.h:419-431
            const float fl = floorf(p);
            pg[d] = (uint32_t)(int)fl;
            fr[d] = p - fl;
        }
#pragma unroll
        for (int corner = 0; corner < (1 << D); ++corner)
        {
            float w = 1.0f;
            uint32_t stride = 1, index = 0;
#pragma unroll
            for (int d = 0; d < D; ++d)
            {
                const bool up = (corner >> d) & 1;
                w *= up ? fr[d] : 1.0f - fr[d];
                if (stride <= size) // grid_index, grid.h:143-160
                {
                    index += (pg[d] + (up ? 1u : 0u)) * stride;
                    stride *= res;
                }
            }
            c.idx[corner] = g.off[level] + (index < size ? index : index % size); // dense levels never wrap: keep the integer division off the hot path
            c.w[corner] = w;
        }
    }

#if defined(__CUDACC__)
    // Encoded features of one point -> enc[KE] (fp32 sums, rounded to fp16 by the caller when it packs the operand).
    // KE >= g.k0 is the compile-time width of enc: the level loop is fully unrolled so that enc stays in registers.
    template <int D, int KE>
    WOST_DEV void encode_point(const GuideNetView &g, const Vec<D> &x, float *enc)
    {
        float xn[D];
        normalize_coord<D>(g, x, xn);
#pragma unroll
        for (int l = 0; l < KE / kGridFeat; ++l)
        {
            float f[kGridFeat] = {0.0f, 0.0f, 0.0f, 0.0f};
            if (l < g.n_levels)
            {
                GridCell<D> c;
                grid_cell<D>(g, l, xn, c);
#pragma unroll
                for (int corner = 0; corner < (1 << D); ++corner)
                {
                    const uint2 raw = __ldg(reinterpret_cast<const uint2 *>(g.grid) + c.idx[corner]);
                    const __half2 h01 = *reinterpret_cast<const __half2 *>(&raw.x), h23 = *reinterpret_cast<const __half2 *>(&raw.y);
                    f[0] += c.w[corner] * __low2float(h01);
                    f[1] += c.w[corner] * __high2float(h01);
                    f[2] += c.w[corner] * __low2float(h23);
                    f[3] += c.w[corner] * __high2float(h23);
                }
            }
#pragma unroll
            for (int k = 0; k < kGridFeat; ++k)
                enc[l * kGridFeat + k] = f[k];
        }
    }

    // SIMT evaluation of the MLP for ONE row by ONE thread, weights read from global/L1 in the canonical tile layout.
    // Used by wost_query_network and as the cross-check of the tcgen05 path (same rounding points: fp16 operands,
    // fp32 accumulation). in/out: float arrays; act[l] (optional) receives the post-ReLU fp16-rounded activations.
    WOST_DEV float round_half(float v) { return __half2float(__float2half_rn(v)); }
    WOST_DEV void mlp_forward_row_simt(const GuideNetView &g, const float *enc, float *out)
    {
        float a[kHidden], b[kHidden];
        for (int k = 0; k < g.k0; ++k)
            a[k] = round_half(enc[k]);
        for (int l = 0; l < 4; ++l)
        {
            const int K = mlp_layer_in(g, l), N = mlp_layer_out(l);
            const __half *W = g.mlp + mlp_layer_offset(g.k0, l);
            for (int o = 0; o < N; ++o)
            {
                float s = 0.0f;
                for (int k = 0; k < K; ++k)
                    s += a[k] * __half2float(__ldg(W + tile_off(N, o, k)));
                b[o] = s;
            }
            if (l == 3)
            {
                for (int o = 0; o < N; ++o)
                    out[o] = b[o];
            }
            else
            {
                for (int o = 0; o < N; ++o)
                    a[o] = round_half(fmaxf(b[o], 0.0f));
            }
        }
    }
#endif
} // namespace wost
