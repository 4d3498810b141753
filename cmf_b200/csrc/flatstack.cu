// Whole-stack kernel for the small flat configs: the complete z -> x push of the
// primal and all d tangent columns through every affine coupling layer INCLUDING its MLP conditioner,
// and the per-sample metric G = J^T J, in ONE launch -- and one launch for the adjoint of all of it.
//
// Replaces, for `flat-realnvp` models whose conditioners are small tanh MLPs (the 2-D configs of
// config/two_d.py: D <= 8, hidden width 10): per layer the F.linear / tanh / JVP-rule ops of
// networks.py:206-224 + jvp_layers.py:38-64 on the primal and the tangents, the coupling of
// acl.py:125-146, the tail pad of non_square.py:397-410 and the stack + bmm of non_square.py:307-308 --
// ~30 launches of microsecond kernels per layer in the layer program (1.2-4.6 KB of
// algorithmic traffic per sample against ~0.01 % of the HBM roofline).  Here the whole model (840-1092
// parameters) sits in shared memory, a sample's state (x: D floats, V: d x D floats) in registers /
// local memory, and J and the layer activations never touch HBM: per sample the kernel reads z (4 d
// bytes) and writes G and x (4 d^2 + 4 D bytes).
//
// One thread per sample, one warp per CTA.  Every Linear is zero-padded to 16 x 16 in shared memory, so
// all matrix-vector loops have compile-time bounds and a column's activations live in registers; the
// tangent columns go through the conditioner one at a time.  The adjoint kernel re-runs the forward per
// sample keeping each layer's INPUT state, then walks the layers backwards recomputing one conditioner
// (and one tangent column) at a time.  Parameter gradients: a reduce-scatter over the warp's 32 samples
// (32 entries per 31 shuffles, fixed tree), each lane adding its entry to the CTA's gradient image -- no
// atomics, so every partial and their host-side sum are run-to-run reproducible.
#include "common.cuh"

namespace cmf {
namespace {

constexpr int kFsD = CMF_FLAT_MAX_D;        // max data dimension
constexpr int kFsK = CMF_FLAT_MAX_LATENT;   // max latent dimension (tangent columns)
constexpr int kFsH = CMF_FLAT_MAX_WIDTH;    // max width of any MLP layer (input, hidden, output = 2 M): 16
constexpr int kFsNH = CMF_FLAT_MAX_HIDDEN;  // max hidden layers of the MLP
constexpr int kFsL = CMF_FLAT_MAX_LAYERS;   // max coupling layers
constexpr int kFsThreads = 32;              // ONE warp per CTA: 32 samples, parameter-gradient partial per CTA in a fixed order
constexpr int kFsHalf = kFsH / 2;           // padded last layer: shift rows at [0, 8), log-scale rows at [8, 16)
constexpr int kFsMat = kFsH * kFsH + kFsH;  // one padded Linear in shared memory: W[16][16] then bias[16]
static_assert(kFsD <= kFsHalf, "the padded output layout needs n_mod <= 8");

struct Desc {
    int D, d, L, nh, P;
    int pos[kFsK];                 // logical position of latent k after the tail pad (non_square.py:397-404)
    int n_keep[kFsL], n_mod[kFsL];
    int keep[kFsL][kFsD], mod[kFsL][kFsD];      // logical positions, conditioner-input / -output order
    int width[kFsL][kFsNH + 2];    // widths of the MLP: width[l][0] = n_keep, ..., width[l][nh + 1] = 2 n_mod
    int w_off[kFsL][kFsNH + 1], b_off[kFsL][kFsNH + 1];   // offsets into the packed parameter buffer
};
static_assert(sizeof(Desc) <= 4000, "flat-stack descriptor must fit the kernel parameter space");

// Shared-memory image of the model: every Linear zero-padded to 16 x 16 (+ 16 biases), so all inner
// loops have compile-time bounds and a sample's activations stay in REGISTERS.  The last Linear of a
// conditioner is re-ordered: shift rows at [0, n_mod), log-scale rows at [8, 8 + n_mod).
__device__ __forceinline__ int padded_row(const Desc& M, int l, int j, int o) {
    if (j < M.nh) return o;
    const int nm = M.n_mod[l];
    return o < nm ? o : kFsHalf + (o - nm);
}
__device__ __forceinline__ void load_padded(const Desc& M, const float* __restrict__ params, float* sw) {
    const int total = M.L * (M.nh + 1) * kFsMat;
    for (int i = threadIdx.x; i < total; i += kFsThreads) sw[i] = 0.f;
    __syncwarp();
    for (int l = 0; l < M.L; ++l)
        for (int j = 0; j <= M.nh; ++j) {
            float* dst = sw + (l * (M.nh + 1) + j) * kFsMat;
            const int nin = M.width[l][j], nout = M.width[l][j + 1];
            for (int e = threadIdx.x; e < nin * nout; e += kFsThreads) {
                const int o = e / nin, i = e - o * nin;
                dst[padded_row(M, l, j, o) * kFsH + i] = params[M.w_off[l][j] + e];
            }
            for (int o = threadIdx.x; o < nout; o += kFsThreads) dst[kFsH * kFsH + padded_row(M, l, j, o)] = params[M.b_off[l][j] + o];
        }
    __syncwarp();
}

// y = W a (+ bias): 16 x 16, fully unrolled, everything in registers
__device__ __forceinline__ void matvec(const float* __restrict__ W, const float (&a)[kFsH], float (&y)[kFsH], bool with_bias) {
#pragma unroll
    for (int o = 0; o < kFsH; ++o) {
        float acc = with_bias ? W[kFsH * kFsH + o] : 0.f;
#pragma unroll
        for (int i = 0; i < kFsH; ++i) acc = fmaf(W[o * kFsH + i], a[i], acc);
        y[o] = acc;
    }
}
// y = W^T g
__device__ __forceinline__ void matvec_t(const float* __restrict__ W, const float (&g)[kFsH], float (&y)[kFsH]) {
#pragma unroll
    for (int i = 0; i < kFsH; ++i) y[i] = 0.f;
#pragma unroll
    for (int o = 0; o < kFsH; ++o)
#pragma unroll
        for (int i = 0; i < kFsH; ++i) y[i] = fmaf(W[o * kFsH + i], g[o], y[i]);
}

// primal conditioner: h[j] = input of Linear j (h[0] = passthrough half), out = last Linear's output
__device__ __forceinline__ void primal_mlp(const Desc& M, int l, const float* __restrict__ sw, const float* x,
                                           float (&h)[kFsNH + 1][kFsH], float (&out)[kFsH]) {
#pragma unroll
    for (int i = 0; i < kFsH; ++i) h[0][i] = i < M.n_keep[l] ? x[M.keep[l][i]] : 0.f;
#pragma unroll
    for (int j = 0; j <= kFsNH; ++j) {
        if (j > M.nh) break;
        float y[kFsH];
        matvec(sw + (l * (M.nh + 1) + j) * kFsMat, h[j], y, true);
        if (j < M.nh) {
#pragma unroll
            for (int o = 0; o < kFsH; ++o) h[j + 1][o] = tanhf(y[o]);
        } else {
#pragma unroll
            for (int o = 0; o < kFsH; ++o) out[o] = y[o];
        }
    }
}
// one tangent column through the same conditioner (jvp_layers.py:38-64: no bias; tanh -> (1 - y^2) v):
// a[j] = input of Linear j, pre[j] = W_j a[j] before the tanh factor, outT = last Linear's output
__device__ __forceinline__ void tangent_mlp(const Desc& M, int l, const float* __restrict__ sw, const float* Vk,
                                            const float (&h)[kFsNH + 1][kFsH], float (&a)[kFsNH + 1][kFsH],
                                            float (&pre)[kFsNH][kFsH], float (&outT)[kFsH]) {
#pragma unroll
    for (int i = 0; i < kFsH; ++i) a[0][i] = i < M.n_keep[l] ? Vk[M.keep[l][i]] : 0.f;
#pragma unroll
    for (int j = 0; j <= kFsNH; ++j) {
        if (j > M.nh) break;
        float y[kFsH];
        matvec(sw + (l * (M.nh + 1) + j) * kFsMat, a[j], y, false);
        if (j < M.nh) {
#pragma unroll
            for (int o = 0; o < kFsH; ++o) {
                pre[j][o] = y[o];
                a[j + 1][o] = (1.f - h[j + 1][o] * h[j + 1][o]) * y[o];
            }
        } else {
#pragma unroll
            for (int o = 0; o < kFsH; ++o) outT[o] = y[o];
        }
    }
}

// x, V: the sample's state (logical order), in local memory (D + d D <= 72 floats, dynamic indices)
__device__ __forceinline__ void init_state(const Desc& M, const float* __restrict__ z, float* x, float (*V)[kFsD]) {
    for (int i = 0; i < M.D; ++i) {
        x[i] = 0.f;
        for (int k = 0; k < M.d; ++k) V[k][i] = 0.f;
    }
    for (int k = 0; k < M.d; ++k) {                      // tail pad + inverse permutation; V0 = one-hot columns
        x[M.pos[k]] = z[k];
        V[k][M.pos[k]] = 1.f;
    }
}

__device__ __forceinline__ void layer_forward(const Desc& M, int l, const float* __restrict__ sw, float* x, float (*V)[kFsD]) {
    float h[kFsNH + 1][kFsH], out[kFsH];
    primal_mlp(M, l, sw, x, h, out);
    const int nm = M.n_mod[l];
    float e[kFsHalf], zq[kFsHalf];
#pragma unroll
    for (int j = 0; j < kFsHalf; ++j) {
        e[j] = expf(-out[kFsHalf + j]);
        zq[j] = j < nm ? x[M.mod[l][j]] : 0.f;
    }
    for (int k = 0; k < M.d; ++k) {                      // tangents one at a time: their activations stay in registers
        float a[kFsNH + 1][kFsH], pre[kFsNH][kFsH], outT[kFsH];
        tangent_mlp(M, l, sw, V[k], h, a, pre, outT);
#pragma unroll
        for (int j = 0; j < kFsHalf; ++j)
            if (j < nm) {                                // acl.py:125-146
                const int q = M.mod[l][j];
                V[k][q] = e[j] * (V[k][q] - zq[j] * outT[kFsHalf + j]) - outT[j];
            }
    }
#pragma unroll
    for (int j = 0; j < kFsHalf; ++j)
        if (j < nm) x[M.mod[l][j]] = zq[j] * e[j] - out[j];
}

__global__ void __launch_bounds__(kFsThreads)
flatstack_fwd_kernel(const Desc M, int B, const float* __restrict__ z, const float* __restrict__ params,
                     float* __restrict__ G, float* __restrict__ xout) {
    extern __shared__ float sw[];
    load_padded(M, params, sw);
    const int b = blockIdx.x * kFsThreads + threadIdx.x;
    if (b >= B) return;
    float x[kFsD];
    float V[kFsK][kFsD];
    init_state(M, z + (size_t)b * M.d, x, V);
    for (int l = 0; l < M.L; ++l) layer_forward(M, l, sw, x, V);
    for (int i = 0; i < M.D; ++i) xout[(size_t)b * M.D + i] = x[i];
    for (int k = 0; k < M.d; ++k)
        for (int c = 0; c <= k; ++c) {
            float g = 0.f;
            for (int i = 0; i < M.D; ++i) g = fmaf(V[k][i], V[c][i], g);
            G[((size_t)b * M.d + k) * M.d + c] = g;
            G[((size_t)b * M.d + c) * M.d + k] = g;
        }
}

// Reduce-scatter over the warp: every lane brings 32 values; afterwards lane L holds in v[0] the sum over
// all 32 lanes (= samples) of their v[L].  31 shuffles for 32 sums (a butterfly per value would take
// 160), a fixed summation tree, registers only.
__device__ __forceinline__ void warp_reduce_scatter32(float (&v)[32]) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        const bool upper = (lane & s) != 0;
#pragma unroll
        for (int i = 0; i < s; ++i) {
            const float send = upper ? v[i] : v[i + s];
            const float keep = upper ? v[i + s] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
        }
    }
}
// g_W += g (x) a and g_b += g for one padded Linear (g = adjoint of its output, a = its input), summed
// over the warp's 32 samples: two rows of 16 products per reduce-scatter, each lane then owns one entry.
// Only rows that exist are visited (padding carries no gradient): rows [0, lo) and [8, hi).
__device__ __forceinline__ void accumulate_linear(float* sg, const float (&g)[kFsH], const float (&a)[kFsH], bool with_bias,
                                                  int lo, int hi) {
    const int lane = threadIdx.x & 31;
    auto exists = [&](int o) { return o < lo || (o >= kFsHalf && o < hi); };
#pragma unroll
    for (int o = 0; o < kFsH; o += 2) {
        if (!exists(o) && !exists(o + 1)) continue;      // warp-uniform
        float v[32];
#pragma unroll
        for (int i = 0; i < kFsH; ++i) {
            v[i] = g[o] * a[i];
            v[kFsH + i] = g[o + 1] * a[i];
        }
        warp_reduce_scatter32(v);
        const int row = o + (lane >> 4);
        if (exists(row)) sg[row * kFsH + (lane & 15)] += v[0];
    }
    if (with_bias) {
        float v[32];
#pragma unroll
        for (int i = 0; i < kFsH; ++i) {
            v[i] = g[i];
            v[kFsH + i] = 0.f;
        }
        warp_reduce_scatter32(v);
        if (lane < kFsH && exists(lane)) sg[kFsH * kFsH + lane] += v[0];
    }
    __syncwarp();
}

__global__ void __launch_bounds__(kFsThreads)
flatstack_bwd_kernel(const Desc M, int B, const float* __restrict__ z, const float* __restrict__ params,
                     const float* __restrict__ gG, const float* __restrict__ gxout,
                     float* __restrict__ gz, float* __restrict__ gparams_partial) {
    extern __shared__ float smem[];
    const int img = M.L * (M.nh + 1) * kFsMat;
    float* sw = smem;
    float* sg = smem + img;                               // gradient image, same padded layout
    load_padded(M, params, sw);
    for (int i = threadIdx.x; i < img; i += kFsThreads) sg[i] = 0.f;
    __syncwarp();
    const int b = blockIdx.x * kFsThreads + threadIdx.x;
    const bool live = b < B;                              // dead lanes run with zero cotangents (warp-wide reductions)
    float x[kFsD];
    float V[kFsK][kFsD];
    float xs[kFsL][kFsD];                                 // layer inputs
    float Vs[kFsL][kFsK][kFsD];
    init_state(M, z + (size_t)(live ? b : 0) * M.d, x, V);
    for (int l = 0; l < M.L; ++l) {
        for (int i = 0; i < M.D; ++i) {
            xs[l][i] = x[i];
            for (int k = 0; k < M.d; ++k) Vs[l][k][i] = V[k][i];
        }
        layer_forward(M, l, sw, x, V);
    }
    // adjoint of G = V V^T and of the primal output
    float gx[kFsD];
    float gV[kFsK][kFsD];
    for (int i = 0; i < M.D; ++i) gx[i] = (live && gxout) ? gxout[(size_t)b * M.D + i] : 0.f;
    for (int k = 0; k < M.d; ++k)
        for (int i = 0; i < M.D; ++i) {
            float a = 0.f;
            if (live && gG)
                for (int c = 0; c < M.d; ++c)
                    a = fmaf(gG[((size_t)b * M.d + k) * M.d + c] + gG[((size_t)b * M.d + c) * M.d + k], V[c][i], a);
            gV[k][i] = a;
        }
    for (int l = M.L - 1; l >= 0; --l) {
        const int nm = M.n_mod[l], p = M.n_keep[l];
        float h[kFsNH + 1][kFsH], out[kFsH];
        primal_mlp(M, l, sw, xs[l], h, out);
        float e[kFsHalf], zq[kFsHalf], gxq[kFsHalf], Asum[kFsHalf], Csum[kFsHalf];
#pragma unroll
        for (int j = 0; j < kFsHalf; ++j) {
            e[j] = expf(-out[kFsHalf + j]);
            zq[j] = j < nm ? xs[l][M.mod[l][j]] : 0.f;
            gxq[j] = j < nm ? gx[M.mod[l][j]] : 0.f;
            Asum[j] = 0.f; Csum[j] = 0.f;
        }
        // gh[j]: adjoint of the primal activations h[j] collected from the tangents' tanh factors
        float gh[kFsNH + 1][kFsH];
#pragma unroll
        for (int j = 0; j <= kFsNH; ++j)
#pragma unroll
            for (int i = 0; i < kFsH; ++i) gh[j][i] = 0.f;
        for (int k = 0; k < M.d; ++k) {
            float a[kFsNH + 1][kFsH], pre[kFsNH][kFsH], outT[kFsH];
            tangent_mlp(M, l, sw, Vs[l][k], h, a, pre, outT);
            // coupling adjoint of column k (the formulas of coupling.cu): g = [g_T_k | g_S_k]
            float g[kFsH];
#pragma unroll
            for (int j = 0; j < kFsHalf; ++j) {
                const float gk = j < nm ? gV[k][M.mod[l][j]] : 0.f;
                Asum[j] = fmaf(outT[kFsHalf + j], gk, Asum[j]);
                Csum[j] = fmaf(j < nm ? Vs[l][k][M.mod[l][j]] : 0.f, gk, Csum[j]);
                g[j] = -gk;
                g[kFsHalf + j] = -e[j] * zq[j] * gk;
                if (j < nm) gV[k][M.mod[l][j]] = e[j] * gk;
            }
            // back through the tangent chain, last Linear first
#pragma unroll
            for (int j = kFsNH; j >= 0; --j) {
                if (j > M.nh) continue;
                const float* W = sw + (l * (M.nh + 1) + j) * kFsMat;
                float* sgj = sg + (l * (M.nh + 1) + j) * kFsMat;
                const bool last = j == M.nh;
                accumulate_linear(sgj, g, a[j], false, last ? nm : M.width[l][j + 1], last ? kFsHalf + nm : 0);
                float gin[kFsH];
                matvec_t(W, g, gin);
                if (j > 0) {
#pragma unroll
                    for (int i = 0; i < kFsH; ++i) {     // a[j] = (1 - h[j]^2) pre[j-1]
                        const float hh = h[j][i];
                        gh[j][i] = fmaf(-2.f * hh * pre[j - 1][i], gin[i], gh[j][i]);
                        g[i] = (1.f - hh * hh) * gin[i];
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < kFsH; ++i)
                        if (i < p) gV[k][M.keep[l][i]] += gin[i];
                }
            }
        }
        // primal: coupling adjoint, then back through the conditioner with the collected gh
        float g[kFsH];
#pragma unroll
        for (int j = 0; j < kFsHalf; ++j) {
            g[j] = -gxq[j];
            g[kFsHalf + j] = -zq[j] * e[j] * gxq[j] - e[j] * (Csum[j] - zq[j] * Asum[j]);
            if (j < nm) gx[M.mod[l][j]] = e[j] * (gxq[j] - Asum[j]);
        }
#pragma unroll
        for (int j = kFsNH; j >= 0; --j) {
            if (j > M.nh) continue;
            const float* W = sw + (l * (M.nh + 1) + j) * kFsMat;
            float* sgj = sg + (l * (M.nh + 1) + j) * kFsMat;
            const bool last = j == M.nh;
            accumulate_linear(sgj, g, h[j], true, last ? nm : M.width[l][j + 1], last ? kFsHalf + nm : 0);
            float gin[kFsH];
            matvec_t(W, g, gin);
            if (j > 0) {
#pragma unroll
                for (int i = 0; i < kFsH; ++i) {         // h[j] = tanh(.): (1 - h^2) (direct + via the tangents' factors)
                    const float hh = h[j][i];
                    g[i] = (1.f - hh * hh) * (gin[i] + gh[j][i]);
                }
            } else {
#pragma unroll
                for (int i = 0; i < kFsH; ++i)
                    if (i < p) gx[M.keep[l][i]] += gin[i];
            }
        }
    }
    if (live)
        for (int k = 0; k < M.d; ++k) gz[(size_t)b * M.d + k] = gx[M.pos[k]];
    __syncwarp();
    // un-pad the CTA's gradient image into its partial (packed parameter order)
    float* outp = gparams_partial + (size_t)blockIdx.x * M.P;
    for (int l = 0; l < M.L; ++l)
        for (int j = 0; j <= M.nh; ++j) {
            const float* src = sg + (l * (M.nh + 1) + j) * kFsMat;
            const int nin = M.width[l][j], nout = M.width[l][j + 1];
            for (int e2 = threadIdx.x; e2 < nin * nout; e2 += kFsThreads) {
                const int o = e2 / nin, i = e2 - o * nin;
                outp[M.w_off[l][j] + e2] = src[padded_row(M, l, j, o) * kFsH + i];
            }
            for (int o = threadIdx.x; o < nout; o += kFsThreads) outp[M.b_off[l][j] + o] = src[kFsH * kFsH + padded_row(M, l, j, o)];
        }
}

}  // namespace
}  // namespace cmf

using namespace cmf;

static int fill_desc(Desc& M, const int32_t* desc, int n) {
    // host layout (int32): D, d, L, nh, P, pos[d], then per layer: n_keep, n_mod, keep[n_keep], mod[n_mod],
    // width[nh + 2], w_off[nh + 1], b_off[nh + 1]
    int c = 0;
    auto next = [&]() { return c < n ? desc[c++] : -1; };
    M.D = next(); M.d = next(); M.L = next(); M.nh = next(); M.P = next();
    CMF_REQUIRE(M.D >= 1 && M.D <= kFsD && M.d >= 1 && M.d <= kFsK && M.d <= M.D, "flat_stack: D = %d, d = %d outside the kernel's range (D <= %d, d <= %d)",
                M.D, M.d, kFsD, kFsK);
    CMF_REQUIRE(M.L >= 1 && M.L <= kFsL && M.nh >= 1 && M.nh <= kFsNH, "flat_stack: %d layers / %d hidden layers outside the kernel's range (%d / %d)",
                M.L, M.nh, kFsL, kFsNH);
    CMF_REQUIRE(M.P >= 1 && M.P <= CMF_FLAT_MAX_PARAMS, "flat_stack: %d parameters (max %d)", M.P, CMF_FLAT_MAX_PARAMS);
    for (int k = 0; k < M.d; ++k) {
        M.pos[k] = next();
        CMF_REQUIRE(M.pos[k] >= 0 && M.pos[k] < M.D, "flat_stack: bad tail position");
    }
    for (int l = 0; l < M.L; ++l) {
        M.n_keep[l] = next(); M.n_mod[l] = next();
        CMF_REQUIRE(M.n_keep[l] >= 1 && M.n_mod[l] >= 1 && M.n_keep[l] + M.n_mod[l] == M.D, "flat_stack: bad partition in layer %d", l);
        for (int i = 0; i < M.n_keep[l]; ++i) { M.keep[l][i] = next(); CMF_REQUIRE(M.keep[l][i] >= 0 && M.keep[l][i] < M.D, "flat_stack: bad index"); }
        for (int i = 0; i < M.n_mod[l]; ++i) { M.mod[l][i] = next(); CMF_REQUIRE(M.mod[l][i] >= 0 && M.mod[l][i] < M.D, "flat_stack: bad index"); }
        for (int j = 0; j <= M.nh + 1; ++j) {
            M.width[l][j] = next();
            CMF_REQUIRE(M.width[l][j] >= 1 && M.width[l][j] <= kFsH, "flat_stack: MLP width %d outside the kernel's range (<= %d)", M.width[l][j], kFsH);
        }
        CMF_REQUIRE(M.width[l][0] == M.n_keep[l] && M.width[l][M.nh + 1] == 2 * M.n_mod[l], "flat_stack: MLP input / output widths do not match the partition");
        for (int j = 0; j <= M.nh; ++j) M.w_off[l][j] = next();
        for (int j = 0; j <= M.nh; ++j) M.b_off[l][j] = next();
        for (int j = 0; j <= M.nh; ++j)
            CMF_REQUIRE(M.w_off[l][j] >= 0 && M.w_off[l][j] + M.width[l][j] * M.width[l][j + 1] <= M.P && M.b_off[l][j] >= 0 &&
                            M.b_off[l][j] + M.width[l][j + 1] <= M.P, "flat_stack: parameter offsets out of range");
    }
    CMF_REQUIRE(c == n, "flat_stack: descriptor has %d entries, %d consumed", n, c);
    return 0;
}

extern "C" int cmf_flat_stack_usable(int D, int d, int L, int nh, int max_width, int P) {
    return D >= 1 && D <= kFsD && d >= 1 && d <= kFsK && d <= D && L >= 1 && L <= kFsL && nh >= 1 && nh <= kFsNH &&
           max_width >= 1 && max_width <= kFsH && P >= 1 && P <= CMF_FLAT_MAX_PARAMS &&
           (size_t)2 * L * (nh + 1) * kFsMat * sizeof(float) <= 48 * 1024;          // weight + gradient images of the adjoint
}

extern "C" int cmf_flat_stack_fwd(const int32_t* desc, int desc_len, int B, const float* z, const float* params,
                                  float* G, float* x, cmf_stream_t stream) {
    Desc M;
    CMF_REQUIRE(desc && desc_len > 5, "flat_stack: null descriptor");
    const int rc = fill_desc(M, desc, desc_len);
    if (rc != 0) return rc;
    CMF_REQUIRE(B >= 0, "flat_stack: negative batch");
    if (B == 0) return 0;
    CMF_REQUIRE(z && params && G && x, "flat_stack: null pointer");
    const size_t smem = (size_t)M.L * (M.nh + 1) * kFsMat * sizeof(float);
    CMF_REQUIRE(smem <= 48 * 1024, "flat_stack: %d layers x %d Linears need %zu bytes of shared memory (max 48 KB)", M.L, M.nh + 1, smem);
    flatstack_fwd_kernel<<<(unsigned)ceil_div(B, kFsThreads), kFsThreads, smem, (cudaStream_t)stream>>>(M, B, z, params, G, x);
    CMF_CHECK_LAUNCH();
    return 0;
}

extern "C" int cmf_flat_stack_bwd_partials(int B) { return B <= 0 ? 0 : (int)ceil_div(B, kFsThreads); }

extern "C" int cmf_flat_stack_bwd(const int32_t* desc, int desc_len, int B, const float* z, const float* params,
                                  const float* g_G, const float* g_x, float* g_z, float* g_params_partial,
                                  cmf_stream_t stream) {
    Desc M;
    CMF_REQUIRE(desc && desc_len > 5, "flat_stack: null descriptor");
    const int rc = fill_desc(M, desc, desc_len);
    if (rc != 0) return rc;
    CMF_REQUIRE(B >= 0, "flat_stack: negative batch");
    if (B == 0) return 0;
    CMF_REQUIRE(z && params && g_z && g_params_partial, "flat_stack_bwd: null pointer");
    const size_t smem = (size_t)2 * M.L * (M.nh + 1) * kFsMat * sizeof(float);
    CMF_REQUIRE(smem <= 48 * 1024, "flat_stack_bwd: %d layers x %d Linears need %zu bytes of shared memory (max 48 KB)", M.L, M.nh + 1, smem);
    flatstack_bwd_kernel<<<(unsigned)ceil_div(B, kFsThreads), kFsThreads, smem, (cudaStream_t)stream>>>(
        M, B, z, params, g_G, g_x, g_z, g_params_partial);
    CMF_CHECK_LAUNCH();
    return 0;
}
