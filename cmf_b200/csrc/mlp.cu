// Tangent-batched tanh-MLP conditioner of the flat (tabular) couplers: the WHOLE MLP -- every Linear,
// tanh and JVP mask product, primal + d tangent columns -- in one launch, and one launch (+ a small
// reduction) for its adjoint including the weight / bias gradients.
//
// Replaces, per coupling layer and direction of the flow (k-flouris/cmf @ df4521aa):
//   cmf/models/components/networks.py:206-224  get_mlp (Linear / Tanh stack) and its forward
//   cmf/models/components/networks.py:24-32    NN_Sequential_JVP.jvp
//   cmf/models/components/jvp_layers.py:38-47  get_linear_jvp (bias-free Linear on the tangent),
//                                              get_jvp_from_activation (tanh' of the primal)
// which the reference runs as one F.linear + elementwise kernels per Linear, per tangent column,
// and whose adjoint autograd derives.  SURVEY 8f N1(ii): "per-layer MLP weights in shared memory".
//
// Schedule.  Rows of the GEMMs are vectors: the S = floor(128 / (1 + d)) samples of a tile contribute
// their S*d tangent rows (row s*d + k) followed by their S primal rows; a CTA keeps the tile's
// activations in shared memory feature-major ([feature][row], pitch 132) through all Linears, the
// weights of the current Linear beside it (128 x 128, zero padded; the next Linear's weights arrive
// by cp.async while the current GEMM runs).  The GEMM is exact FP32 FMA on 8 x 8 register tiles
// (the reference computes these Linears in FP32 with TF32 off; widths <= 128 keep a whole layer in
// one tile, so there is nothing to stream but the activations saved for the adjoint).
// The adjoint walks the Linears in reverse per tile: weight gradient G^T A (contraction over the
// tile's rows, added into the CTA's own L2-resident partial), data gradient G W, then the tanh / mask
// adjoint including the term that flows from the tangents' masks into the primal.  All reductions
// have a fixed order: results are bitwise reproducible.
#include "common.cuh"

namespace cmf {
namespace {

constexpr int kT = 128;                       // rows of a tile = padded width of a Linear
constexpr int kP = 132;                       // shared-memory pitch of an activation tile ([feature][row])
constexpr int kMlpThreads = 256;
constexpr int kMaxLin = CMF_MLP_MAX_LINEARS;
constexpr int kWElems = kT * kT;
constexpr int kPackPerLin = 2 * kWElems + kT;  // W (out-major) | W^T (in-major) | bias, zero padded
constexpr int kPartPerLin = kWElems + kT;      // weight-gradient tile | bias gradient

struct MlpParams {
    int B, d, S, ntiles, nlin;
    int width[kMaxLin + 1];
    const float* x;        // (B, width[0])
    const float* V;        // (B, d, width[0]) or null when d == 0
    const float* wpack;    // nlin x kPackPerLin
    float* st;             // (B, width[nlin])
    float* ST;             // (B, d, width[nlin])
    float* save_T;         // (ntiles, nlin - 1, 128, 128): t (tangent rows) / a (primal rows), feature-major; or null
    float* save_U;         // same shape: unmasked tangents u; or null
    const float* g_st;
    const float* g_ST;
    float* g_x;
    float* g_V;
    float* partial;        // (gridDim.x, nlin, kPartPerLin)
};

struct PackParams {
    int nlin;
    int width[kMaxLin + 1];
    const float* W[kMaxLin];
    const float* b[kMaxLin];
    float* wpack;
};

__device__ __forceinline__ void cp_async16(float* dst, const float* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// rows [0, nrows) of a 128-wide global matrix into shared memory with the same pitch
__device__ __forceinline__ void load_w_async(float* dst, const float* src, int nrows, int tid) {
    for (int i = tid; i < nrows * (kT / 4); i += kMlpThreads) cp_async16(dst + i * 4, src + i * 4);
}

// the tile's input vectors, feature-major; rows past the batch and padding rows are zero
__device__ __forceinline__ void load_inputs(float* T, const MlpParams& P, int b0, int ns, int nfeat, int tid) {
    const int d = P.d, nt = P.S * d, nin = P.width[0];
    for (int i = tid; i < nfeat * kT; i += kMlpThreads) {
        const int k = i / kT, r = i - k * kT;
        float v = 0.f;
        if (k < nin) {
            if (r < nt) {
                if (r / d < ns) v = __ldg(P.V + ((long long)b0 * d + r) * nin + k);
            } else if (r - nt < ns) {
                v = __ldg(P.x + (long long)(b0 + r - nt) * nin + k);
            }
        }
        T[k * kP + r] = v;
    }
}

__device__ __forceinline__ int row_of(int ty, int i) { return (i < 4 ? 0 : 60) + ty * 4 + i; }   // i >= 4: 64 + ty*4 + (i-4)
__device__ __forceinline__ int col_of(int tx, int j) { return (j < 4 ? 0 : 60) + tx * 4 + j; }

// acc[i][j] += sum_{k < K} A[k][row_i] * Bm[k][col_j]     (A pitch kP, Bm pitch kT; NG column groups of 64)
template <int NG>
__device__ __forceinline__ void gemm_rows(const float* __restrict__ A, const float* __restrict__ Bm, int K, int ty, int tx,
                                          float (&acc)[8][8]) {
    const float* ap = A + ty * 4;
    const float* bp = Bm + tx * 4;
    // operands of step k + 1 are in flight while the 64 FMAs of step k issue (two warps per scheduler:
    // the loads must be hidden inside the warp's own instruction stream)
    float4 a0 = *reinterpret_cast<const float4*>(ap), a1 = *reinterpret_cast<const float4*>(ap + 64);
    float4 b0 = *reinterpret_cast<const float4*>(bp), b1 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (NG == 2) b1 = *reinterpret_cast<const float4*>(bp + 64);
#pragma unroll 4
    for (int k = 0; k < K; ++k) {
        const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
        // the last trip reads row K: inside the shared-memory allocation (one spare row after the last
        // buffer), never used
        a0 = *reinterpret_cast<const float4*>(ap + (k + 1) * kP);
        a1 = *reinterpret_cast<const float4*>(ap + (k + 1) * kP + 64);
        b0 = *reinterpret_cast<const float4*>(bp + (k + 1) * kT);
        if (NG == 2) b1 = *reinterpret_cast<const float4*>(bp + (k + 1) * kT + 64);
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4 * NG; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
}

// acc[i][ii] += sum_{r < nrows} G[jg + 16 i][r] * A[kg + 16 ii][r]      (both pitch kP; nrows % 4 == 0)
template <int NI, int NK>
__device__ __forceinline__ void gemm_contract_rows(const float* __restrict__ G, const float* __restrict__ A, int nrows, int jg, int kg,
                                                   float (&acc)[8][8]) {
    const float* gp = G + jg * kP;
    const float* ap = A + kg * kP;
#pragma unroll 1
    for (int r = 0; r < nrows; r += 4) {
        float4 g[NI], a[NK];
#pragma unroll
        for (int i = 0; i < NI; ++i) g[i] = *reinterpret_cast<const float4*>(gp + i * 16 * kP + r);
#pragma unroll
        for (int i = 0; i < NK; ++i) a[i] = *reinterpret_cast<const float4*>(ap + i * 16 * kP + r);
#pragma unroll
        for (int i = 0; i < NI; ++i)
#pragma unroll
            for (int ii = 0; ii < NK; ++ii) {
                float s = acc[i][ii];
                s = fmaf(g[i].x, a[ii].x, s);
                s = fmaf(g[i].y, a[ii].y, s);
                s = fmaf(g[i].z, a[ii].z, s);
                s = fmaf(g[i].w, a[ii].w, s);
                acc[i][ii] = s;
            }
    }
}

__global__ void mlp_pack_kernel(PackParams P) {
    const int l = blockIdx.x, j = blockIdx.y, k = threadIdx.x;
    const int K = P.width[l], N = P.width[l + 1];
    const float v = (j < N && k < K) ? __ldg(P.W[l] + (long long)j * K + k) : 0.f;
    float* base = P.wpack + (long long)l * kPackPerLin;
    base[j * kT + k] = v;                       // W, out-major (the data-gradient operand)
    base[kWElems + k * kT + j] = v;             // W^T, in-major (the forward operand)
    if (k == 0) base[2 * kWElems + j] = (j < N && P.b[l]) ? __ldg(P.b[l] + j) : 0.f;
}

// the thread's 8 x 8 register tile into a [feature][row] buffer, 128 bits along the rows; features >= nvalid
// up to npad are written as zero, the rest not at all
__device__ __forceinline__ void store_tile(float* T, const float (&acc)[8][8], int ty, int tx, int nvalid, int npad) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = col_of(tx, j);
        if (c >= npad) continue;
        const bool z = c >= nvalid;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float4 v = make_float4(acc[4 * h][j], acc[4 * h + 1][j], acc[4 * h + 2][j], acc[4 * h + 3][j]);
            if (z) v = make_float4(0.f, 0.f, 0.f, 0.f);
            *reinterpret_cast<float4*>(T + c * kP + 64 * h + ty * 4) = v;
        }
    }
}

__device__ __forceinline__ float& f4(float4& v, int e) { return e == 0 ? v.x : e == 1 ? v.y : e == 2 ? v.z : v.w; }

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kMlpThreads, 1) mlp_fwd_kernel(MlpParams P) {
    extern __shared__ __align__(16) float smem[];
    __shared__ unsigned char srow[kT];         // tangent row -> sample of the tile
    float* T = smem;
    float* const Wbase = smem + kT * kP;       // two weight buffers (+ one spare row, see gemm_rows)
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int d = P.d, S = P.S, nt = S * d, nlin = P.nlin, nhid = nlin - 1;
    const int ngroups = (nt + S + 3) >> 2;     // 4-row groups that hold vectors
    if (tid < kT) srow[tid] = (unsigned char)(tid < nt ? tid / d : 0);
    for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x) {
        const int b0 = tile * S;
        const int ns = min(S, P.B - b0);
        load_w_async(Wbase, P.wpack + kWElems, P.width[0], tid);
        load_inputs(T, P, b0, ns, P.width[0], tid);
        cp_async_wait_all();
        __syncthreads();
        for (int l = 0; l < nlin; ++l) {
            const int K = P.width[l], N = P.width[l + 1];
            if (l + 1 < nlin) load_w_async(Wbase + ((l + 1) & 1) * kWElems, P.wpack + (long long)(l + 1) * kPackPerLin + kWElems, N, tid);
            float acc[8][8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
            if (N <= 64) gemm_rows<1>(T, Wbase + (l & 1) * kWElems, K, ty, tx, acc);
            else gemm_rows<2>(T, Wbase + (l & 1) * kWElems, K, ty, tx, acc);
            __syncthreads();                                   // every read of T is done
            const float* bias = P.wpack + (long long)l * kPackPerLin + 2 * kWElems;
            if (l + 1 < nlin) {
                // raw products back into the tile; then, all threads together: tanh on the primal rows, the JVP
                // mask on the tangent rows (rows of dead samples and padding rows are zero products and stay zero)
                store_tile(T, acc, ty, tx, N, N);
                __syncthreads();
                for (int i = tid; i < N * S; i += kMlpThreads) {
                    const int c = i / S, sidx = i - c * S;
                    float* p = T + c * kP + nt + sidx;
                    *p = sidx < ns ? tanhf(*p + __ldg(bias + c)) : 0.f;
                }
                __syncthreads();
                float* sT = P.save_T ? P.save_T + ((long long)tile * nhid + l) * kWElems : nullptr;
                float* sU = P.save_U ? P.save_U + ((long long)tile * nhid + l) * kWElems : nullptr;
                for (int i = tid; i < N * 32; i += kMlpThreads) {
                    const int c = i >> 5, q = i & 31, r0 = q * 4;
                    if (q >= ngroups) continue;
                    float* p = T + c * kP + r0;
                    float4 v = *reinterpret_cast<const float4*>(p);
                    if (r0 < nt) {
                        if (sU) *reinterpret_cast<float4*>(sU + c * kT + r0) = v;
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            if (r0 + e < nt) {
                                const float a = T[c * kP + nt + srow[r0 + e]];
                                f4(v, e) *= 1.f - a * a;
                            }
                        *reinterpret_cast<float4*>(p) = v;
                    }
                    if (sT) *reinterpret_cast<float4*>(sT + c * kT + r0) = v;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = row_of(ty, i);
                    const bool tangent = r < nt;
                    const int sidx = tangent ? srow[r] : r - nt;
                    if (sidx >= ns) continue;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int c = col_of(tx, j);
                        if (c >= N) continue;
                        if (tangent) P.ST[((long long)b0 * d + r) * N + c] = acc[i][j];
                        else P.st[(long long)(b0 + sidx) * N + c] = acc[i][j] + __ldg(bias + c);
                    }
                }
            }
            cp_async_wait_all();
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------------------------------------
template <int NI, int NK>
__device__ __forceinline__ void wgrad_tile(const float* G, const float* A, int nrows, int N, int K, float* part, bool first, int tid) {
    const int jg = tid >> 4, kg = tid & 15;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    gemm_contract_rows<NI, NK>(G, A, nrows, jg, kg, acc);
    // all loads of the partial first, then the stores (a load -> add -> store chain per element would
    // expose 64 L2 round trips)
    if (!first) {
        float old[NI][NK];
#pragma unroll
        for (int i = 0; i < NI; ++i)
#pragma unroll
            for (int ii = 0; ii < NK; ++ii) {
                const int j = jg + 16 * i, k = kg + 16 * ii;
                old[i][ii] = (j < N && k < K) ? __ldcg(part + j * kT + k) : 0.f;
            }
#pragma unroll
        for (int i = 0; i < NI; ++i)
#pragma unroll
            for (int ii = 0; ii < NK; ++ii) acc[i][ii] += old[i][ii];
    }
#pragma unroll
    for (int i = 0; i < NI; ++i)
#pragma unroll
        for (int ii = 0; ii < NK; ++ii) {
            const int j = jg + 16 * i, k = kg + 16 * ii;
            if (j < N && k < K) part[j * kT + k] = acc[i][ii];
        }
}

__global__ void __launch_bounds__(kMlpThreads, 1) mlp_bwd_kernel(MlpParams P) {
    extern __shared__ __align__(16) float smem[];
    __shared__ unsigned char srow[kT];
    float* G = smem;                       // gradient w.r.t. the current Linear's output, [feature][row]
    float* A = smem + kT * kP;             // the current Linear's input vectors, [feature][row]
    float* W = smem + 2 * kT * kP;         // the current Linear's weights, out-major (+ one spare row)
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int d = P.d, S = P.S, nt = S * d, nlin = P.nlin, nhid = nlin - 1;
    const int ngroups = (nt + S + 3) >> 2;
    const int nrows4 = ngroups * 4;
    if (tid < kT) srow[tid] = (unsigned char)(tid < nt ? tid / d : 0);
    float* part = P.partial + (long long)blockIdx.x * nlin * kPartPerLin;
    bool first = true;
    for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x, first = false) {
        const int b0 = tile * S;
        const int ns = min(S, P.B - b0);
        __syncthreads();
        {   // gradient of the last Linear's output
            const int N = P.width[nlin];
            const int npad = N <= 32 ? 32 : kT;
            for (int i = tid; i < npad * kT; i += kMlpThreads) {
                const int j = i / kT, r = i - j * kT;
                float v = 0.f;
                if (j < N) {
                    if (r < nt) {
                        if (srow[r] < ns) v = __ldg(P.g_ST + ((long long)b0 * d + r) * N + j);
                    } else if (r - nt < ns) {
                        v = __ldg(P.g_st + (long long)(b0 + r - nt) * N + j);
                    }
                }
                G[j * kP + r] = v;
            }
        }
        for (int l = nlin - 1; l >= 0; --l) {
            const int K = P.width[l], N = P.width[l + 1];
            const int kpad = K <= 32 ? 32 : kT;
            load_w_async(W, P.wpack + (long long)l * kPackPerLin, N, tid);
            if (l == 0) {
                load_inputs(A, P, b0, ns, kpad, tid);
            } else {
                const float* src = P.save_T + ((long long)tile * nhid + (l - 1)) * kWElems;
                for (int i = tid; i < K * 32; i += kMlpThreads) {
                    const int c = i >> 5, q = i & 31;
                    if (q < ngroups) cp_async16(A + c * kP + q * 4, src + c * kT + q * 4);
                }
                for (int i = tid; i < (kpad - K) * kT; i += kMlpThreads) A[(K + i / kT) * kP + (i % kT)] = 0.f;
            }
            cp_async_wait_all();
            __syncthreads();
            // weight gradient: G^T A over the rows of the tile, into this CTA's partial
            float* pl = part + (long long)l * kPartPerLin;
            if (N <= 32 && K <= 32) wgrad_tile<2, 2>(G, A, nrows4, N, K, pl, first, tid);
            else if (N <= 32) wgrad_tile<2, 8>(G, A, nrows4, N, K, pl, first, tid);
            else if (K <= 32) wgrad_tile<8, 2>(G, A, nrows4, N, K, pl, first, tid);
            else wgrad_tile<8, 8>(G, A, nrows4, N, K, pl, first, tid);
            if (tid < N) {                                     // bias gradient: the primal rows only
                float sum = 0.f;
                for (int r = nt; r < nt + ns; ++r) sum += G[tid * kP + r];
                float* p = pl + kWElems + tid;
                *p = first ? sum : *p + sum;
            }
            // data gradient: G W
            float acc[8][8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
            if (K <= 64) gemm_rows<1>(G, W, N, ty, tx, acc);
            else gemm_rows<2>(G, W, N, ty, tx, acc);
            __syncthreads();                                   // every read of G, of A as an operand and of W is done
            if (l == 0) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = row_of(ty, i);
                    const bool tangent = r < nt;
                    const int sidx = tangent ? srow[r] : r - nt;
                    if (sidx >= ns) continue;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int c = col_of(tx, j);
                        if (c >= K) continue;
                        if (tangent) P.g_V[((long long)b0 * d + r) * K + c] = acc[i][j];
                        else P.g_x[(long long)(b0 + sidx) * K + c] = acc[i][j];
                    }
                }
            } else {
                // acc = gradient w.r.t. the outputs of hidden layer l - 1 (t on tangent rows, a on primal rows); A
                // still holds those outputs.  Tangent rows: t = u (1 - a^2)  =>  g_u = g_t (1 - a^2), and the mask's
                // gradient g_t u goes to the primal row of the same sample (parked in A, summed below).
                store_tile(G, acc, ty, tx, K, kpad);
                __syncthreads();
                const float* sU = P.save_U + ((long long)tile * nhid + (l - 1)) * kWElems;
                for (int i = tid; i < K * 32; i += kMlpThreads) {
                    const int c = i >> 5, q = i & 31, r0 = q * 4;
                    if (r0 >= nt) continue;
                    float4 g = *reinterpret_cast<const float4*>(G + c * kP + r0);
                    const float4 u = __ldg(reinterpret_cast<const float4*>(sU + c * kT + r0));
                    float4 pu = make_float4(g.x * u.x, g.y * u.y, g.z * u.z, g.w * u.w);
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        if (r0 + e < nt) {
                            const float a = A[c * kP + nt + srow[r0 + e]];
                            f4(g, e) *= 1.f - a * a;
                        }
                    *reinterpret_cast<float4*>(G + c * kP + r0) = g;
                    if (r0 + 4 <= nt) {
                        *reinterpret_cast<float4*>(A + c * kP + r0) = pu;
                    } else {
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            if (r0 + e < nt) A[c * kP + r0 + e] = f4(pu, e);
                    }
                }
                __syncthreads();
                for (int i = tid; i < K * S; i += kMlpThreads) {
                    const int c = i / S, sidx = i - c * S;
                    const float a = A[c * kP + nt + sidx];
                    float gm = 0.f;
                    for (int k = 0; k < d; ++k) gm += A[c * kP + sidx * d + k];
                    float* p = G + c * kP + nt + sidx;
                    *p = sidx < ns ? (*p - 2.f * a * gm) * (1.f - a * a) : 0.f;
                }
            }
            __syncthreads();
        }
    }
}

// g[l] = sum over the launch's CTAs of their partials, in CTA order
__global__ void mlp_reduce_kernel(int nparts, int nlin, const float* __restrict__ partial, float* __restrict__ gpack) {
    const int l = blockIdx.x, j = blockIdx.y, k = threadIdx.x;      // j == 128: the bias row
    const long long off = (long long)l * kPartPerLin + (long long)j * kT + k;
    float s = 0.f;
    for (int c = 0; c < nparts; ++c) s += partial[(long long)c * nlin * kPartPerLin + off];
    gpack[off] = s;
}

int check_shape(const char* who, int B, int d, int nlin, const int* width) {
    CMF_REQUIRE(B >= 0 && d >= 0 && d < kT, "%s: B = %d, d = %d (need 0 <= d <= %d)", who, B, d, kT - 1);
    CMF_REQUIRE(nlin >= 1 && nlin <= kMaxLin && width, "%s: %d Linears (max %d)", who, nlin, kMaxLin);
    for (int l = 0; l <= nlin; ++l)
        CMF_REQUIRE(width[l] >= 1 && width[l] <= kT, "%s: width[%d] = %d outside 1..%d", who, l, width[l], kT);
    return 0;
}

void fill(MlpParams& P, int B, int d, int nlin, const int* width) {
    P.B = B; P.d = d; P.nlin = nlin;
    P.S = kT / (1 + d);
    P.ntiles = (int)ceil_div(B, P.S);
    for (int l = 0; l <= nlin; ++l) P.width[l] = width[l];
}

DeviceOnce g_fwd_attr, g_bwd_attr;
constexpr size_t kFwdSmem = (size_t)(kT * kP + 2 * kWElems + kT) * sizeof(float);
constexpr size_t kBwdSmem = (size_t)(2 * kT * kP + kWElems + kT) * sizeof(float);

}  // namespace
}  // namespace cmf

using namespace cmf;

extern "C" {

int cmf_mlp_usable(int nlin, const int* width, int d) {
    if (nlin < 1 || nlin > kMaxLin || !width || d < 0 || d >= kT) return 0;
    for (int l = 0; l <= nlin; ++l)
        if (width[l] < 1 || width[l] > kT) return 0;
    return 1;
}

long long cmf_mlp_pack_floats(int nlin) { return (long long)nlin * kPackPerLin; }
int cmf_mlp_tiles(int B, int d) { return d < 0 || d >= kT || B < 0 ? 0 : (int)ceil_div(B, kT / (1 + d)); }
long long cmf_mlp_save_floats(int B, int d, int nlin) { return (long long)cmf_mlp_tiles(B, d) * (nlin - 1) * kWElems; }
int cmf_mlp_partials(int B, int d) { const int t = cmf_mlp_tiles(B, d); return t < kNumSMs ? t : kNumSMs; }
long long cmf_mlp_partial_floats(int B, int d, int nlin) { return (long long)cmf_mlp_partials(B, d) * nlin * kPartPerLin; }

int cmf_mlp_pack(int nlin, const int* width, const float* const* W, const float* const* b, float* wpack, cmf_stream_t stream) {
    if (int rc = check_shape("mlp_pack", 0, 0, nlin, width)) return rc;
    CMF_REQUIRE(W && b && wpack, "mlp_pack: null pointer");
    PackParams P;
    P.nlin = nlin;
    P.wpack = wpack;
    for (int l = 0; l <= nlin; ++l) P.width[l] = width[l];
    for (int l = 0; l < nlin; ++l) {
        CMF_REQUIRE(W[l], "mlp_pack: null weight %d", l);
        P.W[l] = W[l];
        P.b[l] = b[l];
    }
    mlp_pack_kernel<<<dim3(nlin, kT), kT, 0, (cudaStream_t)stream>>>(P);
    CMF_CHECK_LAUNCH();
    return 0;
}

int cmf_mlp_jvp_fwd(int B, int d, int nlin, const int* width, const float* x, const float* V, const float* wpack, float* st,
                    float* ST, float* save_T, float* save_U, cmf_stream_t stream) {
    if (int rc = check_shape("mlp_jvp_fwd", B, d, nlin, width)) return rc;
    if (B == 0) return 0;
    CMF_REQUIRE(x && wpack && st && (d == 0 || (V && ST)), "mlp_jvp_fwd: null pointer");
    CMF_REQUIRE((save_T == nullptr) == (save_U == nullptr), "mlp_jvp_fwd: give both activation buffers or neither");
    CMF_REQUIRE(aligned16(wpack) && aligned16(save_T) && aligned16(save_U), "mlp_jvp_fwd: wpack / saved activations must be 16-byte aligned");
    MlpParams P{};
    fill(P, B, d, nlin, width);
    P.x = x; P.V = V; P.wpack = wpack; P.st = st; P.ST = ST; P.save_T = save_T; P.save_U = save_U;
    if (!g_fwd_attr.done()) {
        cudaError_t e = cudaFuncSetAttribute(mlp_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFwdSmem);
        CMF_REQUIRE(e == cudaSuccess, "mlp_jvp_fwd: %s", cudaGetErrorString(e));
        g_fwd_attr.mark();
    }
    const int grid = P.ntiles < kNumSMs ? P.ntiles : kNumSMs;
    mlp_fwd_kernel<<<grid, kMlpThreads, kFwdSmem, (cudaStream_t)stream>>>(P);
    CMF_CHECK_LAUNCH();
    return 0;
}

int cmf_mlp_jvp_bwd(int B, int d, int nlin, const int* width, const float* x, const float* V, const float* wpack,
                    const float* save_T, const float* save_U, const float* g_st, const float* g_ST, float* g_x, float* g_V,
                    float* partial, float* gpack, cmf_stream_t stream) {
    if (int rc = check_shape("mlp_jvp_bwd", B, d, nlin, width)) return rc;
    CMF_REQUIRE(gpack, "mlp_jvp_bwd: null gradient buffer");
    const int nparts = cmf_mlp_partials(B, d);
    if (B > 0) {
        CMF_REQUIRE(x && wpack && g_st && g_x && partial && (d == 0 || (V && g_ST && g_V)), "mlp_jvp_bwd: null pointer");
        CMF_REQUIRE(nlin == 1 || (save_T && (d == 0 || save_U)), "mlp_jvp_bwd: the forward launch's saved activations are required");
        CMF_REQUIRE(aligned16(wpack) && aligned16(save_T) && aligned16(save_U), "mlp_jvp_bwd: wpack / saved activations must be 16-byte aligned");
        MlpParams P{};
        fill(P, B, d, nlin, width);
        P.x = x; P.V = V; P.wpack = wpack; P.save_T = const_cast<float*>(save_T); P.save_U = const_cast<float*>(save_U);
        P.g_st = g_st; P.g_ST = g_ST; P.g_x = g_x; P.g_V = g_V; P.partial = partial;
        if (!g_bwd_attr.done()) {
            cudaError_t e = cudaFuncSetAttribute(mlp_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBwdSmem);
            CMF_REQUIRE(e == cudaSuccess, "mlp_jvp_bwd: %s", cudaGetErrorString(e));
            g_bwd_attr.mark();
        }
        mlp_bwd_kernel<<<nparts, kMlpThreads, kBwdSmem, (cudaStream_t)stream>>>(P);
        CMF_CHECK_LAUNCH();
    }
    mlp_reduce_kernel<<<dim3(nlin, kT + 1), kT, 0, (cudaStream_t)stream>>>(nparts, nlin, partial, gpack);
    CMF_CHECK_LAUNCH();
    return 0;
}

}  // extern "C"
