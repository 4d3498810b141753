// Rational-quadratic spline coupling (inverse direction = the z -> x push of the injective flow),
// primal + all d tangent columns in one pass, and its adjoint.
//
// Reference arithmetic: gitmodules/nsf/nde/transforms/splines/rational_quadratic.py:13-52 (linear
// tails: identity outside [-tail_bound, tail_bound]), :54-134 (inverse branch), bin search
// gitmodules/nsf/utils/torchutils.py:133-138  (sum(x >= knots) - 1 after knots[-1] += 1e-6).
// The reference's cmf wrapper (cmf/models/components/bijections/nsf.py:19-42) has no `_jvp`
// (SURVEY D4): pushing tangents through the spline is a capability this kernel adds; its oracle is
// autograd over the nsf arithmetic (oracle/rqs_oracle.py).
//
// Contract.  The knot tables are built by the host with the reference's own torch ops (so the
// fp32 knots, and therefore the bin indices, are bit-identical to the reference's):
//   cw, ch (B, M, K+1)  cumulative widths / heights in [-tail_bound, tail_bound]
//   dv     (B, M, K+1)  knot derivatives (min_derivative + softplus, end knots pinned to 1)
//   CW, CH, DV (B, d, M, K+1) their tangents.
// Per element, with b = bin(z) in the HEIGHT knots:  x = f(z; cw_b, cw_b+1, ch_b, ch_b+1, dv_b, dv_b+1),
//   x_dot_k = f_z z_dot_k + sum_q f_q q_dot_k      (7 partials per element, shared by all k)
// Bin search: a warp owns 32 consecutive elements; for each of them lane l compares against knot l
// and the bin is popc(ballot) - 1 -- literally the reference's sum of comparisons.
// Adjoint: g_q = g_x f_q + (H r)_q with r = sum_k g_xdot_k q_dot_k (ONE Hessian-vector product per
// element, by linearity), g_qdot_k = g_xdot_k f_q.  f is written once, generically, and instantiated
// on forward-mode jets (gradient) and jets over dual numbers (Hessian-vector product).
#include <math.h>

#include "common.cuh"

namespace cmf {

// ---------------------------------------------------------------- tiny forward-mode AD
struct Dual {           // value + one directional derivative
    float v, e;
};
__device__ __forceinline__ Dual operator+(Dual a, Dual b) { return {a.v + b.v, a.e + b.e}; }
__device__ __forceinline__ Dual operator-(Dual a, Dual b) { return {a.v - b.v, a.e - b.e}; }
__device__ __forceinline__ Dual operator-(Dual a) { return {-a.v, -a.e}; }
__device__ __forceinline__ Dual operator*(Dual a, Dual b) { return {a.v * b.v, a.v * b.e + a.e * b.v}; }
__device__ __forceinline__ Dual operator/(Dual a, Dual b) {
    const float q = a.v / b.v;
    return {q, (a.e - q * b.e) / b.v};
}
__device__ __forceinline__ Dual ad_sqrt(Dual a) {
    const float s = sqrtf(a.v);
    return {s, a.e / (2.f * s)};
}
__device__ __forceinline__ float ad_sqrt(float a) { return sqrtf(a); }
__device__ __forceinline__ Dual ad_const(Dual, float c) { return {c, 0.f}; }
__device__ __forceinline__ float ad_const(float, float c) { return c; }

constexpr int kNP = 7;   // z, cw0, cw1, ch0, ch1, dv0, dv1
template <class S>
struct Jet {             // value + gradient w.r.t. the kNP inputs, over scalar type S
    S v;
    S g[kNP];
};
template <class S> __device__ __forceinline__ Jet<S> operator+(const Jet<S>& a, const Jet<S>& b) {
    Jet<S> r; r.v = a.v + b.v;
#pragma unroll
    for (int i = 0; i < kNP; ++i) r.g[i] = a.g[i] + b.g[i];
    return r;
}
template <class S> __device__ __forceinline__ Jet<S> operator-(const Jet<S>& a, const Jet<S>& b) {
    Jet<S> r; r.v = a.v - b.v;
#pragma unroll
    for (int i = 0; i < kNP; ++i) r.g[i] = a.g[i] - b.g[i];
    return r;
}
template <class S> __device__ __forceinline__ Jet<S> operator-(const Jet<S>& a) {
    Jet<S> r; r.v = -a.v;
#pragma unroll
    for (int i = 0; i < kNP; ++i) r.g[i] = -a.g[i];
    return r;
}
template <class S> __device__ __forceinline__ Jet<S> operator*(const Jet<S>& a, const Jet<S>& b) {
    Jet<S> r; r.v = a.v * b.v;
#pragma unroll
    for (int i = 0; i < kNP; ++i) r.g[i] = a.v * b.g[i] + a.g[i] * b.v;
    return r;
}
template <class S> __device__ __forceinline__ Jet<S> operator/(const Jet<S>& a, const Jet<S>& b) {
    Jet<S> r; r.v = a.v / b.v;
#pragma unroll
    for (int i = 0; i < kNP; ++i) r.g[i] = (a.g[i] - r.v * b.g[i]) / b.v;
    return r;
}
template <class S> __device__ __forceinline__ Jet<S> ad_sqrt(const Jet<S>& a) {
    Jet<S> r; r.v = ad_sqrt(a.v);
    const S two_s = r.v + r.v;
#pragma unroll
    for (int i = 0; i < kNP; ++i) r.g[i] = a.g[i] / two_s;
    return r;
}
template <class S> __device__ __forceinline__ Jet<S> ad_const(const Jet<S>& like, float c) {
    Jet<S> r; r.v = ad_const(like.v, c);
#pragma unroll
    for (int i = 0; i < kNP; ++i) r.g[i] = ad_const(like.v, 0.f);
    return r;
}

// The inverse rational-quadratic map inside one bin, in the reference's operation order
// (rational_quadratic.py:96-112 with input_heights = ch1 - ch0, input_bin_widths = cw1 - cw0).
template <class T>
__device__ __forceinline__ T rqs_inverse_bin(T z, T c0, T c1, T h0, T h1, T d0, T d1) {
    const T w = c1 - c0;
    const T h = h1 - h0;
    const T delta = h / w;
    const T y = z - h0;
    const T two = ad_const(z, 2.f), four = ad_const(z, 4.f);
    const T t = d0 + d1 - two * delta;
    const T a = y * t + h * (delta - d0);
    const T b = h * d0 - y * t;
    const T c = -(delta * y);
    const T disc = b * b - four * a * c;
    const T root = (two * c) / (-b - ad_sqrt(disc));
    return root * w + c0;
}

template <class S>
__device__ __forceinline__ Jet<S> jet_input(S v, int which, S zero, S one) {
    Jet<S> r; r.v = v;
#pragma unroll
    for (int i = 0; i < kNP; ++i) r.g[i] = (i == which) ? one : zero;
    return r;
}

constexpr int kRqsThreads = 128;

// Shared-memory staging of the knot tables (north_star: "knots staged via shared memory"): the rows of
// a warp's 32 consecutive elements are one contiguous block of 32 (K+1) floats, so the warp copies the
// block with consecutive lanes on consecutive floats (K+1 full-line loads instead of 32 partial-row
// loads), then searches / picks out of the tile.
__device__ __forceinline__ void warp_stage_rows(float* tile, const float* __restrict__ table, long long warp_base,
                                                int n, int K, int lane) {
    const int kp = K + 1;
    __syncwarp();
    const float* src = table + warp_base * kp;
    for (int pos = lane; pos < n * kp; pos += 32) tile[pos] = __ldg(src + pos);
    __syncwarp();
}
// bin = popc(ballot(z >= knot_l)) - 1 over the staged height knots (torchutils.py:133-138)
__device__ __forceinline__ int warp_bin_search_tile(const float* tile, int n, int K, float z, int lane) {
    int bin = -1;
    const int kp = K + 1;
    for (int i = 0; i < n; ++i) {
        const float zi = __shfl_sync(0xffffffffu, z, i);
        float knot = 0.f;
        if (lane <= K) knot = tile[i * kp + lane];
        if (lane == K) knot = __fadd_rn(knot, 1e-6f);            // bin_locations[..., -1] += eps
        const unsigned bits = __ballot_sync(0xffffffffu, lane <= K && zi >= knot);
        if (lane == i) bin = __popc(bits) - 1;
    }
    return bin;
}

struct RqsElem {
    bool inside;
    int bin;
    float p[kNP];        // z, cw0, cw1, ch0, ch1, dv0, dv1
};

// the same element record with all three tables read through the shared-memory tile (all lanes of the
// warp must call this: it stages cooperatively)
__device__ __forceinline__ RqsElem rqs_load_elem_staged(float* tile, bool valid, long long warp_base, int n, int K, float tb,
                                                        float z, int lane, const float* __restrict__ cw,
                                                        const float* __restrict__ ch, const float* __restrict__ dv) {
    RqsElem E;
    const int kp = K + 1;
    warp_stage_rows(tile, ch, warp_base, n, K, lane);
    const int bin = warp_bin_search_tile(tile, n, K, z, lane);
    E.inside = valid && z >= -tb && z <= tb;
    E.bin = E.inside ? min(max(bin, 0), K - 1) : -1;
    E.p[0] = z;
#pragma unroll
    for (int q = 1; q < kNP; ++q) E.p[q] = 0.f;
    const int o = lane * kp + (E.inside ? E.bin : 0);
    if (E.inside) { E.p[3] = tile[o]; E.p[4] = tile[o + 1]; }
    warp_stage_rows(tile, cw, warp_base, n, K, lane);
    if (E.inside) { E.p[1] = tile[o]; E.p[2] = tile[o + 1]; }
    warp_stage_rows(tile, dv, warp_base, n, K, lane);
    if (E.inside) { E.p[5] = tile[o]; E.p[6] = tile[o + 1]; }
    __syncwarp();
    return E;
}

__device__ __forceinline__ Jet<float> rqs_eval_grad(const RqsElem& E) {
    Jet<float> in[kNP];
#pragma unroll
    for (int q = 0; q < kNP; ++q) in[q] = jet_input<float>(E.p[q], q, 0.f, 1.f);
    return rqs_inverse_bin(in[0], in[1], in[2], in[3], in[4], in[5], in[6]);
}

__global__ void __launch_bounds__(kRqsThreads)
rqs_coupling_fwd_kernel(int B, int d, int M, int K, float tb,
                        const float* __restrict__ z_m, const float* __restrict__ V_m,
                        const float* __restrict__ cw, const float* __restrict__ ch, const float* __restrict__ dv,
                        const float* __restrict__ CW, const float* __restrict__ CH, const float* __restrict__ DV,
                        float* __restrict__ x_m, float* __restrict__ V_out, int32_t* __restrict__ bin_out) {
    __shared__ float tiles[kRqsThreads / 32][32 * 32];          // K + 1 <= 32 floats per row
    const long long total = (long long)B * M;
    const long long e = (long long)blockIdx.x * kRqsThreads + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool valid = e < total;
    const long long e0 = e - lane;
    if (e0 >= total) return;                                      // warp-uniform
    const int n = (int)min((long long)32, total - e0);
    const float z = valid ? __ldg(z_m + e) : 0.f;
    const RqsElem E = rqs_load_elem_staged(tiles[threadIdx.x >> 5], valid, e0, n, K, tb, z, lane, cw, ch, dv);
    if (!valid) return;
    Jet<float> f;
    if (E.inside) f = rqs_eval_grad(E);
    x_m[e] = E.inside ? f.v : z;
    if (bin_out) bin_out[e] = E.bin;
    if (d == 0) return;
    const int b = (int)(e / M), j = (int)(e - (long long)b * M);
    for (int k = 0; k < d; ++k) {
        const long long ek = ((long long)b * d + k) * M + j;
        const float zd = __ldg(V_m + ek);
        float xd = zd;
        if (E.inside) {
            const long long o = ek * (K + 1) + E.bin;
            xd = f.g[0] * zd;
            xd = fmaf(f.g[1], __ldg(CW + o), xd); xd = fmaf(f.g[2], __ldg(CW + o + 1), xd);
            xd = fmaf(f.g[3], __ldg(CH + o), xd); xd = fmaf(f.g[4], __ldg(CH + o + 1), xd);
            xd = fmaf(f.g[5], __ldg(DV + o), xd); xd = fmaf(f.g[6], __ldg(DV + o + 1), xd);
        }
        V_out[ek] = xd;
    }
}

__device__ __forceinline__ void rqs_store_row(float* row, int K, int bin, float v0, float v1) {
    for (int l = 0; l <= K; ++l) row[l] = (l == bin) ? v0 : ((l == bin + 1) ? v1 : 0.f);
}

// The adjoint writes DENSE rows of K+1 floats per element (zero outside the two selected knots).  A
// thread storing its own row puts 32 lanes on 32 different sectors per instruction (stride K+1 floats:
// the 35 %-of-HBM finding of round 1).  The rows of a warp's 32 consecutive elements are ONE contiguous
// block of 32 (K+1) floats, so each lane writes its row into a shared-memory tile and the warp copies
// the tile out with consecutive lanes on consecutive floats (full 128-byte lines).
// `n` = rows of this warp that exist, `dst` = global address of row 0 of the block.
__device__ __forceinline__ void rqs_store_rows_coalesced(float* tile, float* dst, int K, int lane, int n,
                                                         int bin, float v0, float v1) {
    const int kp = K + 1;
    float* mine = tile + lane * kp;
    for (int l = 0; l < kp; ++l) mine[l] = (l == bin) ? v0 : ((l == bin + 1) ? v1 : 0.f);
    __syncwarp();
    for (int pos = lane; pos < n * kp; pos += 32) __stcs(dst + pos, tile[pos]);
    __syncwarp();
}

__global__ void __launch_bounds__(kRqsThreads)
rqs_coupling_bwd_kernel(int B, int d, int M, int K, float tb,
                        const float* __restrict__ z_m, const float* __restrict__ V_m,
                        const float* __restrict__ cw, const float* __restrict__ ch, const float* __restrict__ dv,
                        const float* __restrict__ CW, const float* __restrict__ CH, const float* __restrict__ DV,
                        const float* __restrict__ g_x, const float* __restrict__ g_V,
                        float* __restrict__ g_z, float* __restrict__ g_Vin,
                        float* __restrict__ g_cw, float* __restrict__ g_ch, float* __restrict__ g_dv,
                        float* __restrict__ g_CW, float* __restrict__ g_CH, float* __restrict__ g_DV) {
    __shared__ float tiles[kRqsThreads / 32][32 * 32];          // K + 1 <= 32 floats per row
    const long long total = (long long)B * M;
    const long long e = (long long)blockIdx.x * kRqsThreads + threadIdx.x;
    const int lane = threadIdx.x & 31;
    float* tile = tiles[threadIdx.x >> 5];
    const bool valid = e < total;
    const long long e0 = e - lane;                                // the warp's first element
    if (e0 >= total) return;                                      // whole warp past the end (warp-uniform)
    const int n = (int)min((long long)32, total - e0);            // rows of this warp that exist
    const float z = valid ? __ldg(z_m + e) : 0.f;
    const RqsElem E = rqs_load_elem_staged(tile, valid, e0, n, K, tb, z, lane, cw, ch, dv);
    const bool inside = valid && E.inside;
    Jet<float> f;
    if (inside) f = rqs_eval_grad(E);
    const float gx = (valid && g_x) ? __ldg(g_x + e) : 0.f;
    float r[kNP];
#pragma unroll
    for (int q = 0; q < kNP; ++q) r[q] = 0.f;
    const int b = valid ? (int)(e / M) : 0, j = valid ? (int)(e - (long long)b * M) : 0;
    // fast path: the warp's rows lie in ONE sample, so for every tangent k they are one contiguous block
    const int b0 = (int)(e0 / M);
    const bool one_sample = (e0 + n - 1) / M == b0;
    const int j0 = (int)(e0 - (long long)b0 * M);
    const int sbin = inside ? E.bin : -2;                         // -2: no knot selected, an all-zero row
    for (int k = 0; k < d; ++k) {
        const long long ek = ((long long)b * d + k) * M + j;
        const float gk = valid ? __ldg(g_V + ek) : 0.f;
        const long long row = ek * (K + 1);
        float a[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (inside) {
            const long long o = row + E.bin;
            r[0] = fmaf(gk, __ldg(V_m + ek), r[0]);
            r[1] = fmaf(gk, __ldg(CW + o), r[1]); r[2] = fmaf(gk, __ldg(CW + o + 1), r[2]);
            r[3] = fmaf(gk, __ldg(CH + o), r[3]); r[4] = fmaf(gk, __ldg(CH + o + 1), r[4]);
            r[5] = fmaf(gk, __ldg(DV + o), r[5]); r[6] = fmaf(gk, __ldg(DV + o + 1), r[6]);
#pragma unroll
            for (int q = 0; q < 6; ++q) a[q] = gk * f.g[q + 1];
        }
        if (valid) g_Vin[ek] = inside ? gk * f.g[0] : gk;
        if (one_sample) {
            const long long blk = (((long long)b0 * d + k) * M + j0) * (K + 1);
            rqs_store_rows_coalesced(tile, g_CW + blk, K, lane, n, sbin, a[0], a[1]);
            rqs_store_rows_coalesced(tile, g_CH + blk, K, lane, n, sbin, a[2], a[3]);
            rqs_store_rows_coalesced(tile, g_DV + blk, K, lane, n, sbin, a[4], a[5]);
        } else if (valid) {                                       // a warp straddling two samples (M < 32 or a sample end)
            rqs_store_row(g_CW + row, K, sbin, a[0], a[1]);
            rqs_store_row(g_CH + row, K, sbin, a[2], a[3]);
            rqs_store_row(g_DV + row, K, sbin, a[4], a[5]);
        }
    }
    float a[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (inside) {
        // Hessian-vector product H r: the gradient jet evaluated over duals in direction r
        Jet<Dual> in[kNP];
#pragma unroll
        for (int q = 0; q < kNP; ++q) in[q] = jet_input<Dual>(Dual{E.p[q], r[q]}, q, Dual{0.f, 0.f}, Dual{1.f, 0.f});
        const Jet<Dual> h = rqs_inverse_bin(in[0], in[1], in[2], in[3], in[4], in[5], in[6]);
        g_z[e] = fmaf(gx, f.g[0], h.g[0].e);
#pragma unroll
        for (int q = 0; q < 6; ++q) a[q] = fmaf(gx, f.g[q + 1], h.g[q + 1].e);
    } else if (valid) {
        g_z[e] = gx;
    }
    const long long blk = e0 * (K + 1);                           // primal rows: always one contiguous block
    rqs_store_rows_coalesced(tile, g_cw + blk, K, lane, n, sbin, a[0], a[1]);
    rqs_store_rows_coalesced(tile, g_ch + blk, K, lane, n, sbin, a[2], a[3]);
    rqs_store_rows_coalesced(tile, g_dv + blk, K, lane, n, sbin, a[4], a[5]);
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_rqs_coupling_jvp_fwd(int B, int d, int M, int K, float tail_bound,
                                        const float* z_m, const float* V_m,
                                        const float* cw, const float* ch, const float* dv,
                                        const float* CW, const float* CH, const float* DV,
                                        float* x_m, float* V_out, int32_t* bin, cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && d >= 0 && M >= 0, "rqs_fwd: negative size");
    CMF_REQUIRE(K >= 1 && K <= 31, "rqs_fwd: need 1 <= num_bins <= 31 (one knot per lane), got %d", K);
    CMF_REQUIRE(tail_bound > 0.f, "rqs_fwd: tail_bound must be positive");
    if (B == 0 || M == 0) return 0;
    CMF_REQUIRE(z_m && cw && ch && dv && x_m, "rqs_fwd: null primal pointer");
    CMF_REQUIRE(d == 0 || (V_m && CW && CH && DV && V_out), "rqs_fwd: null tangent pointer");
    const long long total = (long long)B * M;
    rqs_coupling_fwd_kernel<<<(unsigned)ceil_div(total, kRqsThreads), kRqsThreads, 0, (cudaStream_t)stream>>>(
        B, d, M, K, tail_bound, z_m, V_m, cw, ch, dv, CW, CH, DV, x_m, V_out, bin);
    CMF_CHECK_LAUNCH();
    return 0;
}

extern "C" int cmf_rqs_coupling_jvp_bwd(int B, int d, int M, int K, float tail_bound,
                                        const float* z_m, const float* V_m,
                                        const float* cw, const float* ch, const float* dv,
                                        const float* CW, const float* CH, const float* DV,
                                        const float* g_x, const float* g_V,
                                        float* g_z, float* g_Vin, float* g_cw, float* g_ch, float* g_dv,
                                        float* g_CW, float* g_CH, float* g_DV, cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && d >= 0 && M >= 0, "rqs_bwd: negative size");
    CMF_REQUIRE(K >= 1 && K <= 31, "rqs_bwd: need 1 <= num_bins <= 31 (one knot per lane), got %d", K);
    CMF_REQUIRE(tail_bound > 0.f, "rqs_bwd: tail_bound must be positive");
    if (B == 0 || M == 0) return 0;
    CMF_REQUIRE(z_m && cw && ch && dv && g_z && g_cw && g_ch && g_dv, "rqs_bwd: null primal pointer");
    CMF_REQUIRE(d == 0 || (V_m && CW && CH && DV && g_V && g_Vin && g_CW && g_CH && g_DV), "rqs_bwd: null tangent pointer");
    const long long total = (long long)B * M;
    rqs_coupling_bwd_kernel<<<(unsigned)ceil_div(total, kRqsThreads), kRqsThreads, 0, (cudaStream_t)stream>>>(
        B, d, M, K, tail_bound, z_m, V_m, cw, ch, dv, CW, CH, DV, g_x, g_V, g_z, g_Vin, g_cw, g_ch, g_dv,
        g_CW, g_CH, g_DV);
    CMF_CHECK_LAUNCH();
    return 0;
}
