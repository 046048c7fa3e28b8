// 1-D integer transforms of the AV1 residual path as fully unrolled register networks.
//
// Every function works in place on a compile-time sized array that lives in registers; all
// indices, cosine weights and rounding constants fold to immediates after unrolling, so a
// half-butterfly is two IMADs and a shift.  Arithmetic is 32-bit wrapping, i.e. what the
// encoder executes on an AVX2 host (mullo/add/add/srai, reference
// Source/Lib/Common/ASM_AVX2/highbd_fwd_txfm_avx2.c:216-227) and identical to the C definition
// (Source/Lib/Common/Codec/EbTransforms.c:1288-1299) on every legal input.
//
// Network definitions follow Source/Lib/Common/Codec/EbTransforms.c: forward DCT :1314-2762,
// forward ADST :2764-3183, forward identity :3620-3659; inverse DCT :5465-5745 and :6938-7716,
// inverse ADST :6097-6502, inverse identity :7717-7755, clamp_value :5458.  They are produced
// here by the recursion that generates them (even half = DCT-N/2 of the input butterflies, odd
// half = alternating rotation / butterfly stages), not written out stage by stage.
//
// The header also compiles as plain C++ (SVT_HD empty) so that tests can run the very same
// templates on the host against the oracle without a GPU.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define SVT_HD __host__ __device__ __forceinline__
#define SVT_TABLE __device__
#define SVT_NOINLINE static __host__ __device__ __noinline__
#else
#define SVT_HD inline __attribute__((always_inline))
#define SVT_TABLE
#define SVT_NOINLINE inline
#endif

namespace svt {

// ---- trig tables, built at compile time -------------------------------------------------------
// cospi[b][j] = round(cos(pi*j/128) * 2^(10+b)), EbTransforms.c:1241
constexpr double kPi = 3.14159265358979323846264338327950288;
constexpr double ccos(double x) {  // Taylor, |x| <= pi/2: error < 1e-17
    double t = 1.0, s = 1.0;
    for (int k = 1; k <= 14; ++k) {
        t = -t * x * x / double((2 * k - 1) * (2 * k));
        s += t;
    }
    return s;
}
struct CosTab { int v[4][64]; };
constexpr CosTab make_cos() {
    CosTab t{};
    for (int b = 0; b < 4; ++b)
        for (int j = 0; j < 64; ++j) {
            double x = ccos(kPi * j / 128.0) * double(1 << (10 + b));
            t.v[b][j] = int(x + 0.5);
        }
    return t;
}
SVT_TABLE constexpr CosTab kCos = make_cos();
// sinpi for ADST4 at 10..13 bits, EbTransforms.c:1303-1309 (data: the 13-bit row is not plain
// rounding, the reference keeps [1]+[2]==[4])
SVT_TABLE constexpr int kSin[4][5] = { { 0, 330, 621, 836, 951 }, { 0, 660, 1241, 1672, 1901 },
                                       { 0, 1321, 2482, 3344, 3803 }, { 0, 2642, 4964, 6689, 7606 } };

SVT_HD constexpr int brev(int v, int bits) {
    int r = 0;
    for (int i = 0; i < bits; ++i) r |= ((v >> i) & 1) << (bits - 1 - i);
    return r;
}
SVT_HD constexpr int ilog2(int n) { int l = 0; while ((1 << l) < n) ++l; return l; }

// half_btf with wrapping 32-bit arithmetic
template <int BIT>
SVT_HD int hb(int w0, int a, int w1, int b) {
    const unsigned r = (unsigned(w0) * unsigned(a) + (1u << (BIT - 1))) + unsigned(w1) * unsigned(b);
    return int(r) >> BIT;
}
template <int BIT>
SVT_HD int rshift(int v) { return int(unsigned(v) + (1u << (BIT - 1))) >> BIT; }
// half_btf followed by round_shift(., POST), as one rounding:
//   ((z >> BIT) + 2^(POST-1)) >> POST  ==  (z + 2^(BIT+POST-1)) >> (BIT + POST)   (floor of a floor)
// with z = w0*a + w1*b + 2^(BIT-1).  Exact as long as z + 2^(BIT+POST-1) does not wrap, which legal
// residuals are far from (|half_btf output| < 2^17 on the forward path, the wrap is at 2^18).
// Used for the last stage of the forward transforms, whose outputs go straight into the
// round_shift of av1_round_shift_array_c (EbTransforms.c:3682, :3747, :3768).
// round_shift(., BIT) followed by round_shift(., POST), as one rounding
template <int BIT, int POST>
SVT_HD int rshift2(int v) {
    if constexpr (POST == 0) return rshift<BIT>(v);
    else return int(unsigned(v) + ((1u << (BIT - 1)) + (1u << (BIT + POST - 1)))) >> (BIT + POST);
}
template <int BIT, int POST>
SVT_HD int hbp(int w0, int a, int w1, int b) {
    if constexpr (POST == 0) return hb<BIT>(w0, a, w1, b);
    else {
        const unsigned r = (unsigned(w0) * unsigned(a) + ((1u << (BIT - 1)) + (1u << (BIT + POST - 1)))) +
                           unsigned(w1) * unsigned(b);
        return int(r) >> (BIT + POST);
    }
}

// Clamp policy: forward transforms do not clamp; inverse transforms clamp every add/sub to the
// stage range (16 or 18 bits, av1_gen_inv_stage_range EbTransforms.c:5404).
struct NoClamp {
    SVT_HD int operator()(int v) const { return v; }
    SVT_HD int add(int a, int b) const { return int(unsigned(a) + unsigned(b)); }
    SVT_HD int sub(int a, int b) const { return int(unsigned(a) - unsigned(b)); }
};
struct Clamp {
    int lo, hi;
    SVT_HD explicit Clamp(int bits) : lo(-(1 << (bits - 1))), hi((1 << (bits - 1)) - 1) {}
    SVT_HD int operator()(int v) const { return v < lo ? lo : (v > hi ? hi : v); }
    // clamp(a +- b): on sm_100a the add and the first bound fuse into one VIADDMNMX
    SVT_HD int add(int a, int b) const {
#if defined(__CUDA_ARCH__)
        return min(__viaddmax_s32(a, b, lo), hi);
#else
        return (*this)(a + b);
#endif
    }
    SVT_HD int sub(int a, int b) const {
#if defined(__CUDA_ARCH__)
        return min(__viaddmax_s32(a, -b, lo), hi);
#else
        return (*this)(a - b);
#endif
    }
};

#define SVT_C(j) (kCos.v[BIT - 10][(j)])

// ---- DCT ---------------------------------------------------------------------------------------
// x is in "position" order (the reference's bf[] index before its final bit-reversal stage).
template <int M, int W, class K>
SVT_HD void odd_bfly(int* y, K k) {
#pragma unroll
    for (int g = 0; g < M / W; ++g)
#pragma unroll
        for (int i = 0; i < W / 2; ++i) {
            const int ia = g * W + i, ib = g * W + W - 1 - i;
            const int a = y[ia], b = y[ib];
            if (g & 1) { y[ia] = k.sub(b, a); y[ib] = k.add(b, a); }
            else       { y[ia] = k.add(a, b); y[ib] = k.sub(a, b); }
        }
}
template <int M, int W, int BIT, bool INV>
SVT_HD void odd_rot(int* y) {
    constexpr int G = M / W;
#pragma unroll
    for (int g = 0; g < G / 2; ++g) {
        const int ca = (32 / G) * (1 + 4 * brev(g, ilog2(G / 2))), cb = 64 - ca;
#pragma unroll
        for (int j = W / 4; j < 3 * W / 4; ++j) {
            const int p = g * W + j, q = M - 1 - p;
            const int P = y[p], Q = y[q];
            if (j < W / 2) {
                y[p] = hb<BIT>(-SVT_C(ca), P, SVT_C(cb), Q);
                y[q] = INV ? hb<BIT>(SVT_C(cb), P, SVT_C(ca), Q) : hb<BIT>(SVT_C(ca), Q, SVT_C(cb), P);
            } else {
                y[p] = hb<BIT>(-SVT_C(cb), P, -SVT_C(ca), Q);
                y[q] = INV ? hb<BIT>(-SVT_C(ca), P, SVT_C(cb), Q) : hb<BIT>(SVT_C(cb), Q, -SVT_C(ca), P);
            }
        }
    }
}
template <int M, int BIT>
SVT_HD void mid_rot45(int* y) {  // pi/4 rotation of the middle half of the odd part
#pragma unroll
    for (int p = M / 4; p < M / 2; ++p) {
        const int q = M - 1 - p;
        const int P = y[p], Q = y[q];
        // forward: (-c32*P + c32*Q, c32*Q + c32*P); the inverse is the same (symmetric) matrix
        y[p] = hb<BIT>(-SVT_C(32), P, SVT_C(32), Q);
        y[q] = hb<BIT>(SVT_C(32), P, SVT_C(32), Q);
    }
}
template <int N, int BIT, int POST = 0>
SVT_HD void end_rot_fwd(int* y) {
    constexpr int M = N / 2;
#pragma unroll
    for (int i = 0; i < M / 2; ++i) {
        const int b = (64 / N) * (1 + 4 * brev(i, ilog2(M / 2))), a = 64 - b;
        const int lo = y[i], hi = y[M - 1 - i];
        y[i] = hbp<BIT, POST>(SVT_C(a), lo, SVT_C(b), hi);
        y[M - 1 - i] = hbp<BIT, POST>(SVT_C(a), hi, -SVT_C(b), lo);
    }
}
template <int N, int BIT>
SVT_HD void end_rot_inv(int* y) {
    constexpr int M = N / 2;
#pragma unroll
    for (int i = 0; i < M / 2; ++i) {
        const int b = (64 / N) * (1 + 4 * brev(i, ilog2(M / 2))), a = 64 - b;
        const int lo = y[i], hi = y[M - 1 - i];
        y[i] = hb<BIT>(SVT_C(a), lo, -SVT_C(b), hi);
        y[M - 1 - i] = hb<BIT>(SVT_C(b), lo, SVT_C(a), hi);
    }
}

template <int M, int W, int BIT>
SVT_HD void fwd_odd_stages(int* y) {  // W runs M/2, M/4, ..., 2
    if constexpr (W >= 2) {
        odd_bfly<M, W>(y, NoClamp());
        if constexpr (W >= 4) odd_rot<M, W, BIT, false>(y);
        fwd_odd_stages<M, W / 2, BIT>(y);
    }
}
template <int M, int W, int BIT, class K>
SVT_HD void inv_odd_stages(int* y, K k) {  // W runs 2, 4, ..., M/2
    if constexpr (W <= M / 2) {
        if constexpr (W >= 4) odd_rot<M, W, BIT, true>(y);
        odd_bfly<M, W>(y, k);
        inv_odd_stages<M, 2 * W, BIT>(y, k);
    }
}

template <int N, int BIT> struct Dct {
    template <int POST = 0> SVT_HD static void fwd_pos(int* x) {
#pragma unroll
        for (int i = 0; i < N / 2; ++i) {
            const int a = x[i], b = x[N - 1 - i];
            x[i] = a + b;
            x[N - 1 - i] = a - b;
        }
        Dct<N / 2, BIT>::template fwd_pos<POST>(x);
        int* y = x + N / 2;
        constexpr int M = N / 2;
        if constexpr (M >= 4) mid_rot45_fwd(y);
        fwd_odd_stages<M, M / 2, BIT>(y);
        end_rot_fwd<N, BIT, POST>(y);  // every forward DCT output leaves through a last-stage half_btf
    }
    SVT_HD static void mid_rot45_fwd(int* y) {
        constexpr int M = N / 2;
#pragma unroll
        for (int p = M / 4; p < M / 2; ++p) {
            const int q = M - 1 - p;
            const int lo = y[p], hi = y[q];
            y[p] = hb<BIT>(-SVT_C(32), lo, SVT_C(32), hi);
            y[q] = hb<BIT>(SVT_C(32), hi, SVT_C(32), lo);
        }
    }
    template <class K> SVT_HD static void inv_pos(int* x, K k) {
        Dct<N / 2, BIT>::inv_pos(x, k);
        int* y = x + N / 2;
        constexpr int M = N / 2;
        end_rot_inv<N, BIT>(y);
        inv_odd_stages<M, 2, BIT>(y, k);
        if constexpr (M >= 4) mid_rot45<M, BIT>(y);
#pragma unroll
        for (int i = 0; i < N / 2; ++i) {
            const int a = x[i], b = x[N - 1 - i];
            x[i] = k.add(a, b);
            x[N - 1 - i] = k.sub(a, b);
        }
    }
};
template <int BIT> struct Dct<2, BIT> {
    template <int POST = 0> SVT_HD static void fwd_pos(int* x) {
        const int a = x[0], b = x[1];
        x[0] = hbp<BIT, POST>(SVT_C(32), a, SVT_C(32), b);
        x[1] = hbp<BIT, POST>(-SVT_C(32), b, SVT_C(32), a);
    }
    template <class K> SVT_HD static void inv_pos(int* x, K) {
        const int a = x[0], b = x[1];
        x[0] = hb<BIT>(SVT_C(32), a, SVT_C(32), b);
        x[1] = hb<BIT>(SVT_C(32), a, -SVT_C(32), b);
    }
};

// POST: a round_shift(., POST) of every output folded into the last stage (see hbp)
template <int N, int BIT, int POST = 0>
SVT_HD void fdct(int (&x)[N]) {
    Dct<N, BIT>::template fwd_pos<POST>(x);
    int t[N];
#pragma unroll
    for (int k = 0; k < N; ++k) t[k] = x[brev(k, ilog2(N))];
#pragma unroll
    for (int k = 0; k < N; ++k) x[k] = t[k];
}
template <int N, int BIT, class K>
SVT_HD void idct(int (&x)[N], K k) {
    int t[N];
#pragma unroll
    for (int p = 0; p < N; ++p) t[p] = x[brev(p, ilog2(N))];
    Dct<N, BIT>::inv_pos(t, k);
#pragma unroll
    for (int p = 0; p < N; ++p) x[p] = t[p];
}

// ---- ADST 8 / 16 ---------------------------------------------------------------------------------
// fadst input permutation with signs (EbTransforms.c:2872-2879, :2986-3001); the inverse uses the
// same table for its signed output permutation (:6285-6292, :6484-6499).
SVT_HD constexpr int adst_perm(int n, int i) {
    constexpr int p8[8] = { 0, -7, -3, 4, -1, 6, 2, -5 };
    constexpr int p16[16] = { 0, -15, -7, 8, -3, 12, 4, -11, -1, 14, 6, -9, 2, -13, -5, 10 };
    return n == 8 ? p8[i & 7] : p16[i & 15];
}
template <int N, int GS, int BIT>
SVT_HD void adst_rot_group(int* x) {
    constexpr int NP = GS / 4, NA = NP > 1 ? NP / 2 : 1;
#pragma unroll
    for (int g = 0; g < N; g += GS)
#pragma unroll
        for (int k = 0; k < NP; ++k) {
            const int ia = g + GS / 2 + 2 * k, ib = ia + 1;
            const int a = x[ia], b = x[ib];
            const int u = (128 / GS) * (1 + 4 * brev(k % NA, ilog2(NA)));
            if (k < NA) {
                x[ia] = hb<BIT>(SVT_C(u), a, SVT_C(64 - u), b);
                x[ib] = hb<BIT>(SVT_C(64 - u), a, -SVT_C(u), b);
            } else {
                x[ia] = hb<BIT>(-SVT_C(64 - u), a, SVT_C(u), b);
                x[ib] = hb<BIT>(SVT_C(u), a, SVT_C(64 - u), b);
            }
        }
}
template <int N, int S, class K>
SVT_HD void adst_bfly(int* x, K k) {
#pragma unroll
    for (int g = 0; g < N; g += 2 * S)
#pragma unroll
        for (int i = 0; i < S; ++i) {
            const int a = x[g + i], b = x[g + i + S];
            x[g + i] = k.add(a, b);
            x[g + i + S] = k.sub(a, b);
        }
}
template <int N, int BIT, int POST = 0>
SVT_HD void adst_last_rot(int* x) {
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
        const int u = 32 / N + (128 / N) * i;
        const int a = x[2 * i], b = x[2 * i + 1];
        x[2 * i] = hbp<BIT, POST>(SVT_C(u), a, SVT_C(64 - u), b);
        x[2 * i + 1] = hbp<BIT, POST>(SVT_C(64 - u), a, -SVT_C(u), b);
    }
}
template <int N, int GS, int BIT>
SVT_HD void fadst_stages(int* x) {  // GS runs 4, 8, ..., N
    if constexpr (GS <= N) {
        adst_rot_group<N, GS, BIT>(x);
        adst_bfly<N, GS / 2>(x, NoClamp());
        fadst_stages<N, GS * 2, BIT>(x);
    }
}
template <int N, int GS, int BIT, class K>
SVT_HD void iadst_stages(int* x, K k) {  // GS runs N, N/2, ..., 4
    if constexpr (GS >= 4) {
        adst_bfly<N, GS / 2>(x, k);
        adst_rot_group<N, GS, BIT>(x);
        iadst_stages<N, GS / 2, BIT>(x, k);
    }
}

// ADST4, EbTransforms.c:2764-2854 / :6097-6185 (no clamps; all-zero early-out is a no-op here
// because the arithmetic maps zeros to zeros)
template <int BIT, int POST = 0>
SVT_HD void fadst4(int (&x)[4]) {
    const unsigned s1 = kSin[BIT - 10][1], s2 = kSin[BIT - 10][2], s3 = kSin[BIT - 10][3],
                   s4 = kSin[BIT - 10][4];
    const unsigned x0 = x[0], x1 = x[1], x2 = x[2], x3 = x[3];
    const unsigned a0 = s1 * x0 + s2 * x1 + s4 * x3;
    const unsigned a1 = s3 * (x0 + x1 - x3);
    const unsigned a2 = s4 * x0 - s1 * x1 + s2 * x3;
    const unsigned a3 = s3 * x2;
    x[0] = rshift2<BIT, POST>(int(a0 + a3));
    x[1] = rshift2<BIT, POST>(int(a1));
    x[2] = rshift2<BIT, POST>(int(a2 - a3));
    x[3] = rshift2<BIT, POST>(int(a2 - a0 + a3));
}
template <int BIT>
SVT_HD void iadst4(int (&x)[4]) {
    const unsigned s1 = kSin[BIT - 10][1], s2 = kSin[BIT - 10][2], s3 = kSin[BIT - 10][3],
                   s4 = kSin[BIT - 10][4];
    const unsigned x0 = x[0], x1 = x[1], x2 = x[2], x3 = x[3];
    const unsigned a0 = s1 * x0 + s4 * x2 + s2 * x3;
    const unsigned a1 = s2 * x0 - s1 * x2 - s4 * x3;
    const unsigned a3 = s3 * x1;
    const unsigned a2 = s3 * (x0 - x2 + x3);
    x[0] = rshift<BIT>(int(a0 + a3));
    x[1] = rshift<BIT>(int(a1 + a3));
    x[2] = rshift<BIT>(int(a2));
    x[3] = rshift<BIT>(int(a0 + a1 - a3));
}

template <int N, int BIT, int POST = 0>
SVT_HD void fadst(int (&x)[N]) {
    if constexpr (N == 4) {
        fadst4<BIT, POST>(x);
    } else {
        int t[N];
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const int p = adst_perm(N, i);
            t[i] = p < 0 ? -x[-p] : x[p];
        }
        fadst_stages<N, 4, BIT>(t);
        adst_last_rot<N, BIT, POST>(t);
#pragma unroll
        for (int j = 0; j < N / 2; ++j) {
            x[2 * j] = t[2 * j + 1];
            x[2 * j + 1] = t[N - 2 - 2 * j];
        }
    }
}
template <int N, int BIT, class K>
SVT_HD void iadst(int (&x)[N], K k) {
    if constexpr (N == 4) {
        iadst4<BIT>(x);
    } else {
        int t[N];
#pragma unroll
        for (int i = 0; i < N / 2; ++i) {
            t[2 * i] = x[N - 1 - 2 * i];
            t[2 * i + 1] = x[2 * i];
        }
        adst_last_rot<N, BIT>(t);
        iadst_stages<N, N, BIT>(t, k);
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const int p = adst_perm(N, i);
            if (p < 0) x[-p] = -t[i]; else x[p] = t[i];
        }
    }
}

// ---- identity, EbTransforms.c:3620-3659 / :7717-7755 ---------------------------------------------
// 64-bit product then round_shift(., 12) for the sqrt2 sizes.
SVT_HD int mul_round12(int v, int m) {
    return int((int64_t(v) * m + 2048) >> 12);
}
template <int N>
SVT_HD void identity(int (&x)[N]) {
#pragma unroll
    for (int i = 0; i < N; ++i) {
        if constexpr (N == 4) x[i] = mul_round12(x[i], 5793);
        else if constexpr (N == 8) x[i] = int(unsigned(x[i]) * 2u);
        else if constexpr (N == 16) x[i] = mul_round12(x[i], 2 * 5793);
        else if constexpr (N == 32) x[i] = int(unsigned(x[i]) * 4u);
        else x[i] = mul_round12(x[i], 4 * 5793);
    }
}

#undef SVT_C
}  // namespace svt
