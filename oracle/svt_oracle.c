/*
 * TEST INFRASTRUCTURE ONLY -- see svt_oracle.h.  Plain-C restatement of the reference's
 * residual-coding path (forward 2-D transform, quantize/dequantize/eob, inverse transform +
 * reconstruction).  Parity PINNED against oracle/_ref (the reference compiled from source).
 *
 * The reference spells every butterfly network out stage by stage (EbTransforms.c:1314-3618
 * forward, :5465-6937 inverse).  Here the same networks are produced by the recursion that
 * generates them: DCT-N = input butterflies + DCT-N/2 on the sums + an "odd half" made of
 * alternating rotation / butterfly stages; the ADSTs by their stride-doubling stages.  The
 * rounding points (one round_shift per half_btf, EbTransforms.c:1288-1299) and, for the
 * inverse, the clamp after every add/sub (EbTransforms.c:5458) are what make the result
 * bit-exact, and they are identical to the reference's.
 *
 * Build: gcc -O2 -fwrapv (32-bit products wrap exactly as the reference's do in practice).
 */
#include "svt_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * trigonometric tables
 * cospi[b-10][j] = round(cos(pi*j/128) * 2^b)                    (EbTransforms.c:1241-1284)
 * sinpi[b-10][j] ~ sqrt(2)*sin(j*pi/9)*2/3 * 2^b                 (EbTransforms.c:1301-1312)
 * ---------------------------------------------------------------------------------------- */
static int32_t g_cospi[7][64];
/* Not reproducible by rounding alone: the reference nudged [2] and [4] at 13 and 16 bits so that
 * [1] + [2] == [4] holds with its own choice of which entry yields (EbTransforms.c:1303-1309),
 * so these 35 values are carried as data. */
static const int32_t g_sinpi[7][5] = {
    { 0, 330, 621, 836, 951 },        { 0, 660, 1241, 1672, 1901 },
    { 0, 1321, 2482, 3344, 3803 },    { 0, 2642, 4964, 6689, 7606 },
    { 0, 5283, 9929, 13377, 15212 },  { 0, 10566, 19858, 26755, 30424 },
    { 0, 21133, 39716, 53510, 60849 } };
static int g_trig_ready;

static void trig_init(void) {
    if (g_trig_ready) return;
    for (int b = 0; b < 7; ++b)
        for (int j = 0; j < 64; ++j)
            g_cospi[b][j] = (int32_t)floor(cos(M_PI * j / 128.0) * (double)(1 << (10 + b)) + 0.5);
    g_trig_ready = 1;
}
const int32_t *svto_cospi(int bit) { trig_init(); return g_cospi[bit - 10]; }
const int32_t *svto_sinpi(int bit) { trig_init(); return g_sinpi[bit - 10]; }

/* round_shift / half_btf, EbTransforms.c:1288-1299: 32-bit products, 64-bit sum */
static inline int32_t rshift64(int64_t v, int bit) { return (int32_t)((v + (1LL << (bit - 1))) >> bit); }
static inline int32_t hbtf(int32_t w0, int32_t a, int32_t w1, int32_t b, int bit) {
    const int64_t r = (int64_t)(int32_t)(w0 * a) + (int64_t)(int32_t)(w1 * b);
    return rshift64(r, bit);
}
/* clamp_value, EbTransforms.c:5458 (bits <= 0: no clamp -- the forward direction) */
static inline int32_t clampv(int32_t v, int bits) {
    if (bits <= 0) return v;
    const int64_t hi = (1LL << (bits - 1)) - 1, lo = -(1LL << (bits - 1));
    return (int32_t)(v > hi ? hi : v < lo ? lo : v);
}
static inline int brev(int v, int bits) {
    int r = 0;
    for (int i = 0; i < bits; ++i) r |= ((v >> i) & 1) << (bits - 1 - i);
    return r;
}
static inline int ilog2(int n) { int l = 0; while ((1 << l) < n) ++l; return l; }

/* ------------------------------------------------------------------------------------------
 * DCT.  Working array is in "position" order: the reference's bf[] index just before its last
 * (bit-reversal) stage; output[k] = pos[brev(k)].
 * ---------------------------------------------------------------------------------------- */

/* group butterflies of width W over the M-element odd half: even groups (a+b, a-b), odd
 * groups (b-a, b+a); cf. EbTransforms.c:1505-1512 (forward), :5634-5641 (inverse, clamped) */
static void odd_bfly(int32_t *y, int M, int W, int cb) {
    for (int g = 0; g < M / W; ++g)
        for (int k = 0; k < W / 2; ++k) {
            int32_t *pa = &y[g * W + k], *pb = &y[g * W + W - 1 - k];
            const int32_t a = *pa, b = *pb;
            if (g & 1) { *pa = clampv(b - a, cb); *pb = clampv(b + a, cb); }
            else       { *pa = clampv(a + b, cb); *pb = clampv(a - b, cb); }
        }
}

/* middle rotation stage that follows (forward) / precedes (inverse) the width-W butterflies:
 * cf. EbTransforms.c:1531-1536 (fdct16 stage 4), :1788-1797 & :1850-1861 (fdct32) */
static void odd_rot(int32_t *y, int M, int W, const int32_t *c, int bit, int inverse) {
    const int G = M / W;
    for (int g = 0; g < G / 2; ++g) {
        const int ca = (32 / G) * (1 + 4 * brev(g, ilog2(G / 2))), cb = 64 - ca;
        for (int j = W / 4; j < 3 * W / 4; ++j) {
            const int p = g * W + j, q = M - 1 - p;
            const int32_t P = y[p], Q = y[q];
            if (j < W / 2) { /* "type 1" */
                y[p] = hbtf(-c[ca], P, c[cb], Q, bit);
                y[q] = inverse ? hbtf(c[cb], P, c[ca], Q, bit) : hbtf(c[ca], Q, c[cb], P, bit);
            } else {         /* "type 2" */
                y[p] = hbtf(-c[cb], P, -c[ca], Q, bit);
                y[q] = inverse ? hbtf(-c[ca], P, c[cb], Q, bit) : hbtf(c[cb], Q, -c[ca], P, bit);
            }
        }
    }
}

static void fdct_odd(int32_t *y, int N, const int32_t *c, int bit) {
    const int M = N / 2;
    if (M >= 4) /* pi/4 rotation of the middle half, EbTransforms.c:1485-1488 */
        for (int p = M / 4; p < M / 2; ++p) {
            const int q = M - 1 - p;
            const int32_t lo = y[p], hi = y[q];
            y[p] = hbtf(-c[32], lo, c[32], hi, bit);
            y[q] = hbtf(c[32], hi, c[32], lo, bit);
        }
    for (int W = M / 2; W >= 2; W >>= 1) {
        odd_bfly(y, M, W, 0);
        if (W >= 4) odd_rot(y, M, W, c, bit, 0);
    }
    for (int i = 0; i < M / 2; ++i) { /* last rotations, EbTransforms.c:1577-1584 */
        const int b = (64 / N) * (1 + 4 * brev(i, ilog2(M / 2))), a = 64 - b;
        const int32_t lo = y[i], hi = y[M - 1 - i];
        y[i] = hbtf(c[a], lo, c[b], hi, bit);
        y[M - 1 - i] = hbtf(c[a], hi, -c[b], lo, bit);
    }
}

static void fdct_pos(int32_t *x, int n, const int32_t *c, int bit) {
    if (n == 2) { /* EbTransforms.c:1340-1341 */
        const int32_t a = x[0], b = x[1];
        x[0] = hbtf(c[32], a, c[32], b, bit);
        x[1] = hbtf(-c[32], b, c[32], a, bit);
        return;
    }
    for (int i = 0; i < n / 2; ++i) { /* EbTransforms.c:1329-1332 */
        const int32_t a = x[i], b = x[n - 1 - i];
        x[i] = a + b;
        x[n - 1 - i] = a - b;
    }
    fdct_pos(x, n / 2, c, bit);
    fdct_odd(x + n / 2, n, c, bit);
}

static void idct_odd(int32_t *y, int N, const int32_t *c, int bit, int cbits) {
    const int M = N / 2;
    for (int i = 0; i < M / 2; ++i) { /* first rotations, EbTransforms.c:5617-5624 */
        const int b = (64 / N) * (1 + 4 * brev(i, ilog2(M / 2))), a = 64 - b;
        const int32_t lo = y[i], hi = y[M - 1 - i];
        y[i] = hbtf(c[a], lo, -c[b], hi, bit);
        y[M - 1 - i] = hbtf(c[b], lo, c[a], hi, bit);
    }
    for (int W = 2; W <= M / 2; W <<= 1) {
        if (W >= 4) odd_rot(y, M, W, c, bit, 1);
        odd_bfly(y, M, W, cbits);
    }
    if (M >= 4) /* EbTransforms.c:5706-5709 */
        for (int p = M / 4; p < M / 2; ++p) {
            const int q = M - 1 - p;
            const int32_t P = y[p], Q = y[q];
            y[p] = hbtf(-c[32], P, c[32], Q, bit);
            y[q] = hbtf(c[32], P, c[32], Q, bit);
        }
}

static void idct_pos(int32_t *x, int n, const int32_t *c, int bit, int cbits) {
    if (n == 2) { /* EbTransforms.c:5489-5490 */
        const int32_t a = x[0], b = x[1];
        x[0] = hbtf(c[32], a, c[32], b, bit);
        x[1] = hbtf(c[32], a, -c[32], b, bit);
        return;
    }
    idct_pos(x, n / 2, c, bit, cbits);
    idct_odd(x + n / 2, n, c, bit, cbits);
    for (int i = 0; i < n / 2; ++i) { /* EbTransforms.c:5498-5501 */
        const int32_t a = x[i], b = x[n - 1 - i];
        x[i] = clampv(a + b, cbits);
        x[n - 1 - i] = clampv(a - b, cbits);
    }
}

static void fdct(const int32_t *in, int32_t *out, int n, int bit) {
    int32_t x[64];
    memcpy(x, in, n * sizeof(int32_t));
    fdct_pos(x, n, svto_cospi(bit), bit);
    for (int k = 0; k < n; ++k) out[k] = x[brev(k, ilog2(n))];
}
static void idct(const int32_t *in, int32_t *out, int n, int bit, int cbits) {
    int32_t x[64];
    for (int p = 0; p < n; ++p) x[p] = in[brev(p, ilog2(n))];
    idct_pos(x, n, svto_cospi(bit), bit, cbits);
    memcpy(out, x, n * sizeof(int32_t));
}

/* ------------------------------------------------------------------------------------------
 * ADST 8 / 16 (EbTransforms.c:2856-3183 forward, :6189-6502 inverse): an input permutation
 * with signs, then rotation stages over groups of 4, 8, (16) alternating with stride-2, 4,
 * (8) butterflies, a last rotation of adjacent pairs and an output permutation.  The inverse
 * runs the same rotations in the opposite order with clamped butterflies.
 * ---------------------------------------------------------------------------------------- */
static const int8_t k_adst8_in[8] = { 0, -7, -3, 4, -1, 6, 2, -5 };                /* :2872-2879 */
static const int8_t k_adst16_in[16] = { 0, -15, -7, 8, -3, 12, 4, -11,
                                        -1, 14, 6, -9, 2, -13, -5, 10 };           /* :2986-3001 */

static void adst_rot_group(int32_t *x, int n, int Gs, const int32_t *c, int bit) {
    const int npairs = Gs / 4, na = npairs > 1 ? npairs / 2 : 1;
    for (int g = 0; g < n; g += Gs)
        for (int k = 0; k < npairs; ++k) {
            int32_t *pa = &x[g + Gs / 2 + 2 * k], *pb = pa + 1;
            const int32_t a = *pa, b = *pb;
            const int kk = k % na;
            const int u = (128 / Gs) * (1 + 4 * brev(kk, ilog2(na)));
            if (k < na) { /* cf. :2918-2919 */
                *pa = hbtf(c[u], a, c[64 - u], b, bit);
                *pb = hbtf(c[64 - u], a, -c[u], b, bit);
            } else {      /* cf. :2920-2921 */
                *pa = hbtf(-c[64 - u], a, c[u], b, bit);
                *pb = hbtf(c[u], a, c[64 - u], b, bit);
            }
        }
}
static void adst_bfly(int32_t *x, int n, int S, int cb) {
    for (int g = 0; g < n; g += 2 * S)
        for (int i = 0; i < S; ++i) {
            const int32_t a = x[g + i], b = x[g + i + S];
            x[g + i] = clampv(a + b, cb);
            x[g + i + S] = clampv(a - b, cb);
        }
}
static void adst_last_rot(int32_t *x, int n, const int32_t *c, int bit) {
    for (int i = 0; i < n / 2; ++i) { /* :2946-2953, :6213-6220 */
        const int u = 32 / n + (128 / n) * i;
        const int32_t a = x[2 * i], b = x[2 * i + 1];
        x[2 * i] = hbtf(c[u], a, c[64 - u], b, bit);
        x[2 * i + 1] = hbtf(c[64 - u], a, -c[u], b, bit);
    }
}
static void fadst(const int32_t *in, int32_t *out, int n, int bit) {
    const int32_t *c = svto_cospi(bit);
    const int8_t *perm = n == 8 ? k_adst8_in : k_adst16_in;
    int32_t x[16];
    for (int i = 0; i < n; ++i) x[i] = perm[i] < 0 ? -in[-perm[i]] : in[perm[i]];
    for (int Gs = 4; Gs <= n; Gs <<= 1) {
        adst_rot_group(x, n, Gs, c, bit);
        adst_bfly(x, n, Gs / 2, 0);
    }
    adst_last_rot(x, n, c, bit);
    for (int j = 0; j < n / 2; ++j) { /* :2958-2965 */
        out[2 * j] = x[2 * j + 1];
        out[2 * j + 1] = x[n - 2 - 2 * j];
    }
}
static void iadst(const int32_t *in, int32_t *out, int n, int bit, int cbits) {
    const int32_t *c = svto_cospi(bit);
    const int8_t *perm = n == 8 ? k_adst8_in : k_adst16_in;
    int32_t x[16];
    for (int i = 0; i < n / 2; ++i) { /* :6203-6210 */
        x[2 * i] = in[n - 1 - 2 * i];
        x[2 * i + 1] = in[2 * i];
    }
    adst_last_rot(x, n, c, bit);
    for (int Gs = n; Gs >= 4; Gs >>= 1) {
        adst_bfly(x, n, Gs / 2, cbits);
        adst_rot_group(x, n, Gs, c, bit);
    }
    for (int i = 0; i < n; ++i) { /* :6285-6292 */
        if (perm[i] < 0) out[-perm[i]] = -x[i];
        else out[perm[i]] = x[i];
    }
}

/* ADST4: EbTransforms.c:2764-2854 (forward), :6097-6185 (inverse) */
static void fadst4(const int32_t *in, int32_t *out, int bit) {
    const int32_t *s = svto_sinpi(bit);
    const int32_t x0 = in[0], x1 = in[1], x2 = in[2], x3 = in[3];
    if (!(x0 | x1 | x2 | x3)) { out[0] = out[1] = out[2] = out[3] = 0; return; }
    const int32_t a0 = s[1] * x0 + s[2] * x1 + s[4] * x3;
    const int32_t a1 = s[3] * (x0 + x1 - x3);
    const int32_t a2 = s[4] * x0 - s[1] * x1 + s[2] * x3;
    const int32_t a3 = s[3] * x2;
    out[0] = rshift64(a0 + a3, bit);
    out[1] = rshift64(a1, bit);
    out[2] = rshift64(a2 - a3, bit);
    out[3] = rshift64(a2 - a0 + a3, bit);
}
static void iadst4(const int32_t *in, int32_t *out, int bit) {
    const int32_t *s = svto_sinpi(bit);
    const int32_t x0 = in[0], x1 = in[1], x2 = in[2], x3 = in[3];
    if (!(x0 | x1 | x2 | x3)) { out[0] = out[1] = out[2] = out[3] = 0; return; }
    const int32_t a0 = s[1] * x0 + s[4] * x2 + s[2] * x3;
    const int32_t a1 = s[2] * x0 - s[1] * x2 - s[4] * x3;
    const int32_t a3 = s[3] * x1;
    const int32_t a2 = s[3] * ((x0 - x2) + x3);
    out[0] = rshift64(a0 + a3, bit);
    out[1] = rshift64(a1 + a3, bit);
    out[2] = rshift64(a2, bit);
    out[3] = rshift64(a0 + a1 - a3, bit);
}

/* identity: EbTransforms.c:3620-3659 (forward), :7717-7755 (inverse): x sqrt2, x2, x 2sqrt2,
 * x4, x 4sqrt2 with 64-bit products and one round_shift(.,12) for the sqrt2 ones */
static void identity(const int32_t *in, int32_t *out, int n) {
    for (int i = 0; i < n; ++i) {
        switch (n) {
        case 4: out[i] = rshift64((int64_t)in[i] * 5793, 12); break;
        case 8: out[i] = (int32_t)((int64_t)in[i] * 2); break;
        case 16: out[i] = rshift64((int64_t)in[i] * 2 * 5793, 12); break;
        case 32: out[i] = (int32_t)((int64_t)in[i] * 4); break;
        default: out[i] = rshift64((int64_t)in[i] * 4 * 5793, 12); break;
        }
    }
}

int svto_txfm1d(int kind, int n, int inverse, const int32_t *in, int32_t *out, int cos_bit,
                int range_bits) {
    int32_t tmp[64];
    if (n != 4 && n != 8 && n != 16 && n != 32 && n != 64) return -1;
    memcpy(tmp, in, n * sizeof(int32_t));
    if (kind == 0) {
        if (inverse) idct(tmp, out, n, cos_bit, range_bits); else fdct(tmp, out, n, cos_bit);
    } else if (kind == 1) {
        if (n > 16) return -1;
        if (n == 4) { if (inverse) iadst4(tmp, out, cos_bit); else fadst4(tmp, out, cos_bit); }
        else if (inverse) iadst(tmp, out, n, cos_bit, range_bits);
        else fadst(tmp, out, n, cos_bit);
    } else if (kind == 2) {
        identity(tmp, out, n);
    } else return -1;
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * per-(tx_size, tx_type) configuration: Av1TransformConfig (EbTransforms.c:4329),
 * av1_get_inv_txfm_cfg (:8154), get_flip_cfg (:4263)
 * ---------------------------------------------------------------------------------------- */
static const uint8_t k_wlog2[19] = { 2, 3, 4, 5, 6, 2, 3, 3, 4, 4, 5, 5, 6, 2, 4, 3, 5, 4, 6 };
static const uint8_t k_hlog2[19] = { 2, 3, 4, 5, 6, 3, 2, 4, 3, 5, 4, 6, 5, 4, 2, 5, 3, 6, 4 };
/* fwd_shift_* (EbTransforms.h:121-139), inv_shift_* (:268-286) */
static const int8_t k_fwd_shift[19][3] = {
    { 2, 0, 0 },  { 2, -1, 0 }, { 2, -2, 0 }, { 2, -4, 0 }, { 0, -2, -2 }, { 2, -1, 0 },
    { 2, -1, 0 }, { 2, -2, 0 }, { 2, -2, 0 }, { 2, -4, 0 }, { 2, -4, 0 },  { 0, -2, -2 },
    { 2, -4, -2 }, { 2, -1, 0 }, { 2, -1, 0 }, { 2, -2, 0 }, { 2, -2, 0 }, { 0, -2, 0 },
    { 2, -4, 0 } };
static const int8_t k_inv_shift0[19] = { 0, -1, -2, -2, -2, 0, 0, -1, -1, -1, -1, -1, -1,
                                         -1, -1, -2, -2, -2, -2 };
/* fwd_cos_bit_col / _row [txw_idx][txh_idx] (EbTransforms.h:141-156) */
static const int8_t k_fwd_cos_col[5][5] = { { 13, 13, 13, 0, 0 }, { 13, 13, 13, 12, 0 },
    { 13, 13, 13, 12, 13 }, { 0, 13, 13, 12, 13 }, { 0, 0, 13, 12, 13 } };
static const int8_t k_fwd_cos_row[5][5] = { { 13, 13, 12, 0, 0 }, { 13, 13, 13, 12, 0 },
    { 13, 13, 12, 13, 12 }, { 0, 12, 13, 12, 11 }, { 0, 0, 12, 11, 10 } };
/* vtx_tab / htx_tab (EbTransforms.h:87-98): 0 DCT, 1 ADST, 2 FLIPADST, 3 IDTX */
static const uint8_t k_vtx[16] = { 0, 1, 0, 1, 2, 0, 2, 1, 2, 3, 0, 3, 1, 3, 2, 3 };
static const uint8_t k_htx[16] = { 0, 0, 1, 1, 0, 2, 2, 2, 1, 3, 3, 0, 3, 1, 3, 2 };

int svto_tx_wide(int s) { return 1 << k_wlog2[s]; }
int svto_tx_high(int s) { return 1 << k_hlog2[s]; }
int svto_max_eob(int s) {
    const int w = svto_tx_wide(s), h = svto_tx_high(s);
    return (w > 32 ? 32 : w) * (h > 32 ? 32 : h);
}
int svto_tx_scale(int s) {
    const int pels = svto_tx_wide(s) * svto_tx_high(s);
    return (pels > 256) + (pels > 1024);
}
int svto_is_valid(int s, int t) {
    const int m = k_wlog2[s] > k_hlog2[s] ? k_wlog2[s] : k_hlog2[s];
    if (s < 0 || s > 18 || t < 0 || t > 15) return 0;
    if (m == 6) return t == 0;
    if (m == 5) return t == 0 || t == 9;
    return 1;
}
int svto_scan_class(int s, int t) { (void)s; return t < 10 ? 0 : (t & 1) ? 2 : 1; }

typedef struct Cfg {
    int w, h, ud, lr, kcol, krow; /* 1-D kind: 0 DCT 1 ADST 2 identity */
    int fsh[3], fcos_col, fcos_row, ish0, rect;
} Cfg;
static void get_cfg(int s, int t, Cfg *c) {
    c->w = svto_tx_wide(s); c->h = svto_tx_high(s);
    c->ud = k_vtx[t] == 2; c->lr = k_htx[t] == 2;
    c->kcol = k_vtx[t] == 3 ? 2 : k_vtx[t] != 0;
    c->krow = k_htx[t] == 3 ? 2 : k_htx[t] != 0;
    for (int i = 0; i < 3; ++i) c->fsh[i] = k_fwd_shift[s][i];
    c->fcos_col = k_fwd_cos_col[k_wlog2[s] - 2][k_hlog2[s] - 2];
    c->fcos_row = k_fwd_cos_row[k_wlog2[s] - 2][k_hlog2[s] - 2];
    c->ish0 = k_inv_shift0[s];
    const int d = (int)k_wlog2[s] - (int)k_hlog2[s];
    c->rect = d == 1 || d == -1;
}
void svto_get_cfg(int s, int t, int32_t *o) {
    Cfg c; get_cfg(s, t, &c);
    o[0] = c.w; o[1] = c.h; o[2] = c.fsh[0]; o[3] = c.fsh[1]; o[4] = c.fsh[2];
    o[5] = c.fcos_col; o[6] = c.fcos_row; o[7] = c.ish0; o[8] = -4; o[9] = c.ud; o[10] = c.lr;
    o[11] = c.kcol; o[12] = c.krow; o[13] = svto_tx_scale(s); o[14] = svto_max_eob(s); o[15] = 0;
}

/* av1_round_shift_array_c, EbTransforms.c:3682: bit > 0 rounds down, bit < 0 scales up */
static void shift_array(int32_t *a, int n, int bit) {
    if (bit > 0) for (int i = 0; i < n; ++i) a[i] = rshift64(a[i], bit);
    else if (bit < 0) for (int i = 0; i < n; ++i) a[i] = a[i] * (1 << (-bit));
}

/* Av1TranformTwoDCore_c, EbTransforms.c:3701-3779 */
int svto_fwd_txfm2d(const int16_t *input, uint32_t stride, int32_t *output, int s, int t, int bd) {
    (void)bd; /* only feeds the (unused) stage-range bookkeeping in the reference */
    if (!svto_is_valid(s, t)) return -1;
    Cfg c; get_cfg(s, t, &c);
    const int w = c.w, h = c.h;
    int32_t *buf = (int32_t *)malloc(sizeof(int32_t) * w * h);
    int32_t col[64], tmp[64];
    for (int x = 0; x < w; ++x) {
        for (int r = 0; r < h; ++r) col[r] = input[(c.ud ? h - 1 - r : r) * stride + x];
        shift_array(col, h, -c.fsh[0]);
        svto_txfm1d(c.kcol, h, 0, col, tmp, c.fcos_col, 0);
        shift_array(tmp, h, -c.fsh[1]);
        for (int r = 0; r < h; ++r) buf[r * w + (c.lr ? w - 1 - x : x)] = tmp[r];
    }
    for (int r = 0; r < h; ++r) {
        svto_txfm1d(c.krow, w, 0, buf + r * w, output + r * w, c.fcos_row, 0);
        shift_array(output + r * w, w, -c.fsh[2]);
        if (c.rect) /* x sqrt2, EbTransforms.c:3769-3776 */
            for (int x = 0; x < w; ++x)
                output[r * w + x] = rshift64((int64_t)output[r * w + x] * 5793, 12);
    }
    free(buf);
    return 0;
}

/* av1_estimate_transform, EbTransforms.c:4918-5291; HandleTransform*_c + EnergyComputation
 * :4351-4731; re-pack :4962-4965, :5007-5010, :5169-5173 */
int svto_estimate_transform(const int16_t *residual, uint32_t stride, int32_t *coeff, int s, int t,
                            int bd, uint64_t *energy) {
    if (svto_fwd_txfm2d(residual, stride, coeff, s, t, bd)) return -1;
    const int w = svto_tx_wide(s), h = svto_tx_high(s);
    const int kw = w > 32 ? 32 : w, kh = h > 32 ? 32 : h;
    *energy = 0;
    if (kw == w && kh == h) return 0;
    uint64_t e = 0;
    for (int r = 0; r < h; ++r)
        for (int x = 0; x < w; ++x)
            if (r >= kh || x >= kw) {
                const int64_t v = coeff[r * w + x];
                e += (uint64_t)(v * v);
                coeff[r * w + x] = 0;
            }
    *energy = e;
    if (kw != w)
        for (int r = 1; r < kh; ++r) memmove(coeff + r * kw, coeff + r * w, kw * sizeof(int32_t));
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * quantizer: highbd_quantize_b_helper_c (EbFullLoop.c:239-296); with clamp16 the 8-bit
 * quantize_b_helper_c_II (EbFullLoop.c:44-106), whose trailing-zero pre-scan is equivalent to
 * testing every position.
 * ---------------------------------------------------------------------------------------- */
void svto_quantize(const int32_t *coeff, int n, const int16_t *zbin, const int16_t *round,
                   const int16_t *quant, const int16_t *quant_shift, const int16_t *dequant,
                   int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob_out, const int16_t *scan,
                   int log_scale, const uint8_t *qm, const uint8_t *iqm, int clamp16) {
    const int32_t rnd_ls = log_scale ? (1 << (log_scale - 1)) : 0;
    const int32_t zbins[2] = { (zbin[0] + rnd_ls) >> log_scale, (zbin[1] + rnd_ls) >> log_scale };
    const int32_t rounds[2] = { (round[0] + rnd_ls) >> log_scale, (round[1] + rnd_ls) >> log_scale };
    int eob = -1;
    memset(qcoeff, 0, n * sizeof(int32_t));
    memset(dqcoeff, 0, n * sizeof(int32_t));
    for (int i = 0; i < n; ++i) {
        const int rc = scan[i], ac = rc != 0;
        const int32_t wt = qm ? qm[rc] : 32, iwt = iqm ? iqm[rc] : 32;
        const int32_t cf = coeff[rc];
        const int32_t cw = cf * wt;
        if (!(cw >= zbins[ac] * 32 || cw <= -zbins[ac] * 32)) continue;
        const int32_t sign = cf >> 31;
        const int32_t ab = (cf ^ sign) - sign;
        int64_t t1 = (int64_t)ab + rounds[ac];
        if (clamp16) t1 = t1 > 32767 ? 32767 : t1 < -32768 ? -32768 : t1; /* EbFullLoop.c:86-88 */
        const int64_t tw = t1 * wt;
        const int64_t t2 = ((tw * quant[ac]) >> 16) + tw;
        const int32_t aq = (int32_t)((t2 * quant_shift[ac]) >> (16 - log_scale + 5));
        qcoeff[rc] = (aq ^ sign) - sign;
        const int32_t dq = (dequant[ac] * iwt + 16) >> 5;
        const int32_t adq = (aq * dq) >> log_scale;
        dqcoeff[rc] = (adq ^ sign) - sign;
        if (aq) eob = i;
    }
    *eob_out = (uint16_t)(eob + 1);
}

/* av1_build_quantizer, EbModeDecisionConfigurationProcess.c:429-521 (all deltas 0);
 * invert_quant :322, get_qzbin_factor :300 */
void svto_build_quant_tables(const int16_t *dc_q, const int16_t *ac_q, int bd, int16_t *out) {
    for (int q = 0; q < 256; ++q) {
        const int thr = bd == 8 ? 148 : bd == 10 ? 592 : 2368;
        const int zf = q == 0 ? 64 : (dc_q[q] < thr ? 84 : 80);
        const int rf = q == 0 ? 64 : 48;
        int16_t *o = out + q * 40;
        for (int i = 0; i < 8; ++i) {
            const int d = i == 0 ? dc_q[q] : ac_q[q];
            int l = 0;
            for (uint32_t tt = (uint32_t)d; tt > 1; tt >>= 1) ++l;
            const int m = 1 + (1 << (16 + l)) / d;
            o[i] = (int16_t)((zf * d + 64) >> 7);
            o[8 + i] = (int16_t)((rf * d) >> 7);
            o[16 + i] = (int16_t)(m - (1 << 16));
            o[24 + i] = (int16_t)(1 << (16 - l));
            o[32 + i] = (int16_t)d;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * inverse: inv_txfm2d_add_c (EbTransforms.c:8180-8266); 64-point re-expansion :8299-8426;
 * stage ranges av1_gen_inv_stage_range (:5404): rows 16 (bd 8) / 18 (bd 10), columns 16;
 * pixel add highbd_clip_pixel_add + check_range (:7794-7813)
 * ---------------------------------------------------------------------------------------- */
int svto_inv_txfm2d_add(const int32_t *dq, uint16_t *dst, int32_t stride, int s, int t, int bd) {
    if (!svto_is_valid(s, t)) return -1;
    Cfg c; get_cfg(s, t, &c);
    const int w = c.w, h = c.h;
    const int kw = w > 32 ? 32 : w, kh = h > 32 ? 32 : h;
    const int row_bits = bd == 8 ? 16 : bd == 10 ? 18 : 20;
    const int col_bits = bd == 12 ? 18 : 16;
    const int in_bits = bd + 8, mid_bits = bd + 6 > 16 ? bd + 6 : 16;
    int32_t *buf = (int32_t *)malloc(sizeof(int32_t) * w * h);
    int32_t a[64], b[64];
    for (int r = 0; r < h; ++r) {
        for (int x = 0; x < w; ++x) {
            int32_t v = (r < kh && x < kw) ? dq[r * kw + x] : 0;
            if (c.rect) v = rshift64((int64_t)v * 2896, 12);
            a[x] = clampv(v, in_bits);
        }
        svto_txfm1d(c.krow, w, 1, a, buf + r * w, 12, row_bits);
        shift_array(buf + r * w, w, -c.ish0);
    }
    const int32_t pmax = (1 << bd) - 1;
    const int64_t rmax = (1LL << (7 + bd)) - 1 + (914LL << (bd - 7)), rmin = -rmax - 1;
    for (int x = 0; x < w; ++x) {
        for (int r = 0; r < h; ++r) a[r] = clampv(buf[r * w + (c.lr ? w - 1 - x : x)], mid_bits);
        svto_txfm1d(c.kcol, h, 1, a, b, 12, col_bits);
        shift_array(b, h, 4);
        for (int r = 0; r < h; ++r) {
            int64_t res = b[c.ud ? h - 1 - r : r];
            res = res > rmax ? rmax : res < rmin ? rmin : res;
            int32_t p = (int32_t)dst[r * stride + x] + (int32_t)res;
            dst[r * stride + x] = (uint16_t)(p < 0 ? 0 : p > pmax ? pmax : p);
        }
    }
    free(buf);
    return 0;
}

/* av1_inv_txfm_add_c, EbTransforms.c:8882: widen, run the 16-bit core with bd = 8, narrow */
int svto_inv_txfm2d_add8(const int32_t *dq, uint8_t *dst, int32_t stride, int s, int t) {
    uint16_t tmp[64 * 64];
    const int w = svto_tx_wide(s), h = svto_tx_high(s);
    for (int r = 0; r < h; ++r) for (int x = 0; x < w; ++x) tmp[r * 64 + x] = dst[r * stride + x];
    if (svto_inv_txfm2d_add(dq, tmp, 64, s, t, 8)) return -1;
    for (int r = 0; r < h; ++r) for (int x = 0; x < w; ++x) dst[r * stride + x] = (uint8_t)tmp[r * 64 + x];
    return 0;
}

/* the whole path for a slice of a block list (bench.py cpu_baseline "port") */
void svto_fused_blocks(const SvtoBlock *blk, int first, int last, const int16_t *residual,
                       void *recon, const int16_t *tables, int32_t *qcoeff, int32_t *dqcoeff,
                       uint16_t *eob, uint64_t *energy, const int16_t *const *scans, int clamp16,
                       const uint8_t *wt, const uint8_t *iwt, const int32_t *qm_off) {
    int32_t *coeff = (int32_t *)malloc(sizeof(int32_t) * 4096);
    for (int i = first; i < last; ++i) {
        const SvtoBlock *b = &blk[i];
        const int n = svto_max_eob(b->tx_size);
        const int16_t *t = tables + (size_t)b->qrow * 40;
        const int16_t *scan = scans[b->tx_size * 3 + svto_scan_class(b->tx_size, b->tx_type)];
        svto_estimate_transform(residual + b->res_off, b->res_stride, coeff, b->tx_size,
                                b->tx_type, b->bd, &energy[i]);
        const int use_qm = (b->flags & 2) && wt != NULL;
        const int qo = use_qm ? qm_off[(b->plane > 0) * 19 + b->tx_size] : 0;
        svto_quantize(coeff, n, t, t + 8, t + 16, t + 24, t + 32, qcoeff + b->coeff_off,
                      dqcoeff + b->coeff_off, &eob[i], scan, svto_tx_scale(b->tx_size),
                      use_qm ? wt + qo : NULL, use_qm ? iwt + qo : NULL,
                      (clamp16 || (b->flags & 1)) && b->bd == 8);
        if (eob[i]) {
            if (b->bd == 8)
                svto_inv_txfm2d_add8(dqcoeff + b->coeff_off, (uint8_t *)recon + b->rec_off,
                                     b->rec_stride, b->tx_size, b->tx_type);
            else
                svto_inv_txfm2d_add(dqcoeff + b->coeff_off, (uint16_t *)recon + b->rec_off,
                                    b->rec_stride, b->tx_size, b->tx_type, b->bd);
        }
    }
    free(coeff);
}

/* ---- residual from source - prediction, transform-domain distortion ---------- */
/* residual_kernel_c / residual_kernel16bit (Codec/EbPictureOperators.c:166-191, :134-160):
 * residual = (int16)input - (int16)pred, element by element. */
void svto_residual(const void *input, uint32_t input_stride, const void *pred, uint32_t pred_stride,
                   int16_t *residual, uint32_t residual_stride, int w, int h, int bd) {
    for (int r = 0; r < h; ++r)
        for (int c = 0; c < w; ++c) {
            const int a = bd == 8 ? ((const uint8_t *)input)[(size_t)r * input_stride + c]
                                  : ((const uint16_t *)input)[(size_t)r * input_stride + c];
            const int b = bd == 8 ? ((const uint8_t *)pred)[(size_t)r * pred_stride + c]
                                  : ((const uint16_t *)pred)[(size_t)r * pred_stride + c];
            residual[(size_t)r * residual_stride + c] = (int16_t)((int16_t)a - (int16_t)b);
        }
}

/* The per-TU distortion of the full loop (Codec/EbFullLoop.c:808-835):
 * picture_full_distortion32_bits (Codec/EbPictureOperators.c:349-402) over the coded
 * min(w,32) x min(h,32) region -- full_distortion_kernel32_bits (:283-313) when the block has
 * non-zero coefficients, full_distortion_kernel_cbf_zero32_bits (:318-347) otherwise -- then
 * += three_quad_energy on both entries and RIGHT_SIGNED_SHIFT by (MAX_TX_SCALE - tx_scale) * 2
 * with MAX_TX_SCALE = 1 (Codec/EbTransforms.h:40). dist[0] = DIST_CALC_RESIDUAL,
 * dist[1] = DIST_CALC_PREDICTION. */
void svto_tu_distortion(const int32_t *coeff, const int32_t *recon_coeff, int tx_size, int eob,
                        uint64_t three_quad_energy, uint64_t dist[2]) {
    const int n = svto_max_eob(tx_size);
    uint64_t res = 0, pred = 0;
    for (int i = 0; i < n; ++i) {
        const int64_t c = coeff[i];
        const int64_t d = c - (int64_t)recon_coeff[i];
        res += (uint64_t)(d * d);
        pred += (uint64_t)(c * c);
    }
    if (!eob) res = pred;
    res += three_quad_energy;
    pred += three_quad_energy;
    const int shift = (1 - svto_tx_scale(tx_size)) * 2;
    dist[0] = shift < 0 ? res << -shift : res >> shift;
    dist[1] = shift < 0 ? pred << -shift : pred >> shift;
}

/* svto_fused_blocks with the residual taken from source - prediction (src has the geometry of
 * the prediction / reconstruction planes) and the per-block distortion pair. `recon` holds the
 * prediction on entry, like svto_fused_blocks. */
static void fused_blocks_md(const SvtoBlock *blk, int first, int last, const void *src, void *recon,
                            const int16_t *tables, int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob,
                            uint64_t *energy, uint64_t *dist, const int16_t *const *scans,
                            int clamp16, int do_inverse) {
    int32_t *coeff = (int32_t *)malloc(sizeof(int32_t) * 4096);
    int16_t *res = (int16_t *)malloc(sizeof(int16_t) * 4096);
    for (int i = first; i < last; ++i) {
        const SvtoBlock *b = &blk[i];
        const int w = svto_tx_wide(b->tx_size), h = svto_tx_high(b->tx_size);
        const int n = svto_max_eob(b->tx_size);
        const int16_t *t = tables + (size_t)b->qrow * 40;
        const int16_t *scan = scans[b->tx_size * 3 + svto_scan_class(b->tx_size, b->tx_type)];
        const size_t px = b->bd == 8 ? 1 : 2;
        svto_residual((const char *)src + (size_t)b->rec_off * px, b->rec_stride,
                      (const char *)recon + (size_t)b->rec_off * px, b->rec_stride, res, w, w, h,
                      b->bd);
        svto_estimate_transform(res, w, coeff, b->tx_size, b->tx_type, b->bd, &energy[i]);
        svto_quantize(coeff, n, t, t + 8, t + 16, t + 24, t + 32, qcoeff + b->coeff_off,
                      dqcoeff + b->coeff_off, &eob[i], scan, svto_tx_scale(b->tx_size), NULL, NULL,
                      (clamp16 || (b->flags & 1)) && b->bd == 8);
        svto_tu_distortion(coeff, dqcoeff + b->coeff_off, b->tx_size, eob[i], energy[i],
                           dist + 2 * (size_t)i);
        if (eob[i] && do_inverse) {
            if (b->bd == 8)
                svto_inv_txfm2d_add8(dqcoeff + b->coeff_off, (uint8_t *)recon + b->rec_off,
                                     b->rec_stride, b->tx_size, b->tx_type);
            else
                svto_inv_txfm2d_add(dqcoeff + b->coeff_off, (uint16_t *)recon + b->rec_off,
                                    b->rec_stride, b->tx_size, b->tx_type, b->bd);
        }
    }
    free(coeff);
    free(res);
}
void svto_fused_blocks_md(const SvtoBlock *blk, int first, int last, const void *src, void *recon,
                          const int16_t *tables, int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob,
                          uint64_t *energy, uint64_t *dist, const int16_t *const *scans,
                          int clamp16) {
    fused_blocks_md(blk, first, last, src, recon, tables, qcoeff, dqcoeff, eob, energy, dist, scans,
                    clamp16, 1);
}
/* one iteration of the tx-type loop of ProductFullLoopTxSearch (EbFullLoop.c:984-1075)
 * per listed candidate -- transform, quantize, distortion; the prediction is only read, so any
 * number of candidates may name the same block */
void svto_tx_search_blocks(const SvtoBlock *blk, int first, int last, const void *src, const void *pred,
                           const int16_t *tables, int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob,
                           uint64_t *energy, uint64_t *dist, const int16_t *const *scans,
                           int clamp16) {
    fused_blocks_md(blk, first, last, src, (void *)pred, tables, qcoeff, dqcoeff, eob, energy, dist,
                    scans, clamp16, 0);
}

/* ---- av1_txb_init_levels_c (Source/Lib/Common/Codec/EbRateDistortionCost.c:125-148)
 * on a whole levels_buf: `buf` is what the reference declares as levels_buf[TX_PAD_2D]; the
 * reference's `levels` pointer is buf + TX_PAD_TOP * (width + TX_PAD_HOR) (set_levels, :121).
 * Writes (width + 4) * (height + 6) + 16 bytes. */
void svto_txb_init_levels(const int32_t *coeff, int width, int height, uint8_t *buf) {
    const int stride = width + 4;                         /* TX_PAD_HOR */
    uint8_t *ls = buf;
    for (int i = 0; i < 2 * stride; ++i) *ls++ = 0;       /* TX_PAD_TOP rows */
    for (int i = 0; i < height; ++i) {
        for (int j = 0; j < width; ++j) {
            int32_t a = coeff[i * width + j];
            a = a < 0 ? -a : a;
            *ls++ = (uint8_t)(a > 127 ? 127 : a);         /* clamp(abs, 0, INT8_MAX) */
        }
        for (int j = 0; j < 4; ++j) *ls++ = 0;
    }
    for (int i = 0; i < 4 * stride + 16; ++i) *ls++ = 0;  /* TX_PAD_BOTTOM rows + TX_PAD_END */
}
