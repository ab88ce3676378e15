/*
 * pib_oracle.c -- CPU restatement of the reference arithmetic on PipeInfer's speculation path.
 *
 * TEST INFRASTRUCTURE (see pib_oracle.h).  Plain C, scalar loops.  Where the reference takes an AVX2 branch on the
 * x86 hosts this project runs on (built with -mavx2 -mfma -mf16c, oracle/Makefile), the scalar code below reproduces
 * that branch's *association order* (8 lanes, lane-wise fused multiply-add, then the same horizontal reduction), so
 * that results are bit-identical to the compiled reference and not merely close.  Compiled with -ffp-contract=off;
 * every fused multiply-add the reference performs is written as an explicit fmaf().
 * All paths below are relative to /root/reference.
 */
#define _GNU_SOURCE
#include "pib_oracle.h"

#include <immintrin.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define QK 32
#define QK_K 256

/* ------------------------------------------------------------------------------------------------------------- */
/* fp16 <-> fp32: ggml-impl.h GGML_COMPUTE_FP16_TO_FP32 / FP32_TO_FP16 with F16C (_cvtsh_ss / _cvtss_sh, RNE)      */
float   po_fp16_to_fp32(po_fp16 h) { return _cvtsh_ss(h); }
po_fp16 po_fp32_to_fp16(float f)   { return _cvtss_sh(f, 0); }

static po_fp16 tab_exp[1 << 16], tab_silu[1 << 16], tab_gelu[1 << 16];
static int     tab_ready = 0;

/* ggml.c:1408-1411 ggml_gelu_f32, :1476-1478 ggml_silu_f32; tables filled as in ggml.c:2190-2200 */
static float gelu_f32(float x) {
    const float GELU_COEF_A = 0.044715f, SQRT_2_OVER_PI = 0.79788456080286535587989211986876f;
    return 0.5f * x * (1.0f + tanhf(SQRT_2_OVER_PI * x * (1.0f + GELU_COEF_A * x * x)));
}
static float silu_f32(float x) { return x / (1.0f + expf(-x)); }

void po_tables_init(void) {
    if (tab_ready) return;
    for (int i = 0; i < (1 << 16); ++i) {
        const float f = po_fp16_to_fp32((po_fp16) i);
        tab_gelu[i] = po_fp32_to_fp16(gelu_f32(f));
        tab_silu[i] = po_fp32_to_fp16(silu_f32(f));
        tab_exp[i]  = po_fp32_to_fp16(expf(f));
    }
    tab_ready = 1;
}
const po_fp16 * po_table_exp_f16(void)  { po_tables_init(); return tab_exp; }
const po_fp16 * po_table_silu_f16(void) { po_tables_init(); return tab_silu; }
const po_fp16 * po_table_gelu_f16(void) { po_tables_init(); return tab_gelu; }

/* ------------------------------------------------------------------------------------------------------------- */
/* block formats: ggml-quants.h:10-15 (q4_0), :42-47 (q8_0), :116-122 (q4_K), :161-166 (q8_K)                      */
#pragma pack(push, 1)
typedef struct { po_fp16 d; uint8_t qs[QK / 2]; } blk_q4_0;                                   /* 18 B */
typedef struct { po_fp16 d; int8_t qs[QK]; } blk_q8_0;                                        /* 34 B */
typedef struct { po_fp16 d; po_fp16 dmin; uint8_t scales[12]; uint8_t qs[QK_K / 2]; } blk_q4_K; /* 144 B */
typedef struct { float d; int8_t qs[QK_K]; int16_t bsums[QK_K / 16]; } blk_q8_K;              /* 292 B */
typedef struct { po_fp16 d; po_fp16 dmin; uint8_t scales[12]; uint8_t qh[QK_K / 8]; uint8_t qs[QK_K / 2]; } blk_q5_K;   /* 176 B, ggml-quants.h:138-145 */
typedef struct { uint8_t hmask[QK_K / 8]; uint8_t qs[QK_K / 4]; uint8_t scales[12]; po_fp16 d; } blk_q3_K;                       /* 110 B, ggml-quants.h:95-101 */
typedef struct { uint8_t ql[QK_K / 2]; uint8_t qh[QK_K / 4]; int8_t scales[QK_K / 16]; po_fp16 d; } blk_q6_K;   /* 210 B, ggml-quants.h:152-157 */
#pragma pack(pop)

/* hsum_float_8, ggml-quants.c:67-73: (hi128 + lo128) -> (x + movehl) -> (x + movehdup) */
static float hsum8(const float a[8]) {
    float r[4];
    for (int i = 0; i < 4; i++) r[i] = a[4 + i] + a[i];
    const float s0 = r[0] + r[2], s1 = r[1] + r[3];
    return s0 + s1;
}

/* ggml-quants.c:701-760 (AVX2): d = max/127 stored as fp16, id = 127/max, round-to-nearest-even, saturating packs */
void po_quantize_row_q8_0(const float * x, void * vy, int k) {
    blk_q8_0 * y  = (blk_q8_0 *) vy;
    const int  nb = k / QK;
    for (int i = 0; i < nb; i++) {
        float amax = 0.0f;
        for (int j = 0; j < QK; j++) { const float a = fabsf(x[i * QK + j]); if (a > amax) amax = a; }
        const float d  = amax / 127.f;
        y[i].d         = po_fp32_to_fp16(d);
        const float id = (amax != 0.0f) ? 127.f / amax : 0.0f;
        for (int j = 0; j < QK; j++) {
            const float v = nearbyintf(x[i * QK + j] * id);     /* _mm256_round_ps(NEAREST) then cvtps_epi32 */
            int q = (int) v;
            if (q > 127) q = 127;
            if (q < -128) q = -128;                               /* packs_epi32 / packs_epi16 saturate */
            y[i].qs[j] = (int8_t) q;
        }
    }
}

/* ggml-quants.c:1191-1196 */
static int nearest_int(float fval) {
    float val = fval + 12582912.f;
    int   i;
    memcpy(&i, &val, sizeof(int));
    return (i & 0x007fffff) - 0x00400000;
}

/* ggml-quants.c:2329-2363 */
void po_quantize_row_q8_K(const float * x, void * vy, int k) {
    blk_q8_K * y  = (blk_q8_K *) vy;
    const int  nb = k / QK_K;
    for (int i = 0; i < nb; i++) {
        float max = 0, amax = 0;
        for (int j = 0; j < QK_K; ++j) {
            float ax = fabsf(x[j]);
            if (ax > amax) { amax = ax; max = x[j]; }
        }
        if (!amax) {
            y[i].d = 0;
            memset(y[i].qs, 0, QK_K);
            x += QK_K;
            continue;       /* NB: bsums are left untouched by the reference in this case */
        }
        const float iscale = -128.f / max;
        for (int j = 0; j < QK_K; ++j) {
            int v      = nearest_int(iscale * x[j]);
            y[i].qs[j] = (int8_t) (v < 127 ? v : 127);
        }
        for (int j = 0; j < QK_K / 16; ++j) {
            int sum = 0;
            for (int ii = 0; ii < 16; ++ii) sum += y[i].qs[j * 16 + ii];
            y[i].bsums[j] = (int16_t) sum;
        }
        y[i].d = 1 / iscale;
        x += QK_K;
    }
}

/* ggml-quants.c:1074-1092 */
void po_dequantize_row_q4_0(const void * vx, float * y, int k) {
    const blk_q4_0 * x  = (const blk_q4_0 *) vx;
    const int        nb = k / QK;
    for (int i = 0; i < nb; i++) {
        const float d = po_fp16_to_fp32(x[i].d);
        for (int j = 0; j < QK / 2; ++j) {
            const int x0 = (x[i].qs[j] & 0x0F) - 8;
            const int x1 = (x[i].qs[j] >> 4) - 8;
            y[i * QK + j]          = x0 * d;
            y[i * QK + j + QK / 2] = x1 * d;
        }
    }
}

/* ggml-quants.c:1168-1182 */
void po_dequantize_row_q8_0(const void * vx, float * y, int k) {
    const blk_q8_0 * x  = (const blk_q8_0 *) vx;
    const int        nb = k / QK;
    for (int i = 0; i < nb; i++) {
        const float d = po_fp16_to_fp32(x[i].d);
        for (int j = 0; j < QK; ++j) y[i * QK + j] = x[i].qs[j] * d;
    }
}

/* ggml-quants.c:1446-1454 */
static void get_scale_min_k4(int j, const uint8_t * q, uint8_t * d, uint8_t * m) {
    if (j < 4) {
        *d = q[j] & 63;
        *m = q[j + 4] & 63;
    } else {
        *d = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        *m = (q[j + 4] >> 4) | ((q[j - 0] >> 6) << 4);
    }
}

/* ggml-quants.c:1917-1941 */
void po_dequantize_row_q4_K(const void * vx, float * y, int k) {
    const blk_q4_K * x  = (const blk_q4_K *) vx;
    const int        nb = k / QK_K;
    for (int i = 0; i < nb; i++) {
        const uint8_t * q   = x[i].qs;
        const float     d   = po_fp16_to_fp32(x[i].d);
        const float     min = po_fp16_to_fp32(x[i].dmin);
        int             is  = 0;
        uint8_t         sc, m;
        for (int j = 0; j < QK_K; j += 64) {
            get_scale_min_k4(is + 0, x[i].scales, &sc, &m);
            const float d1 = d * sc, m1 = min * m;
            get_scale_min_k4(is + 1, x[i].scales, &sc, &m);
            const float d2 = d * sc, m2 = min * m;
            for (int l = 0; l < 32; ++l) *y++ = d1 * (q[l] & 0xF) - m1;
            for (int l = 0; l < 32; ++l) *y++ = d2 * (q[l] >> 4) - m2;
            q += 32;
            is += 2;
        }
    }
}

/* ggml-quants.c:2500-2523 (AVX2).  mul_sum_i8_pairs_float (:136-146) yields 8 lanes, lane l = sum of the 4 products
 * of elements 4l..4l+3 (exact in int, converted to float); acc = fma(d, q, acc) lane-wise; hsum_float_8 at the end.
 * bytes_from_nibbles_32 (:108-114) puts the 16 low nibbles in bytes 0..15 and the 16 high nibbles in bytes 16..31. */
void po_vec_dot_q4_0_q8_0(int n, float * s, const void * vx, const void * vy) {
    const blk_q4_0 * x  = (const blk_q4_0 *) vx;
    const blk_q8_0 * y  = (const blk_q8_0 *) vy;
    const int        nb = n / QK;
    float acc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    for (int i = 0; i < nb; ++i) {
        const float d = po_fp16_to_fp32(x[i].d) * po_fp16_to_fp32(y[i].d);
        int8_t bx[QK];
        for (int j = 0; j < QK / 2; j++) {
            bx[j]          = (int8_t) ((x[i].qs[j] & 0x0F) - 8);
            bx[j + QK / 2] = (int8_t) ((x[i].qs[j] >> 4) - 8);
        }
        for (int l = 0; l < 8; l++) {
            int sum = 0;
            for (int t = 0; t < 4; t++) sum += (int) bx[4 * l + t] * (int) y[i].qs[4 * l + t];
            acc[l] = fmaf(d, (float) sum, acc[l]);
        }
    }
    *s = hsum8(acc);
}

/* ggml-quants.c:3587-3607 (AVX2) */
void po_vec_dot_q8_0_q8_0(int n, float * s, const void * vx, const void * vy) {
    const blk_q8_0 * x  = (const blk_q8_0 *) vx;
    const blk_q8_0 * y  = (const blk_q8_0 *) vy;
    const int        nb = n / QK;
    float acc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    for (int i = 0; i < nb; ++i) {
        const float d = po_fp16_to_fp32(x[i].d) * po_fp16_to_fp32(y[i].d);
        for (int l = 0; l < 8; l++) {
            int sum = 0;
            for (int t = 0; t < 4; t++) sum += (int) x[i].qs[4 * l + t] * (int) y[i].qs[4 * l + t];
            acc[l] = fmaf(d, (float) sum, acc[l]);
        }
    }
    *s = hsum8(acc);
}

/* ggml-quants.c:5318-5381 (AVX2).  Per super-block: d = y.d * fp16(x.d), dmin = -y.d * fp16(x.dmin);
 * mins: q8s = hadd(bsums) -> 8 pair sums, prod lanes 0..3 = madd(mins16, q8s) = m[2l]*bs[2l] + m[2l+1]*bs[2l+1],
 * acc_m = fma(dmin, float(prod), acc_m) (4 lanes); scales: 8 int32 lanes, lane l accumulates, over the four 64-value
 * chunks j, sc[2j]*(sum of 4 products of low-nibble elements 4l..4l+3) + sc[2j+1]*(same for high nibbles) -- note
 * maddubs pairs are int16 (no overflow: 15*127*2) and madd by scale is exact in int32; acc = fma(d, float(sumi), acc).
 * Final: hsum_float_8(acc) + (acc_m reduced as (x + movehl) then (x + movehdup)). */
void po_vec_dot_q4_K_q8_K(int n, float * s, const void * vx, const void * vy) {
    const blk_q4_K * x  = (const blk_q4_K *) vx;
    const blk_q8_K * y  = (const blk_q8_K *) vy;
    const int        nb = n / QK_K;
    float acc[8]   = { 0, 0, 0, 0, 0, 0, 0, 0 };
    float acc_m[4] = { 0, 0, 0, 0 };
    for (int i = 0; i < nb; ++i) {
        const float d    = y[i].d * po_fp16_to_fp32(x[i].d);
        const float dmin = -y[i].d * po_fp16_to_fp32(x[i].dmin);
        uint8_t sc[8], mn[8];
        for (int j = 0; j < 8; j++) get_scale_min_k4(j, x[i].scales, &sc[j], &mn[j]);
        for (int l = 0; l < 4; l++) {
            /* _mm_hadd_epi16(lo, hi): q8s[t] = bsums[2t] + bsums[2t+1] (int16 wrap cannot occur: |.| <= 2*16*128) */
            const int q8s0 = (int16_t) (y[i].bsums[4 * l + 0] + y[i].bsums[4 * l + 1]);
            const int q8s1 = (int16_t) (y[i].bsums[4 * l + 2] + y[i].bsums[4 * l + 3]);
            const int prod = (int) mn[2 * l] * q8s0 + (int) mn[2 * l + 1] * q8s1;
            acc_m[l] = fmaf(dmin, (float) prod, acc_m[l]);
        }
        int32_t sumi[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
        const uint8_t * q4 = x[i].qs;
        const int8_t *  q8 = y[i].qs;
        for (int j = 0; j < QK_K / 64; ++j) {
            for (int l = 0; l < 8; l++) {
                int lo = 0, hi = 0;
                for (int t = 0; t < 4; t++) {
                    lo += (int) (q4[4 * l + t] & 0xF) * (int) q8[4 * l + t];
                    hi += (int) (q4[4 * l + t] >> 4) * (int) q8[32 + 4 * l + t];
                }
                sumi[l] += (int) sc[2 * j] * lo + (int) sc[2 * j + 1] * hi;
            }
            q4 += 32;
            q8 += 64;
        }
        for (int l = 0; l < 8; l++) acc[l] = fmaf(d, (float) sumi[l], acc[l]);
    }
    const float m0 = acc_m[0] + acc_m[2], m1 = acc_m[1] + acc_m[3];
    *s = hsum8(acc) + (m0 + m1);
}

/* ggml-quants.c:1917-1960 */
void po_dequantize_row_q5_K(const void * vx, float * y, int k) {
    const blk_q5_K * x = (const blk_q5_K *) vx;
    const int nb = k / QK_K;
    for (int i = 0; i < nb; i++) {
        const uint8_t * ql = x[i].qs;
        const uint8_t * qh = x[i].qh;
        const float d = po_fp16_to_fp32(x[i].d), min = po_fp16_to_fp32(x[i].dmin);
        int     is = 0;
        uint8_t sc, m, u1 = 1, u2 = 2;
        for (int j = 0; j < QK_K; j += 64) {
            get_scale_min_k4(is + 0, x[i].scales, &sc, &m);
            const float d1 = d * sc, m1 = min * m;
            get_scale_min_k4(is + 1, x[i].scales, &sc, &m);
            const float d2 = d * sc, m2 = min * m;
            for (int l = 0; l < 32; ++l) *y++ = d1 * ((ql[l] & 0xF) + (qh[l] & u1 ? 16 : 0)) - m1;
            for (int l = 0; l < 32; ++l) *y++ = d2 * ((ql[l] >> 4) + (qh[l] & u2 ? 16 : 0)) - m2;
            ql += 32; is += 2;
            u1 <<= 2; u2 <<= 2;
        }
    }
}

/* ggml-quants.c:5773-5850 (AVX2).  As q4_K with a fifth bit per value (qh bit 2j for the low nibbles of chunk j, bit 2j+1 for
 * the high nibbles); the mins go through a SCALAR float: summs += dmin * (int) sum_j m_j (bsums[2j] + bsums[2j+1]) (mul then
 * add, no contraction); result = hsum_float_8(acc) + summs. */
void po_vec_dot_q5_K_q8_K(int n, float * s, const void * vx, const void * vy) {
    const blk_q5_K * x  = (const blk_q5_K *) vx;
    const blk_q8_K * y  = (const blk_q8_K *) vy;
    const int        nb = n / QK_K;
    float acc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    float summs  = 0.f;
    for (int i = 0; i < nb; ++i) {
        const float d    = y[i].d * po_fp16_to_fp32(x[i].d);
        const float dmin = -y[i].d * po_fp16_to_fp32(x[i].dmin);
        uint8_t sc[8], mn[8];
        for (int j = 0; j < 8; j++) get_scale_min_k4(j, x[i].scales, &sc[j], &mn[j]);
        int prod = 0;
        for (int j = 0; j < 8; j++) prod += (int) mn[j] * (int) (int16_t) (y[i].bsums[2 * j] + y[i].bsums[2 * j + 1]);
        {
            const float t = dmin * (float) prod;
            summs = summs + t;
        }
        int32_t sumi[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
        const uint8_t * q5 = x[i].qs;
        const int8_t *  q8 = y[i].qs;
        for (int j = 0; j < QK_K / 64; ++j) {
            for (int l = 0; l < 8; l++) {
                int lo = 0, hi = 0;
                for (int t = 0; t < 4; t++) {
                    const int e  = 4 * l + t;
                    const int ql = (q5[e] & 0xF) + (((x[i].qh[e] >> (2 * j)) & 1) << 4);
                    const int qu = (q5[e] >> 4) + (((x[i].qh[e] >> (2 * j + 1)) & 1) << 4);
                    lo += ql * (int) q8[e];
                    hi += qu * (int) q8[32 + e];
                }
                sumi[l] += (int) sc[2 * j] * lo + (int) sc[2 * j + 1] * hi;
            }
            q5 += 32;
            q8 += 64;
        }
        for (int l = 0; l < 8; l++) acc[l] = fmaf(d, (float) sumi[l], acc[l]);
    }
    *s = hsum8(acc) + summs;
}

/* the sixteen 6-bit scales of a Q3_K super-block (ggml-quants.c:1720-1730, :4459-4465), minus 32 */
static void q3k_scales(const uint8_t * packed, int8_t * out) {
    const uint32_t kmask1 = 0x03030303, kmask2 = 0x0f0f0f0f;
    uint32_t aux[4];
    memcpy(aux, packed, 12);
    const uint32_t tmp = aux[2];
    aux[2] = ((aux[0] >> 4) & kmask2) | (((tmp >> 4) & kmask1) << 4);
    aux[3] = ((aux[1] >> 4) & kmask2) | (((tmp >> 6) & kmask1) << 4);
    aux[0] = (aux[0] & kmask2) | (((tmp >> 0) & kmask1) << 4);
    aux[1] = (aux[1] & kmask2) | (((tmp >> 2) & kmask1) << 4);
    for (int i = 0; i < 16; i++) out[i] = (int8_t) (((const uint8_t *) aux)[i] - 32);
}

/* ggml-quants.c:1711-1758 */
void po_dequantize_row_q3_K(const void * vx, float * y, int k) {
    const blk_q3_K * x = (const blk_q3_K *) vx;
    const int nb = k / QK_K;
    for (int i = 0; i < nb; i++) {
        const float     d_all = po_fp16_to_fp32(x[i].d);
        const uint8_t * q  = x[i].qs;
        const uint8_t * hm = x[i].hmask;
        uint8_t m = 1;
        int8_t  scales[16];
        q3k_scales(x[i].scales, scales);
        int is = 0;
        for (int n = 0; n < QK_K; n += 128) {
            int shift = 0;
            for (int j = 0; j < 4; ++j) {
                float dl = d_all * scales[is++];
                for (int l = 0; l < 16; ++l) *y++ = dl * ((int8_t) ((q[l + 0] >> shift) & 3) - ((hm[l + 0] & m) ? 0 : 4));
                dl = d_all * scales[is++];
                for (int l = 0; l < 16; ++l) *y++ = dl * ((int8_t) ((q[l + 16] >> shift) & 3) - ((hm[l + 16] & m) ? 0 : 4));
                shift += 2;
                m <<= 1;
            }
            q += 32;
        }
    }
}

/* ggml-quants.c:4443-4546 (AVX2).  Per super-block d = y.d * fp16(x.d).  The 256 values are 2 halves x 4 groups of 32 (group g of half
 * j: bits 2g, 2g+1 of the 32 qs bytes of the half, bit 4j + g of the hmask bytes, y[128 j + 32 g ..]); value = low2 - (high bit ? 0 : 4).
 * maddubs gives int16 pair sums (no saturation), madd with the group's two scales - 32 (first 16 values / last 16) gives 8 int32
 * lanes, lane m = elements 4m..4m+3 of the group; the lanes add up over the 8 groups, then acc = fma(d, float(sumi), acc) and
 * hsum_float_8. */
void po_vec_dot_q3_K_q8_K(int n, float * s, const void * vx, const void * vy) {
    const blk_q3_K * x  = (const blk_q3_K *) vx;
    const blk_q8_K * y  = (const blk_q8_K *) vy;
    const int        nb = n / QK_K;
    float acc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    for (int i = 0; i < nb; ++i) {
        const float     d  = y[i].d * po_fp16_to_fp32(x[i].d);
        const uint8_t * q3 = x[i].qs;
        const uint8_t * hm = x[i].hmask;
        const int8_t *  q8 = y[i].qs;
        int8_t scales[16];
        q3k_scales(x[i].scales, scales);
        int32_t sumi[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
        for (int j = 0; j < QK_K / 128; ++j) {
            for (int g = 0; g < 4; g++) {
                for (int m = 0; m < 8; m++) {
                    int dot = 0;
                    for (int t = 0; t < 4; t++) {
                        const int l = 4 * m + t;
                        const int q = ((q3[l] >> (2 * g)) & 3) - ((hm[l] & (1 << (4 * j + g))) ? 0 : 4);
                        dot += q * (int) q8[32 * g + l];
                    }
                    sumi[m] += (int) scales[8 * j + 2 * g + (m >= 4)] * dot;
                }
            }
            q3 += 32; q8 += 128;
        }
        for (int l = 0; l < 8; l++) acc[l] = fmaf(d, (float) sumi[l], acc[l]);
    }
    *s = hsum8(acc);
}

/* ggml-quants.c:1994-2040 */
void po_dequantize_row_q6_K(const void * vx, float * y, int k) {
    const blk_q6_K * x = (const blk_q6_K *) vx;
    const int nb = k / QK_K;
    for (int i = 0; i < nb; i++) {
        const float     d  = po_fp16_to_fp32(x[i].d);
        const uint8_t * ql = x[i].ql;
        const uint8_t * qh = x[i].qh;
        const int8_t *  sc = x[i].scales;
        for (int n = 0; n < QK_K; n += 128) {
            for (int l = 0; l < 32; ++l) {
                const int    is = l / 16;
                const int8_t q1 = (int8_t) ((ql[l + 0] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                const int8_t q2 = (int8_t) ((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                const int8_t q3 = (int8_t) ((ql[l + 0] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                const int8_t q4 = (int8_t) ((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                y[l + 0]  = d * sc[is + 0] * q1;
                y[l + 32] = d * sc[is + 2] * q2;
                y[l + 64] = d * sc[is + 4] * q3;
                y[l + 96] = d * sc[is + 6] * q4;
            }
            y += 128; ql += 64; qh += 32; sc += 8;
        }
    }
}

/* ggml-quants.c:6739-6812 (AVX2).  Per super-block d = y.d * fp16(x.d).  The 256 values are 2 halves x 4 groups of 32
 * (group g of half j: ql nibbles low/high of bytes [0,32) / [32,64), 2 bits of qh, y[128 j + 32 g ..]).  maddubs gives int16
 * pair sums q*y (no saturation: 63*127*2), minus 32*(y pair sum); madd with the group's two int8 scales (first 16 values /
 * last 16) gives 8 int32 lanes, lane m = elements 4m..4m+3 of the group; the lanes add up over the 8 groups, then
 * acc = fma(d, float(sumi), acc) and hsum_float_8. */
void po_vec_dot_q6_K_q8_K(int n, float * s, const void * vx, const void * vy) {
    const blk_q6_K * x  = (const blk_q6_K *) vx;
    const blk_q8_K * y  = (const blk_q8_K *) vy;
    const int        nb = n / QK_K;
    float acc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    for (int i = 0; i < nb; ++i) {
        const float     d  = y[i].d * po_fp16_to_fp32(x[i].d);
        const uint8_t * ql = x[i].ql;
        const uint8_t * qh = x[i].qh;
        const int8_t *  q8 = y[i].qs;
        int32_t sumi[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
        for (int j = 0; j < QK_K / 128; ++j) {
            for (int g = 0; g < 4; g++) {
                for (int m = 0; m < 8; m++) {
                    int dot = 0;
                    for (int t = 0; t < 4; t++) {
                        const int l  = 4 * m + t;
                        const int lo = (g & 1) ? ql[32 + l] : ql[l];
                        const int q  = ((g < 2) ? (lo & 0xF) : (lo >> 4)) | (((qh[l] >> (2 * g)) & 3) << 4);
                        dot += (q - 32) * (int) q8[32 * g + l];
                    }
                    sumi[m] += (int) x[i].scales[8 * j + 2 * g + (m >= 4)] * dot;
                }
            }
            ql += 64; qh += 32; q8 += 128;
        }
        for (int l = 0; l < 8; l++) acc[l] = fmaf(d, (float) sumi[l], acc[l]);
    }
    *s = hsum8(acc);
}

/* ggml.c:1200-1235 with GGML_F16_STEP 32, GGML_F16_EPR 8, GGML_F16_ARR 4 (F16C + FMA), reduce macro ggml.c:772-790 */
void po_vec_dot_f16(int n, float * s, const po_fp16 * x, const po_fp16 * y) {
    double    sumf = 0.0;
    const int np   = n & ~31;
    float     sum[4][8];
    memset(sum, 0, sizeof(sum));
    for (int i = 0; i < np; i += 32)
        for (int j = 0; j < 4; j++)
            for (int l = 0; l < 8; l++)
                sum[j][l] = fmaf(po_fp16_to_fp32(x[i + j * 8 + l]), po_fp16_to_fp32(y[i + j * 8 + l]), sum[j][l]);
    for (int l = 0; l < 8; l++) { sum[0][l] += sum[2][l]; sum[1][l] += sum[3][l]; }
    for (int l = 0; l < 8; l++) sum[0][l] += sum[1][l];
    float t0[4];
    for (int l = 0; l < 4; l++) t0[l] = sum[0][l] + sum[0][4 + l];
    const float t1a = t0[0] + t0[1], t1b = t0[2] + t0[3];   /* _mm_hadd_ps(t0, t0) */
    sumf = (double) (t1a + t1b);                             /* second hadd, then float -> ggml_float */
    for (int i = np; i < n; ++i) sumf += (double) (po_fp16_to_fp32(x[i]) * po_fp16_to_fp32(y[i]));
    *s = (float) sumf;
}

size_t po_row_size(int type, int64_t k) {
    switch (type) {
        case PO_TYPE_Q4_0: return (size_t) (k / QK) * sizeof(blk_q4_0);
        case PO_TYPE_Q8_0: return (size_t) (k / QK) * sizeof(blk_q8_0);
        case PO_TYPE_Q4_K: return (size_t) (k / QK_K) * sizeof(blk_q4_K);
        case PO_TYPE_Q3_K: return (size_t) (k / QK_K) * sizeof(blk_q3_K);
        case PO_TYPE_Q5_K: return (size_t) (k / QK_K) * sizeof(blk_q5_K);
        case PO_TYPE_Q6_K: return (size_t) (k / QK_K) * sizeof(blk_q6_K);
        case PO_TYPE_F32:  return (size_t) k * 4;
    }
    return 0;
}

/* ggml.c:9388-9600: INIT quantises every src1 row with type_traits[vec_dot_type].from_float (Q8_0 for Q4_0/Q8_0
 * weights, Q8_K for Q4_K: ggml.c:455-465, 541-548), then dst[i1][ir0] = vec_dot(ne00, src0 row ir0, src1 row i1). */
void po_mul_mat(int type, const void * w, int64_t N, int64_t K, const float * x, int b, float * y) {
    const size_t wrow = po_row_size(type, K);
    const int    kq   = type == PO_TYPE_Q3_K || type == PO_TYPE_Q4_K || type == PO_TYPE_Q5_K || type == PO_TYPE_Q6_K;             /* K-quants multiply with Q8_K activations */
    const size_t arow = kq ? (size_t) (K / QK_K) * sizeof(blk_q8_K) : (size_t) (K / QK) * sizeof(blk_q8_0);
    char * act = (char *) malloc(arow * (size_t) b);
    for (int t = 0; t < b; t++) {
        if (kq) po_quantize_row_q8_K(x + (size_t) t * K, act + arow * t, (int) K);
        else                      po_quantize_row_q8_0(x + (size_t) t * K, act + arow * t, (int) K);
    }
    for (int t = 0; t < b; t++) {
        for (int64_t r = 0; r < N; r++) {
            const char * wr = (const char *) w + wrow * (size_t) r;
            float *      o  = y + (size_t) t * N + r;
            switch (type) {
                case PO_TYPE_Q4_0: po_vec_dot_q4_0_q8_0((int) K, o, wr, act + arow * t); break;
                case PO_TYPE_Q8_0: po_vec_dot_q8_0_q8_0((int) K, o, wr, act + arow * t); break;
                case PO_TYPE_Q4_K: po_vec_dot_q4_K_q8_K((int) K, o, wr, act + arow * t); break;
                case PO_TYPE_Q3_K: po_vec_dot_q3_K_q8_K((int) K, o, wr, act + arow * t); break;
                case PO_TYPE_Q5_K: po_vec_dot_q5_K_q8_K((int) K, o, wr, act + arow * t); break;
                case PO_TYPE_Q6_K: po_vec_dot_q6_K_q8_K((int) K, o, wr, act + arow * t); break;
            }
        }
    }
    free(act);
}

/* ggml.c:9028-9074: sum of float squares in double, mean as float, scale = 1/sqrtf(mean + eps) */
void po_rms_norm(const float * x, float * y, int n, float eps) {
    double sum = 0.0;
    for (int i = 0; i < n; i++) sum += (double) (x[i] * x[i]);
    const float mean  = (float) (sum / n);
    const float scale = 1.0f / sqrtf(mean + eps);
    for (int i = 0; i < n; i++) y[i] = x[i] * scale;
}

/* ggml.c:8959-9005 */
void po_norm(const float * x, float * y, int n, float eps) {
    double sum = 0.0;
    for (int i = 0; i < n; i++) sum += (double) x[i];
    float  mean = (float) (sum / n);
    double sum2 = 0.0;
    for (int i = 0; i < n; i++) {
        float v = x[i] - mean;
        y[i]    = v;
        sum2 += (double) (v * v);
    }
    float       variance = (float) (sum2 / n);
    const float scale    = 1.0f / sqrtf(variance + eps);
    for (int i = 0; i < n; i++) y[i] *= scale;
}

/* ggml.c:11006-11166 with ext_factor = 0, attn_factor = 1, xpos_base = 0 (the values llama.cpp passes for a model
 * without YaRN metadata).  mode 0 (llama): pairs (2i, 2i+1), theta_i = pos * theta_scale^i by repeated multiply,
 * rope_yarn multiplies by freq_scale (:10979).  mode 2 (NeoX / falcon): pairs (i, i + n_dims/2) and theta_base is
 * multiplied by freq_scale *before* the loop as well (:11137).  x: [n_head][head_dim] of one token. */
void po_rope(float * x, int n_head, int head_dim, int n_dims, int pos, int mode, float freq_base, float freq_scale) {
    const float theta_scale = powf(freq_base, -2.0f / n_dims);
    for (int h = 0; h < n_head; h++) {
        float * v          = x + (size_t) h * head_dim;
        float   theta_base = (float) pos;
        if (!(mode & 2)) {
            for (int i0 = 0; i0 < head_dim; i0 += 2) {
                const float theta     = freq_scale * theta_base;
                const float cos_theta = cosf(theta) * 1.0f;
                const float sin_theta = sinf(theta) * 1.0f;
                const float zeta      = 1.0f;
                theta_base *= theta_scale;
                const float x0 = v[i0], x1 = v[i0 + 1];
                v[i0]     = x0 * cos_theta * zeta - x1 * sin_theta * zeta;
                v[i0 + 1] = x0 * sin_theta * zeta + x1 * cos_theta * zeta;
            }
        } else {
            theta_base *= freq_scale;
            for (int ib = 0; ib < head_dim / n_dims; ++ib) {
                for (int ic = 0; ic < n_dims; ic += 2) {
                    const float theta     = freq_scale * theta_base;
                    const float cos_theta = cosf(theta) * 1.0f;
                    const float sin_theta = sinf(theta) * 1.0f;
                    theta_base *= theta_scale;
                    const int   i0 = ib * n_dims + ic / 2;
                    const float x0 = v[i0], x1 = v[i0 + n_dims / 2];
                    /* the reference Makefile builds with -std=c11 (ISO mode => -ffp-contract=off): no fused ops here */
                    v[i0]              = x0 * cos_theta - x1 * sin_theta;
                    v[i0 + n_dims / 2] = x0 * sin_theta + x1 * cos_theta;
                }
            }
        }
    }
}

/* ggml.c:10551-10612: max, exp through the fp16 table on fp16(x - max), sum in double, scale by (float)(1/sum) */
void po_soft_max(float * p, int n) {
    po_tables_init();
    float max = -INFINITY;
    for (int i = 0; i < n; i++) if (p[i] > max) max = p[i];
    double sum = 0.0;
    for (int i = 0; i < n; i++) {
        if (p[i] == -INFINITY) {
            p[i] = 0.0f;
        } else {
            const float val = po_fp16_to_fp32(tab_exp[po_fp32_to_fp16(p[i] - max)]);
            sum += (double) val;
            p[i] = val;
        }
    }
    sum = 1.0 / sum;
    const float sc = (float) sum;
    for (int i = 0; i < n; i++) p[i] *= sc;
}

void po_silu(float * x, int n) {
    po_tables_init();
    for (int i = 0; i < n; i++) x[i] = po_fp16_to_fp32(tab_silu[po_fp32_to_fp16(x[i])]);
}
void po_gelu(float * x, int n) {
    po_tables_init();
    for (int i = 0; i < n; i++) x[i] = po_fp16_to_fp32(tab_gelu[po_fp32_to_fp16(x[i])]);
}

/* ------------------------------------------------------------------------------------------------------------- */
/* KV-cell metadata.  The reference keeps std::set<llama_seq_id> per cell (llama.cpp:1292-1301); membership, insert,
 * erase and empty() are all the scheduler observes, so a bitset is an exact model.                                */
static int  cell_has(const po_kv_meta * kv, uint32_t i, int32_t s) {
    if (s < 0 || s >= kv->n_words * 32) return 0;
    return (kv->bits[(size_t) i * kv->n_words + (s >> 5)] >> (s & 31)) & 1u;
}
static void cell_set(po_kv_meta * kv, uint32_t i, int32_t s) { kv->bits[(size_t) i * kv->n_words + (s >> 5)] |= 1u << (s & 31); }
static void cell_clr(po_kv_meta * kv, uint32_t i, int32_t s) { kv->bits[(size_t) i * kv->n_words + (s >> 5)] &= ~(1u << (s & 31)); }
static int  cell_empty(const po_kv_meta * kv, uint32_t i) {
    for (int w = 0; w < kv->n_words; w++) if (kv->bits[(size_t) i * kv->n_words + w]) return 0;
    return 1;
}
static void cell_clear(po_kv_meta * kv, uint32_t i) { memset(kv->bits + (size_t) i * kv->n_words, 0, 4 * (size_t) kv->n_words); }

po_kv_meta * po_kv_new(uint32_t n_ctx, int32_t max_seq_ids) {
    po_kv_meta * kv = (po_kv_meta *) calloc(1, sizeof(*kv));
    kv->size    = n_ctx;
    kv->n_words = (max_seq_ids + 31) / 32;
    kv->pos     = (int32_t *) malloc(4 * (size_t) n_ctx);
    kv->bits    = (uint32_t *) calloc((size_t) n_ctx * kv->n_words, 4);
    for (uint32_t i = 0; i < n_ctx; i++) kv->pos[i] = -1;
    return kv;
}
void po_kv_free(po_kv_meta * kv) { if (kv) { free(kv->pos); free(kv->bits); free(kv); } }

void po_kv_clear(po_kv_meta * kv) {
    for (uint32_t i = 0; i < kv->size; ++i) { kv->pos[i] = -1; cell_clear(kv, i); }
    kv->head = 0;
    kv->used = 0;
}

/* llama.cpp:1593-1643 */
int po_kv_find_slot(po_kv_meta * kv, int n_tokens_i, const int32_t * pos, const int32_t * n_seq_id, const int32_t * const * seq_id) {
    const uint32_t n_ctx = kv->size, n_tokens = (uint32_t) n_tokens_i;
    if (n_tokens > n_ctx) return 0;
    uint32_t n_tested = 0;
    while (1) {
        if (kv->head + n_tokens > n_ctx) {
            n_tested += n_ctx - kv->head;
            kv->head = 0;
            continue;
        }
        int found = 1;
        for (uint32_t i = 0; i < n_tokens; i++) {
            if (kv->pos[kv->head + i] >= 0) {
                found = 0;
                kv->head += i + 1;
                n_tested += i + 1;
                break;
            }
        }
        if (found) break;
        if (n_tested >= n_ctx) return 0;
    }
    for (uint32_t i = 0; i < n_tokens; i++) {
        kv->pos[kv->head + i] = pos[i];
        for (int32_t j = 0; j < n_seq_id[i]; j++) cell_set(kv, kv->head + i, seq_id[i][j]);
    }
    kv->used += n_tokens;
    return 1;
}

/* llama.cpp:1647-1655 (note: the loop stops at i > 0, so a cache whose only used cell is cell 0 reports 0) */
int32_t po_kv_cell_max(const po_kv_meta * kv) {
    for (uint32_t i = kv->size - 1; i > 0; --i)
        if (kv->pos[i] >= 0 && !cell_empty(kv, i)) return (int32_t) (i + 1);
    return 0;
}

/* llama.cpp:1666-1697 -- PipeInfer's predicate ((pos >= p0 || pos < 0) && pos < p1) and the removed `continue`:
 * every cell in range that is (or becomes) empty is reset, and the first such cell is the new-head candidate. */
void po_kv_seq_rm(po_kv_meta * kv, int32_t seq, int32_t p0, int32_t p1) {
    uint32_t new_head = kv->size;
    if (p0 < 0) p0 = 0;
    if (p1 < 0) p1 = INT32_MAX;
    for (uint32_t i = 0; i < kv->size; ++i) {
        if ((kv->pos[i] >= p0 || kv->pos[i] < 0) && kv->pos[i] < p1) {
            if (seq < 0) cell_clear(kv, i);
            else if (cell_has(kv, i, seq)) cell_clr(kv, i, seq);
            if (cell_empty(kv, i)) {
                if (kv->pos[i] >= 0) kv->used--;
                kv->pos[i] = -1;
                if (new_head == kv->size) new_head = i;
            }
        }
    }
    if (new_head != kv->size && new_head < kv->head) kv->head = new_head;
}

/* llama.cpp:1699-1715 (sets head = 0 unconditionally) */
void po_kv_seq_cp(po_kv_meta * kv, int32_t src, int32_t dst, int32_t p0, int32_t p1) {
    if (p0 < 0) p0 = 0;
    if (p1 < 0) p1 = INT32_MAX;
    kv->head = 0;
    for (uint32_t i = 0; i < kv->size; ++i)
        if (cell_has(kv, i, src) && kv->pos[i] >= p0 && kv->pos[i] < p1) cell_set(kv, i, dst);
}

/* llama.cpp:1717-1734 */
void po_kv_seq_keep(po_kv_meta * kv, int32_t seq) {
    uint32_t new_head = kv->size;
    for (uint32_t i = 0; i < kv->size; ++i) {
        if (!cell_has(kv, i, seq)) {
            if (kv->pos[i] >= 0) kv->used--;
            kv->pos[i] = -1;
            cell_clear(kv, i);
            if (new_head == kv->size) new_head = i;
        } else {
            cell_clear(kv, i);
            cell_set(kv, i, seq);
        }
    }
    if (new_head != kv->size && new_head < kv->head) kv->head = new_head;
}

/* llama.cpp:5200-5240 */
void po_kv_mask(const po_kv_meta * kv, int n_tokens, int n_kv, const int32_t * pos, const int32_t * n_seq_id,
                const int32_t * const * seq_id, uint8_t * mask) {
    for (int j = 0; j < n_tokens; ++j) {
        for (int i = 0; i < n_kv; ++i) {
            int has = 0;
            for (int s = 0; s < n_seq_id[j]; s++) if (cell_has(kv, (uint32_t) i, seq_id[j][s])) { has = 1; break; }
            mask[(size_t) j * n_kv + i] = (uint8_t) (!has || kv->pos[i] > pos[j]);
        }
    }
}

/* ------------------------------------------------------------------------------------------------------------- */
po_context * po_context_new(const po_model * m, uint32_t n_ctx, int32_t max_seq_ids) {
    po_context * c   = (po_context *) calloc(1, sizeof(*c));
    const size_t gqa = (size_t) (m->n_embd / m->n_head) * m->n_head_kv;
    c->model = m;
    c->n_ctx = n_ctx;
    c->kv    = po_kv_new(n_ctx, max_seq_ids);
    c->k     = (po_fp16 *) calloc((size_t) m->n_layer * n_ctx * gqa, 2);
    c->v     = (po_fp16 *) calloc((size_t) m->n_layer * n_ctx * gqa, 2);
    po_tables_init();
    return c;
}
void po_context_free(po_context * c) { if (c) { po_kv_free(c->kv); free(c->k); free(c->v); free(c); } }

static void dequant_row(int type, const void * w, int64_t row, int64_t k, float * out) {
    const char * p = (const char *) w + po_row_size(type, k) * (size_t) row;
    switch (type) {
        case PO_TYPE_Q4_0: po_dequantize_row_q4_0(p, out, (int) k); break;
        case PO_TYPE_Q8_0: po_dequantize_row_q8_0(p, out, (int) k); break;
        case PO_TYPE_Q4_K: po_dequantize_row_q4_K(p, out, (int) k); break;
        case PO_TYPE_Q3_K: po_dequantize_row_q3_K(p, out, (int) k); break;
        case PO_TYPE_Q5_K: po_dequantize_row_q5_K(p, out, (int) k); break;
        case PO_TYPE_Q6_K: po_dequantize_row_q6_K(p, out, (int) k); break;
        case PO_TYPE_F32:  memcpy(out, p, 4 * (size_t) k); break;
    }
}

/* llm_build_norm with LLM_NORM (llama.cpp:3578-3600): ggml_norm, then ggml_mul by the weight, then ggml_add of the bias */
static void layer_norm_rows(const float * x, float * y, int T, int E, float eps, const float * w, const float * b) {
    for (int t = 0; t < T; t++) {
        po_norm(x + (size_t) t * E, y + (size_t) t * E, E, eps);
        for (int i = 0; i < E; i++) y[(size_t) t * E + i] = y[(size_t) t * E + i] * w[i];
        for (int i = 0; i < E; i++) y[(size_t) t * E + i] = y[(size_t) t * E + i] + b[i];
    }
}

/* masked multi-sequence attention of one layer (llm_build_kqv llama.cpp:3698): shared by both architectures */
static void attention_rows(const po_context * c, int il, int T, int n_kv, const uint8_t * mask, const float * q, float * att,
                           float * kq, po_fp16 * q16, po_fp16 * p16) {
    const po_model * m = c->model;
    const int E = m->n_embd, H = m->n_head, HKV = m->n_head_kv, hd = E / H, GQA = hd * HKV, NC = (int) c->n_ctx;
    const po_fp16 * kl = c->k + (size_t) il * NC * GQA;
    const po_fp16 * vl = c->v + (size_t) il * NC * GQA;
    const float kq_scale = 1.0f / sqrtf((float) hd);
    for (int t = 0; t < T; t++) {
        for (int h = 0; h < H; h++) {
            const int hk = h / (H / HKV);
            for (int i = 0; i < hd; i++) q16[i] = po_fp32_to_fp16(q[(size_t) t * E + h * hd + i]);
            for (int i = 0; i < n_kv; i++) {
                float s;
                po_vec_dot_f16(hd, &s, kl + (size_t) i * GQA + hk * hd, q16);
                s *= kq_scale;
                kq[i] = mask[(size_t) t * n_kv + i] ? -INFINITY : s + 0.0f;
            }
            po_soft_max(kq, n_kv);
            for (int i = 0; i < n_kv; i++) p16[i] = po_fp32_to_fp16(kq[i]);
            for (int d = 0; d < hd; d++) {
                float s;
                po_vec_dot_f16(n_kv, &s, vl + (size_t) (hk * hd + d) * NC, p16);
                att[(size_t) t * E + h * hd + d] = s;
            }
        }
    }
}

/* build_falcon (llama.cpp:4106-4225) inside the decode of :5461-5726 */
static int decode_falcon(po_context * c, int T, const int32_t * token, const int32_t * pos, int n_kv, uint32_t kv_head,
                         const uint8_t * mask, float * logits) {
    const po_model * m = c->model;
    const int E = m->n_embd, H = m->n_head, HKV = m->n_head_kv, hd = E / H, GQA = hd * HKV, FF = m->n_ff, NC = (int) c->n_ctx;
    const int NQKV = E + 2 * GQA;
    float * inpL = (float *) malloc(4 * (size_t) T * E);
    float * an   = (float *) malloc(4 * (size_t) T * E);      /* attn_norm output: feeds the MLP (and QKV in the 7B layout) */
    float * an2  = (float *) malloc(4 * (size_t) T * E);
    float * qkv  = (float *) malloc(4 * (size_t) T * NQKV);
    float * q    = (float *) malloc(4 * (size_t) T * E);
    float * kc   = (float *) malloc(4 * (size_t) T * GQA);
    float * att  = (float *) malloc(4 * (size_t) T * E);
    float * ao   = (float *) malloc(4 * (size_t) T * E);
    float * up   = (float *) malloc(4 * (size_t) T * FF);
    float * fo   = (float *) malloc(4 * (size_t) T * E);
    float * kq   = (float *) malloc(4 * (size_t) n_kv);
    po_fp16 * q16 = (po_fp16 *) malloc(2 * (size_t) hd);
    po_fp16 * p16 = (po_fp16 *) malloc(2 * (size_t) n_kv);
    for (int t = 0; t < T; t++) dequant_row(m->wtype, m->tok_embd, token[t], E, inpL + (size_t) t * E);
    for (int il = 0; il < m->n_layer; il++) {
        po_fp16 * kl = c->k + (size_t) il * NC * GQA;
        po_fp16 * vl = c->v + (size_t) il * NC * GQA;
        layer_norm_rows(inpL, an, T, E, m->norm_eps, m->attn_norm[il], m->attn_norm_b[il]);
        const float * qin = an;
        if (m->attn_norm2 && m->attn_norm2[il]) {                                            /* Falcon-40B :4137-4147 */
            layer_norm_rows(inpL, an2, T, E, m->norm_eps, m->attn_norm2[il], m->attn_norm2_b[il]);
            qin = an2;
        }
        po_mul_mat(m->wtype, m->wq[il], NQKV, E, qin, T, qkv);                                /* wqkv :4149 */
        for (int t = 0; t < T; t++) {                                                        /* ggml_cont of the three views */
            memcpy(q + (size_t) t * E, qkv + (size_t) t * NQKV, 4 * (size_t) E);
            memcpy(kc + (size_t) t * GQA, qkv + (size_t) t * NQKV + E, 4 * (size_t) GQA);
            po_rope(q + (size_t) t * E, H, hd, hd, pos[t], 2, m->rope_freq_base, m->rope_freq_scale);      /* NeoX :4163-4173 */
            po_rope(kc + (size_t) t * GQA, HKV, hd, hd, pos[t], 2, m->rope_freq_base, m->rope_freq_scale);
            for (int i = 0; i < GQA; i++) {
                kl[(size_t) (kv_head + t) * GQA + i] = po_fp32_to_fp16(kc[(size_t) t * GQA + i]);
                vl[(size_t) i * NC + kv_head + t]    = po_fp32_to_fp16(qkv[(size_t) t * NQKV + E + GQA + i]);
            }
        }
        attention_rows(c, il, T, n_kv, mask, q, att, kq, q16, p16);
        po_mul_mat(m->wtype, m->wo[il], E, E, att, T, ao);                                    /* kqv_out */
        po_mul_mat(m->wtype, m->ffn_up[il], FF, E, an, T, up);                                /* the MLP reads attn_norm :4187 */
        po_gelu(up, T * FF);
        po_mul_mat(m->wtype, m->ffn_down[il], E, FF, up, T, fo);
        for (size_t i = 0; i < (size_t) T * E; i++) fo[i] = fo[i] + ao[i];                    /* :4197 */
        for (size_t i = 0; i < (size_t) T * E; i++) inpL[i] = fo[i] + inpL[i];                /* :4200 */
    }
    layer_norm_rows(inpL, an, T, E, m->norm_eps, m->output_norm, m->output_norm_b);
    po_mul_mat(m->wtype, m->output, m->n_vocab, E, an, T, logits);
    free(inpL); free(an); free(an2); free(qkv); free(q); free(kc); free(att); free(ao); free(up); free(fo); free(kq); free(q16); free(p16);
    return 0;
}

/* debugging aid: with PO_DUMP_DIR set every intermediate of the llama graph is written to <dir>/<name>.<layer>.bin (float32) */
static void po_dump(const char * name, int il, const float * p, size_t n) {
    const char * dir = getenv("PO_DUMP_DIR");
    if (!dir) return;
    char path[512];
    snprintf(path, sizeof(path), "%s/%s.%d.bin", dir, name, il);
    FILE * f = fopen(path, "wb");
    if (f) { fwrite(p, 4, n, f); fclose(f); }
}

/* llama.cpp:5461-5726 (phase 0) + build_llama :3872-3980 + llm_build_kv_store :3546 + llm_build_kqv :3698 */
int po_decode(po_context * c, int n_tokens, const int32_t * token, const int32_t * pos, const int32_t * n_seq_id,
              const int32_t * const * seq_id, float * logits) {
    const po_model * m  = c->model;
    po_kv_meta *     kv = c->kv;
    const int E = m->n_embd, H = m->n_head, HKV = m->n_head_kv, hd = E / H, GQA = hd * HKV, FF = m->n_ff;
    const int T = n_tokens, NC = (int) c->n_ctx;

    if (kv->head > kv->used + 2 * (uint32_t) T) kv->head = 0;                       /* :5548-5550 */
    if (!po_kv_find_slot(kv, T, pos, n_seq_id, seq_id)) return -1;                  /* :5562-5565 */
    {
        int32_t cm = po_kv_cell_max(kv);                                            /* :5571 */
        int32_t n  = cm > 32 ? cm : 32;
        kv->n      = (uint32_t) (n < NC ? n : NC);
    }
    const int      n_kv    = (int) kv->n;
    const uint32_t kv_head = kv->head;
    uint8_t *      mask    = (uint8_t *) malloc((size_t) T * n_kv);
    po_kv_mask(kv, T, n_kv, pos, n_seq_id, seq_id, mask);
    const float kq_scale = 1.0f / sqrtf((float) hd);                                /* :5189 */
    if (m->arch == 1) {
        const int rc = decode_falcon(c, T, token, pos, n_kv, kv_head, mask, logits);
        kv->head += (uint32_t) T;
        if (kv->head >= kv->size) kv->head = 0;
        free(mask);
        return rc;
    }

    float * inpL = (float *) malloc(4 * (size_t) T * E);
    float * cur  = (float *) malloc(4 * (size_t) T * E);
    float * q    = (float *) malloc(4 * (size_t) T * E);
    float * kc   = (float *) malloc(4 * (size_t) T * GQA);
    float * vc   = (float *) malloc(4 * (size_t) T * GQA);
    float * att  = (float *) malloc(4 * (size_t) T * E);
    float * ffi  = (float *) malloc(4 * (size_t) T * E);
    float * gate = (float *) malloc(4 * (size_t) T * FF);
    float * up   = (float *) malloc(4 * (size_t) T * FF);
    float * kq   = (float *) malloc(4 * (size_t) n_kv);
    po_fp16 * q16 = (po_fp16 *) malloc(2 * (size_t) hd);
    po_fp16 * p16 = (po_fp16 *) malloc(2 * (size_t) n_kv);

#define TT(arr) ((arr) ? (arr)[il] : m->wtype)          /* per-tensor type of a mixed file */
    for (int t = 0; t < T; t++) dequant_row(m->wtype, m->tok_embd, token[t], E, inpL + (size_t) t * E);   /* get_rows */

    for (int il = 0; il < m->n_layer; il++) {
        po_fp16 * kl = c->k + (size_t) il * NC * GQA;
        po_fp16 * vl = c->v + (size_t) il * NC * GQA;
        po_dump("inp", il, inpL, (size_t) T * E);
        for (int t = 0; t < T; t++) {
            po_rms_norm(inpL + (size_t) t * E, cur + (size_t) t * E, E, m->rms_eps);
            for (int i = 0; i < E; i++) cur[(size_t) t * E + i] *= m->attn_norm[il][i];       /* ggml_mul :3593 */
        }
        po_mul_mat(TT(m->t_wq), m->wq[il], E, E, cur, T, q);
        po_mul_mat(TT(m->t_wk), m->wk[il], GQA, E, cur, T, kc);
        po_mul_mat(TT(m->t_wv), m->wv[il], GQA, E, cur, T, vc);
        po_dump("q", il, q, (size_t) T * E); po_dump("k", il, kc, (size_t) T * GQA); po_dump("v", il, vc, (size_t) T * GQA);
        for (int t = 0; t < T; t++) {
            po_rope(q + (size_t) t * E, H, hd, m->n_rot, pos[t], 0, m->rope_freq_base, m->rope_freq_scale);
            po_rope(kc + (size_t) t * GQA, HKV, hd, m->n_rot, pos[t], 0, m->rope_freq_base, m->rope_freq_scale);
        }
        for (int t = 0; t < T; t++) {                                                        /* kv_store */
            for (int i = 0; i < GQA; i++) {
                kl[(size_t) (kv_head + t) * GQA + i] = po_fp32_to_fp16(kc[(size_t) t * GQA + i]);
                vl[(size_t) i * NC + kv_head + t]    = po_fp32_to_fp16(vc[(size_t) t * GQA + i]);
            }
        }
        for (int t = 0; t < T; t++) {                                                        /* kqv */
            for (int h = 0; h < H; h++) {
                const int hk = h / (H / HKV);                                                /* ggml.c:9567 i02 = i12/r2 */
                for (int i = 0; i < hd; i++) q16[i] = po_fp32_to_fp16(q[(size_t) t * E + h * hd + i]);
                for (int i = 0; i < n_kv; i++) {
                    float s;
                    po_vec_dot_f16(hd, &s, kl + (size_t) i * GQA + hk * hd, q16);
                    s *= kq_scale;                                                           /* ggml_scale */
                    kq[i] = mask[(size_t) t * n_kv + i] ? -INFINITY : s + 0.0f;              /* ggml_add mask */
                }
                po_soft_max(kq, n_kv);
                for (int i = 0; i < n_kv; i++) p16[i] = po_fp32_to_fp16(kq[i]);
                for (int d = 0; d < hd; d++) {
                    float s;
                    po_vec_dot_f16(n_kv, &s, vl + (size_t) (hk * hd + d) * NC, p16);
                    att[(size_t) t * E + h * hd + d] = s;
                }
            }
        }
        po_dump("att", il, att, (size_t) T * E);
        po_mul_mat(TT(m->t_wo), m->wo[il], E, E, att, T, cur);
        po_dump("o", il, cur, (size_t) T * E);
        for (size_t i = 0; i < (size_t) T * E; i++) ffi[i] = cur[i] + inpL[i];               /* ffn_inp :3944 */
        po_dump("ffi", il, ffi, (size_t) T * E);
        for (int t = 0; t < T; t++) {
            po_rms_norm(ffi + (size_t) t * E, cur + (size_t) t * E, E, m->rms_eps);
            for (int i = 0; i < E; i++) cur[(size_t) t * E + i] *= m->ffn_norm[il][i];
        }
        po_mul_mat(TT(m->t_up), m->ffn_up[il], FF, E, cur, T, up);                              /* llm_build_ffn :3611 */
        po_mul_mat(TT(m->t_gate), m->ffn_gate[il], FF, E, cur, T, gate);
        po_dump("gate", il, gate, (size_t) T * FF); po_dump("up", il, up, (size_t) T * FF);
        po_silu(gate, T * FF);
        for (size_t i = 0; i < (size_t) T * FF; i++) gate[i] = gate[i] * up[i];
        po_mul_mat(TT(m->t_down), m->ffn_down[il], E, FF, gate, T, cur);
        po_dump("down", il, cur, (size_t) T * E);
        for (size_t i = 0; i < (size_t) T * E; i++) inpL[i] = cur[i] + ffi[i];               /* l_out :3962 */
    }
    for (int t = 0; t < T; t++) {
        po_rms_norm(inpL + (size_t) t * E, cur + (size_t) t * E, E, m->rms_eps);
        for (int i = 0; i < E; i++) cur[(size_t) t * E + i] *= m->output_norm[i];
    }
    po_mul_mat(m->t_output ? m->t_output : m->wtype, m->output, m->n_vocab, E, cur, T, logits);
#undef TT

    kv->head += (uint32_t) T;                                                                /* :5708-5713 */
    if (kv->head >= kv->size) kv->head = 0;

    free(mask); free(inpL); free(cur); free(q); free(kc); free(vc); free(att); free(ffi); free(gate); free(up);
    free(kq); free(q16); free(p16);
    return 0;
}

/* llama.cpp:7496-7510: std::max_element with a.logit < b.logit returns the FIRST maximum */
int32_t po_sample_greedy(const float * logits, int n_vocab) {
    int best = 0;
    for (int i = 1; i < n_vocab; i++) if (logits[best] < logits[i]) best = i;
    return best;
}

typedef struct { float l; int32_t id; } cand_t;
static int cand_cmp(const void * a, const void * b) {
    const cand_t * x = (const cand_t *) a, * y = (const cand_t *) b;
    if (x->l > y->l) return -1;
    if (x->l < y->l) return 1;
    return x->id - y->id;   /* std::sort leaves the order of equal logits unspecified; we pick ascending id */
}
/* llama.cpp:7019-7046: sort descending, p_i = expf(l_i - max), cum_sum accumulated in float in sorted order, p /= cum_sum */
void po_sample_softmax_topk(const float * logits, int n_vocab, int k, int32_t * ids, float * logit, float * p) {
    cand_t * c = (cand_t *) malloc(sizeof(cand_t) * (size_t) n_vocab);
    for (int i = 0; i < n_vocab; i++) { c[i].l = logits[i]; c[i].id = i; }
    qsort(c, (size_t) n_vocab, sizeof(cand_t), cand_cmp);
    const float max_l = c[0].l;
    float cum = 0.0f;
    for (int i = 0; i < n_vocab; i++) cum += expf(c[i].l - max_l);
    for (int i = 0; i < k && i < n_vocab; i++) { ids[i] = c[i].id; logit[i] = c[i].l; p[i] = expf(c[i].l - max_l) / cum; }
    free(c);
}

/* ggml-mpi.c:523-559 with the call of llama.cpp:8933 (start = 0, end = n_layer - 1): range_length = end - start + 1;
 * ranges[i] = [next, MIN(end, next + w_i * range_length + start)] (float -> uint16 truncation), last node gets
 * [next, end].  A stage then runs layers [ranges[i][0], MIN(n_layer, ranges[i][1])) except the last, which runs to
 * the end of the graph (ggml-mpi.c:640-650). */
void po_split_range(int n_nodes, uint16_t start, uint16_t end, const float * weights, uint16_t * ranges) {
    uint16_t range_length = (uint16_t) (end - start + 1);
    uint16_t next_layer   = 0;
    for (int i = 0; i < n_nodes - 1; i++) {
        ranges[2 * i] = next_layer;
        float    hi = ranges[2 * i] + (weights[i] * range_length) + start;
        uint16_t r  = (uint16_t) ((float) end < hi ? (float) end : hi);
        ranges[2 * i + 1] = r;
        next_layer        = r;
    }
    ranges[2 * (n_nodes - 1)]     = next_layer;
    ranges[2 * (n_nodes - 1) + 1] = end;
}
