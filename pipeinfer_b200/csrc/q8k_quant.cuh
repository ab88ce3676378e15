// q8k_quant.cuh -- one Q8_K super-block (256 values) quantised by one warp, lane l holding values 8l .. 8l+7; shared by the
// activation-quantiser kernels (act_quant.cu) and by the GEMV that quantises its own activations in its prologue (gemv_mma.cu),
// so that the two produce the same bytes.  Reference: quantize_row_q8_K (ggml-quants.c:2329-2363): iscale = -128 / (signed value
// of the first max-|x| element), q = min(127, rne(iscale * x)), d = 1 / iscale, 16-wide bsums.
#pragma once
#include "pib_common.cuh"

// Every pointer addresses this super-block's part of its plane (see pib_act_layout) and may be null:
//   qs   [256] int8   quants in element order           qsp [256] int8   the same in B-fragment order of mma.m16n8k32
//   bs   [16]  int16  sums of 16                         bs8 [16]  int8   sub-block (32) sums split as 64 hi + lo: hi[8], lo[8]
//   sbs  [8]   int16  sub-block sums                     qsl [64]  int2   B registers of the lane-exact kernel     d [1] float
//   q6p  [256] int8   quants in the k16 B-fragment order of the Q6_K kernel      b16 [32] int8   sums of 16 split as 64 hi + lo: hi[16], lo[16]
__device__ __forceinline__ void q8k_quantise_sb(const float (&v)[8], int lane, int8_t * qs, int8_t * qsp, int16_t * bs, int8_t * bs8,
                                                int16_t * sbs, int2 * qsl, float * d, int8_t * q6p = nullptr, int8_t * b16 = nullptr) {
    // first element (in index order) with the largest |x|: strict '>' scan, ties keep the lower index
    float amax = 0.0f, mx = 0.0f;
    int   idx  = 1 << 30;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const float ax = fabsf(v[j]);
        if (ax > amax) { amax = ax; mx = v[j]; idx = lane * 8 + j; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oa = __shfl_xor_sync(0xffffffffu, amax, o);
        const float om = __shfl_xor_sync(0xffffffffu, mx, o);
        const int   oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (oa > amax || (oa == amax && oi < idx)) { amax = oa; mx = om; idx = oi; }
    }
    int   q[8];
    float dval = 0.0f;
    if (amax == 0.0f) {
#pragma unroll
        for (int j = 0; j < 8; j++) q[j] = 0;
    } else {
        const float iscale = -128.f / mx;
#pragma unroll
        for (int j = 0; j < 8; j++) q[j] = min(127, __float2int_rn(__fmul_rn(iscale, v[j])));
        dval = 1.0f / iscale;
    }
    int2 packed;
    packed.x = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((q[3] & 0xff) << 24);
    packed.y = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((q[7] & 0xff) << 24);
    if (qs) *reinterpret_cast<int2 *>(qs + lane * 8) = packed;
    // fragment-order copy for the tensor-core GEMV: lane = 8c + 4h + t  ->  offset 64t + 16c + 8h
    if (qsp) *reinterpret_cast<int2 *>(qsp + (lane & 3) * 64 + (lane >> 3) * 16 + ((lane >> 2) & 1) * 8) = packed;
    // lane l holds elements 8l .. 8l+7 = group l / 2, offsets 8 (l & 1) ..: quarters t = 2 (l & 1) and 2 (l & 1) + 1
    if (q6p) {
        *reinterpret_cast<int *>(q6p + (2 * (lane & 1)) * 64 + (lane >> 1) * 4)     = packed.x;
        *reinterpret_cast<int *>(q6p + (2 * (lane & 1) + 1) * 64 + (lane >> 1) * 4) = packed.y;
    }
    const int s8  = q[0] + q[1] + q[2] + q[3] + q[4] + q[5] + q[6] + q[7];
    const int s16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 1);
    if (bs && (lane & 1) == 0) bs[lane >> 1] = (int16_t) s16;
    if (b16 && (lane & 1) == 0) { b16[lane >> 1] = (int8_t) (s16 >> 6); b16[16 + (lane >> 1)] = (int8_t) (s16 & 63); }
    const int s32 = s16 + __shfl_xor_sync(0xffffffffu, s16, 2);        // the 32-element sub-block sum (lanes 4j..4j+3)
    if ((lane & 3) == 0) {
        if (bs8) { bs8[lane >> 2] = (int8_t) (s32 >> 6); bs8[8 + (lane >> 2)] = (int8_t) (s32 & 63); }
        if (sbs) sbs[lane >> 2] = (int16_t) s32;
    }
    // lane-group copy for the lane-exact kernel: this lane holds lane groups 2 (lane % 4) and 2 (lane % 4) + 1 of sub-block lane / 4
    if (qsl) {
        int2 * o = qsl + 2 * (lane & 3) * 8 + (lane >> 2);
        o[0] = make_int2(packed.x, 0);
        o[8] = make_int2(0, packed.y);
    }
    if (d && lane == 0) *d = dval;
}
