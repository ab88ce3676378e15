// gemv_dp4a.cu -- variant 1 of the quantised matmul (K1): one warp per output row, 128-bit streaming weight loads,
// int8 dot products with dp4a, fp32 block-scale FMA, warp-shuffle reduction.
//
// Arithmetic follows the reference CPU path (reference ggml-quants.c:2426 q4_0.q8_0, :3521 q8_0.q8_0, :5216 q4_K.q8_K):
// exact integer sums per 32-element block, scaled by fp16(d_w) * d_x in fp32.  Only the order in which the fp32 block
// terms are added differs (per-lane partial sums + shuffle tree instead of 8 AVX lanes), i.e. results agree with the
// reference to fp32 rounding, not bit for bit.
//
// HBM-bound: every weight byte is read exactly once (ld.global.nc.L1::no_allocate); the quantised activations
// (b*K bytes) are re-read by every warp but stay L1/L2 resident.  This is the simple, always-correct kernel; the
// TMA-staged tensor-core kernel in gemv_mma.cu is the fast one for b > 1.
#include "pib_common.cuh"

#define GV_WARPS 4

__device__ __forceinline__ int dot8(const int4 & w0, const int4 & w1, const int4 & a0, const int4 & a1) {
    int s = __dp4a(w0.x, a0.x, 0);
    s     = __dp4a(w0.y, a0.y, s);
    s     = __dp4a(w0.z, a0.z, s);
    s     = __dp4a(w0.w, a0.w, s);
    s     = __dp4a(w1.x, a1.x, s);
    s     = __dp4a(w1.y, a1.y, s);
    s     = __dp4a(w1.z, a1.z, s);
    s     = __dp4a(w1.w, a1.w, s);
    return s;
}

template <int B>
__global__ void __launch_bounds__(GV_WARPS * 32) gemv_q8_0_kernel(const int8_t * __restrict__ wqs, const uint16_t * __restrict__ wd,
                                                                  const int8_t * __restrict__ aqs, const float * __restrict__ ad,
                                                                  float * __restrict__ y, int N, int K, const int * __restrict__ cancel) {
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;
    const int lane = threadIdx.x & 31;
    const int row  = blockIdx.x * GV_WARPS + (threadIdx.x >> 5);
    if (row >= N) return;
    const int        nb   = K / 32;
    const int4 *     wrow = reinterpret_cast<const int4 *>(wqs + (size_t) row * K);
    const uint16_t * drow = wd + (size_t) row * nb;
    float            acc[B];
#pragma unroll
    for (int t = 0; t < B; t++) acc[t] = 0.f;
#pragma unroll 2
    for (int blk = lane; blk < nb; blk += 32) {
        const int4  w0 = ld_stream_v4(wrow + blk * 2);
        const int4  w1 = ld_stream_v4(wrow + blk * 2 + 1);
        const float dw = pib_h2f(drow[blk]);
#pragma unroll
        for (int t = 0; t < B; t++) {
            const int4 * a  = reinterpret_cast<const int4 *>(aqs + (size_t) t * K + (size_t) blk * 32);
            const int4   a0 = __ldg(a), a1 = __ldg(a + 1);
            const int    s  = dot8(w0, w1, a0, a1);
            acc[t]          = fmaf(__fmul_rn(dw, ad[(size_t) t * nb + blk]), (float) s, acc[t]);
        }
    }
#pragma unroll
    for (int t = 0; t < B; t++) {
        const float v = warp_sum(acc[t]);
        if (lane == 0) y[(size_t) t * N + row] = v;
    }
}

template <int B>
__global__ void __launch_bounds__(GV_WARPS * 32) gemv_q4_0_kernel(const uint8_t * __restrict__ wqs, const uint16_t * __restrict__ wd,
                                                                  const int8_t * __restrict__ aqs, const float * __restrict__ ad,
                                                                  float * __restrict__ y, int N, int K, const int * __restrict__ cancel) {
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;
    const int lane = threadIdx.x & 31;
    const int row  = blockIdx.x * GV_WARPS + (threadIdx.x >> 5);
    if (row >= N) return;
    const int        nb   = K / 32;
    const int4 *     wrow = reinterpret_cast<const int4 *>(wqs + (size_t) row * (K / 2));
    const uint16_t * drow = wd + (size_t) row * nb;
    float            acc[B];
#pragma unroll
    for (int t = 0; t < B; t++) acc[t] = 0.f;
#pragma unroll 2
    for (int blk = lane; blk < nb; blk += 32) {
        const int4  w  = ld_stream_v4(wrow + blk);
        const float dw = pib_h2f(drow[blk]);
        // byte j: low nibble = element j, high nibble = element j + 16 (ggml-quants.c:2715-2727); offset by -8
        int4 lo, hi;
        lo.x = __vsub4(w.x & 0x0F0F0F0F, 0x08080808);  hi.x = __vsub4((w.x >> 4) & 0x0F0F0F0F, 0x08080808);
        lo.y = __vsub4(w.y & 0x0F0F0F0F, 0x08080808);  hi.y = __vsub4((w.y >> 4) & 0x0F0F0F0F, 0x08080808);
        lo.z = __vsub4(w.z & 0x0F0F0F0F, 0x08080808);  hi.z = __vsub4((w.z >> 4) & 0x0F0F0F0F, 0x08080808);
        lo.w = __vsub4(w.w & 0x0F0F0F0F, 0x08080808);  hi.w = __vsub4((w.w >> 4) & 0x0F0F0F0F, 0x08080808);
#pragma unroll
        for (int t = 0; t < B; t++) {
            const int4 * a  = reinterpret_cast<const int4 *>(aqs + (size_t) t * K + (size_t) blk * 32);
            const int4   a0 = __ldg(a), a1 = __ldg(a + 1);
            const int    s  = dot8(lo, hi, a0, a1);
            acc[t]          = fmaf(__fmul_rn(dw, ad[(size_t) t * nb + blk]), (float) s, acc[t]);
        }
    }
#pragma unroll
    for (int t = 0; t < B; t++) {
        const float v = warp_sum(acc[t]);
        if (lane == 0) y[(size_t) t * N + row] = v;
    }
}

// 6-bit scale / min of sub-block j from the 12 packed bytes (reference get_scale_min_k4, ggml-quants.c:1446-1454)
__device__ __forceinline__ void scale_min_k4(int j, const uint8_t * q, int & sc, int & m) {
    if (j < 4) {
        sc = q[j] & 63;
        m  = q[j + 4] & 63;
    } else {
        sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        m  = (q[j + 4] >> 4) | ((q[j] >> 6) << 4);
    }
}

template <int B>
__global__ void __launch_bounds__(GV_WARPS * 32) gemv_q4_K_kernel(const uint8_t * __restrict__ wqs, const uint8_t * __restrict__ whdr,
                                                                  const int8_t * __restrict__ aqs, const float * __restrict__ ad,
                                                                  const int16_t * __restrict__ absum, float * __restrict__ y,
                                                                  int N, int K, const int * __restrict__ cancel) {
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;
    const int lane = threadIdx.x & 31;
    const int row  = blockIdx.x * GV_WARPS + (threadIdx.x >> 5);
    if (row >= N) return;
    const int    nsb  = K / 256;
    const int    nu   = K / 64;       // 64-element chunks: 32 bytes of nibbles, two 32-element sub-blocks
    const int4 * wrow = reinterpret_cast<const int4 *>(wqs + (size_t) row * (K / 2));
    const int4 * hrow = reinterpret_cast<const int4 *>(whdr + (size_t) row * nsb * 16);
    float        acc[B];
#pragma unroll
    for (int t = 0; t < B; t++) acc[t] = 0.f;
    for (int u = lane; u < nu; u += 32) {
        const int  sb = u >> 2, c = u & 3;
        const int4 w0 = ld_stream_v4(wrow + u * 2);
        const int4 w1 = ld_stream_v4(wrow + u * 2 + 1);
        const int4 h  = __ldg(hrow + sb);
        const float dw   = pib_h2f((uint16_t) (h.x & 0xffff));
        const float dmin = pib_h2f((uint16_t) ((uint32_t) h.x >> 16));
        uint8_t     scb[12];
        *reinterpret_cast<int *>(scb)     = h.y;
        *reinterpret_cast<int *>(scb + 4) = h.z;
        *reinterpret_cast<int *>(scb + 8) = h.w;
        int sc0, m0, sc1, m1;
        scale_min_k4(2 * c, scb, sc0, m0);
        scale_min_k4(2 * c + 1, scb, sc1, m1);
        int4 lo0, lo1, hi0, hi1;    // low nibbles = elements 0..31 of the chunk, high nibbles = elements 32..63
        lo0.x = w0.x & 0x0F0F0F0F; lo0.y = w0.y & 0x0F0F0F0F; lo0.z = w0.z & 0x0F0F0F0F; lo0.w = w0.w & 0x0F0F0F0F;
        lo1.x = w1.x & 0x0F0F0F0F; lo1.y = w1.y & 0x0F0F0F0F; lo1.z = w1.z & 0x0F0F0F0F; lo1.w = w1.w & 0x0F0F0F0F;
        hi0.x = (w0.x >> 4) & 0x0F0F0F0F; hi0.y = (w0.y >> 4) & 0x0F0F0F0F; hi0.z = (w0.z >> 4) & 0x0F0F0F0F; hi0.w = (w0.w >> 4) & 0x0F0F0F0F;
        hi1.x = (w1.x >> 4) & 0x0F0F0F0F; hi1.y = (w1.y >> 4) & 0x0F0F0F0F; hi1.z = (w1.z >> 4) & 0x0F0F0F0F; hi1.w = (w1.w >> 4) & 0x0F0F0F0F;
#pragma unroll
        for (int t = 0; t < B; t++) {
            const int4 * a  = reinterpret_cast<const int4 *>(aqs + (size_t) t * K + (size_t) u * 64);
            const int4   a0 = __ldg(a), a1 = __ldg(a + 1), a2 = __ldg(a + 2), a3 = __ldg(a + 3);
            const int    i0 = dot8(lo0, lo1, a0, a1);
            const int    i1 = dot8(hi0, hi1, a2, a3);
            const int2   bs = __ldg(reinterpret_cast<const int2 *>(absum + (size_t) t * (K / 16) + (size_t) u * 4));
            const int    b0 = (int) (short) (bs.x & 0xffff) + (int) (short) ((uint32_t) bs.x >> 16);
            const int    b1 = (int) (short) (bs.y & 0xffff) + (int) (short) ((uint32_t) bs.y >> 16);
            const float  dx = ad[(size_t) t * nsb + sb];
            acc[t] = fmaf(__fmul_rn(dx, dw), (float) (sc0 * i0 + sc1 * i1), acc[t]);
            acc[t] = fmaf(__fmul_rn(-dx, dmin), (float) (m0 * b0 + m1 * b1), acc[t]);
        }
    }
#pragma unroll
    for (int t = 0; t < B; t++) {
        const float v = warp_sum(acc[t]);
        if (lane == 0) y[(size_t) t * N + row] = v;
    }
}

template <int B>
static void launch_b(const pib_weight * w, const uint8_t * act, const pib_act_layout & L, int t0, int b_total, float * y, const int * cancel, cudaStream_t st) {
    const int     N = (int) w->N, K = (int) w->K;
    const dim3    grid((N + GV_WARPS - 1) / GV_WARPS), block(GV_WARPS * 32);
    const int8_t * aqs = (const int8_t *) (act + L.qs_off) + (size_t) t0 * K;
    float *        yo  = y + (size_t) t0 * N;
    (void) b_total;
    if (w->type == PIB_TYPE_Q8_0) {
        const float * ad = (const float *) (act + L.d_off) + (size_t) t0 * (K / 32);
        pib_launch(gemv_q8_0_kernel<B>, grid, block, 0, st, (const int8_t *) w->qs, w->d, aqs, ad, yo, N, K, cancel);
    } else if (w->type == PIB_TYPE_Q4_0) {
        const float * ad = (const float *) (act + L.d_off) + (size_t) t0 * (K / 32);
        pib_launch(gemv_q4_0_kernel<B>, grid, block, 0, st, w->qs, w->d, aqs, ad, yo, N, K, cancel);
    } else {
        const float *   ad = (const float *) (act + L.d_off) + (size_t) t0 * (K / 256);
        const int16_t * bs = (const int16_t *) (act + L.bs_off) + (size_t) t0 * (K / 16);
        pib_launch(gemv_q4_K_kernel<B>, grid, block, 0, st, w->qs, (const uint8_t *) w->d, aqs, ad, bs, yo, N, K, cancel);
    }
}

int pib_gemv_dp4a(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (w->type == PIB_TYPE_Q6_K || w->type == PIB_TYPE_Q5_K) { pib_set_error("pib_gemv: Q5_K / Q6_K have their own kernels"); return PIB_ERR_UNSUPP; }
    if (!w->qs) { pib_set_error("pib_gemv: variant 1 needs the plane form, which this matrix dropped (set PIB_KEEP_PLANES=1 before loading the model)"); return PIB_ERR_UNSUPP; }
    const pib_act_layout L   = pib_act_layout_of(w->type, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    for (int t0 = 0; t0 < b; t0 += 8) {
        const int nb = b - t0 < 8 ? b - t0 : 8;
        switch (nb) {
            case 1: launch_b<1>(w, act, L, t0, b, y_dev, cancel, st); break;
            case 2: launch_b<2>(w, act, L, t0, b, y_dev, cancel, st); break;
            case 3: launch_b<3>(w, act, L, t0, b, y_dev, cancel, st); break;
            case 4: launch_b<4>(w, act, L, t0, b, y_dev, cancel, st); break;
            case 5: launch_b<5>(w, act, L, t0, b, y_dev, cancel, st); break;
            case 6: launch_b<6>(w, act, L, t0, b, y_dev, cancel, st); break;
            case 7: launch_b<7>(w, act, L, t0, b, y_dev, cancel, st); break;
            default: launch_b<8>(w, act, L, t0, b, y_dev, cancel, st); break;
        }
    }
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}
