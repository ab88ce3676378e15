// gemv_q5k.cu -- dequant-GEMV for Q5_K weights (5.5 bits per weight) at verification batch 1..16.
//
// Q4_K_M files of models with 4 or more query heads per KV head (the 70B class) keep attn_v in Q5_K (reference
// llama.cpp:7909-7933).  Arithmetic of ggml_vec_dot_q5_K_q8_K (ggml-quants.c:5680, AVX2 path): as Q4_K with a fifth bit per
// value -- per super-block eight sub-blocks of 32 with 6-bit scale and min, exact integer sums
// sum_j sc_j * sum(q * y) and sum_j m_j * bsum_j, scaled by fp16(d) * d_x and -fp16(dmin) * d_x in fp32.
//
// Layout (weights.cu): per row and super-block 160 bytes {qs[128] low nibbles, qh[32] fifth bits} and the 16-byte header
// {d, dmin, scales[12]} of Q4_K.  One warp per row, half a warp per super-block: lane = (64-value chunk j, eight byte
// positions 8p..8p+7): two sub-blocks' worth of bytes assembled with SIMD-in-a-word logic, dp4a.
#include "pib_common.cuh"

#define Q5_WARPS 4

// 6-bit scale / min of sub-block j from the 12 packed bytes (reference get_scale_min_k4, ggml-quants.c:1446-1454)
__device__ __forceinline__ void q5_scale_min(int j, const uint8_t * q, int & sc, int & m) {
    if (j < 4) { sc = q[j] & 63; m = q[j + 4] & 63; }
    else { sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

template <int B>
__global__ void __launch_bounds__(Q5_WARPS * 32) gemv_q5_K_kernel(const uint8_t * __restrict__ wqs, const uint8_t * __restrict__ whdr,
                                                                  const int8_t * __restrict__ aqs, const float * __restrict__ ad,
                                                                  const int16_t * __restrict__ absum, float * __restrict__ y, int N, int K,
                                                                  const int * __restrict__ cancel) {
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;
    const int lane = threadIdx.x & 31;
    const int row  = blockIdx.x * Q5_WARPS + (threadIdx.x >> 5);
    if (row >= N) return;
    const int nsb = K / 256;
    const int j = (lane >> 2) & 3, p = lane & 3;
    float acc[B];
#pragma unroll
    for (int t = 0; t < B; t++) acc[t] = 0.f;
    for (int sb = lane >> 4; sb < nsb; sb += 2) {
        const uint8_t * base = wqs + ((size_t) row * nsb + sb) * 160;
        const uint8_t * hdr  = whdr + ((size_t) row * nsb + sb) * 16;
        const uint2 w = *reinterpret_cast<const uint2 *>(base + 32 * j + 8 * p);
        const uint2 h = *reinterpret_cast<const uint2 *>(base + 128 + 8 * p);
        const int lo0 = (int) ((w.x & 0x0F0F0F0Fu) | (((h.x >> (2 * j)) & 0x01010101u) << 4));
        const int lo1 = (int) ((w.y & 0x0F0F0F0Fu) | (((h.y >> (2 * j)) & 0x01010101u) << 4));
        const int hi0 = (int) (((w.x >> 4) & 0x0F0F0F0Fu) | (((h.x >> (2 * j + 1)) & 0x01010101u) << 4));
        const int hi1 = (int) (((w.y >> 4) & 0x0F0F0F0Fu) | (((h.y >> (2 * j + 1)) & 0x01010101u) << 4));
        int sc0, m0, sc1, m1;
        q5_scale_min(2 * j, hdr + 4, sc0, m0);
        q5_scale_min(2 * j + 1, hdr + 4, sc1, m1);
        const float dw   = pib_h2f(*reinterpret_cast<const uint16_t *>(hdr));
        const float dmin = pib_h2f(*reinterpret_cast<const uint16_t *>(hdr + 2));
#pragma unroll
        for (int t = 0; t < B; t++) {
            const int8_t * a  = aqs + (size_t) t * K + (size_t) sb * 256 + 64 * j + 8 * p;
            const int2     yl = *reinterpret_cast<const int2 *>(a), yh = *reinterpret_cast<const int2 *>(a + 32);
            const int      sum = sc0 * __dp4a(lo1, yl.y, __dp4a(lo0, yl.x, 0)) + sc1 * __dp4a(hi1, yh.y, __dp4a(hi0, yh.x, 0));
            const float    dx  = ad[(size_t) t * nsb + sb];
            acc[t] = fmaf(__fmul_rn(dx, dw), (float) sum, acc[t]);
            if (p == 0) {                                   // the min term of sub-blocks 2j, 2j+1: bsums are sums of 16
                const int16_t * bs = absum + (size_t) t * (K / 16) + (size_t) sb * 16 + 4 * j;
                const int mins = m0 * ((int) bs[0] + (int) bs[1]) + m1 * ((int) bs[2] + (int) bs[3]);
                acc[t] = fmaf(__fmul_rn(-dx, dmin), (float) mins, acc[t]);
            }
        }
    }
#pragma unroll
    for (int t = 0; t < B; t++) {
        const float v = warp_sum(acc[t]);
        if (lane == 0) y[(size_t) t * N + row] = v;
    }
}

template <int B>
static cudaError_t launch_q5(const pib_weight * w, const uint8_t * act, const pib_act_layout & L, int t0, float * y, const int * cancel,
                             cudaStream_t st) {
    const int       N = (int) w->N, K = (int) w->K;
    const int8_t *  aqs = (const int8_t *) (act + L.qs_off) + (size_t) t0 * K;
    const float *   ad  = (const float *) (act + L.d_off) + (size_t) t0 * (K / 256);
    const int16_t * bs  = (const int16_t *) (act + L.bs_off) + (size_t) t0 * (K / 16);
    return pib_launch(gemv_q5_K_kernel<B>, dim3((N + Q5_WARPS - 1) / Q5_WARPS), dim3(Q5_WARPS * 32), 0, st, w->qs, (const uint8_t *) w->d, aqs, ad,
                      bs, y + (size_t) t0 * N, N, K, cancel);
}

int pib_gemv_q5k(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!w || w->type != PIB_TYPE_Q5_K || !w->qs) { pib_set_error("pib_gemv: not a resident Q5_K matrix"); return PIB_ERR_UNSUPP; }
    const pib_act_layout L   = pib_act_layout_of(PIB_TYPE_Q4_K, w->K, b);          // Q8_K family activations
    const uint8_t *      act = (const uint8_t *) act_dev;
    for (int t0 = 0; t0 < b; t0 += 8) {
        const int   nb = b - t0 < 8 ? b - t0 : 8;
        cudaError_t e  = cudaSuccess;
        switch (nb) {
            case 1: e = launch_q5<1>(w, act, L, t0, y_dev, cancel, st); break;
            case 2: e = launch_q5<2>(w, act, L, t0, y_dev, cancel, st); break;
            case 3: e = launch_q5<3>(w, act, L, t0, y_dev, cancel, st); break;
            case 4: e = launch_q5<4>(w, act, L, t0, y_dev, cancel, st); break;
            case 5: e = launch_q5<5>(w, act, L, t0, y_dev, cancel, st); break;
            case 6: e = launch_q5<6>(w, act, L, t0, y_dev, cancel, st); break;
            case 7: e = launch_q5<7>(w, act, L, t0, y_dev, cancel, st); break;
            default: e = launch_q5<8>(w, act, L, t0, y_dev, cancel, st); break;
        }
        PIB_CHECK(e);
    }
    return PIB_OK;
}
