// gemv_q6k.cu -- dequant-GEMV for Q6_K weights (6.5625 bits per weight) at verification batch 1..16.
//
// Q4_K_M files -- what the paper's models are distributed as -- keep output.weight and part of attn_v / ffn_down in Q6_K
// (reference llama.cpp:7909-7957).  Arithmetic of the reference's ggml_vec_dot_q6_K_q8_K (ggml-quants.c:6594, AVX2 :6739):
// per 256-value super-block, sixteen groups of 16 values with an int8 scale each, q = 6-bit value - 32, exact integer sums
// sum_g scale_g * sum (q * y), scaled by fp16(d_w) * d_x in fp32.  Only the order of the fp32 super-block terms differs.
//
// Layout (weights.cu): per row and super-block 192 bytes {ql[128] low nibbles, qh[64] two high bits} in the reference's
// order, and a 32-byte header {int8 scales[16], fp16 d, pad}.  One warp per row; half a warp per super-block: lane =
// (half j of the super-block, 4 positions l = 4p..4p+3) assembles the four groups' bytes with SIMD-in-a-word logic and
// multiplies with dp4a.  This is the simple kernel of the type (HBM-bound, 30-40 % of peak): the lm_head pass of a Q4_K_M
// model is 215 MB, the layers stay on the tensor-core kernels of their own types.
#include "pib_common.cuh"

#define Q6_WARPS 4

template <int B>
__global__ void __launch_bounds__(Q6_WARPS * 32) gemv_q6_K_kernel(const uint8_t * __restrict__ wqs, const uint8_t * __restrict__ whdr,
                                                                  const int8_t * __restrict__ aqs, const float * __restrict__ ad,
                                                                  float * __restrict__ y, int N, int K, const int * __restrict__ cancel) {
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;
    const int lane = threadIdx.x & 31;
    const int row  = blockIdx.x * Q6_WARPS + (threadIdx.x >> 5);
    if (row >= N) return;
    const int nsb = K / 256;
    const int j = (lane >> 3) & 1, p = lane & 7;
    float acc[B];
#pragma unroll
    for (int t = 0; t < B; t++) acc[t] = 0.f;
    for (int sb = lane >> 4; sb < nsb; sb += 2) {
        const uint8_t * base = wqs + ((size_t) row * nsb + sb) * 192;
        const uint8_t * hdr  = whdr + ((size_t) row * nsb + sb) * 32;
        const uint32_t  w1 = *reinterpret_cast<const uint32_t *>(base + 64 * j + 4 * p);
        const uint32_t  w2 = *reinterpret_cast<const uint32_t *>(base + 64 * j + 32 + 4 * p);
        const uint32_t  h  = *reinterpret_cast<const uint32_t *>(base + 128 + 32 * j + 4 * p);
        int q[4];
        q[0] = (int) ((w1 & 0x0F0F0F0Fu) | ((h & 0x03030303u) << 4));
        q[1] = (int) ((w2 & 0x0F0F0F0Fu) | (((h >> 2) & 0x03030303u) << 4));
        q[2] = (int) (((w1 >> 4) & 0x0F0F0F0Fu) | (((h >> 4) & 0x03030303u) << 4));
        q[3] = (int) (((w2 >> 4) & 0x0F0F0F0Fu) | (((h >> 6) & 0x03030303u) << 4));
        int sc[4];
#pragma unroll
        for (int g = 0; g < 4; g++) sc[g] = (int) (int8_t) hdr[8 * j + 2 * g + (p >> 2)];
        const float dw = pib_h2f(*reinterpret_cast<const uint16_t *>(hdr + 16));
#pragma unroll
        for (int t = 0; t < B; t++) {
            const int8_t * a = aqs + (size_t) t * K + (size_t) sb * 256 + 128 * j + 4 * p;
            int sum = 0;
#pragma unroll
            for (int g = 0; g < 4; g++) {
                const int yv = *reinterpret_cast<const int *>(a + 32 * g);
                sum += sc[g] * (__dp4a(q[g], yv, 0) - 32 * __dp4a(0x01010101, yv, 0));
            }
            acc[t] = fmaf(__fmul_rn(ad[(size_t) t * nsb + sb], dw), (float) sum, acc[t]);
        }
    }
#pragma unroll
    for (int t = 0; t < B; t++) {
        const float v = warp_sum(acc[t]);
        if (lane == 0) y[(size_t) t * N + row] = v;
    }
}

template <int B>
static cudaError_t launch_q6(const pib_weight * w, const uint8_t * act, const pib_act_layout & L, int t0, float * y, const int * cancel,
                             cudaStream_t st) {
    const int      N = (int) w->N, K = (int) w->K;
    const int8_t * aqs = (const int8_t *) (act + L.qs_off) + (size_t) t0 * K;
    const float *  ad  = (const float *) (act + L.d_off) + (size_t) t0 * (K / 256);
    return pib_launch(gemv_q6_K_kernel<B>, dim3((N + Q6_WARPS - 1) / Q6_WARPS), dim3(Q6_WARPS * 32), 0, st, w->qs, (const uint8_t *) w->d, aqs, ad,
                      y + (size_t) t0 * N, N, K, cancel);
}

int pib_gemv_q6k(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!w || w->type != PIB_TYPE_Q6_K || !w->qs) { pib_set_error("pib_gemv: not a resident Q6_K matrix"); return PIB_ERR_UNSUPP; }
    // the activations are in the Q8_K family layout, the one Q4_K uses
    const pib_act_layout L   = pib_act_layout_of(PIB_TYPE_Q4_K, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    for (int t0 = 0; t0 < b; t0 += 8) {
        const int   nb = b - t0 < 8 ? b - t0 : 8;
        cudaError_t e  = cudaSuccess;
        switch (nb) {
            case 1: e = launch_q6<1>(w, act, L, t0, y_dev, cancel, st); break;
            case 2: e = launch_q6<2>(w, act, L, t0, y_dev, cancel, st); break;
            case 3: e = launch_q6<3>(w, act, L, t0, y_dev, cancel, st); break;
            case 4: e = launch_q6<4>(w, act, L, t0, y_dev, cancel, st); break;
            case 5: e = launch_q6<5>(w, act, L, t0, y_dev, cancel, st); break;
            case 6: e = launch_q6<6>(w, act, L, t0, y_dev, cancel, st); break;
            case 7: e = launch_q6<7>(w, act, L, t0, y_dev, cancel, st); break;
            default: e = launch_q6<8>(w, act, L, t0, y_dev, cancel, st); break;
        }
        PIB_CHECK(e);
    }
    return PIB_OK;
}
