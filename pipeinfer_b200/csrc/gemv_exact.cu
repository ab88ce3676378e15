// gemv_exact.cu -- variant 4 of the quantised matmul (K1): the reference's summation ORDER restated literally, one thread per
// (output row, token), from the plane form of the weights.  Slow (no tensor cores, no coalescing effort); it exists to pin
// down what "the same bits as the CPU" means for the fast kernels: the reference's AVX2 dot products keep eight fp32 lane
// accumulators per output -- lane l collects, block after block, fma(d_block, float(sum of the four products of elements
// 4l .. 4l+3), acc[l]) -- and add the lanes at the end as ((a0+a4)+(a2+a6)) + ((a1+a5)+(a3+a7))
// (ggml-quants.c:67-73 hsum_float_8, :2500-2530 q4_0, :3587-3607 q8_0, :5318-5381 q4_K with its 4-lane acc_m for the mins).
// A kernel that forms the same integers and applies the same fmas in the same order returns the same bits.
#include "pib_common.cuh"

__device__ __forceinline__ float hsum8_ref(const float (&a)[8]) {
    const float r0 = __fadd_rn(a[4], a[0]), r1 = __fadd_rn(a[5], a[1]), r2 = __fadd_rn(a[6], a[2]), r3 = __fadd_rn(a[7], a[3]);
    return __fadd_rn(__fadd_rn(r0, r2), __fadd_rn(r1, r3));
}

template <int TYPE>
__global__ void __launch_bounds__(128) gemv_exact_ref_kernel(const uint8_t * __restrict__ wqs, const uint16_t * __restrict__ wd, int N, int K,
                                                             const int8_t * __restrict__ aqs, const float * __restrict__ ad,
                                                             const int16_t * __restrict__ absum, float * __restrict__ y,
                                                             const int * __restrict__ cancel) {
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;
    const int row = blockIdx.x * blockDim.x + threadIdx.x, tok = blockIdx.y;
    if (row >= N) return;
    float acc[8] = { 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f };
    if (TYPE == PIB_TYPE_Q8_0 || TYPE == PIB_TYPE_Q4_0) {
        const int nb = K / 32;
        for (int blk = 0; blk < nb; blk++) {
            const float   d = __fmul_rn(pib_h2f(wd[(size_t) row * nb + blk]), ad[(size_t) tok * nb + blk]);
            const int *   a = reinterpret_cast<const int *>(aqs + (size_t) tok * K + (size_t) blk * 32);
            if (TYPE == PIB_TYPE_Q8_0) {
                const int * w = reinterpret_cast<const int *>(wqs + (size_t) row * K + (size_t) blk * 32);
#pragma unroll
                for (int l = 0; l < 8; l++) acc[l] = fmaf(d, (float) __dp4a(w[l], a[l], 0), acc[l]);
            } else {
                // byte j of the block: low nibble = element j, high nibble = element j + 16 (bytes_from_nibbles_32, :108-114)
                const uint32_t * w = reinterpret_cast<const uint32_t *>(wqs + (size_t) row * (K / 2) + (size_t) blk * 16);
#pragma unroll
                for (int l = 0; l < 4; l++) {
                    const uint32_t v  = w[l];
                    const int      lo = (int) ((((v & 0x0F0F0F0Fu) | 0x80808080u) - 0x08080808u) ^ 0x80808080u);         // q - 8 per byte
                    const int      hi = (int) (((((v >> 4) & 0x0F0F0F0Fu) | 0x80808080u) - 0x08080808u) ^ 0x80808080u);
                    acc[l]     = fmaf(d, (float) __dp4a(lo, a[l], 0), acc[l]);
                    acc[l + 4] = fmaf(d, (float) __dp4a(hi, a[l + 4], 0), acc[l + 4]);
                }
            }
        }
        y[(size_t) tok * N + row] = hsum8_ref(acc);
        return;
    }
    // Q4_K x Q8_K
    const int nsb = K / 256;
    float accm[4] = { 0.f, 0.f, 0.f, 0.f };
    for (int s = 0; s < nsb; s++) {
        const uint4    h  = *reinterpret_cast<const uint4 *>(reinterpret_cast<const uint8_t *>(wd) + ((size_t) row * nsb + s) * 16);
        const float    dx = ad[(size_t) tok * nsb + s];
        const float    d  = __fmul_rn(dx, pib_h2f((uint16_t) (h.x & 0xffff)));
        const float    dm = __fmul_rn(-dx, pib_h2f((uint16_t) (h.x >> 16)));
        // get_scale_min_k4 for the eight sub-blocks (ggml-quants.c:1446)
        const uint32_t sc_lo = h.y & 0x3f3f3f3fu, m_lo = h.z & 0x3f3f3f3fu;
        const uint32_t sc_hi = (h.w & 0x0f0f0f0fu) | (((h.y >> 6) & 0x03030303u) << 4);
        const uint32_t m_hi  = ((h.w >> 4) & 0x0f0f0f0fu) | (((h.z >> 6) & 0x03030303u) << 4);
        const int16_t * bs = absum + ((size_t) tok * nsb + s) * 16;
#pragma unroll
        for (int l = 0; l < 4; l++) {
            const int q0 = (int16_t) (bs[4 * l + 0] + bs[4 * l + 1]), q1 = (int16_t) (bs[4 * l + 2] + bs[4 * l + 3]);
            const uint32_t mm = l < 2 ? m_lo : m_hi;
            const int m0 = (int) ((mm >> (16 * (l & 1))) & 0xff), m1 = (int) ((mm >> (16 * (l & 1) + 8)) & 0xff);
            accm[l] = fmaf(dm, (float) (m0 * q0 + m1 * q1), accm[l]);
        }
        int sumi[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
        const uint32_t * q4 = reinterpret_cast<const uint32_t *>(wqs + (size_t) row * (K / 2) + (size_t) s * 128);
        const int *      q8 = reinterpret_cast<const int *>(aqs + (size_t) tok * K + (size_t) s * 256);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t scw = j < 2 ? sc_lo : sc_hi;
            const int s0 = (int) ((scw >> (16 * (j & 1))) & 0xff), s1 = (int) ((scw >> (16 * (j & 1) + 8)) & 0xff);
#pragma unroll
            for (int l = 0; l < 8; l++) {
                const uint32_t v = q4[j * 8 + l];
                sumi[l] += s0 * __dp4a((int) (v & 0x0F0F0F0Fu), q8[j * 16 + l], 0) + s1 * __dp4a((int) ((v >> 4) & 0x0F0F0F0Fu), q8[j * 16 + 8 + l], 0);
            }
        }
#pragma unroll
        for (int l = 0; l < 8; l++) acc[l] = fmaf(d, (float) sumi[l], acc[l]);
    }
    y[(size_t) tok * N + row] = __fadd_rn(hsum8_ref(acc), __fadd_rn(__fadd_rn(accm[0], accm[2]), __fadd_rn(accm[1], accm[3])));
}

int pib_gemv_exact_ref(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!w->qs || !w->d) { pib_set_error("pib_gemv: variant 4 needs the plane form of the weights (PIB_KEEP_PLANES=1 inside the engine)"); return PIB_ERR_UNSUPP; }
    const pib_act_layout L   = pib_act_layout_of(w->type, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    const dim3           grid((unsigned) ((w->N + 127) / 128), b);
    const int8_t *       aqs = (const int8_t *) (act + L.qs_off);
    const float *        ad  = (const float *) (act + L.d_off);
    const int16_t *      bs  = (const int16_t *) (act + L.bs_off);
    switch (w->type) {
        case PIB_TYPE_Q8_0: pib_launch(gemv_exact_ref_kernel<PIB_TYPE_Q8_0>, grid, dim3(128), 0, st, w->qs, w->d, (int) w->N, (int) w->K, aqs, ad, bs, y_dev, cancel); break;
        case PIB_TYPE_Q4_0: pib_launch(gemv_exact_ref_kernel<PIB_TYPE_Q4_0>, grid, dim3(128), 0, st, w->qs, w->d, (int) w->N, (int) w->K, aqs, ad, bs, y_dev, cancel); break;
        case PIB_TYPE_Q4_K: pib_launch(gemv_exact_ref_kernel<PIB_TYPE_Q4_K>, grid, dim3(128), 0, st, w->qs, w->d, (int) w->N, (int) w->K, aqs, ad, bs, y_dev, cancel); break;
        default: pib_set_error("pib_gemv: variant 4 supports Q4_0, Q8_0, Q4_K"); return PIB_ERR_UNSUPP;
    }
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}
