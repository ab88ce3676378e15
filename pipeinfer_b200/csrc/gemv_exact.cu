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

// ---- the other K-quants (Q4_K_M / Q4_K_S / Q3_K_M files), same idea: the AVX2 lane accumulators of ggml_vec_dot_q{3,5,6}_K_q8_K restated
// literally from the plane forms (ggml-quants.c:4443 q3_K, :5773 q5_K, :6739 q6_K; oracle/pib_oracle.c po_vec_dot_q*_K_q8_K) ------------------
__device__ __forceinline__ void kq_scale_min_k4(int j, const uint8_t * q, int & sc, int & m) {
    if (j < 4) { sc = q[j] & 63; m = q[j + 4] & 63; }
    else { sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

template <int TYPE>
__global__ void __launch_bounds__(128) gemv_exact_kq_kernel(const uint8_t * __restrict__ wqs, const uint8_t * __restrict__ whdr, int N, int K,
                                                            const int8_t * __restrict__ aqs, const float * __restrict__ ad,
                                                            const int16_t * __restrict__ absum, float * __restrict__ y, const int * __restrict__ cancel) {
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;
    const int row = blockIdx.x * blockDim.x + threadIdx.x, tok = blockIdx.y;
    if (row >= N) return;
    const int nsb = K / 256;
    float acc[8] = { 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f };
    float summs = 0.f;
    for (int s = 0; s < nsb; s++) {
        const float    dx = ad[(size_t) tok * nsb + s];
        const int8_t * q8 = aqs + (size_t) tok * K + (size_t) s * 256;
        int sumi[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
        float d;
        if (TYPE == PIB_TYPE_Q6_K) {
            const uint8_t * ql = wqs + ((size_t) row * nsb + s) * 192, * qh = ql + 128;
            const uint8_t * hd = whdr + ((size_t) row * nsb + s) * 32;
            const int8_t *  sc = reinterpret_cast<const int8_t *>(hd);
            d = __fmul_rn(dx, pib_h2f(*reinterpret_cast<const uint16_t *>(hd + 16)));
            for (int j = 0; j < 2; j++) {
                for (int g = 0; g < 4; g++)
                    for (int m = 0; m < 8; m++) {
                        int dot = 0;
                        for (int t = 0; t < 4; t++) {
                            const int l  = 4 * m + t;
                            const int lo = (g & 1) ? ql[32 + l] : ql[l];
                            const int q  = ((g < 2) ? (lo & 0xF) : (lo >> 4)) | (((qh[l] >> (2 * g)) & 3) << 4);
                            dot += (q - 32) * (int) q8[32 * g + l];
                        }
                        sumi[m] += (int) sc[8 * j + 2 * g + (m >= 4)] * dot;
                    }
                ql += 64; qh += 32; q8 += 128;
            }
        } else if (TYPE == PIB_TYPE_Q5_K) {
            const uint8_t * q5 = wqs + ((size_t) row * nsb + s) * 160, * qh = q5 + 128;
            const uint8_t * hd = whdr + ((size_t) row * nsb + s) * 16;
            d = __fmul_rn(dx, pib_h2f(*reinterpret_cast<const uint16_t *>(hd)));
            const float dmin = __fmul_rn(-dx, pib_h2f(*reinterpret_cast<const uint16_t *>(hd + 2)));
            int sc[8], mn[8];
            for (int j = 0; j < 8; j++) kq_scale_min_k4(j, hd + 4, sc[j], mn[j]);
            const int16_t * bs = absum + ((size_t) tok * nsb + s) * 16;
            int prod = 0;
            for (int j = 0; j < 8; j++) prod += mn[j] * (int) (int16_t) (bs[2 * j] + bs[2 * j + 1]);
            summs = __fadd_rn(summs, __fmul_rn(dmin, (float) prod));
            for (int j = 0; j < 4; j++) {
                for (int l = 0; l < 8; l++) {
                    int lo = 0, hi = 0;
                    for (int t = 0; t < 4; t++) {
                        const int e  = 4 * l + t;
                        const int ql = (q5[e] & 0xF) + (((qh[e] >> (2 * j)) & 1) << 4);
                        const int qu = (q5[e] >> 4) + (((qh[e] >> (2 * j + 1)) & 1) << 4);
                        lo += ql * (int) q8[e];
                        hi += qu * (int) q8[32 + e];
                    }
                    sumi[l] += sc[2 * j] * lo + sc[2 * j + 1] * hi;
                }
                q5 += 32; q8 += 64;
            }
        } else {   // Q3_K
            const uint8_t * q3 = wqs + ((size_t) row * nsb + s) * 96, * hm = q3 + 64;
            const uint8_t * hd = whdr + ((size_t) row * nsb + s) * 16;
            d = __fmul_rn(dx, pib_h2f(*reinterpret_cast<const uint16_t *>(hd + 12)));
            // the sixteen 6-bit scales, minus 32 (ggml-quants.c:4459-4465)
            uint32_t aux[4];
            for (int i = 0; i < 3; i++) aux[i] = (uint32_t) hd[4 * i] | ((uint32_t) hd[4 * i + 1] << 8) | ((uint32_t) hd[4 * i + 2] << 16) | ((uint32_t) hd[4 * i + 3] << 24);
            const uint32_t kmask1 = 0x03030303u, kmask2 = 0x0f0f0f0fu, tmp = aux[2];
            aux[2] = ((aux[0] >> 4) & kmask2) | (((tmp >> 4) & kmask1) << 4);
            aux[3] = ((aux[1] >> 4) & kmask2) | (((tmp >> 6) & kmask1) << 4);
            aux[0] = (aux[0] & kmask2) | (((tmp >> 0) & kmask1) << 4);
            aux[1] = (aux[1] & kmask2) | (((tmp >> 2) & kmask1) << 4);
            for (int j = 0; j < 2; j++) {
                for (int g = 0; g < 4; g++)
                    for (int m = 0; m < 8; m++) {
                        int dot = 0;
                        for (int t = 0; t < 4; t++) {
                            const int l = 4 * m + t;
                            const int q = ((q3[l] >> (2 * g)) & 3) - ((hm[l] & (1 << (4 * j + g))) ? 0 : 4);
                            dot += q * (int) q8[32 * g + l];
                        }
                        const int is = 8 * j + 2 * g + (m >= 4);
                        sumi[m] += ((int) ((aux[is >> 2] >> (8 * (is & 3))) & 0xff) - 32) * dot;
                    }
                q3 += 32; q8 += 128;
            }
        }
#pragma unroll
        for (int l = 0; l < 8; l++) acc[l] = fmaf(d, (float) sumi[l], acc[l]);
    }
    const float r = hsum8_ref(acc);
    y[(size_t) tok * N + row] = TYPE == PIB_TYPE_Q5_K ? __fadd_rn(r, summs) : r;
}

int pib_gemv_exact_ref(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!w->qs || !w->d) { pib_set_error("pib_gemv: variant 4 needs the plane form of the weights (PIB_KEEP_PLANES=1 inside the engine)"); return PIB_ERR_UNSUPP; }
    const pib_act_layout L   = pib_act_layout_of(pib_act_family(w->type), w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    const dim3           grid((unsigned) ((w->N + 127) / 128), b);
    const int8_t *       aqs = (const int8_t *) (act + L.qs_off);
    const float *        ad  = (const float *) (act + L.d_off);
    const int16_t *      bs  = (const int16_t *) (act + L.bs_off);
    switch (w->type) {
        case PIB_TYPE_Q8_0: pib_launch(gemv_exact_ref_kernel<PIB_TYPE_Q8_0>, grid, dim3(128), 0, st, w->qs, w->d, (int) w->N, (int) w->K, aqs, ad, bs, y_dev, cancel); break;
        case PIB_TYPE_Q4_0: pib_launch(gemv_exact_ref_kernel<PIB_TYPE_Q4_0>, grid, dim3(128), 0, st, w->qs, w->d, (int) w->N, (int) w->K, aqs, ad, bs, y_dev, cancel); break;
        case PIB_TYPE_Q4_K: pib_launch(gemv_exact_ref_kernel<PIB_TYPE_Q4_K>, grid, dim3(128), 0, st, w->qs, w->d, (int) w->N, (int) w->K, aqs, ad, bs, y_dev, cancel); break;
        case PIB_TYPE_Q6_K: pib_launch(gemv_exact_kq_kernel<PIB_TYPE_Q6_K>, grid, dim3(128), 0, st, (const uint8_t *) w->qs, (const uint8_t *) w->d, (int) w->N, (int) w->K, aqs, ad, bs, y_dev, cancel); break;
        case PIB_TYPE_Q5_K: pib_launch(gemv_exact_kq_kernel<PIB_TYPE_Q5_K>, grid, dim3(128), 0, st, (const uint8_t *) w->qs, (const uint8_t *) w->d, (int) w->N, (int) w->K, aqs, ad, bs, y_dev, cancel); break;
        case PIB_TYPE_Q3_K: pib_launch(gemv_exact_kq_kernel<PIB_TYPE_Q3_K>, grid, dim3(128), 0, st, (const uint8_t *) w->qs, (const uint8_t *) w->d, (int) w->N, (int) w->K, aqs, ad, bs, y_dev, cancel); break;
        default: pib_set_error("pib_gemv: variant 4 supports Q4_0, Q8_0, Q3_K, Q4_K, Q5_K, Q6_K"); return PIB_ERR_UNSUPP;
    }
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}
