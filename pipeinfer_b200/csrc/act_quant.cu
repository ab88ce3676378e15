// act_quant.cu -- activation quantisation (K2), fused with the producer that precedes each matmul (K4 / K8).
//
// The reference's CPU matmul first quantises every activation row to the weight type's "vec_dot type" (reference
// ggml.c:9493-9509): Q8_0 blocks for Q4_0 / Q8_0 weights (ggml-quants.c:632; on x86 the AVX2 branch :701-760:
// d = max|x| / 127 stored as fp16, q = round-half-even(x * (127 / max|x|))) and Q8_K super-blocks for Q4_K weights
// (ggml-quants.c:2329-2363: iscale = -128 / (signed value of the first max-|x| element), q = min(127, rne(iscale*x)),
// d = 1 / iscale, 16-wide bsums).  Greedy parity needs exactly this integer arithmetic, so it is reproduced bit for
// bit; only the layout differs (planes instead of blocks, see pib_act_layout in pib_common.cuh).
//
// One CTA per token row.  Prologues (template PRO):
//   PRO_PLAIN     v = x
//   PRO_RMSNORM   v = (x * 1/sqrtf(mean(x^2) + eps)) * w        (ggml.c:9028 with the double accumulator, then ggml_mul)
//   PRO_SILU_MUL  v = silu_f16lut(x) * x2                        (ggml.c:1490 table look-up, then ggml_mul; llama.cpp:3680)
//   PRO_GELU      v = gelu_f16lut(x)                             (ggml.c:1433; falcon MLP)
#include "pib_common.cuh"
#include "q8k_quant.cuh"

enum { PRO_PLAIN = 0, PRO_RMSNORM = 1, PRO_SILU_MUL = 2, PRO_GELU = 3, PRO_LAYERNORM = 4 };

#define AQ_THREADS 256

template <int PRO>
__device__ __forceinline__ float aq_value(const float * xr, const float * __restrict__ x2r, float scale,
                                          const uint16_t * __restrict__ lut, int64_t i, float mean = 0.f,
                                          const float * __restrict__ biasr = nullptr) {
    const float x = xr[i];
    if (PRO == PRO_RMSNORM) return __fmul_rn(__fmul_rn(x, scale), x2r[i]);
    // reference norm_f32 (ggml.c:8959-9020): y = (x - mean) * scale, then ggml_mul by the weight, ggml_add of the bias
    if (PRO == PRO_LAYERNORM) return __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(x, mean), scale), x2r[i]), biasr[i]);
    if (PRO == PRO_SILU_MUL) return __fmul_rn(pib_h2f(lut[pib_f2h(x)]), x2r[i]);
    if (PRO == PRO_GELU) return pib_h2f(lut[pib_f2h(x)]);
    return x;
}

// eight consecutive values (32-byte aligned) with two 128-bit loads per operand
template <int PRO>
__device__ __forceinline__ void aq_value8(const float * xr, const float * __restrict__ x2r, float scale,
                                          const uint16_t * __restrict__ lut, int64_t i0, float * v, float mean = 0.f,
                                          const float * __restrict__ biasr = nullptr) {
    const float4 xa = *reinterpret_cast<const float4 *>(xr + i0), xb = *reinterpret_cast<const float4 *>(xr + i0 + 4);
    const float  x[8] = { xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w };
    float        w[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    if (PRO == PRO_RMSNORM || PRO == PRO_SILU_MUL || PRO == PRO_LAYERNORM) {
        const float4 wa = *reinterpret_cast<const float4 *>(x2r + i0), wb = *reinterpret_cast<const float4 *>(x2r + i0 + 4);
        w[0] = wa.x; w[1] = wa.y; w[2] = wa.z; w[3] = wa.w; w[4] = wb.x; w[5] = wb.y; w[6] = wb.z; w[7] = wb.w;
    }
#pragma unroll
    for (int j = 0; j < 8; j++) {
        if (PRO == PRO_RMSNORM)       v[j] = __fmul_rn(__fmul_rn(x[j], scale), w[j]);
        else if (PRO == PRO_LAYERNORM) v[j] = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(x[j], mean), scale), w[j]), biasr[i0 + j]);
        else if (PRO == PRO_SILU_MUL) v[j] = __fmul_rn(pib_h2f(lut[pib_f2h(x[j])]), w[j]);
        else if (PRO == PRO_GELU)     v[j] = pib_h2f(lut[pib_f2h(x[j])]);
        else                          v[j] = x[j];
    }
}

template <bool Q8K, int PRO>
__global__ void __launch_bounds__(AQ_THREADS) act_quant_kernel(const float * x, const float * __restrict__ x2,
                                                               int x2_per_row, float eps, int64_t K,
                                                               int8_t * __restrict__ qs, float * __restrict__ dplane,
                                                               void * __restrict__ bsplane, int8_t * __restrict__ bs8plane,
                                                               int8_t * __restrict__ qsp, int2 * __restrict__ qsl,
                                                               int16_t * __restrict__ sbs, int8_t * __restrict__ q6p, int8_t * __restrict__ b16,
                                                               float * __restrict__ vout,
                                                               const uint16_t * __restrict__ lut,
                                                               const float * __restrict__ addend, float * sum_out,
                                                               const int * __restrict__ cancel, int64_t xstride,
                                                               const float * __restrict__ bias, const float * __restrict__ addend2) {
    const int     row  = blockIdx.x;
    const float * xr   = x + (size_t) row * xstride;
    const float * x2r  = x2 ? (x2_per_row ? x2 + (size_t) row * xstride : x2) : nullptr;
    const int     lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int     nwarp = AQ_THREADS / 32;
    // RMSNorm into Q8_K with one super-block per warp (every llama shape here: K / 256 == 8 * gridDim.y): the lane's eight norm
    // weights are immutable, so they are fetched before the wait on the producer of x, and the lane's eight x values of pass 2
    // are kept from pass 1 instead of being read again
    const bool own = Q8K && PRO == PRO_RMSNORM && K / PIB_QK_K == (int64_t) nwarp * gridDim.y && !x2_per_row;
    float wk[8] = { 0, 0, 0, 0, 0, 0, 0, 0 }, keep[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    if (own) {
        const float4 * w4 = reinterpret_cast<const float4 *>(x2r + ((int64_t) blockIdx.y * nwarp + warp) * PIB_QK_K + lane * 8);
        const float4 wa = w4[0], wb = w4[1];
        wk[0] = wa.x; wk[1] = wa.y; wk[2] = wa.z; wk[3] = wa.w; wk[4] = wb.x; wk[5] = wb.y; wk[6] = wb.z; wk[7] = wb.w;
    }
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;

    float scale = 1.0f, mean = 0.0f;
    if (own) {
        __shared__ double red1[AQ_THREADS / 32];
        double s = 0.0;
        const float4 * __restrict__ x4 = reinterpret_cast<const float4 *>(xr);
        const float4 * __restrict__ a4 = addend ? reinterpret_cast<const float4 *>(addend + (size_t) row * K) : nullptr;
        float4 *                    o4 = sum_out ? reinterpret_cast<float4 *>(sum_out + (size_t) row * K) : nullptr;
        // warp w reads super-blocks w, w + 8, ... (a lane: the same eight elements of each); the one of index blockIdx.y is its own
#pragma unroll 4
        for (int j = 0; j < (int) gridDim.y; j++) {
            const int i = (j * nwarp + warp) * (PIB_QK_K / 4) + lane * 2;
            float4 v0 = x4[i], v1 = x4[i + 1];
            if (a4) {
                const float4 a0 = a4[i], a1 = a4[i + 1];
                v0.x = __fadd_rn(v0.x, a0.x); v0.y = __fadd_rn(v0.y, a0.y); v0.z = __fadd_rn(v0.z, a0.z); v0.w = __fadd_rn(v0.w, a0.w);
                v1.x = __fadd_rn(v1.x, a1.x); v1.y = __fadd_rn(v1.y, a1.y); v1.z = __fadd_rn(v1.z, a1.z); v1.w = __fadd_rn(v1.w, a1.w);
                if (j == (int) blockIdx.y) { o4[i] = v0; o4[i + 1] = v1; }        // every CTA of the row writes its own quarter of the sums
            }
            if (j == (int) blockIdx.y) { keep[0] = v0.x; keep[1] = v0.y; keep[2] = v0.z; keep[3] = v0.w; keep[4] = v1.x; keep[5] = v1.y; keep[6] = v1.z; keep[7] = v1.w; }
            s += (double) __fmul_rn(v0.x, v0.x); s += (double) __fmul_rn(v0.y, v0.y); s += (double) __fmul_rn(v0.z, v0.z); s += (double) __fmul_rn(v0.w, v0.w);
            s += (double) __fmul_rn(v1.x, v1.x); s += (double) __fmul_rn(v1.y, v1.y); s += (double) __fmul_rn(v1.z, v1.z); s += (double) __fmul_rn(v1.w, v1.w);
        }
        s = warp_sum_d(s);
        if (lane == 0) red1[warp] = s;
        __syncthreads();
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < nwarp; w++) t += red1[w];                           // every thread forms the same total
        const float ms = (float) (t / (double) K);
        scale          = 1.0f / sqrtf(ms + eps);
    } else if (PRO == PRO_RMSNORM || PRO == PRO_LAYERNORM) {
        __shared__ double red[AQ_THREADS / 32];
        __shared__ float  s_scale;
        double            s = 0.0;
        // optional residual add fused in front (reference ggml_add, llama.cpp:3944 / :3962): sum_out = x + addend.
        // (with several CTAs per row each one recomputes the whole row's sum -- K floats from L2 -- and all of them
        //  write the same sums, so pass 2 never depends on another CTA).  128-bit loads, four in flight per thread.
        const float4 * __restrict__ x4 = reinterpret_cast<const float4 *>(xr);
        const float4 * __restrict__ a4 = addend ? reinterpret_cast<const float4 *>(addend + (size_t) row * K) : nullptr;
        const float4 * __restrict__ b4 = addend2 ? reinterpret_cast<const float4 *>(addend2 + (size_t) row * K) : nullptr;   // (x + addend) + addend2
        float4 *                    o4 = sum_out ? reinterpret_cast<float4 *>(sum_out + (size_t) row * K) : nullptr;
        const int n4 = (int) (K / 4);
#pragma unroll 4
        for (int i = threadIdx.x; i < n4; i += AQ_THREADS) {
            float4 v = x4[i];
            if (a4) {
                const float4 a = a4[i];
                v.x = __fadd_rn(v.x, a.x); v.y = __fadd_rn(v.y, a.y); v.z = __fadd_rn(v.z, a.z); v.w = __fadd_rn(v.w, a.w);
                if (b4) {
                    const float4 c2 = b4[i];
                    v.x = __fadd_rn(v.x, c2.x); v.y = __fadd_rn(v.y, c2.y); v.z = __fadd_rn(v.z, c2.z); v.w = __fadd_rn(v.w, c2.w);
                }
                o4[i] = v;
            }
            if (PRO == PRO_LAYERNORM) {
                s += (double) v.x; s += (double) v.y; s += (double) v.z; s += (double) v.w;
            } else {
                s += (double) __fmul_rn(v.x, v.x);
                s += (double) __fmul_rn(v.y, v.y);
                s += (double) __fmul_rn(v.z, v.z);
                s += (double) __fmul_rn(v.w, v.w);
            }
        }
        if (a4) xr = sum_out + (size_t) row * K;     // pass 2 re-reads the sums (visible CTA-wide after the barrier below)
        s = warp_sum_d(s);
        if (lane == 0) red[warp] = s;
        __syncthreads();
        if (PRO == PRO_RMSNORM) {
            if (threadIdx.x == 0) {
                double t = 0.0;
                for (int w = 0; w < nwarp; w++) t += red[w];
                const float ms = (float) (t / (double) K);
                s_scale        = 1.0f / sqrtf(ms + eps);
            }
            __syncthreads();
            scale = s_scale;
        } else {
            // LayerNorm (reference norm_f32, ggml.c:8959-9020): mean from the double sum, then the double sum of the float
            // squares of (x - mean), scale = 1 / sqrtf(variance + eps)
            __shared__ float s_mean;
            if (threadIdx.x == 0) {
                double t = 0.0;
                for (int w = 0; w < nwarp; w++) t += red[w];
                s_mean = (float) (t / (double) K);
            }
            __syncthreads();
            mean = s_mean;
            double s2 = 0.0;
            const float4 * v4 = reinterpret_cast<const float4 *>(xr);
#pragma unroll 4
            for (int i = threadIdx.x; i < n4; i += AQ_THREADS) {
                const float4 v = v4[i];
                const float  d0 = __fsub_rn(v.x, mean), d1 = __fsub_rn(v.y, mean), d2 = __fsub_rn(v.z, mean), d3 = __fsub_rn(v.w, mean);
                s2 += (double) __fmul_rn(d0, d0); s2 += (double) __fmul_rn(d1, d1); s2 += (double) __fmul_rn(d2, d2); s2 += (double) __fmul_rn(d3, d3);
            }
            s2 = warp_sum_d(s2);
            __syncthreads();
            if (lane == 0) red[warp] = s2;
            __syncthreads();
            if (threadIdx.x == 0) {
                double t = 0.0;
                for (int w = 0; w < nwarp; w++) t += red[w];
                const float var = (float) (t / (double) K);
                s_scale         = 1.0f / sqrtf(var + eps);
            }
            __syncthreads();
            scale = s_scale;
        }
    }
    const float * biasr = bias;

    if (!Q8K) {
        // one warp per 32-element block, lane == element.  Blocks nb .. nbp - 1 are the zero padding of the last group of
        // eight (K not a multiple of 256): only their fragment-order copies exist.
        const int64_t nb = K / PIB_QK, nbp = ((K + 255) / 256) * 8;
        __half *      dxp = reinterpret_cast<__half *>(bs8plane);
        for (int64_t blk = (int64_t) blockIdx.y * nwarp + warp; blk < nbp; blk += (int64_t) gridDim.y * nwarp) {
            int   q = 0;
            float d = 0.0f;
            if (blk < nb) {
                const int64_t i    = blk * PIB_QK + lane;
                const float   v    = aq_value<PRO>(xr, x2r, scale, lut, i, mean, biasr);
                const float   amax = warp_max(fabsf(v));
                d                  = amax / 127.f;
                const float   id   = (amax != 0.0f) ? 127.f / amax : 0.0f;
                q                  = __float2int_rn(__fmul_rn(v, id));
                q                  = max(-128, min(127, q));
                qs[(size_t) row * K + i] = (int8_t) q;
                if (vout) vout[(size_t) row * K + i] = v;
                int sum = q;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                if (lane == 0) {
                    dplane[(size_t) row * nb + blk]            = pib_h2f(pib_f2h(d));   // the reference stores d as fp16
                    ((int32_t *) bsplane)[(size_t) row * nb + blk] = sum;
                }
            }
            // fragment-order copy for the tensor-core GEMV (gemv_blk32.cu): element e = 16 h + 4 t + i of block j of a group
            // of eight goes to 64 t + 8 j + 4 h + i, so that a lane's B operands of a whole group are 64 contiguous bytes
            qsp[(size_t) row * nbp * 32 + (blk >> 3) * 256 + ((lane & 15) >> 2) * 64 + (blk & 7) * 8 + (lane >> 4) * 4 + (lane & 3)] = (int8_t) q;
            if (lane == 0) dxp[(size_t) row * nbp + blk] = __float2half_rn(d);
        }
    } else {
        // one warp per 256-element super-block, lane owns 8 consecutive elements
        const int64_t nsb = K / PIB_QK_K;
        for (int64_t sb = (int64_t) blockIdx.y * nwarp + warp; sb < nsb; sb += (int64_t) gridDim.y * nwarp) {
            const int64_t i0 = sb * PIB_QK_K + lane * 8;
            float v[8];
            if (own) {
#pragma unroll
                for (int j = 0; j < 8; j++) v[j] = __fmul_rn(__fmul_rn(keep[j], scale), wk[j]);
            } else {
                aq_value8<PRO>(xr, x2r, scale, lut, i0, v, mean, biasr);
            }
            if (vout) {
                float4 * vo = reinterpret_cast<float4 *>(vout + (size_t) row * K + i0);
                vo[0] = make_float4(v[0], v[1], v[2], v[3]);
                vo[1] = make_float4(v[4], v[5], v[6], v[7]);
            }
            const size_t rs = (size_t) row * nsb + sb;
            q8k_quantise_sb(v, lane, qs + (size_t) row * K + sb * PIB_QK_K, qsp + (size_t) row * K + sb * PIB_QK_K, (int16_t *) bsplane + rs * 16, bs8plane + rs * 16,
                            sbs + rs * 8, qsl + rs * 64, dplane + rs, q6p + (size_t) row * K + sb * PIB_QK_K, b16 + rs * 32);
        }
    }
}

template <int PRO>
static int launch_aq(int wtype, const float * x, const float * x2, int x2_per_row, float eps, int b, int64_t K, void * act,
                     float * vout, const uint16_t * lut, cudaStream_t st, const float * addend = nullptr,
                     float * sum_out = nullptr, const int * cancel = nullptr, int64_t xstride = 0, const float * bias = nullptr,
                     const float * addend2 = nullptr) {
    if (xstride == 0) xstride = K;
    const pib_act_layout L = pib_act_layout_of(wtype, K, b);
    int8_t * qs = (int8_t *) act + L.qs_off;
    float *  d  = (float *) ((uint8_t *) act + L.d_off);
    void *   bs = (uint8_t *) act + L.bs_off;
    int8_t * b8 = (int8_t *) act + L.bs8_off;
    int8_t * qp = (int8_t *) act + L.qsp_off;
    int2 *    ql = (int2 *) ((uint8_t *) act + L.qsl_off);
    int16_t * sb = (int16_t *) ((uint8_t *) act + L.sbs_off);
    int8_t *  q6 = (int8_t *) act + L.q6p_off;
    int8_t *  b6 = (int8_t *) act + L.b16_off;
    const int64_t units = wtype == PIB_TYPE_Q4_K ? K / PIB_QK_K : K / PIB_QK / 4;        // super-blocks, or groups of 4 blocks
    const int     parts = (int) max((int64_t) 1, min((int64_t) 32, units / (AQ_THREADS / 32)));
    const dim3    grid(b, parts);
    if (wtype == PIB_TYPE_Q4_K) pib_launch(act_quant_kernel<true, PRO>, grid, dim3(AQ_THREADS), 0, st, x, x2, x2_per_row, eps, K, qs, d, bs, b8, qp, ql, sb, q6, b6, vout, lut, addend, sum_out, cancel, xstride, bias, addend2);
    else                        pib_launch(act_quant_kernel<false, PRO>, grid, dim3(AQ_THREADS), 0, st, x, x2, x2_per_row, eps, K, qs, d, bs, b8, qp, ql, sb, q6, b6, vout, lut, addend, sum_out, cancel, xstride, bias, addend2);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

static int check_args(int wtype, int b, int64_t K) {
    const int blk = wtype == PIB_TYPE_Q4_K ? PIB_QK_K : PIB_QK;
    if ((wtype != PIB_TYPE_Q4_0 && wtype != PIB_TYPE_Q8_0 && wtype != PIB_TYPE_Q4_K) || b < 1 || K < blk || K % blk != 0) {
        pib_set_error("act quantise: bad arguments (type %d, b %d, K %lld)", wtype, b, (long long) K);
        return PIB_ERR_ARG;
    }
    return PIB_OK;
}

extern "C" size_t pib_act_bytes(int wtype, int64_t K, int b) { return pib_act_layout_of(wtype, K, b).total; }

extern "C" int pib_quantize_act(int wtype, const float * x_dev, int b, int64_t K, void * act_dev, pib_stream s) {
    if (check_args(wtype, b, K) != PIB_OK) return PIB_ERR_ARG;
    return launch_aq<PRO_PLAIN>(wtype, x_dev, nullptr, 0, 0.f, b, K, act_dev, nullptr, nullptr, (cudaStream_t) s);
}

extern "C" int pib_norm_quantize_act(int wtype, int norm_kind, const float * x_dev, const float * norm_w_dev, float eps, int b,
                                     int64_t K, void * act_dev, float * xnorm_dev, pib_stream s) {
    if (check_args(wtype, b, K) != PIB_OK) return PIB_ERR_ARG;
    if (norm_kind != 1) { pib_set_error("pib_norm_quantize_act: only RMSNorm (1) is implemented"); return PIB_ERR_UNSUPP; }
    return launch_aq<PRO_RMSNORM>(wtype, x_dev, norm_w_dev, 0, eps, b, K, act_dev, xnorm_dev, nullptr, (cudaStream_t) s);
}

// ---- internal (engine) entry points ---------------------------------------------------------------------------------
// silu(gate) * up -> quantised activations for the down projection; gate_up: [b][2*K] rows = {gate[K], up[K]}
int pib_silu_mul_quantize_act(int wtype, const float * gate_dev, const float * up_dev, int64_t row_stride, int b, int64_t K,
                              void * act_dev, const int * cancel, cudaStream_t st) {
    if (check_args(wtype, b, K) != PIB_OK) return PIB_ERR_ARG;
    return launch_aq<PRO_SILU_MUL>(wtype, gate_dev, up_dev, 1, 0.f, b, K, act_dev, nullptr, pib_get_luts().silu_f16, st,
                                   nullptr, nullptr, cancel, row_stride);
}

int pib_quantize_act_c(int wtype, const float * x_dev, int b, int64_t K, void * act_dev, const int * cancel, cudaStream_t st) {
    if (check_args(wtype, b, K) != PIB_OK) return PIB_ERR_ARG;
    return launch_aq<PRO_PLAIN>(wtype, x_dev, nullptr, 0, 0.f, b, K, act_dev, nullptr, nullptr, st, nullptr, nullptr, cancel);
}

// [sum_out = x + addend;] act = quantise(rmsnorm(sum or x) * w)
int pib_add_rmsnorm_quantize_act(int wtype, const float * x_dev, const float * addend_dev, float * sum_out_dev,
                                 const float * norm_w_dev, float eps, int b, int64_t K, void * act_dev, const int * cancel,
                                 cudaStream_t st) {
    if (check_args(wtype, b, K) != PIB_OK) return PIB_ERR_ARG;
    return launch_aq<PRO_RMSNORM>(wtype, x_dev, norm_w_dev, 0, eps, b, K, act_dev, nullptr, nullptr, st, addend_dev, sum_out_dev,
                                  cancel);
}

// falcon: [sum_out = (x + addend) + addend2;] act = quantise(layernorm(sum or x) * w + bias)
int pib_add_layernorm_quantize_act(int wtype, const float * x_dev, const float * addend_dev, const float * addend2_dev,
                                   float * sum_out_dev, const float * norm_w_dev, const float * norm_b_dev, float eps, int b,
                                   int64_t K, void * act_dev, const int * cancel, cudaStream_t st) {
    if (check_args(wtype, b, K) != PIB_OK) return PIB_ERR_ARG;
    return launch_aq<PRO_LAYERNORM>(wtype, x_dev, norm_w_dev, 0, eps, b, K, act_dev, nullptr, nullptr, st, addend_dev, sum_out_dev,
                                    cancel, 0, norm_b_dev, addend2_dev);
}

// falcon: act = quantise(gelu(x)) through the reference's fp16 table (ggml.c:8676)
int pib_gelu_quantize_act(int wtype, const float * x_dev, int b, int64_t K, void * act_dev, const int * cancel, cudaStream_t st) {
    if (check_args(wtype, b, K) != PIB_OK) return PIB_ERR_ARG;
    return launch_aq<PRO_GELU>(wtype, x_dev, nullptr, 0, 0.f, b, K, act_dev, nullptr, pib_get_luts().gelu_f16, st, nullptr, nullptr, cancel);
}

extern "C" int pib_act_to_blocks(int wtype, const void * act_dev, int b, int64_t K, void * out) {
    if (check_args(wtype, b, K) != PIB_OK) return PIB_ERR_ARG;
    const pib_act_layout L = pib_act_layout_of(wtype, K, b);
    uint8_t * h = (uint8_t *) malloc(L.total);
    cudaError_t e = cudaMemcpy(h, act_dev, L.total, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { free(h); PIB_CHECK(e); }
    uint8_t * o = (uint8_t *) out;
    if (wtype == PIB_TYPE_Q4_K) {
        const int64_t nsb = K / 256;
        for (int t = 0; t < b; t++)
            for (int64_t s = 0; s < nsb; s++) {
                uint8_t * blk = o + ((size_t) t * nsb + s) * 292;
                memcpy(blk, h + L.d_off + ((size_t) t * nsb + s) * 4, 4);
                memcpy(blk + 4, h + L.qs_off + (size_t) t * K + s * 256, 256);
                memcpy(blk + 260, h + L.bs_off + ((size_t) t * (K / 16) + s * 16) * 2, 32);
            }
    } else {
        const int64_t nb = K / 32;
        for (int t = 0; t < b; t++)
            for (int64_t s = 0; s < nb; s++) {
                uint8_t * blk = o + ((size_t) t * nb + s) * 34;
                float     d;
                memcpy(&d, h + L.d_off + ((size_t) t * nb + s) * 4, 4);
                const __half hd = __float2half_rn(d);
                memcpy(blk, &hd, 2);
                memcpy(blk + 2, h + L.qs_off + (size_t) t * K + s * 32, 32);
            }
    }
    free(h);
    return PIB_OK;
}
