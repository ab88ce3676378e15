// gemv_q5k.cu -- dequant-GEMV for Q5_K weights (5.5 bits per weight) at verification batch 1..16.
//
// Q4_K_M files of models with 4 or more query heads per KV head (the 70B class) keep attn_v in Q5_K (reference
// llama.cpp:7909-7933).  Arithmetic of ggml_vec_dot_q5_K_q8_K (ggml-quants.c:5680, AVX2 path): as Q4_K with a fifth bit per
// value -- per super-block eight sub-blocks of 32 with 6-bit scale and min, exact integer sums
// sum_j sc_j * sum(q * y) and sum_j m_j * bsum_j, scaled by fp16(d) * d_x and -fp16(dmin) * d_x in fp32.
//
// Layout (weights.cu): per row and super-block 160 bytes {qs[128] low nibbles, qh[32] fifth bits} and the 16-byte header
// {d, dmin, scales[12]} of Q4_K.  One warp per row, half a warp per super-block: lane = (64-value chunk j, eight byte
// positions 8p..8p+7): two sub-blocks' worth of bytes assembled with SIMD-in-a-word logic, dp4a.  That was round 1's kernel (kept as
// PIB_Q5K_TC=0); the default since round 2 is the tensor-core kernel on fragment-major records further down.
#include "pib_common.cuh"

#include <stdlib.h>

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

// ---- tensor-core kernel (default when the matrix has its tiles) ----------------------------------------------------------------------
// The Q4_K ring kernel (gemv_mma.cu) with a fifth bit: a record is the Q4_K record of the (16-row tile, super-block) -- 2048 bytes of
// nibbles in mma.m16n8k32 A-fragment order + 16 row headers -- followed by 32 lanes x 16 bytes of fifth bits, 2816 bytes = the 16 x 176
// of the reference blocks.  Lane (g, t) word c holds the bits of chunk c: in byte k, bit s belongs to value k of set
// s = 4 sub + 2 lhalf + rowsel (sub: low / high-nibble sub-block 2c / 2c+1; lhalf: chunk positions 8t.. / 8t+4..; rowsel: row g / g+8),
// so every A register is (nibbles & 0x0f0f0f0f) | ((word >> s) & 0x01010101) << 4.  Integer sums exact, fp32 terms per super-block.
#define Q5T_BYTES  2816
#define Q5T_WARPS  4
#define Q5T_STAGES 4

__device__ __forceinline__ void q5_mma_u8s8(int (&c)[4], int a0, int a1, int a2, int a3, int b0, int b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void q5_mma_s8s8(int (&c)[4], int a0, int a1, int a2, int a3, int b0, int b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// one thread per (record, lane)
__global__ void q5k_tiles_kernel(const uint8_t * __restrict__ qs, const uint8_t * __restrict__ hdr, int64_t N, int64_t K, uint8_t * __restrict__ tiles) {
    const int64_t nsb = K / 256, n_tiles = (N + 15) / 16;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tiles * nsb * 32) return;
    const int     lane = (int) (i & 31), g = lane >> 2, t = lane & 3;
    const int64_t rec = i >> 5, T = rec / nsb, S = rec % nsb;
    uint8_t *     out = tiles + rec * Q5T_BYTES;
    const int64_t r0 = T * 16 + g, r1 = r0 + 8;
    const uint8_t * b0 = r0 < N ? qs + (r0 * nsb + S) * 160 : nullptr;
    const uint8_t * b1 = r1 < N ? qs + (r1 * nsb + S) * 160 : nullptr;
    uint32_t H[4];
    for (int c = 0; c < 4; c++) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (b0) { const uint2 a = *reinterpret_cast<const uint2 *>(b0 + c * 32 + t * 8); v.x = a.x; v.y = a.y; }
        if (b1) { const uint2 a = *reinterpret_cast<const uint2 *>(b1 + c * 32 + t * 8); v.z = a.x; v.w = a.y; }
        *reinterpret_cast<uint4 *>(out + c * 512 + lane * 16) = v;
        uint32_t h = 0;
        for (int r = 0; r < 2; r++) {
            const uint8_t * blk = r ? b1 : b0;
            if (!blk) continue;
            for (int lh = 0; lh < 2; lh++)
                for (int k = 0; k < 4; k++) {
                    const uint32_t q = blk[128 + 8 * t + 4 * lh + k];                  // qh[l]: bit j = fifth bit of element 32 j + l
                    h |= ((q >> (2 * c)) & 1u) << (8 * k + 2 * lh + r);
                    h |= ((q >> (2 * c + 1)) & 1u) << (8 * k + 4 + 2 * lh + r);
                }
        }
        H[c] = h;
    }
    *reinterpret_cast<uint4 *>(out + 2304 + lane * 16) = make_uint4(H[0], H[1], H[2], H[3]);
    if (lane < 16) {
        const int64_t r = T * 16 + lane;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (r < N) v = *reinterpret_cast<const uint4 *>(hdr + (r * nsb + S) * 16);     // rows beyond N: d = dmin = 0
        *reinterpret_cast<uint4 *>(out + 2048 + lane * 16) = v;
    }
}

int pib_weight_build_tiles_q5k(pib_weight * w) {
    if (w->type != PIB_TYPE_Q5_K || w->tiles) return PIB_OK;
    const int64_t nsb = w->K / 256, n_tiles = (w->N + 15) / 16;
    const size_t  bytes = (size_t) n_tiles * nsb * Q5T_BYTES;
    if (cudaMalloc(&w->tiles, bytes) != cudaSuccess) { cudaGetLastError(); pib_set_error("q5_K tiles: cudaMalloc of %zu bytes failed", bytes); w->tiles = nullptr; return PIB_ERR_NOMEM; }
    w->tile_bytes = bytes;
    const int64_t n = n_tiles * nsb * 32;
    q5k_tiles_kernel<<<(unsigned) ((n + 127) / 128), 128>>>(w->qs, (const uint8_t *) w->d, w->N, w->K, w->tiles);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

template <int NT>
__global__ void __launch_bounds__(Q5T_WARPS * 32, NT == 1 ? 4 : 3)
gemv_q5k_tc_kernel(const uint8_t * __restrict__ tiles, int nsb, int N, int K, const int8_t * __restrict__ aqs, const float * __restrict__ ad,
                   const int8_t * __restrict__ abs8, int b, float * __restrict__ y, const int * __restrict__ cancel) {
    if (PIB_CANCELLED(cancel)) { pib_pdl_wait(); return; }
    extern __shared__ __align__(16) uint8_t q5_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int T = blockIdx.x;
    const uint8_t * tbase  = tiles + (size_t) T * nsb * Q5T_BYTES + lane * 16;
    uint8_t *       ring   = q5_smem + warp * (Q5T_STAGES * Q5T_BYTES);
    const uint32_t  ring_a = (uint32_t) __cvta_generic_to_shared(ring) + lane * 16;
    int ps = warp;
    auto issue = [&](int stage) {
        if (ps < nsb) {
            const uint8_t * src = tbase + (size_t) ps * Q5T_BYTES;
            const uint32_t  dst = ring_a + stage * Q5T_BYTES;
#pragma unroll
            for (int c = 0; c < 4; c++) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + c * 512), "l"(src + c * 512) : "memory");
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 2304), "l"(src + 2304) : "memory");
            if (lane < 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 2048), "l"(src + 2048) : "memory");
            ps += Q5T_WARPS;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int i = 0; i < Q5T_STAGES - 1; i++) issue(i);        // weights are immutable: before the wait on the producer of the activations

    float acc[NT][4];
#pragma unroll
    for (int n = 0; n < NT; n++) { acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f; }
    const int8_t * a_qs[NT];
    const int8_t * a_b8[NT];
    const float *  a_dx[NT][2];
#pragma unroll
    for (int n = 0; n < NT; n++) {
        const int tok = min(n * 8 + g, b - 1);
        a_qs[n] = aqs + (size_t) tok * K + t * 64;
        a_b8[n] = abs8 + (size_t) tok * nsb * 16 + (t & 1) * 4;
#pragma unroll
        for (int e = 0; e < 2; e++) a_dx[n][e] = ad + (size_t) min(n * 8 + 2 * t + e, b - 1) * nsb;
    }
    pib_pdl_wait();
    pib_pdl_trigger();
    int stage = 0;
    for (int s = warp; s < nsb; s += Q5T_WARPS) {
        __syncwarp();
        issue((stage + Q5T_STAGES - 1) % Q5T_STAGES);
        asm volatile("cp.async.wait_group %0;" ::"n"(Q5T_STAGES - 1) : "memory");
        __syncwarp();
        const uint8_t * rs = ring + stage * Q5T_BYTES;
        uint4 wq[4];
#pragma unroll
        for (int c = 0; c < 4; c++) wq[c] = *reinterpret_cast<const uint4 *>(rs + c * 512 + lane * 16);
        const uint4 hb = *reinterpret_cast<const uint4 *>(rs + 2304 + lane * 16);
        const uint4 h0 = *reinterpret_cast<const uint4 *>(rs + 2048 + g * 16), h1 = *reinterpret_cast<const uint4 *>(rs + 2048 + (g + 8) * 16);
        // headers of rows g and g + 8: get_scale_min_k4 (ggml-quants.c:1446) for all eight sub-blocks at once
        const float    d0 = pib_h2f((uint16_t) (h0.x & 0xffff)), dm0 = pib_h2f((uint16_t) (h0.x >> 16));
        const float    d1 = pib_h2f((uint16_t) (h1.x & 0xffff)), dm1 = pib_h2f((uint16_t) (h1.x >> 16));
        const uint32_t sc0[2] = { h0.y & 0x3f3f3f3fu, (h0.w & 0x0f0f0f0fu) | (((h0.y >> 6) & 0x03030303u) << 4) };
        const uint32_t sc1[2] = { h1.y & 0x3f3f3f3fu, (h1.w & 0x0f0f0f0fu) | (((h1.y >> 6) & 0x03030303u) << 4) };
        const uint32_t m0[2]  = { h0.z & 0x3f3f3f3fu, ((h0.w >> 4) & 0x0f0f0f0fu) | (((h0.z >> 6) & 0x03030303u) << 4) };
        const uint32_t m1[2]  = { h1.z & 0x3f3f3f3fu, ((h1.w >> 4) & 0x0f0f0f0fu) | (((h1.z >> 6) & 0x03030303u) << 4) };
        const uint32_t hw[4]  = { hb.x, hb.y, hb.z, hb.w };
#pragma unroll
        for (int n = 0; n < NT; n++) {
            const int4 * ap = reinterpret_cast<const int4 *>(a_qs[n] + (size_t) s * 256);
            const int *  bp = reinterpret_cast<const int *>(a_b8[n] + (size_t) s * 16);
            const int    amh = __ldg(bp), aml = __ldg(bp + 2);
            const float  dx0 = __ldg(a_dx[n][0] + s), dx1 = __ldg(a_dx[n][1] + s);
            int sum[4] = { 0, 0, 0, 0 };
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int4     aq = __ldg(ap + c);
                const uint32_t wx = wq[c].x, wy = wq[c].y, wz = wq[c].z, ww = wq[c].w, H = hw[c];
                int c_lo[4] = { 0, 0, 0, 0 }, c_hi[4] = { 0, 0, 0, 0 };
                q5_mma_u8s8(c_lo, (int) ((wx & 0x0F0F0F0Fu) | ((H << 4) & 0x10101010u)), (int) ((wz & 0x0F0F0F0Fu) | ((H << 3) & 0x10101010u)),
                            (int) ((wy & 0x0F0F0F0Fu) | ((H << 2) & 0x10101010u)), (int) ((ww & 0x0F0F0F0Fu) | ((H << 1) & 0x10101010u)), aq.x, aq.y);
                q5_mma_u8s8(c_hi, (int) (((wx >> 4) & 0x0F0F0F0Fu) | (H & 0x10101010u)), (int) (((wz >> 4) & 0x0F0F0F0Fu) | ((H >> 1) & 0x10101010u)),
                            (int) (((wy >> 4) & 0x0F0F0F0Fu) | ((H >> 2) & 0x10101010u)), (int) (((ww >> 4) & 0x0F0F0F0Fu) | ((H >> 3) & 0x10101010u)), aq.z, aq.w);
                const int j0 = 2 * c, j1 = 2 * c + 1;
                const int s00 = (int) ((sc0[j0 >> 2] >> (8 * (j0 & 3))) & 0xff), s01 = (int) ((sc0[j1 >> 2] >> (8 * (j1 & 3))) & 0xff);
                const int s10 = (int) ((sc1[j0 >> 2] >> (8 * (j0 & 3))) & 0xff), s11 = (int) ((sc1[j1 >> 2] >> (8 * (j1 & 3))) & 0xff);
                sum[0] += s00 * c_lo[0] + s01 * c_hi[0]; sum[1] += s00 * c_lo[1] + s01 * c_hi[1];
                sum[2] += s10 * c_lo[2] + s11 * c_hi[2]; sum[3] += s10 * c_lo[3] + s11 * c_hi[3];
            }
            // the min term: sum_j m_j * bsum_j with the sub-block sums split as 64 hi + lo (bs8 plane), mins as the A operand
            const int am0 = t == 0 ? (int) m0[0] : t == 1 ? (int) m0[1] : 0, am1 = t == 0 ? (int) m1[0] : t == 1 ? (int) m1[1] : 0;
            const int bmh = t < 2 ? amh : 0, bml = t < 2 ? aml : 0;
            int c_mh[4] = { 0, 0, 0, 0 }, c_ml[4] = { 0, 0, 0, 0 };
            q5_mma_s8s8(c_mh, am0, am1, 0, 0, bmh, 0);
            q5_mma_s8s8(c_ml, am0, am1, 0, 0, bml, 0);
            acc[n][0] = fmaf(__fmul_rn(dx0, d0), (float) sum[0], acc[n][0]);  acc[n][0] = fmaf(__fmul_rn(-dx0, dm0), (float) (64 * c_mh[0] + c_ml[0]), acc[n][0]);
            acc[n][1] = fmaf(__fmul_rn(dx1, d0), (float) sum[1], acc[n][1]);  acc[n][1] = fmaf(__fmul_rn(-dx1, dm0), (float) (64 * c_mh[1] + c_ml[1]), acc[n][1]);
            acc[n][2] = fmaf(__fmul_rn(dx0, d1), (float) sum[2], acc[n][2]);  acc[n][2] = fmaf(__fmul_rn(-dx0, dm1), (float) (64 * c_mh[2] + c_ml[2]), acc[n][2]);
            acc[n][3] = fmaf(__fmul_rn(dx1, d1), (float) sum[3], acc[n][3]);  acc[n][3] = fmaf(__fmul_rn(-dx1, dm1), (float) (64 * c_mh[3] + c_ml[3]), acc[n][3]);
        }
        stage = stage + 1 == Q5T_STAGES ? 0 : stage + 1;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                                            // the rings are done with: their memory becomes the reduction buffer
    float (*red)[NT * 8][16] = reinterpret_cast<float (*)[NT * 8][16]>(q5_smem);
#pragma unroll
    for (int n = 0; n < NT; n++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
            red[warp][n * 8 + 2 * t + e][g]     = acc[n][e];
            red[warp][n * 8 + 2 * t + e][g + 8] = acc[n][2 + e];
        }
    __syncthreads();
    for (int i = threadIdx.x; i < NT * 8 * 16; i += Q5T_WARPS * 32) {
        const int tok = i / 16, r = i % 16, row = T * 16 + r;
        if (tok < b && row < N) {
            float v = red[0][tok][r];
#pragma unroll
            for (int w2 = 1; w2 < Q5T_WARPS; w2++) v += red[w2][tok][r];
            y[(size_t) tok * N + row] = v;
        }
    }
}

template <int NT>
static int launch_q5_tc(const pib_weight * w, const uint8_t * act, const pib_act_layout & L, int b, float * y, const int * cancel, cudaStream_t st) {
    const size_t smem = (size_t) Q5T_WARPS * Q5T_STAGES * Q5T_BYTES;
    static bool configured = false;
    if (!configured) {
        PIB_CHECK(cudaFuncSetAttribute(gemv_q5k_tc_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        PIB_CHECK(cudaFuncSetAttribute(gemv_q5k_tc_kernel<NT>, cudaFuncAttributePreferredSharedMemoryCarveout, (int) cudaSharedmemCarveoutMaxShared));
        configured = true;
    }
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    PIB_CHECK(pib_launch(gemv_q5k_tc_kernel<NT>, dim3(n_tiles), dim3(Q5T_WARPS * 32), smem, st, w->tiles, nsb, (int) w->N, (int) w->K,
                         (const int8_t *) (act + L.qsp_off), (const float *) (act + L.d_off), (const int8_t *) (act + L.bs8_off), b, y, cancel));
    return PIB_OK;
}

int pib_gemv_q5k(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!w || w->type != PIB_TYPE_Q5_K || (!w->qs && !w->tiles)) { pib_set_error("pib_gemv: not a resident Q5_K matrix"); return PIB_ERR_UNSUPP; }
    const pib_act_layout L   = pib_act_layout_of(PIB_TYPE_Q4_K, w->K, b);          // Q8_K family activations
    const uint8_t *      act = (const uint8_t *) act_dev;
    static const int use_tc = getenv("PIB_Q5K_TC") ? atoi(getenv("PIB_Q5K_TC")) : 1;
    if (w->tiles && (use_tc || !w->qs) && b <= PIB_MAX_BATCH)
        return b <= 8 ? launch_q5_tc<1>(w, act, L, b, y_dev, cancel, st) : launch_q5_tc<2>(w, act, L, b, y_dev, cancel, st);
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

// embedding rows (token_embd of Q5_K_S / Q5_K_M files): dequantize_row_q5_K (ggml-quants.c:1930-1960), y = (d * sc) * q - min * m
__global__ void q5k_get_rows_kernel(const uint8_t * __restrict__ qs, const uint8_t * __restrict__ hdr, int64_t K, const int32_t * __restrict__ ids,
                                    float * __restrict__ out) {
    const int64_t row = ids[blockIdx.x], nsb = K / 256;
    for (int64_t e = threadIdx.x; e < K; e += blockDim.x) {
        const int64_t sb = e >> 8;
        const int     ee = (int) (e & 255), j = ee >> 6, r = ee & 63, l = r & 31, hi = r >> 5;
        const uint8_t * blk = qs + (row * nsb + sb) * 160;
        const uint8_t * h   = hdr + (row * nsb + sb) * 16;
        const float d = pib_h2f(*reinterpret_cast<const uint16_t *>(h)), mn = pib_h2f(*reinterpret_cast<const uint16_t *>(h + 2));
        int sc, m;
        q5_scale_min(2 * j + hi, h + 4, sc, m);
        const int q = (hi ? (blk[32 * j + l] >> 4) : (blk[32 * j + l] & 0xF)) + (((blk[128 + l] >> (2 * j + hi)) & 1) ? 16 : 0);
        const float d1 = __fmul_rn(d, (float) sc), m1 = __fmul_rn(mn, (float) m);
        out[(size_t) blockIdx.x * K + e] = __fsub_rn(__fmul_rn(d1, (float) q), m1);
    }
}
int pib_get_rows_q5k(const pib_weight * w, const int32_t * ids_dev, int n, float * out_dev, cudaStream_t st) {
    if (!w->qs) { pib_set_error("pib_get_rows: this matrix keeps only its tensor-core tiles"); return PIB_ERR_UNSUPP; }
    PIB_CHECK(pib_launch(q5k_get_rows_kernel, dim3(n), dim3(256), 0, st, (const uint8_t *) w->qs, (const uint8_t *) w->d, w->K, ids_dev, out_dev));
    return PIB_OK;
}
