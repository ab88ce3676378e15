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
// multiplies with dp4a.  That was round 1's kernel (30-40 % of the HBM peak; kept as PIB_Q6K_TC=0); the default since round 2 is the
// tensor-core kernel on fragment-major records further down (0.90 of peak on the 215 MB lm_head of a 70B Q4_K_M model).
#include "pib_common.cuh"

#include <stdlib.h>

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

// ---- tensor-core kernel (default when the matrix has its tiles) ----------------------------------------------------------------------
// Same structure as the Q4_K ring kernel (gemv_mma.cu): the matrix is stored as one 3360-byte record per (16-row tile, super-block)
// -- exactly the 16 x 210 bytes of the reference blocks, regrouped so that a lane's registers ARE the A operand:
//   [0, 2048)     4 chunks x 32 lanes x 16 B   low nibbles; chunk c, lane (g, t): {W[2c][0], W[2c][1], W[2c+1][0], W[2c+1][1]}, where
//                                              W[p][r] holds, for row g + 8r, byte i = lo4(e = 32p + 4t + i) | lo4(e + 16) << 4:
//                                              the two 16-element scale groups 2p, 2p+1 of the pair share a word
//   [2048, 3072)  2 chunks x 32 lanes x 16 B   the two high bits; chunk h word i <-> pair p = 4h + i; in byte b, bits 2s, 2s+1 belong
//                                              to value b of set s: 0 = group 2p row g, 1 = 2p row g+8, 2 = 2p+1 row g, 3 = 2p+1 row g+8
//   [3072, 3328)  16 rows x 16 int8 scales     [3328, 3360)  16 x fp16 d
// One mma.m16n8k16 (u8 . s8) per scale group gives sum q.y for 16 rows x 8 tokens; the group's int8 scale is applied to the int32
// result.  The -32 offset (q6 = q - 32) is 32 * sum_j scale_j * bsum_j: the activations' 16-sums come split as 64 hi + lo (b16 plane),
// so it is two more MMAs with the SCALES as the A operand (k = the 16 groups).  Exact integer sums, fp32 fma(d_w * d_x, sum, acc) per
// super-block as the reference; only the order of the super-block terms differs.  Records are streamed through per-warp cp.async rings.
#define Q6T_BYTES  3360
#define Q6T_WARPS  4
#define Q6T_STAGES 4

__device__ __forceinline__ void q6_mma_u8s8_k16(int (&c)[4], int a0, int a1, int b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ void q6_mma_s8s8_k16(int (&c)[4], int a0, int a1, int b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ int q6_sbyte(uint32_t w, int k) { return (int) (int8_t) (w >> (8 * k)); }

// the 6-bit value of element e of a row's super-block in the plane form (reference order, dequantize_row_q6_K ggml-quants.c:2242)
__device__ __forceinline__ uint32_t q6_value(const uint8_t * blk192, int e) {
    const int n = e >> 7, ep = e & 127, l = ep & 31, quarter = ep >> 5;
    const uint8_t ql = blk192[64 * n + l + 32 * (quarter & 1)];
    const uint8_t qh = blk192[128 + 32 * n + l];
    return (uint32_t) ((quarter >= 2 ? ql >> 4 : ql & 0xF) | (((qh >> (2 * quarter)) & 3) << 4));
}

// one thread per (record, lane): the lane's 6 x 16 bytes of quants; lanes 0..17 also copy the scales / d block
__global__ void q6k_tiles_kernel(const uint8_t * __restrict__ qs, const uint8_t * __restrict__ hdr, int64_t N, int64_t K, uint8_t * __restrict__ tiles) {
    const int64_t nsb = K / 256, n_tiles = (N + 15) / 16;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tiles * nsb * 32) return;
    const int     lane = (int) (i & 31), g = lane >> 2, t = lane & 3;
    const int64_t rec = i >> 5, T = rec / nsb, sb = rec % nsb;
    uint8_t *     out = tiles + rec * Q6T_BYTES;
    const int64_t r0 = T * 16 + g, r1 = r0 + 8;
    const uint8_t * b0 = r0 < N ? qs + (r0 * nsb + sb) * 192 : nullptr;
    const uint8_t * b1 = r1 < N ? qs + (r1 * nsb + sb) * 192 : nullptr;
    uint32_t W[8][2], H[8];
    for (int p = 0; p < 8; p++) {
        H[p] = 0;
        for (int r = 0; r < 2; r++) {
            const uint8_t * blk = r ? b1 : b0;
            uint32_t w = 0;
            for (int k = 0; k < 4; k++) {
                const uint32_t qa = blk ? q6_value(blk, 32 * p + 4 * t + k) : 32u, qb = blk ? q6_value(blk, 32 * p + 16 + 4 * t + k) : 32u;   // padding rows: q - 32 = 0
                w |= ((qa & 0xF) | ((qb & 0xF) << 4)) << (8 * k);
                H[p] |= ((qa >> 4) << (2 * r) | (qb >> 4) << (2 * (2 + r))) << (8 * k);
            }
            W[p][r] = w;
        }
    }
    for (int c = 0; c < 4; c++) *reinterpret_cast<uint4 *>(out + c * 512 + lane * 16) = make_uint4(W[2 * c][0], W[2 * c][1], W[2 * c + 1][0], W[2 * c + 1][1]);
    for (int h = 0; h < 2; h++) *reinterpret_cast<uint4 *>(out + 2048 + h * 512 + lane * 16) = make_uint4(H[4 * h], H[4 * h + 1], H[4 * h + 2], H[4 * h + 3]);
    if (lane < 16) {
        const int64_t r = T * 16 + lane;
        uint4 sc = make_uint4(0, 0, 0, 0);
        uint16_t d = 0;
        if (r < N) { sc = *reinterpret_cast<const uint4 *>(hdr + (r * nsb + sb) * 32); d = *reinterpret_cast<const uint16_t *>(hdr + (r * nsb + sb) * 32 + 16); }
        *reinterpret_cast<uint4 *>(out + 3072 + lane * 16) = sc;
        *reinterpret_cast<uint16_t *>(out + 3328 + lane * 2) = d;
    }
}

int pib_weight_build_tiles_q6k(pib_weight * w) {
    if (w->type != PIB_TYPE_Q6_K || w->tiles) return PIB_OK;
    const int64_t nsb = w->K / 256, n_tiles = (w->N + 15) / 16;
    const size_t  bytes = (size_t) n_tiles * nsb * Q6T_BYTES;
    if (cudaMalloc(&w->tiles, bytes) != cudaSuccess) { cudaGetLastError(); pib_set_error("q6_K tiles: cudaMalloc of %zu bytes failed", bytes); w->tiles = nullptr; return PIB_ERR_NOMEM; }
    w->tile_bytes = bytes;
    const int64_t n = n_tiles * nsb * 32;
    q6k_tiles_kernel<<<(unsigned) ((n + 127) / 128), 128>>>(w->qs, (const uint8_t *) w->d, w->N, w->K, w->tiles);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

template <int NT>
__global__ void __launch_bounds__(Q6T_WARPS * 32, NT == 1 ? 4 : 3)
gemv_q6k_tc_kernel(const uint8_t * __restrict__ tiles, int nsb, int N, int K, const int8_t * __restrict__ aq6, const float * __restrict__ ad,
                   const int8_t * __restrict__ ab16, int b, float * __restrict__ y, const int * __restrict__ cancel) {
    if (PIB_CANCELLED(cancel)) { pib_pdl_wait(); return; }
    extern __shared__ __align__(16) uint8_t q6_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int T = blockIdx.x;
    const uint8_t * tbase  = tiles + (size_t) T * nsb * Q6T_BYTES + lane * 16;
    uint8_t *       ring   = q6_smem + warp * (Q6T_STAGES * Q6T_BYTES);
    const uint32_t  ring_a = (uint32_t) __cvta_generic_to_shared(ring) + lane * 16;
    int ps = warp;
    auto issue = [&](int stage) {
        if (ps < nsb) {
            const uint8_t * src = tbase + (size_t) ps * Q6T_BYTES;
            const uint32_t  dst = ring_a + stage * Q6T_BYTES;
#pragma unroll
            for (int c = 0; c < 6; c++) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + c * 512), "l"(src + c * 512) : "memory");
            if (lane < 18) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 3072), "l"(src + 3072) : "memory");
            ps += Q6T_WARPS;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int i = 0; i < Q6T_STAGES - 1; i++) issue(i);        // weights are immutable: before the wait on the producer of the activations

    float acc[NT][4];
#pragma unroll
    for (int n = 0; n < NT; n++) { acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f; }
    const int8_t * a_q[NT];
    const int8_t * a_b[NT];
    const float *  a_dx[NT][2];
#pragma unroll
    for (int n = 0; n < NT; n++) {
        const int tok = min(n * 8 + g, b - 1);
        a_q[n] = aq6 + (size_t) tok * K + t * 64;
        a_b[n] = ab16 + (size_t) tok * nsb * 32 + 4 * t;
#pragma unroll
        for (int e = 0; e < 2; e++) a_dx[n][e] = ad + (size_t) min(n * 8 + 2 * t + e, b - 1) * nsb;
    }
    pib_pdl_wait();
    pib_pdl_trigger();
    int stage = 0;
    for (int s = warp; s < nsb; s += Q6T_WARPS) {
        __syncwarp();
        issue((stage + Q6T_STAGES - 1) % Q6T_STAGES);
        asm volatile("cp.async.wait_group %0;" ::"n"(Q6T_STAGES - 1) : "memory");
        __syncwarp();
        const uint8_t * rs = ring + stage * Q6T_BYTES;
        uint4 lo[4], hi[2];
#pragma unroll
        for (int c = 0; c < 4; c++) lo[c] = *reinterpret_cast<const uint4 *>(rs + c * 512 + lane * 16);
#pragma unroll
        for (int h = 0; h < 2; h++) hi[h] = *reinterpret_cast<const uint4 *>(rs + 2048 + h * 512 + lane * 16);
        const uint4 sc0 = *reinterpret_cast<const uint4 *>(rs + 3072 + g * 16), sc1 = *reinterpret_cast<const uint4 *>(rs + 3072 + (g + 8) * 16);
        const float d0 = pib_h2f(*reinterpret_cast<const uint16_t *>(rs + 3328 + g * 2)), d1 = pib_h2f(*reinterpret_cast<const uint16_t *>(rs + 3328 + (g + 8) * 2));
        const uint32_t s0w[4] = { sc0.x, sc0.y, sc0.z, sc0.w }, s1w[4] = { sc1.x, sc1.y, sc1.z, sc1.w };
        const uint32_t s0t = t == 0 ? sc0.x : t == 1 ? sc0.y : t == 2 ? sc0.z : sc0.w, s1t = t == 0 ? sc1.x : t == 1 ? sc1.y : t == 2 ? sc1.z : sc1.w;
        const uint32_t low[4][4] = { { lo[0].x, lo[0].y, lo[0].z, lo[0].w }, { lo[1].x, lo[1].y, lo[1].z, lo[1].w }, { lo[2].x, lo[2].y, lo[2].z, lo[2].w }, { lo[3].x, lo[3].y, lo[3].z, lo[3].w } };
        const uint32_t hiw[8] = { hi[0].x, hi[0].y, hi[0].z, hi[0].w, hi[1].x, hi[1].y, hi[1].z, hi[1].w };
#pragma unroll
        for (int n = 0; n < NT; n++) {
            const int4 * ap = reinterpret_cast<const int4 *>(a_q[n] + (size_t) s * 256);
            int4 bq[4];
#pragma unroll
            for (int c = 0; c < 4; c++) bq[c] = __ldg(ap + c);
            const int bw[16] = { bq[0].x, bq[0].y, bq[0].z, bq[0].w, bq[1].x, bq[1].y, bq[1].z, bq[1].w, bq[2].x, bq[2].y, bq[2].z, bq[2].w, bq[3].x, bq[3].y, bq[3].z, bq[3].w };
            const int bh = __ldg(reinterpret_cast<const int *>(a_b[n] + (size_t) s * 32)), bl = __ldg(reinterpret_cast<const int *>(a_b[n] + (size_t) s * 32 + 16));
            const float dx0 = __ldg(a_dx[n][0] + s), dx1 = __ldg(a_dx[n][1] + s);
            int sum[4] = { 0, 0, 0, 0 };
#pragma unroll
            for (int p = 0; p < 8; p++) {
                const uint32_t W0 = low[p >> 1][2 * (p & 1)], W1 = low[p >> 1][2 * (p & 1) + 1], H = hiw[p];
                int cA[4] = { 0, 0, 0, 0 }, cB[4] = { 0, 0, 0, 0 };
                q6_mma_u8s8_k16(cA, (int) ((W0 & 0x0F0F0F0Fu) | ((H << 4) & 0x30303030u)), (int) ((W1 & 0x0F0F0F0Fu) | ((H << 2) & 0x30303030u)), bw[2 * p]);
                q6_mma_u8s8_k16(cB, (int) (((W0 >> 4) & 0x0F0F0F0Fu) | (H & 0x30303030u)), (int) (((W1 >> 4) & 0x0F0F0F0Fu) | ((H >> 2) & 0x30303030u)), bw[2 * p + 1]);
                const int jA = 2 * p, jB = 2 * p + 1;
                const int a0 = q6_sbyte(s0w[jA >> 2], jA & 3), a1 = q6_sbyte(s1w[jA >> 2], jA & 3), c0 = q6_sbyte(s0w[jB >> 2], jB & 3), c1 = q6_sbyte(s1w[jB >> 2], jB & 3);
                sum[0] += a0 * cA[0] + c0 * cB[0]; sum[1] += a0 * cA[1] + c0 * cB[1];
                sum[2] += a1 * cA[2] + c1 * cB[2]; sum[3] += a1 * cA[3] + c1 * cB[3];
            }
            // the -32 offset: 32 * sum_j scale_j * bsum_j, bsum = 64 hi + lo; A = this lane's four scales of rows g, g + 8
            int ch[4] = { 0, 0, 0, 0 }, cl[4] = { 0, 0, 0, 0 };
            q6_mma_s8s8_k16(ch, (int) s0t, (int) s1t, bh);
            q6_mma_s8s8_k16(cl, (int) s0t, (int) s1t, bl);
            acc[n][0] = fmaf(__fmul_rn(dx0, d0), (float) (sum[0] - 32 * (64 * ch[0] + cl[0])), acc[n][0]);
            acc[n][1] = fmaf(__fmul_rn(dx1, d0), (float) (sum[1] - 32 * (64 * ch[1] + cl[1])), acc[n][1]);
            acc[n][2] = fmaf(__fmul_rn(dx0, d1), (float) (sum[2] - 32 * (64 * ch[2] + cl[2])), acc[n][2]);
            acc[n][3] = fmaf(__fmul_rn(dx1, d1), (float) (sum[3] - 32 * (64 * ch[3] + cl[3])), acc[n][3]);
        }
        stage = stage + 1 == Q6T_STAGES ? 0 : stage + 1;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                                            // the rings are done with: their memory becomes the reduction buffer
    float (*red)[NT * 8][16] = reinterpret_cast<float (*)[NT * 8][16]>(q6_smem);
#pragma unroll
    for (int n = 0; n < NT; n++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
            red[warp][n * 8 + 2 * t + e][g]     = acc[n][e];
            red[warp][n * 8 + 2 * t + e][g + 8] = acc[n][2 + e];
        }
    __syncthreads();
    for (int i = threadIdx.x; i < NT * 8 * 16; i += Q6T_WARPS * 32) {
        const int tok = i / 16, r = i % 16, row = T * 16 + r;
        if (tok < b && row < N) {
            float v = red[0][tok][r];
#pragma unroll
            for (int w2 = 1; w2 < Q6T_WARPS; w2++) v += red[w2][tok][r];
            y[(size_t) tok * N + row] = v;
        }
    }
}

template <int NT>
static int launch_q6_tc(const pib_weight * w, const uint8_t * act, const pib_act_layout & L, int b, float * y, const int * cancel, cudaStream_t st) {
    const size_t smem = (size_t) Q6T_WARPS * Q6T_STAGES * Q6T_BYTES;
    static bool configured = false;
    if (!configured) {
        PIB_CHECK(cudaFuncSetAttribute(gemv_q6k_tc_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        PIB_CHECK(cudaFuncSetAttribute(gemv_q6k_tc_kernel<NT>, cudaFuncAttributePreferredSharedMemoryCarveout, (int) cudaSharedmemCarveoutMaxShared));
        configured = true;
    }
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    PIB_CHECK(pib_launch(gemv_q6k_tc_kernel<NT>, dim3(n_tiles), dim3(Q6T_WARPS * 32), smem, st, w->tiles, nsb, (int) w->N, (int) w->K,
                         (const int8_t *) (act + L.q6p_off), (const float *) (act + L.d_off), (const int8_t *) (act + L.b16_off), b, y, cancel));
    return PIB_OK;
}

int pib_gemv_q6k(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!w || w->type != PIB_TYPE_Q6_K || (!w->qs && !w->tiles)) { pib_set_error("pib_gemv: not a resident Q6_K matrix"); return PIB_ERR_UNSUPP; }
    // the activations are in the Q8_K family layout, the one Q4_K uses
    const pib_act_layout L   = pib_act_layout_of(PIB_TYPE_Q4_K, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    static const int use_tc = getenv("PIB_Q6K_TC") ? atoi(getenv("PIB_Q6K_TC")) : 1;
    if (w->tiles && (use_tc || !w->qs) && b <= PIB_MAX_BATCH)
        return b <= 8 ? launch_q6_tc<1>(w, act, L, b, y_dev, cancel, st) : launch_q6_tc<2>(w, act, L, b, y_dev, cancel, st);
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

// embedding rows (token_embd of Q6_K files): dequantize_row_q6_K (ggml-quants.c:1994-2040), y = (d * scale) * q, one thread per element
__global__ void q6k_get_rows_kernel(const uint8_t * __restrict__ qs, const uint8_t * __restrict__ hdr, int64_t K, const int32_t * __restrict__ ids,
                                    float * __restrict__ out) {
    const int64_t row = ids[blockIdx.x], nsb = K / 256;
    for (int64_t e = threadIdx.x; e < K; e += blockDim.x) {
        const int64_t sb = e >> 8;
        const int     ee = (int) (e & 255), n = ee >> 7, ep = ee & 127, quarter = ep >> 5, l = ep & 31;
        const uint8_t * blk = qs + (row * nsb + sb) * 192;
        const uint8_t * h   = hdr + (row * nsb + sb) * 32;
        const int q = (int) (int8_t) q6_value(blk, ee) - 32;
        const float d  = pib_h2f(*reinterpret_cast<const uint16_t *>(h + 16));
        const float sc = (float) reinterpret_cast<const int8_t *>(h)[8 * n + (l >> 4) + 2 * quarter];
        out[(size_t) blockIdx.x * K + e] = __fmul_rn(__fmul_rn(d, sc), (float) q);
    }
}
int pib_get_rows_q6k(const pib_weight * w, const int32_t * ids_dev, int n, float * out_dev, cudaStream_t st) {
    if (!w->qs) { pib_set_error("pib_get_rows: this matrix keeps only its tensor-core tiles"); return PIB_ERR_UNSUPP; }
    PIB_CHECK(pib_launch(q6k_get_rows_kernel, dim3(n), dim3(256), 0, st, (const uint8_t *) w->qs, (const uint8_t *) w->d, w->K, ids_dev, out_dev));
    return PIB_OK;
}
