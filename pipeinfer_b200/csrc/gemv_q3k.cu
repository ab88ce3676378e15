// gemv_q3k.cu -- dequant-GEMV for Q3_K weights (3.4375 bits per weight) at verification batch 1..16, on the tensor cores.
//
// Q3_K_M files -- the type of the authors' 70B target (examples/speculative/scripts/dolphin_tinyllama_8_nodes.sh) -- keep token_embd,
// attn_q / attn_k and ffn_gate / ffn_up in Q3_K (reference llama.cpp:7900-7970).  Arithmetic of ggml_vec_dot_q3_K_q8_K
// (ggml-quants.c:4314, AVX2 :4443): per 256-value super-block sixteen groups of 16 values with a 6-bit scale (- 32) each,
// value = (two low bits) - (high bit ? 0 : 4), exact integer sums sum_g scale_g * sum(q * y), scaled by fp16(d_w) * d_x in fp32.
// Only the order of the fp32 super-block terms differs from the CPU's.
//
// Same structure as the Q6_K kernel (gemv_q6k.cu): one 1824-byte record per (16-row tile, super-block), regrouped so that a lane's
// registers become the A operand of mma.m16n8k16 with two logic ops per register:
//   [0, 1024)     2 chunks x 32 lanes x 16 B   low bit pairs; lane (g, t) word W[r][q] (row g + 8r, groups 4q .. 4q+3): in byte b, bits
//                                              2i, 2i+1 belong to element 16 (4q + i) + 4t + b.  Chunk c = {W[0][2c], W[1][2c], W[0][2c+1], W[1][2c+1]}
//   [1024, 1536)  32 lanes x 16 B              high bits; word H[r][h] (groups 8h .. 8h+7): in byte b, bit i belongs to element
//                                              16 (8h + i) + 4t + b.  {H[0][0], H[1][0], H[0][1], H[1][1]}
//   [1536, 1792)  16 rows x 16 int8            the scales, already unpacked and with the 32 subtracted (the reference packs them in 12 bytes:
//                                              1824 instead of 16 x 110 = 1760 bytes per record, + 3.6 %)
//   [1792, 1824)  16 x fp16 d
// u = low | high << 2 is 0..7 and the value is u - 4: the offset is 4 * sum_j scale_j * bsum_j, two more MMAs of the scales against
// the activations' 16-sums (b16 plane), as for Q6_K's - 32.
#include "pib_common.cuh"

#include <stdlib.h>

#define Q3T_BYTES  1824
#define Q3T_WARPS  4
#define Q3T_STAGES 4

__device__ __forceinline__ void q3_mma_u8s8_k16(int (&c)[4], int a0, int a1, int b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ void q3_mma_s8s8_k16(int (&c)[4], int a0, int a1, int b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ int q3_sbyte(uint32_t w, int k) { return (int) (int8_t) (w >> (8 * k)); }

// element e of a row's super-block in the plane form {qs[64], hmask[32]} (reference order, dequantize_row_q3_K ggml-quants.c:1711)
__device__ __forceinline__ void q3_bits(const uint8_t * blk96, int e, uint32_t & low, uint32_t & high) {
    const int n = e >> 7, ep = e & 127, j = ep >> 5, l = ep & 31;
    low  = (blk96[32 * n + l] >> (2 * j)) & 3u;
    high = (blk96[64 + l] >> (4 * n + j)) & 1u;
}
// the sixteen scales of a super-block, minus 32 (ggml-quants.c:1720-1730)
__device__ __forceinline__ void q3_scales(const uint8_t * packed12, int8_t * out16) {
    uint32_t aux[4];
    for (int i = 0; i < 3; i++) aux[i] = (uint32_t) packed12[4 * i] | ((uint32_t) packed12[4 * i + 1] << 8) | ((uint32_t) packed12[4 * i + 2] << 16) | ((uint32_t) packed12[4 * i + 3] << 24);
    const uint32_t kmask1 = 0x03030303u, kmask2 = 0x0f0f0f0fu, tmp = aux[2];
    aux[2] = ((aux[0] >> 4) & kmask2) | (((tmp >> 4) & kmask1) << 4);
    aux[3] = ((aux[1] >> 4) & kmask2) | (((tmp >> 6) & kmask1) << 4);
    aux[0] = (aux[0] & kmask2) | (((tmp >> 0) & kmask1) << 4);
    aux[1] = (aux[1] & kmask2) | (((tmp >> 2) & kmask1) << 4);
    for (int i = 0; i < 16; i++) out16[i] = (int8_t) ((int) ((aux[i >> 2] >> (8 * (i & 3))) & 0xff) - 32);
}

// one thread per (record, lane): the lane's three 16-byte pieces of bits; lanes 0..15 also write their row's scales and d
__global__ void q3k_tiles_kernel(const uint8_t * __restrict__ qs, const uint8_t * __restrict__ hdr, int64_t N, int64_t K, uint8_t * __restrict__ tiles) {
    const int64_t nsb = K / 256, n_tiles = (N + 15) / 16;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tiles * nsb * 32) return;
    const int     lane = (int) (i & 31), g = lane >> 2, t = lane & 3;
    const int64_t rec = i >> 5, T = rec / nsb, sb = rec % nsb;
    uint8_t *     out = tiles + rec * Q3T_BYTES;
    uint32_t W[2][4], H[2][2];
    for (int r = 0; r < 2; r++) {
        const int64_t row = T * 16 + g + 8 * r;
        const uint8_t * blk = row < N ? qs + (row * nsb + sb) * 96 : nullptr;
        for (int q = 0; q < 4; q++) W[r][q] = 0;
        H[r][0] = H[r][1] = 0;
        for (int j = 0; j < 16; j++)
            for (int b = 0; b < 4; b++) {
                uint32_t lo = 0, hi = 1;                                    // padding rows: value 0 = u 4 (the scales are 0 anyway)
                if (blk) q3_bits(blk, 16 * j + 4 * t + b, lo, hi);
                W[r][j >> 2] |= lo << (8 * b + 2 * (j & 3));
                H[r][j >> 3] |= hi << (8 * b + (j & 7));
            }
    }
    for (int c = 0; c < 2; c++) *reinterpret_cast<uint4 *>(out + c * 512 + lane * 16) = make_uint4(W[0][2 * c], W[1][2 * c], W[0][2 * c + 1], W[1][2 * c + 1]);
    *reinterpret_cast<uint4 *>(out + 1024 + lane * 16) = make_uint4(H[0][0], H[1][0], H[0][1], H[1][1]);
    if (lane < 16) {
        const int64_t r = T * 16 + lane;
        int8_t   sc[16];
        uint16_t d = 0;
        for (int k = 0; k < 16; k++) sc[k] = 0;
        if (r < N) { q3_scales(hdr + (r * nsb + sb) * 16, sc); d = *reinterpret_cast<const uint16_t *>(hdr + (r * nsb + sb) * 16 + 12); }
        for (int k = 0; k < 16; k++) out[1536 + lane * 16 + k] = (uint8_t) sc[k];
        *reinterpret_cast<uint16_t *>(out + 1792 + lane * 2) = d;
    }
}

int pib_weight_build_tiles_q3k(pib_weight * w) {
    if (w->type != PIB_TYPE_Q3_K || w->tiles) return PIB_OK;
    const int64_t nsb = w->K / 256, n_tiles = (w->N + 15) / 16;
    const size_t  bytes = (size_t) n_tiles * nsb * Q3T_BYTES;
    if (cudaMalloc(&w->tiles, bytes) != cudaSuccess) { cudaGetLastError(); pib_set_error("q3_K tiles: cudaMalloc of %zu bytes failed", bytes); w->tiles = nullptr; return PIB_ERR_NOMEM; }
    w->tile_bytes = bytes;
    const int64_t n = n_tiles * nsb * 32;
    q3k_tiles_kernel<<<(unsigned) ((n + 127) / 128), 128>>>(w->qs, (const uint8_t *) w->d, w->N, w->K, w->tiles);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

template <int NT>
__global__ void __launch_bounds__(Q3T_WARPS * 32, NT == 1 ? 4 : 3)
gemv_q3k_tc_kernel(const uint8_t * __restrict__ tiles, int nsb, int N, int K, const int8_t * __restrict__ aq6, const float * __restrict__ ad,
                   const int8_t * __restrict__ ab16, int b, float * __restrict__ y, const int * __restrict__ cancel) {
    if (PIB_CANCELLED(cancel)) { pib_pdl_wait(); return; }
    extern __shared__ __align__(16) uint8_t q3_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int T = blockIdx.x;
    const uint8_t * tbase  = tiles + (size_t) T * nsb * Q3T_BYTES + lane * 16;
    uint8_t *       ring   = q3_smem + warp * (Q3T_STAGES * Q3T_BYTES);
    const uint32_t  ring_a = (uint32_t) __cvta_generic_to_shared(ring) + lane * 16;
    int ps = warp;
    auto issue = [&](int stage) {
        if (ps < nsb) {
            const uint8_t * src = tbase + (size_t) ps * Q3T_BYTES;
            const uint32_t  dst = ring_a + stage * Q3T_BYTES;
#pragma unroll
            for (int c = 0; c < 3; c++) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + c * 512), "l"(src + c * 512) : "memory");
            if (lane < 18) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 1536), "l"(src + 1536) : "memory");
            ps += Q3T_WARPS;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int i = 0; i < Q3T_STAGES - 1; i++) issue(i);        // weights are immutable: before the wait on the producer of the activations

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
    for (int s = warp; s < nsb; s += Q3T_WARPS) {
        __syncwarp();
        issue((stage + Q3T_STAGES - 1) % Q3T_STAGES);
        asm volatile("cp.async.wait_group %0;" ::"n"(Q3T_STAGES - 1) : "memory");
        __syncwarp();
        const uint8_t * rs = ring + stage * Q3T_BYTES;
        const uint4 l0 = *reinterpret_cast<const uint4 *>(rs + lane * 16), l1 = *reinterpret_cast<const uint4 *>(rs + 512 + lane * 16);
        const uint4 hh = *reinterpret_cast<const uint4 *>(rs + 1024 + lane * 16);
        const uint4 sc0 = *reinterpret_cast<const uint4 *>(rs + 1536 + g * 16), sc1 = *reinterpret_cast<const uint4 *>(rs + 1536 + (g + 8) * 16);
        const float d0 = pib_h2f(*reinterpret_cast<const uint16_t *>(rs + 1792 + g * 2)), d1 = pib_h2f(*reinterpret_cast<const uint16_t *>(rs + 1792 + (g + 8) * 2));
        const uint32_t W0[4] = { l0.x, l0.z, l1.x, l1.z }, W1[4] = { l0.y, l0.w, l1.y, l1.w };      // rows g / g + 8, quads 0..3
        const uint32_t H0[2] = { hh.x, hh.z }, H1[2] = { hh.y, hh.w };
        const uint32_t s0w[4] = { sc0.x, sc0.y, sc0.z, sc0.w }, s1w[4] = { sc1.x, sc1.y, sc1.z, sc1.w };
        const uint32_t s0t = t == 0 ? sc0.x : t == 1 ? sc0.y : t == 2 ? sc0.z : sc0.w, s1t = t == 0 ? sc1.x : t == 1 ? sc1.y : t == 2 ? sc1.z : sc1.w;
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
            for (int j = 0; j < 16; j++) {
                const int q = j >> 2, i = j & 3, h = j >> 3, ih = j & 7;
                int c4[4] = { 0, 0, 0, 0 };
                q3_mma_u8s8_k16(c4, (int) (((W0[q] >> (2 * i)) & 0x03030303u) | (((H0[h] >> ih) & 0x01010101u) << 2)),
                                (int) (((W1[q] >> (2 * i)) & 0x03030303u) | (((H1[h] >> ih) & 0x01010101u) << 2)), bw[j]);
                const int a0 = q3_sbyte(s0w[j >> 2], j & 3), a1 = q3_sbyte(s1w[j >> 2], j & 3);
                sum[0] += a0 * c4[0]; sum[1] += a0 * c4[1]; sum[2] += a1 * c4[2]; sum[3] += a1 * c4[3];
            }
            // the - 4 offset: 4 * sum_j scale_j * bsum_j, bsum = 64 hi + lo; A = this lane's four scales of rows g, g + 8
            int ch[4] = { 0, 0, 0, 0 }, cl[4] = { 0, 0, 0, 0 };
            q3_mma_s8s8_k16(ch, (int) s0t, (int) s1t, bh);
            q3_mma_s8s8_k16(cl, (int) s0t, (int) s1t, bl);
            acc[n][0] = fmaf(__fmul_rn(dx0, d0), (float) (sum[0] - 4 * (64 * ch[0] + cl[0])), acc[n][0]);
            acc[n][1] = fmaf(__fmul_rn(dx1, d0), (float) (sum[1] - 4 * (64 * ch[1] + cl[1])), acc[n][1]);
            acc[n][2] = fmaf(__fmul_rn(dx0, d1), (float) (sum[2] - 4 * (64 * ch[2] + cl[2])), acc[n][2]);
            acc[n][3] = fmaf(__fmul_rn(dx1, d1), (float) (sum[3] - 4 * (64 * ch[3] + cl[3])), acc[n][3]);
        }
        stage = stage + 1 == Q3T_STAGES ? 0 : stage + 1;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                                            // the rings are done with: their memory becomes the reduction buffer
    float (*red)[NT * 8][16] = reinterpret_cast<float (*)[NT * 8][16]>(q3_smem);
#pragma unroll
    for (int n = 0; n < NT; n++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
            red[warp][n * 8 + 2 * t + e][g]     = acc[n][e];
            red[warp][n * 8 + 2 * t + e][g + 8] = acc[n][2 + e];
        }
    __syncthreads();
    for (int i = threadIdx.x; i < NT * 8 * 16; i += Q3T_WARPS * 32) {
        const int tok = i / 16, r = i % 16, row = T * 16 + r;
        if (tok < b && row < N) {
            float v = red[0][tok][r];
#pragma unroll
            for (int w2 = 1; w2 < Q3T_WARPS; w2++) v += red[w2][tok][r];
            y[(size_t) tok * N + row] = v;
        }
    }
}

template <int NT>
static int launch_q3_tc(const pib_weight * w, const uint8_t * act, const pib_act_layout & L, int b, float * y, const int * cancel, cudaStream_t st) {
    const size_t smem = (size_t) Q3T_WARPS * Q3T_STAGES * Q3T_BYTES;
    static bool configured = false;
    if (!configured) {
        PIB_CHECK(cudaFuncSetAttribute(gemv_q3k_tc_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        PIB_CHECK(cudaFuncSetAttribute(gemv_q3k_tc_kernel<NT>, cudaFuncAttributePreferredSharedMemoryCarveout, (int) cudaSharedmemCarveoutMaxShared));
        configured = true;
    }
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    PIB_CHECK(pib_launch(gemv_q3k_tc_kernel<NT>, dim3(n_tiles), dim3(Q3T_WARPS * 32), smem, st, w->tiles, nsb, (int) w->N, (int) w->K,
                         (const int8_t *) (act + L.q6p_off), (const float *) (act + L.d_off), (const int8_t *) (act + L.b16_off), b, y, cancel));
    return PIB_OK;
}

int pib_gemv_q3k(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!w || w->type != PIB_TYPE_Q3_K || !w->tiles || b < 1 || b > PIB_MAX_BATCH) { pib_set_error("pib_gemv: not a resident Q3_K matrix (or batch > %d)", PIB_MAX_BATCH); return PIB_ERR_UNSUPP; }
    const pib_act_layout L   = pib_act_layout_of(PIB_TYPE_Q4_K, w->K, b);          // Q8_K family activations
    const uint8_t *      act = (const uint8_t *) act_dev;
    return b <= 8 ? launch_q3_tc<1>(w, act, L, b, y_dev, cancel, st) : launch_q3_tc<2>(w, act, L, b, y_dev, cancel, st);
}

// embedding rows: dequantize_row_q3_K (ggml-quants.c:1711-1758), one thread per element, same float operations in the same order
__global__ void q3k_get_rows_kernel(const uint8_t * __restrict__ qs, const uint8_t * __restrict__ hdr, int64_t K, const int32_t * __restrict__ ids,
                                    float * __restrict__ out) {
    const int64_t row = ids[blockIdx.x], nsb = K / 256;
    for (int64_t e = threadIdx.x; e < K; e += blockDim.x) {
        const int64_t sb = e >> 8;
        const int     ee = (int) (e & 255);
        const uint8_t * blk = qs + (row * nsb + sb) * 96;
        const uint8_t * h   = hdr + (row * nsb + sb) * 16;
        int8_t sc[16];
        q3_scales(h, sc);
        uint32_t lo, hi;
        q3_bits(blk, ee, lo, hi);
        const float d_all = pib_h2f(*reinterpret_cast<const uint16_t *>(h + 12));
        const float dl    = __fmul_rn(d_all, (float) sc[ee >> 4]);
        out[(size_t) blockIdx.x * K + e] = __fmul_rn(dl, (float) ((int) lo - (hi ? 0 : 4)));
    }
}

int pib_get_rows_q3k(const pib_weight * w, const int32_t * ids_dev, int n, float * out_dev, cudaStream_t st) {
    if (!w->qs) { pib_set_error("pib_get_rows: this matrix keeps only its tensor-core tiles"); return PIB_ERR_UNSUPP; }
    PIB_CHECK(pib_launch(q3k_get_rows_kernel, dim3(n), dim3(256), 0, st, (const uint8_t *) w->qs, (const uint8_t *) w->d, w->K, ids_dev, out_dev));
    return PIB_OK;
}
