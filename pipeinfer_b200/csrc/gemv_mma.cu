// gemv_mma.cu -- variant 2 of the quantised matmul (K1) for Q4_K weights: fragment-major weight tiles streamed with
// coalesced 128-bit loads, int8 tensor-core dot products (mma.sync m16n8k32 s8), per-sub-block integer scaling,
// fp32 super-block accumulation, cross-warp + (optional) cross-CTA reduction in a fixed order.
//
// Arithmetic is the reference's (ggml-quants.c:5216 ggml_vec_dot_q4_K_q8_K): per super-block
//     sum_j sc_j * <q_j, x_j>   (exact int32)      times   d_w * d_x
//   - sum_j m_j * bsum_j        (exact int32)      times   dmin_w * d_x
// accumulated in fp32; only the order of the fp32 additions differs from the CPU.
//
// Why tensor cores on an HBM-bound kernel: at b tokens every weight byte feeds 2*b int8 MACs.  dp4a gives 64 lanes/clk/SM
// -- at the bandwidth we want (~35 B/clk/SM) that is enough for b <= 4 only; one mma.sync covers 16 rows x 8 tokens x 32
// weights, so b = 1..8 cost the same issue slots and b = 9..16 twice that, both far below the issue limit.
//
// Layout ("tiles", built once at load from the planes of weights.cu; values untouched): for row tile T (16 rows) and
// super-block S the 16 x 128 bytes of nibbles are stored as [chunk c = 0..3][lane 0..31][16 B], where lane (g = lane/4,
// t = lane%4) gets {row g: bytes [8t, 8t+8) of chunk c, row g+8: the same bytes} -- exactly its A fragments for the two
// 32-element sub-blocks 2c (low nibbles) and 2c+1 (high nibbles) under the k-permutation "mma k-slot 4t+i <-> element 8t+i,
// k-slot 16+4t+i <-> element 8t+4+i" (the activations use the same permutation, so every B fragment is one contiguous
// 8-byte load).  Then the 16 row headers {fp16 d, fp16 dmin, scales[12]}.  2304 bytes per (T, S), consecutive S
// contiguous: every warp-wide load instruction reads 512 contiguous bytes.
#include "pib_common.cuh"

#define MM_WARPS   4
#define MM_THREADS (MM_WARPS * 32)
#define MM_TILE_BYTES 2304          // 2048 nibble bytes + 16 x 16 header bytes per (row tile, super-block)

// ---- repack: planes -> tiles ------------------------------------------------------------------------------------------
__global__ void q4k_tiles_kernel(const uint8_t * __restrict__ qs, const uint8_t * __restrict__ hdr, int64_t N, int64_t K,
                                 uint8_t * __restrict__ tiles) {
    const int64_t nsb = K / 256;
    const int64_t n_tiles = (N + 15) / 16;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;     // one thread per 16-byte piece
    const int64_t per_ts = MM_TILE_BYTES / 16;                              // 144 pieces
    if (i >= n_tiles * nsb * per_ts) return;
    const int64_t ts = i / per_ts;
    const int     p  = (int) (i % per_ts);
    const int64_t T = ts / nsb, S = ts % nsb;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (p < 128) {
        const int c = p / 32, lane = p % 32, g = lane / 4, t = lane % 4;
        const int64_t r0 = T * 16 + g, r1 = r0 + 8;
        if (r0 < N) { const uint2 a = *reinterpret_cast<const uint2 *>(qs + r0 * (K / 2) + S * 128 + c * 32 + t * 8); v.x = a.x; v.y = a.y; }
        if (r1 < N) { const uint2 a = *reinterpret_cast<const uint2 *>(qs + r1 * (K / 2) + S * 128 + c * 32 + t * 8); v.z = a.x; v.w = a.y; }
    } else {
        const int64_t r = T * 16 + (p - 128);
        if (r < N) v = *reinterpret_cast<const uint4 *>(hdr + (r * nsb + S) * 16);     // rows beyond N: d = dmin = 0
    }
    *reinterpret_cast<uint4 *>(tiles + i * 16) = v;
}

int pib_weight_build_tiles(pib_weight * w) {
    if (w->type != PIB_TYPE_Q4_K || w->tiles) return PIB_OK;
    const int64_t nsb = w->K / 256, n_tiles = (w->N + 15) / 16;
    const size_t  bytes = (size_t) n_tiles * nsb * MM_TILE_BYTES;
    if (cudaMalloc(&w->tiles, bytes) != cudaSuccess) {
        cudaGetLastError();
        pib_set_error("gemv tiles: cudaMalloc of %zu bytes failed", bytes);
        w->tiles = nullptr;
        return PIB_ERR_NOMEM;
    }
    w->tile_bytes = bytes;
    const int64_t n = (int64_t) (bytes / 16);
    q4k_tiles_kernel<<<(unsigned) ((n + 255) / 256), 256>>>(w->qs, (const uint8_t *) w->d, w->N, w->K, w->tiles);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

// ---- kernel -------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void mma_s8(int (&c)[4], int a0, int a1, int a2, int a3, int b0, int b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

struct q4k_hdr {          // one row's decoded super-block header
    float    d, dmin;
    uint32_t sc_lo, sc_hi;   // sc0..3 | sc4..7 as bytes
    uint32_t m_lo, m_hi;     // m0..3  | m4..7
};

__device__ __forceinline__ q4k_hdr decode_hdr(const uint4 h) {
    q4k_hdr r;
    r.d    = pib_h2f((uint16_t) (h.x & 0xffff));
    r.dmin = pib_h2f((uint16_t) (h.x >> 16));
    // reference get_scale_min_k4 (ggml-quants.c:1446) for all eight sub-blocks at once (the utmp trick of :5329-5334)
    r.sc_lo = h.y & 0x3f3f3f3fu;
    r.m_lo  = h.z & 0x3f3f3f3fu;
    r.sc_hi = (h.w & 0x0f0f0f0fu) | (((h.y >> 6) & 0x03030303u) << 4);
    r.m_hi  = ((h.w >> 4) & 0x0f0f0f0fu) | (((h.z >> 6) & 0x03030303u) << 4);
    return r;
}

// NT = number of 8-token n-tiles (1: b <= 8, 2: b <= 16).  grid = (row tiles, k splits); each warp of a CTA takes every
// MM_WARPS-th super-block of the CTA's k range.  partial != nullptr: write fp32 partial sums [split][b][N] instead of y.
template <int NT>
__global__ void __launch_bounds__(MM_THREADS, 4)
gemv_q4k_mma_kernel(const uint8_t * __restrict__ tiles, int nsb, int N, int K, const int8_t * __restrict__ aqs,
                    const float * __restrict__ ad, const int16_t * __restrict__ absum, int b, float * __restrict__ y,
                    float * __restrict__ partial, unsigned int * __restrict__ counters, const int * __restrict__ cancel) {
    if (cancel && *cancel) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int T = blockIdx.x;
    const int nsplit = gridDim.y, split = blockIdx.y;
    const int sb_per = (nsb + nsplit - 1) / nsplit;
    const int sb0 = split * sb_per, sb1 = min(nsb, sb0 + sb_per);
    const uint8_t * tbase = tiles + (size_t) T * nsb * MM_TILE_BYTES;

    float acc[NT][4];
#pragma unroll
    for (int n = 0; n < NT; n++) { acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f; }

    // tokens this lane supplies B fragments for (column g of each n-tile) and owns C columns of (2t, 2t+1)
    int  tokB[NT];
    bool okB[NT];
#pragma unroll
    for (int n = 0; n < NT; n++) { tokB[n] = n * 8 + g; okB[n] = tokB[n] < b; }

    uint4 wq[4], hq[2];
    int   s = sb0 + warp;
    if (s < sb1) {
        const uint8_t * p = tbase + (size_t) s * MM_TILE_BYTES;
#pragma unroll
        for (int c = 0; c < 4; c++) wq[c] = *reinterpret_cast<const uint4 *>(p + c * 512 + lane * 16);   // streamed below via ld_stream
        hq[0] = *reinterpret_cast<const uint4 *>(p + 2048 + g * 16);
        hq[1] = *reinterpret_cast<const uint4 *>(p + 2048 + (g + 8) * 16);
    }
    for (; s < sb1; s += MM_WARPS) {
        // prefetch the next super-block of this warp while this one is consumed
        uint4 wn[4], hn[2];
        const int sn = s + MM_WARPS;
        if (sn < sb1) {
            const uint8_t * p = tbase + (size_t) sn * MM_TILE_BYTES;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int4 v = ld_stream_v4(p + c * 512 + lane * 16);
                wn[c] = make_uint4((uint32_t) v.x, (uint32_t) v.y, (uint32_t) v.z, (uint32_t) v.w);
            }
            const int4 h0 = ld_stream_v4(p + 2048 + g * 16), h1 = ld_stream_v4(p + 2048 + (g + 8) * 16);
            hn[0] = make_uint4((uint32_t) h0.x, (uint32_t) h0.y, (uint32_t) h0.z, (uint32_t) h0.w);
            hn[1] = make_uint4((uint32_t) h1.x, (uint32_t) h1.y, (uint32_t) h1.z, (uint32_t) h1.w);
        }
        const q4k_hdr H0 = decode_hdr(hq[0]), H1 = decode_hdr(hq[1]);     // rows g and g + 8
        int sumi[NT][4];
#pragma unroll
        for (int n = 0; n < NT; n++) { sumi[n][0] = sumi[n][1] = sumi[n][2] = sumi[n][3] = 0; }
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const uint4 w = wq[c];            // {row g: x, y ; row g+8: z, w}
            const int a_lo[4] = { (int) (w.x & 0x0F0F0F0Fu), (int) (w.z & 0x0F0F0F0Fu), (int) (w.y & 0x0F0F0F0Fu), (int) (w.w & 0x0F0F0F0Fu) };
            const int a_hi[4] = { (int) ((w.x >> 4) & 0x0F0F0F0Fu), (int) ((w.z >> 4) & 0x0F0F0F0Fu), (int) ((w.y >> 4) & 0x0F0F0F0Fu),
                                  (int) ((w.w >> 4) & 0x0F0F0F0Fu) };
            const int j0 = 2 * c, j1 = 2 * c + 1;
            const int sc0_lo = (int) (((j0 < 4 ? H0.sc_lo : H0.sc_hi) >> (8 * (j0 & 3))) & 0xff);
            const int sc1_lo = (int) (((j0 < 4 ? H1.sc_lo : H1.sc_hi) >> (8 * (j0 & 3))) & 0xff);
            const int sc0_hi = (int) (((j1 < 4 ? H0.sc_lo : H0.sc_hi) >> (8 * (j1 & 3))) & 0xff);
            const int sc1_hi = (int) (((j1 < 4 ? H1.sc_lo : H1.sc_hi) >> (8 * (j1 & 3))) & 0xff);
#pragma unroll
            for (int n = 0; n < NT; n++) {
                int2 b_lo = make_int2(0, 0), b_hi = make_int2(0, 0);
                if (okB[n]) {
                    const int8_t * ap = aqs + (size_t) tokB[n] * K + (size_t) s * 256 + c * 64 + t * 8;
                    b_lo = __ldg(reinterpret_cast<const int2 *>(ap));            // sub-block 2c   : elements 8t .. 8t+7
                    b_hi = __ldg(reinterpret_cast<const int2 *>(ap + 32));       // sub-block 2c+1
                }
                int c_lo[4] = { 0, 0, 0, 0 }, c_hi[4] = { 0, 0, 0, 0 };
                mma_s8(c_lo, a_lo[0], a_lo[1], a_lo[2], a_lo[3], b_lo.x, b_lo.y);
                mma_s8(c_hi, a_hi[0], a_hi[1], a_hi[2], a_hi[3], b_hi.x, b_hi.y);
                sumi[n][0] += sc0_lo * c_lo[0] + sc0_hi * c_hi[0];
                sumi[n][1] += sc0_lo * c_lo[1] + sc0_hi * c_hi[1];
                sumi[n][2] += sc1_lo * c_lo[2] + sc1_hi * c_hi[2];
                sumi[n][3] += sc1_lo * c_lo[3] + sc1_hi * c_hi[3];
            }
        }
        // mins: sum_j m_j * (bsums[2j] + bsums[2j+1]) with dp2a on the int16 pairs exactly as block_q8_K stores them
        const int m0_01 = (int) __byte_perm(H0.m_lo, 0, 0x1100), m0_23 = (int) __byte_perm(H0.m_lo, 0, 0x3322);
        const int m0_45 = (int) __byte_perm(H0.m_hi, 0, 0x1100), m0_67 = (int) __byte_perm(H0.m_hi, 0, 0x3322);
        const int m1_01 = (int) __byte_perm(H1.m_lo, 0, 0x1100), m1_23 = (int) __byte_perm(H1.m_lo, 0, 0x3322);
        const int m1_45 = (int) __byte_perm(H1.m_hi, 0, 0x1100), m1_67 = (int) __byte_perm(H1.m_hi, 0, 0x3322);
#pragma unroll
        for (int n = 0; n < NT; n++) {
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int tok = n * 8 + 2 * t + e;
                if (tok < b) {
                    const int4 * bp = reinterpret_cast<const int4 *>(absum + (size_t) tok * (K / 16) + (size_t) s * 16);
                    const int4  q0 = __ldg(bp), q1 = __ldg(bp + 1);            // int16 pairs for sub-blocks 0..3 | 4..7
                    int mi0 = __dp2a_lo(q0.x, m0_01, 0); mi0 = __dp2a_hi(q0.y, m0_01, mi0);
                    mi0 = __dp2a_lo(q0.z, m0_23, mi0);   mi0 = __dp2a_hi(q0.w, m0_23, mi0);
                    mi0 = __dp2a_lo(q1.x, m0_45, mi0);   mi0 = __dp2a_hi(q1.y, m0_45, mi0);
                    mi0 = __dp2a_lo(q1.z, m0_67, mi0);   mi0 = __dp2a_hi(q1.w, m0_67, mi0);
                    int mi1 = __dp2a_lo(q0.x, m1_01, 0); mi1 = __dp2a_hi(q0.y, m1_01, mi1);
                    mi1 = __dp2a_lo(q0.z, m1_23, mi1);   mi1 = __dp2a_hi(q0.w, m1_23, mi1);
                    mi1 = __dp2a_lo(q1.x, m1_45, mi1);   mi1 = __dp2a_hi(q1.y, m1_45, mi1);
                    mi1 = __dp2a_lo(q1.z, m1_67, mi1);   mi1 = __dp2a_hi(q1.w, m1_67, mi1);
                    const float dx = __ldg(ad + (size_t) tok * nsb + s);
                    acc[n][e]     = fmaf(__fmul_rn(dx, H0.d), (float) sumi[n][e], acc[n][e]);
                    acc[n][e]     = fmaf(__fmul_rn(-dx, H0.dmin), (float) mi0, acc[n][e]);
                    acc[n][2 + e] = fmaf(__fmul_rn(dx, H1.d), (float) sumi[n][2 + e], acc[n][2 + e]);
                    acc[n][2 + e] = fmaf(__fmul_rn(-dx, H1.dmin), (float) mi1, acc[n][2 + e]);
                }
            }
        }
        if (sn < sb1) {
#pragma unroll
            for (int c = 0; c < 4; c++) wq[c] = wn[c];
            hq[0] = hn[0]; hq[1] = hn[1];
        }
    }

    // cross-warp reduction in warp order, then (k-split) cross-CTA reduction in split order by the last CTA to arrive
    __shared__ float red[MM_WARPS][NT * 8][16];
#pragma unroll
    for (int n = 0; n < NT; n++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
            red[warp][n * 8 + 2 * t + e][g]     = acc[n][e];
            red[warp][n * 8 + 2 * t + e][g + 8] = acc[n][2 + e];
        }
    __syncthreads();
    __shared__ bool s_last;
    for (int i = threadIdx.x; i < NT * 8 * 16; i += MM_THREADS) {
        const int tok = i / 16, r = i % 16, row = T * 16 + r;
        if (tok < b && row < N) {
            float v = red[0][tok][r];
#pragma unroll
            for (int w2 = 1; w2 < MM_WARPS; w2++) v += red[w2][tok][r];
            if (nsplit == 1) y[(size_t) tok * N + row] = v;
            else partial[((size_t) split * b + tok) * N + row] = v;
        }
    }
    if (nsplit == 1) return;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&counters[T], 1u) == (unsigned) nsplit - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int i = threadIdx.x; i < NT * 8 * 16; i += MM_THREADS) {
        const int tok = i / 16, r = i % 16, row = T * 16 + r;
        if (tok < b && row < N) {
            float v = 0.f;
            for (int sp = 0; sp < nsplit; sp++) v += __ldcg(partial + ((size_t) sp * b + tok) * N + row);
            y[(size_t) tok * N + row] = v;
        }
    }
    if (threadIdx.x == 0) counters[T] = 0;          // self-cleaning: ready for the next launch
}

// ---- launcher -----------------------------------------------------------------------------------------------------------
// scratch for the k-split path: partial sums [split][b][N] + one counter per row tile; owned by the caller (the engine
// keeps one per context; the stateless C ABI allocates it per device on first use)
struct mm_scratch { float * partial; unsigned int * counters; size_t partial_floats; size_t n_counters; };
static mm_scratch g_scratch[16];

static int ensure_scratch(size_t floats, size_t counters, mm_scratch ** out) {
    int dev = 0;
    cudaGetDevice(&dev);
    mm_scratch & s = g_scratch[dev & 15];
    if (s.partial_floats < floats) {
        if (s.partial) cudaFree(s.partial);
        s.partial = nullptr; s.partial_floats = 0;
        if (cudaMalloc(&s.partial, floats * sizeof(float)) != cudaSuccess) { cudaGetLastError(); return PIB_ERR_NOMEM; }
        s.partial_floats = floats;
    }
    if (s.n_counters < counters) {
        if (s.counters) cudaFree(s.counters);
        s.counters = nullptr; s.n_counters = 0;
        if (cudaMalloc(&s.counters, counters * sizeof(unsigned int)) != cudaSuccess) { cudaGetLastError(); return PIB_ERR_NOMEM; }
        cudaMemset(s.counters, 0, counters * sizeof(unsigned int));
        s.n_counters = counters;
    }
    *out = &s;
    return PIB_OK;
}

bool pib_gemv_mma_supported(const pib_weight * w, int b) {
    return w && w->type == PIB_TYPE_Q4_K && w->tiles != nullptr && b >= 1 && b <= PIB_MAX_BATCH;
}

// number of k splits: enough CTAs to fill the machine when there are few row tiles (small N: draft model, Wk/Wv)
static int choose_split(int n_tiles, int nsb) {
    const int target = 4 * pib_num_sms();
    int split = 1;
    while (n_tiles * split < target && split * 2 * MM_WARPS <= nsb && split < 16) split *= 2;
    return split;
}

int pib_gemv_mma_launch(const pib_weight * w, const void * act_dev, int b, float * y_dev, int split, float * partial,
                        unsigned int * counters, const int * cancel, cudaStream_t st) {
    const pib_act_layout L   = pib_act_layout_of(w->type, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    const int            nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    const dim3           grid(n_tiles, split);
    const int8_t *       aqs = (const int8_t *) (act + L.qs_off);
    const float *        ad  = (const float *) (act + L.d_off);
    const int16_t *      bs  = (const int16_t *) (act + L.bs_off);
    if (b <= 8) gemv_q4k_mma_kernel<1><<<grid, MM_THREADS, 0, st>>>(w->tiles, nsb, (int) w->N, (int) w->K, aqs, ad, bs, b, y_dev, partial, counters, cancel);
    else        gemv_q4k_mma_kernel<2><<<grid, MM_THREADS, 0, st>>>(w->tiles, nsb, (int) w->N, (int) w->K, aqs, ad, bs, b, y_dev, partial, counters, cancel);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

int pib_gemv_mma(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!pib_gemv_mma_supported(w, b)) { pib_set_error("pib_gemv: variant 2 needs Q4_K weights with tiles"); return PIB_ERR_UNSUPP; }
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    const int split = choose_split(n_tiles, nsb);
    mm_scratch * s = nullptr;
    if (split > 1) {
        if (ensure_scratch((size_t) split * b * w->N, (size_t) n_tiles, &s) != PIB_OK) { pib_set_error("pib_gemv: scratch allocation failed"); return PIB_ERR_NOMEM; }
    }
    return pib_gemv_mma_launch(w, act_dev, b, y_dev, split, s ? s->partial : nullptr, s ? s->counters : nullptr, cancel, st);
}
