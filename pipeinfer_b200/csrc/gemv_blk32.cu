// gemv_blk32.cu -- tensor-core dequant-GEMV for the 32-element block formats (Q4_0, Q8_0) at verification batch 1..16.
//
// Same arithmetic as the reference (ggml_vec_dot_q4_0_q8_0 ggml-quants.c:2426, q8_0_q8_0 :3521): per 32-element block an
// exact integer dot product of (q - 8) resp. q with the int8 activations, scaled by fp16(d_w) * d_x in fp32 and summed
// over the blocks.  Only the order of the fp32 block terms differs (per-warp partial sums, then a fixed-order reduction).
//
// Layout.  Eight consecutive blocks (256 columns) of a 16-row tile form one record, in mma.m16n8k32 A-fragment order so
// that a lane's 128-bit loads ARE its tensor-core operands:
//   Q4_0  2304 B: [pair p = 0..3][lane][16 B] = { row g blk 2p, row g+8 blk 2p, row g blk 2p+1, row g+8 blk 2p+1 } -- of each
//                 block the 4 bytes 4t..4t+3, whose low nibbles are columns 4t..4t+3 and high nibbles columns 16+4t..16+4t+3 --
//                 then [row 0..15][8 x fp16 d].
//   Q8_0  4352 B: [blk j = 0..7][lane][16 B] = { row g k 4t.., row g+8 k 4t.., row g k 16+4t.., row g+8 k 16+4t.. }, then the
//                 same 256 B of scales.
// K is padded to a multiple of 256 with zero blocks (d = 0), e.g. Falcon-7B's 4544 columns.  The activations come in a
// matching fragment order (act_quant.cu: qsp / dxp planes), zero-padded the same way.
#include <map>
#include <mutex>
#include <utility>

#include <math.h>

#include "pib_common.cuh"

#define BK_WARPS   4
#define BK_THREADS (BK_WARPS * 32)

__host__ __device__ static inline int bk_rec_bytes(int type) { return type == PIB_TYPE_Q4_0 ? 2304 : 4352; }

// ---- repack: planes -> records ------------------------------------------------------------------------------------------
__global__ void blk32_tiles_kernel(int type, const uint8_t * __restrict__ qs, const uint16_t * __restrict__ d, int64_t N, int64_t K,
                                   uint8_t * __restrict__ tiles) {
    const int64_t nb = K / 32, ngr = (K + 255) / 256, n_tiles = (N + 15) / 16;
    const int     rec = bk_rec_bytes(type), per = rec / 16, nq = per - 16;          // 16-byte pieces: quants, then 16 scale rows
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tiles * ngr * per) return;
    const int64_t ts = i / per;
    const int     p  = (int) (i % per);
    const int64_t T = ts / ngr, G = ts % ngr;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (p < nq) {
        const int     lane = p % 32, g = lane / 4, t = lane % 4;
        const int64_t r0 = T * 16 + g, r1 = r0 + 8;
        if (type == PIB_TYPE_Q4_0) {
            const int     pr = p / 32;                                               // block pair
            const int64_t b0 = G * 8 + 2 * pr, b1 = b0 + 1;
            auto word = [&](int64_t r, int64_t b) -> uint32_t {
                return (r < N && b < nb) ? *reinterpret_cast<const uint32_t *>(qs + r * (K / 2) + b * 16 + t * 4) : 0x88888888u;   // (8 - 8) = 0
            };
            v = make_uint4(word(r0, b0), word(r1, b0), word(r0, b1), word(r1, b1));
        } else {
            const int     j = p / 32;
            const int64_t b = G * 8 + j;
            auto word = [&](int64_t r, int off) -> uint32_t {
                return (r < N && b < nb) ? *reinterpret_cast<const uint32_t *>(qs + r * K + b * 32 + off + t * 4) : 0u;
            };
            v = make_uint4(word(r0, 0), word(r1, 0), word(r0, 16), word(r1, 16));
        }
    } else {
        const int64_t r = T * 16 + (p - nq);
        uint16_t h[8];
        for (int j = 0; j < 8; j++) { const int64_t b = G * 8 + j; h[j] = (r < N && b < nb) ? d[r * nb + b] : (uint16_t) 0; }
        v = make_uint4(h[0] | ((uint32_t) h[1] << 16), h[2] | ((uint32_t) h[3] << 16), h[4] | ((uint32_t) h[5] << 16), h[6] | ((uint32_t) h[7] << 16));
    }
    *reinterpret_cast<uint4 *>(tiles + i * 16) = v;
}

int pib_weight_build_tiles_blk32(pib_weight * w) {
    if ((w->type != PIB_TYPE_Q4_0 && w->type != PIB_TYPE_Q8_0) || w->tiles) return PIB_OK;
    const int64_t ngr = (w->K + 255) / 256, n_tiles = (w->N + 15) / 16;
    const size_t  bytes = (size_t) n_tiles * ngr * bk_rec_bytes(w->type);
    if (cudaMalloc(&w->tiles, bytes) != cudaSuccess) {
        cudaGetLastError();
        pib_set_error("gemv tiles: cudaMalloc of %zu bytes failed", bytes);
        w->tiles = nullptr;
        return PIB_ERR_NOMEM;
    }
    w->tile_bytes = bytes;
    const int64_t n = (int64_t) (bytes / 16);
    blk32_tiles_kernel<<<(unsigned) ((n + 255) / 256), 256>>>(w->type, w->qs, w->d, w->N, w->K, w->tiles);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

// ---- kernel -------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void bk_mma_s8(int (&c)[4], int a0, int a1, int a2, int a3, int b0, int b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// four nibbles 0..15 in the low half of each byte -> four int8 (q - 8): the forced high bit keeps the per-byte subtraction from
// borrowing across bytes, the final xor turns 0x78..0x87 into two's complement -8..7
__device__ __forceinline__ int bk_sub8(uint32_t nib) { return (int) (((nib | 0x80808080u) - 0x08080808u) ^ 0x80808080u); }

// grid (row tiles, k splits), BK_WARPS warps: warp w takes records sb0 + w, sb0 + w + BK_WARPS, ... of the CTA's k range;
// NT = token groups of 8 (1: b <= 8, 2: b <= 16).  The next record is prefetched into registers while the current one is
// multiplied; partial sums of the warps are reduced through shared memory in warp order; with k splits the last CTA to
// arrive adds the splits' partial sums in split order (deterministic), exactly as gemv_q4k_direct_kernel does.
template <int WT, int NT>
__global__ void __launch_bounds__(BK_THREADS, NT == 1 ? 4 : 3)
gemv_blk32_kernel(const uint8_t * __restrict__ tiles, int ngr, int N, const int8_t * __restrict__ aqs, const __half * __restrict__ adx,
                  int b, float * __restrict__ y, float * __restrict__ partial, unsigned int * __restrict__ counters,
                  const int * __restrict__ cancel) {
    if (PIB_CANCELLED(cancel)) {
        pib_pdl_wait();
        // cancelled: no math, but the arrival protocol of the k split still runs -- the flag can flip while the grid is starting,
        // and a tile whose splits did not all arrive would leave its counter non-zero for the next launch
        if (gridDim.y > 1 && threadIdx.x == 0 && atomicAdd(&counters[blockIdx.x], 1u) == gridDim.y - 1) counters[blockIdx.x] = 0;
        return;
    }
    constexpr int REC = WT == PIB_TYPE_Q4_0 ? 2304 : 4352;
    constexpr int NQ  = WT == PIB_TYPE_Q4_0 ? 4 : 8;            // 128-bit quant loads per lane and record
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int T = blockIdx.x;
    const int nsplit = gridDim.y, split = blockIdx.y;
    const int per = (ngr + nsplit - 1) / nsplit;
    const int s0 = split * per, s1 = min(ngr, s0 + per);
    const uint8_t * qbase = tiles + (size_t) T * ngr * REC + lane * 16;
    const uint8_t * dbase = tiles + (size_t) T * ngr * REC + (REC - 256) + g * 16;

    float acc[NT][4];
#pragma unroll
    for (int n = 0; n < NT; n++) { acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f; }
    const int8_t * a_q[NT];
    const __half * a_d[NT][2];
#pragma unroll
    for (int n = 0; n < NT; n++) {
        a_q[n] = aqs + (size_t) min(n * 8 + g, b - 1) * ngr * 256 + t * 64;
#pragma unroll
        for (int e = 0; e < 2; e++) a_d[n][e] = adx + (size_t) min(n * 8 + 2 * t + e, b - 1) * ngr * 8;
    }

    int4 wq[NQ], dq[2];
    int  s = s0 + warp;
    if (s < s1) {
#pragma unroll
        for (int c = 0; c < NQ; c++) wq[c] = ld_stream_v4(qbase + (size_t) s * REC + c * 512);
        dq[0] = ld_stream_v4(dbase + (size_t) s * REC);
        dq[1] = ld_stream_v4(dbase + (size_t) s * REC + 128);
    }
    // the weights above are immutable, so their loads were issued before waiting for the kernel that produces the activations
    pib_pdl_wait();
    pib_pdl_trigger();
    for (; s < s1; s += BK_WARPS) {
        int4 wn[NQ], dn[2];
        const int sn = s + BK_WARPS;
        if (sn < s1) {
#pragma unroll
            for (int c = 0; c < NQ; c++) wn[c] = ld_stream_v4(qbase + (size_t) sn * REC + c * 512);
            dn[0] = ld_stream_v4(dbase + (size_t) sn * REC);
            dn[1] = ld_stream_v4(dbase + (size_t) sn * REC + 128);
        }
        int4 aq[NT][4];
        int4 ax[NT][2];
#pragma unroll
        for (int n = 0; n < NT; n++) {
            const int4 * ap = reinterpret_cast<const int4 *>(a_q[n] + (size_t) s * 256);
#pragma unroll
            for (int c = 0; c < 4; c++) aq[n][c] = __ldg(ap + c);
#pragma unroll
            for (int e = 0; e < 2; e++) ax[n][e] = __ldg(reinterpret_cast<const int4 *>(a_d[n][e] + (size_t) s * 8));
        }
        const uint32_t dw0[4] = { (uint32_t) dq[0].x, (uint32_t) dq[0].y, (uint32_t) dq[0].z, (uint32_t) dq[0].w };   // row g:   d of blocks 2i | 2i+1
        const uint32_t dw1[4] = { (uint32_t) dq[1].x, (uint32_t) dq[1].y, (uint32_t) dq[1].z, (uint32_t) dq[1].w };   // row g+8
#pragma unroll
        for (int j = 0; j < 8; j++) {
            int a0, a1, a2, a3;
            if (WT == PIB_TYPE_Q4_0) {
                const int4     w  = wq[j >> 1];
                const uint32_t w0 = (uint32_t) ((j & 1) ? w.z : w.x), w1 = (uint32_t) ((j & 1) ? w.w : w.y);
                a0 = bk_sub8(w0 & 0x0F0F0F0Fu); a2 = bk_sub8((w0 >> 4) & 0x0F0F0F0Fu);
                a1 = bk_sub8(w1 & 0x0F0F0F0Fu); a3 = bk_sub8((w1 >> 4) & 0x0F0F0F0Fu);
            } else {
                const int4 w = wq[j & (NQ - 1)];
                a0 = w.x; a1 = w.y; a2 = w.z; a3 = w.w;
            }
            const float d0 = __half2float(__ushort_as_half((unsigned short) ((j & 1) ? dw0[j >> 1] >> 16 : dw0[j >> 1] & 0xffffu)));
            const float d1 = __half2float(__ushort_as_half((unsigned short) ((j & 1) ? dw1[j >> 1] >> 16 : dw1[j >> 1] & 0xffffu)));
#pragma unroll
            for (int n = 0; n < NT; n++) {
                const int4 bq = aq[n][j >> 1];
                int c[4] = { 0, 0, 0, 0 };
                bk_mma_s8(c, a0, a1, a2, a3, (j & 1) ? bq.z : bq.x, (j & 1) ? bq.w : bq.y);
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const int4     xv = ax[n][e];
                    const uint32_t xw = (uint32_t) ((j >> 1) == 0 ? xv.x : (j >> 1) == 1 ? xv.y : (j >> 1) == 2 ? xv.z : xv.w);
                    const float    dx = __half2float(__ushort_as_half((unsigned short) ((j & 1) ? xw >> 16 : xw & 0xffffu)));
                    acc[n][e]     = fmaf(__fmul_rn(d0, dx), (float) c[e], acc[n][e]);
                    acc[n][2 + e] = fmaf(__fmul_rn(d1, dx), (float) c[2 + e], acc[n][2 + e]);
                }
            }
        }
        if (sn < s1) {
#pragma unroll
            for (int c = 0; c < NQ; c++) wq[c] = wn[c];
            dq[0] = dn[0]; dq[1] = dn[1];
        }
    }

    __shared__ float red[BK_WARPS][NT * 8][16];
    __shared__ bool  s_last;
#pragma unroll
    for (int n = 0; n < NT; n++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
            red[warp][n * 8 + 2 * t + e][g]     = acc[n][e];
            red[warp][n * 8 + 2 * t + e][g + 8] = acc[n][2 + e];
        }
    __syncthreads();
    for (int i = threadIdx.x; i < NT * 8 * 16; i += BK_THREADS) {
        const int tok = i / 16, r = i % 16, row = T * 16 + r;
        if (tok < b && row < N) {
            float v = red[0][tok][r];
#pragma unroll
            for (int w2 = 1; w2 < BK_WARPS; w2++) v += red[w2][tok][r];
            if (nsplit == 1) y[(size_t) tok * N + row] = v;
            else partial[((size_t) split * b + tok) * N + row] = v;
        }
    }
    if (nsplit == 1) return;
    __syncthreads();
    if (threadIdx.x == 0) { __threadfence(); s_last = atomicAdd(&counters[T], 1u) == (unsigned) nsplit - 1; }
    __syncthreads();
    if (!s_last) return;
    for (int i = threadIdx.x; i < NT * 8 * 16; i += BK_THREADS) {
        const int tok = i / 16, r = i % 16, row = T * 16 + r;
        if (tok < b && row < N) {
            float v = 0.f;
            for (int sp = 0; sp < nsplit; sp++) v += __ldcg(partial + ((size_t) sp * b + tok) * N + row);
            y[(size_t) tok * N + row] = v;
        }
    }
    if (threadIdx.x == 0) counters[T] = 0;
}

// ---- launcher -----------------------------------------------------------------------------------------------------------
struct bk_scratch { float * partial = nullptr; unsigned int * counters = nullptr; };
static std::map<std::pair<int, cudaStream_t>, bk_scratch> g_bk_scratch;
static std::mutex                                        g_bk_mu;

// per (device, stream), sized once for every shape the split rule can produce (see gemv_mma.cu: ensure_scratch)
static int bk_ensure_scratch(cudaStream_t st, bk_scratch ** out) {
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lock(g_bk_mu);
    bk_scratch & s = g_bk_scratch[std::make_pair(dev, st)];
    if (!s.partial) {
        if (cudaMalloc(&s.partial, (size_t) 32 << 20) != cudaSuccess || cudaMalloc(&s.counters, 8192 * sizeof(unsigned int)) != cudaSuccess) {
            cudaGetLastError();
            return PIB_ERR_NOMEM;
        }
        cudaMemset(s.counters, 0, 8192 * sizeof(unsigned int));
    }
    *out = &s;
    return PIB_OK;
}

int pib_split_fill_target();         // gemv_mma.cu: CTA count below which a k split is worth its second phase

bool pib_gemv_blk32_supported(const pib_weight * w, int b) {
    return w && (w->type == PIB_TYPE_Q4_0 || w->type == PIB_TYPE_Q8_0) && w->tiles != nullptr && b >= 1 && b <= PIB_MAX_BATCH;
}

int pib_gemv_blk32(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!pib_gemv_blk32_supported(w, b)) { pib_set_error("pib_gemv: the tensor-core variant needs Q4_0 / Q8_0 weights with tiles"); return PIB_ERR_UNSUPP; }
    const pib_act_layout L   = pib_act_layout_of(w->type, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    const int            ngr = (int) ((w->K + 255) / 256), n_tiles = (int) ((w->N + 15) / 16);
    // k splits: enough CTAs to fill the machine when there are few row tiles; at least two records per warp
    int split = 1;
    while (n_tiles * split < pib_split_fill_target() && split * 2 * BK_WARPS <= ngr && split < 16) split *= 2;     // see gemv_mma.cu
    while (split > 1 && ((size_t) split * b * w->N * sizeof(float) > ((size_t) 32 << 20) || n_tiles > 8192)) split /= 2;
    bk_scratch * s = nullptr;
    if (split > 1 && bk_ensure_scratch(st, &s) != PIB_OK) { pib_set_error("pib_gemv: scratch allocation failed"); return PIB_ERR_NOMEM; }
    const dim3     grid(n_tiles, split);
    const int8_t * aqs = (const int8_t *) (act + L.qsp_off);
    const __half * adx = (const __half *) (act + L.bs8_off);          // dxp plane of the 32-element block formats
    float *        pp  = s ? s->partial : nullptr;
    unsigned int * cc  = s ? s->counters : nullptr;
    cudaError_t    e;
    if (w->type == PIB_TYPE_Q4_0) {
        if (b <= 8) e = pib_launch(gemv_blk32_kernel<PIB_TYPE_Q4_0, 1>, grid, dim3(BK_THREADS), 0, st, w->tiles, ngr, (int) w->N, aqs, adx, b, y_dev, pp, cc, cancel);
        else        e = pib_launch(gemv_blk32_kernel<PIB_TYPE_Q4_0, 2>, grid, dim3(BK_THREADS), 0, st, w->tiles, ngr, (int) w->N, aqs, adx, b, y_dev, pp, cc, cancel);
    } else {
        if (b <= 8) e = pib_launch(gemv_blk32_kernel<PIB_TYPE_Q8_0, 1>, grid, dim3(BK_THREADS), 0, st, w->tiles, ngr, (int) w->N, aqs, adx, b, y_dev, pp, cc, cancel);
        else        e = pib_launch(gemv_blk32_kernel<PIB_TYPE_Q8_0, 2>, grid, dim3(BK_THREADS), 0, st, w->tiles, ngr, (int) w->N, aqs, adx, b, y_dev, pp, cc, cancel);
    }
    PIB_CHECK(e);
    return PIB_OK;
}
