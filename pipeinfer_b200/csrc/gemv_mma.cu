// gemv_mma.cu -- variant 2 of the quantised matmul (K1) for Q4_K weights: fragment-major weight tiles streamed with
// coalesced 128-bit loads, int8 tensor-core dot products (mma.sync m16n8k32 s8), per-sub-block integer scaling,
// fp32 super-block accumulation, cross-warp + (optional) cross-CTA reduction in a fixed order.  Persistent CTAs; every warp
// keeps a private ring of tiles in shared memory filled by TMA bulk copies (cp.async.bulk + mbarrier).
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
#include <map>
#include <mutex>
#include <utility>
#include "pib_common.cuh"
#include "q8k_quant.cuh"

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

int pib_weight_build_tiles_blk32(pib_weight * w);     // gemv_blk32.cu
int pib_weight_build_tiles_q6k(pib_weight * w);       // gemv_q6k.cu
int pib_weight_build_tiles_q5k(pib_weight * w);       // gemv_q5k.cu
int pib_weight_build_tiles_q3k(pib_weight * w);       // gemv_q3k.cu

int pib_weight_build_tiles(pib_weight * w) {
    if (w->type == PIB_TYPE_Q4_0 || w->type == PIB_TYPE_Q8_0) return pib_weight_build_tiles_blk32(w);
    if (w->type == PIB_TYPE_Q6_K) return pib_weight_build_tiles_q6k(w);
    if (w->type == PIB_TYPE_Q5_K) return pib_weight_build_tiles_q5k(w);
    if (w->type == PIB_TYPE_Q3_K) return pib_weight_build_tiles_q3k(w);
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

// A unsigned: the high nibbles are used in place (q << 4, 0..240); the factor 16 is shifted out of the scaled sum
__device__ __forceinline__ void mma_u8s8(int (&c)[4], int a0, int a1, int a2, int a3, int b0, int b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
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

// ---- TMA bulk copy + mbarrier helpers (sm_90+ PTX; SASS: UBLKCP / SYNCS) ----------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_LOOP:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra WAIT_DONE;\n\tbra WAIT_LOOP;\n\tWAIT_DONE:\n\t}"
                 ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void * src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

#define MM_STAGES 3            // ring stages per warp
#define MM_CHUNK  2            // consecutive super-blocks per stage = per TMA bulk copy (2 x 2304 B)
#define MM_STAGE_BYTES (MM_CHUNK * MM_TILE_BYTES)

template <int NT>
struct act_frag {           // everything a lane needs from the activations for one super-block
    int4  q[NT][4];                 // chunk c: B fragments {lo.x, lo.y, hi.x, hi.y} of sub-blocks 2c / 2c+1, token column g
    int   mh[NT], ml[NT];           // B fragments of the min term: hi / lo bytes of the eight 32-wide block sums
    float dx[NT][2];                // activation scale of the tokens of C columns 2t, 2t+1
};

// Persistent kernel, warp-granular work.  Work item = (row tile T, k split p) = sb_per consecutive super-blocks of 16 rows;
// item i belongs to warp (i mod total warps): no cross-warp reduction, no CTA barrier in the main loop.  Every warp owns a
// private ring of MM_STAGES tiles in shared memory that one lane fills with 1-D TMA bulk copies (cp.async.bulk -> mbarrier
// complete_tx); the ring runs MM_STAGES super-blocks ahead across item boundaries and the activation fragments one ahead,
// so weight bytes keep streaming while results are written.  NT = number of 8-token n-tiles (1: b <= 8, 2: b <= 16).
// nsplit (a power of two, dividing nsb) > 1: fp32 partial sums [split][b][N] go through `partial`; the last warp to arrive
// for a row tile (per-tile counter) adds them in split order -- deterministic -- and resets the counter.
template <int NT>
__global__ void __launch_bounds__(MM_THREADS, NT == 1 ? 4 : 3)
gemv_q4k_mma_kernel(const uint8_t * __restrict__ tiles, int nsb, int n_tiles, int lg_split, int N, int K,
                    const int8_t * __restrict__ aqs, const float * __restrict__ ad, const int8_t * __restrict__ abs8, int b,
                    float * __restrict__ y, float * __restrict__ partial, unsigned int * __restrict__ counters,
                    const int * __restrict__ cancel) {
    if (PIB_CANCELLED(cancel)) {
        // A cancelled warp skips the math but still arrives for every item it owns: the flag can flip while the grid is
        // starting, and a tile whose splits did not all arrive would leave its counter non-zero for the next launch.
        if (lg_split > 0 && (threadIdx.x & 31) == 0) {
            const unsigned nsp = 1u << lg_split;
            for (int item = (int) blockIdx.x * MM_WARPS + (int) (threadIdx.x >> 5); item < (n_tiles << lg_split); item += (int) gridDim.x * MM_WARPS)
                if (atomicAdd(&counters[item >> lg_split], 1u) == nsp - 1) counters[item >> lg_split] = 0;
        }
        return;
    }
    extern __shared__ __align__(128) uint8_t dsm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    uint8_t *      ring = dsm + (size_t) warp * MM_STAGES * MM_STAGE_BYTES;
    const uint32_t ring_a = smem_u32(ring);
    const uint32_t bar_a  = smem_u32(dsm + (size_t) MM_WARPS * MM_STAGES * MM_STAGE_BYTES) + (uint32_t) warp * MM_STAGES * 8;
    const int nsplit = 1 << lg_split, pmask = nsplit - 1;
    const int n_items = n_tiles << lg_split;
    const int sb_per  = nsb >> lg_split;
    const int gw = (int) blockIdx.x * MM_WARPS + warp, wtot = (int) gridDim.x * MM_WARPS;

    if (lane == 0) {
        for (int i = 0; i < MM_STAGES; i++) mbar_init(bar_a + i * 8, 1);
        fence_async_smem();
    }
    __syncwarp();

    // three cursors over this warp's sequence of super-blocks: weights (TMA), activations (registers), compute
    struct cur_t { int item, sl, s; };           // s: super-block index inside the row, sl: inside the item
    auto cur_init = [&](cur_t & c) { c.item = gw; c.sl = 0; c.s = (gw & pmask) * sb_per; };
    auto cur_next = [&](cur_t & c) {
        c.sl++; c.s++;
        if (c.sl == sb_per) { c.sl = 0; c.item += wtot; c.s = (c.item & pmask) * sb_per; }
    };
    cur_t wc, ac, cc;
    cur_init(wc); cur_init(ac); cur_init(cc);
    uint32_t issued = 0, consumed = 0;
    // the weight cursor moves in chunks of up to MM_CHUNK consecutive super-blocks of one item: one bulk copy each
    auto issue = [&]() {
        const int      csz   = min(MM_CHUNK, sb_per - wc.sl);
        const uint32_t bytes = (uint32_t) csz * MM_TILE_BYTES;
        if (lane == 0) {
            const uint32_t st = issued % MM_STAGES;
            mbar_expect_tx(bar_a + st * 8, bytes);
            bulk_g2s(ring_a + st * MM_STAGE_BYTES, tiles + ((size_t) (wc.item >> lg_split) * nsb + wc.s) * MM_TILE_BYTES, bytes,
                     bar_a + st * 8);
        }
        issued++;
        for (int i = 0; i < csz; i++) cur_next(wc);
    };
    for (int i = 0; i < MM_STAGES && wc.item < n_items; i++) issue();

    // per-lane activation base pointers (token column g of each n-tile; C columns 2t, 2t+1).  Token indices are clamped
    // to b - 1: padded columns compute on a duplicate of the last token and are never stored.
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
    auto load_act = [&](act_frag<NT> & f, int s) {
#pragma unroll
        for (int n = 0; n < NT; n++) {
            const int4 * ap = reinterpret_cast<const int4 *>(a_qs[n] + (size_t) s * 256);
#pragma unroll
            for (int c = 0; c < 4; c++) f.q[n][c] = __ldg(ap + c);
            const int * bp = reinterpret_cast<const int *>(a_b8[n] + (size_t) s * 16);
            f.mh[n] = __ldg(bp);
            f.ml[n] = __ldg(bp + 2);
#pragma unroll
            for (int e = 0; e < 2; e++) f.dx[n][e] = __ldg(a_dx[n][e] + s);
        }
    };
    act_frag<NT> A0;
    if (ac.item < n_items) { load_act(A0, ac.s); cur_next(ac); }

    float acc[NT][4];
#pragma unroll
    for (int n = 0; n < NT; n++) { acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f; }

    while (cc.item < n_items) {
        act_frag<NT> A1;
        const bool more = ac.item < n_items;
        if (more) { load_act(A1, ac.s); cur_next(ac); }         // activations of the next super-block: in flight during this one

        const uint32_t st  = consumed % MM_STAGES;
        const int      cin = cc.sl % MM_CHUNK;                   // position inside the chunk
        if (cin == 0) mbar_wait(bar_a + st * 8, (consumed / MM_STAGES) & 1);
        const uint8_t * p = ring + (size_t) st * MM_STAGE_BYTES + (size_t) cin * MM_TILE_BYTES;
        uint4 wq[4];
#pragma unroll
        for (int c = 0; c < 4; c++) wq[c] = *reinterpret_cast<const uint4 *>(p + c * 512 + lane * 16);
        const q4k_hdr H0 = decode_hdr(*reinterpret_cast<const uint4 *>(p + 2048 + g * 16));
        const q4k_hdr H1 = decode_hdr(*reinterpret_cast<const uint4 *>(p + 2048 + (g + 8) * 16));
        if (cin == MM_CHUNK - 1 || cc.sl == sb_per - 1) {   // last super-block of the chunk: the stage is drained
            __syncwarp();                                   // every lane has read its part
            if (wc.item < n_items) { if (lane == 0) fence_async_smem(); issue(); }
            consumed++;
        }

        int sumi[NT][4], sumh[NT][4];        // low-nibble sub-blocks; high-nibble sub-blocks (x 16)
#pragma unroll
        for (int n = 0; n < NT; n++)
#pragma unroll
            for (int i = 0; i < 4; i++) { sumi[n][i] = 0; sumh[n][i] = 0; }
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const uint4 w = wq[c];            // {row g: x, y ; row g+8: z, w}
            const int j0 = 2 * c, j1 = 2 * c + 1;
            const int sc0_lo = (int) (((j0 < 4 ? H0.sc_lo : H0.sc_hi) >> (8 * (j0 & 3))) & 0xff);
            const int sc1_lo = (int) (((j0 < 4 ? H1.sc_lo : H1.sc_hi) >> (8 * (j0 & 3))) & 0xff);
            const int sc0_hi = (int) (((j1 < 4 ? H0.sc_lo : H0.sc_hi) >> (8 * (j1 & 3))) & 0xff);
            const int sc1_hi = (int) (((j1 < 4 ? H1.sc_lo : H1.sc_hi) >> (8 * (j1 & 3))) & 0xff);
#pragma unroll
            for (int n = 0; n < NT; n++) {
                int c_lo[4] = { 0, 0, 0, 0 }, c_hi[4] = { 0, 0, 0, 0 };
                mma_u8s8(c_lo, (int) (w.x & 0x0F0F0F0Fu), (int) (w.z & 0x0F0F0F0Fu), (int) (w.y & 0x0F0F0F0Fu), (int) (w.w & 0x0F0F0F0Fu),
                         A0.q[n][c].x, A0.q[n][c].y);
                mma_u8s8(c_hi, (int) (w.x & 0xF0F0F0F0u), (int) (w.z & 0xF0F0F0F0u), (int) (w.y & 0xF0F0F0F0u), (int) (w.w & 0xF0F0F0F0u),
                         A0.q[n][c].z, A0.q[n][c].w);
                sumi[n][0] += sc0_lo * c_lo[0]; sumh[n][0] += sc0_hi * c_hi[0];
                sumi[n][1] += sc0_lo * c_lo[1]; sumh[n][1] += sc0_hi * c_hi[1];
                sumi[n][2] += sc1_lo * c_lo[2]; sumh[n][2] += sc1_hi * c_hi[2];
                sumi[n][3] += sc1_lo * c_lo[3]; sumh[n][3] += sc1_hi * c_hi[3];
            }
        }
        // min term sum_j m_j * bsum_j on the tensor core as well: bsum_j = 64 * hi_j + lo_j (bytes prepared by the
        // activation quantiser); k-slots 0..7 carry j, the rest are zero
        const int am0 = t == 0 ? (int) H0.m_lo : t == 1 ? (int) H0.m_hi : 0;
        const int am1 = t == 0 ? (int) H1.m_lo : t == 1 ? (int) H1.m_hi : 0;
#pragma unroll
        for (int n = 0; n < NT; n++) {
            int c_mh[4] = { 0, 0, 0, 0 }, c_ml[4] = { 0, 0, 0, 0 };
            const int bmh = t < 2 ? A0.mh[n] : 0, bml = t < 2 ? A0.ml[n] : 0;
            mma_s8(c_mh, am0, am1, 0, 0, bmh, 0);
            mma_s8(c_ml, am0, am1, 0, 0, bml, 0);
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const float dx = A0.dx[n][e];
                acc[n][e]     = fmaf(__fmul_rn(dx, H0.d), (float) (sumi[n][e] + (sumh[n][e] >> 4)), acc[n][e]);
                acc[n][e]     = fmaf(__fmul_rn(-dx, H0.dmin), (float) (64 * c_mh[e] + c_ml[e]), acc[n][e]);
                acc[n][2 + e] = fmaf(__fmul_rn(dx, H1.d), (float) (sumi[n][2 + e] + (sumh[n][2 + e] >> 4)), acc[n][2 + e]);
                acc[n][2 + e] = fmaf(__fmul_rn(-dx, H1.dmin), (float) (64 * c_mh[2 + e] + c_ml[2 + e]), acc[n][2 + e]);
            }
        }
        if (more) A0 = A1;

        if (cc.sl == sb_per - 1) {                          // the item is complete: 16 rows x b tokens of (partial) sums
            const int T = cc.item >> lg_split, split = cc.item & pmask;
            const int r0 = T * 16 + g, r1 = r0 + 8;
#pragma unroll
            for (int n = 0; n < NT; n++)
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const int tok = n * 8 + 2 * t + e;
                    if (tok < b) {
                        float * o = nsplit == 1 ? y + (size_t) tok * N : partial + ((size_t) split * b + tok) * N;
                        if (r0 < N) o[r0] = acc[n][e];
                        if (r1 < N) o[r1] = acc[n][2 + e];
                    }
                    acc[n][e] = 0.f; acc[n][2 + e] = 0.f;
                }
            if (nsplit > 1) {
                __threadfence();
                __syncwarp();
                unsigned int old = 0;
                if (lane == 0) old = atomicAdd(&counters[T], 1u);
                old = __shfl_sync(0xffffffffu, old, 0);
                if (old == (unsigned) nsplit - 1) {         // last to arrive: add the splits in order
                    __threadfence();
#pragma unroll
                    for (int n = 0; n < NT; n++)
#pragma unroll
                        for (int e = 0; e < 2; e++) {
                            const int tok = n * 8 + 2 * t + e;
                            if (tok < b) {
                                float v0 = 0.f, v1 = 0.f;
                                for (int sp = 0; sp < nsplit; sp++) {
                                    const float * q = partial + ((size_t) sp * b + tok) * N;
                                    if (r0 < N) v0 += __ldcg(q + r0);
                                    if (r1 < N) v1 += __ldcg(q + r1);
                                }
                                if (r0 < N) y[(size_t) tok * N + r0] = v0;
                                if (r1 < N) y[(size_t) tok * N + r1] = v1;
                            }
                        }
                    __syncwarp();
                    if (lane == 0) counters[T] = 0;         // self-cleaning: ready for the next launch
                }
            }
        }
        cur_next(cc);
    }
}

// the arithmetic of one (16-row tile, super-block) record for NT n-tiles of 8 tokens, shared by the kernels below
// (ACT_SMEM: the activation planes live in shared memory -- the prologue-fused kernel -- and are read with ordinary loads)
template <int NT, bool ACT_SMEM = false>
__device__ __forceinline__ void q4k_record(const int4 (&wq)[4], const int4 (&hq)[2], int s, int t, const int8_t * const (&a_qs)[NT],
                                           const int8_t * const (&a_b8)[NT], const float * const (&a_dx)[NT][2], float (&acc)[NT][4]) {
    const q4k_hdr H0 = decode_hdr(make_uint4((uint32_t) hq[0].x, (uint32_t) hq[0].y, (uint32_t) hq[0].z, (uint32_t) hq[0].w));
    const q4k_hdr H1 = decode_hdr(make_uint4((uint32_t) hq[1].x, (uint32_t) hq[1].y, (uint32_t) hq[1].z, (uint32_t) hq[1].w));
    int4  aq[NT][4];
    int   amh[NT], aml[NT];
    float adx[NT][2];
#pragma unroll
    for (int n = 0; n < NT; n++) {
        const int4 * ap = reinterpret_cast<const int4 *>(a_qs[n] + (size_t) s * 256);
#pragma unroll
        for (int c = 0; c < 4; c++) aq[n][c] = ACT_SMEM ? ap[c] : __ldg(ap + c);
        const int * bp = reinterpret_cast<const int *>(a_b8[n] + (size_t) s * 16);
        amh[n] = ACT_SMEM ? bp[0] : __ldg(bp);
        aml[n] = ACT_SMEM ? bp[2] : __ldg(bp + 2);
#pragma unroll
        for (int e = 0; e < 2; e++) adx[n][e] = ACT_SMEM ? a_dx[n][e][s] : __ldg(a_dx[n][e] + s);
    }
    int sumi[NT][4], sumh[NT][4];
#pragma unroll
    for (int n = 0; n < NT; n++)
#pragma unroll
        for (int i = 0; i < 4; i++) { sumi[n][i] = 0; sumh[n][i] = 0; }
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint32_t wx = (uint32_t) wq[c].x, wy = (uint32_t) wq[c].y, wz = (uint32_t) wq[c].z, ww = (uint32_t) wq[c].w;
        const int j0 = 2 * c, j1 = 2 * c + 1;
        const int sc0_lo = (int) __byte_perm(j0 < 4 ? H0.sc_lo : H0.sc_hi, 0, 0x4440 + (j0 & 3));
        const int sc1_lo = (int) __byte_perm(j0 < 4 ? H1.sc_lo : H1.sc_hi, 0, 0x4440 + (j0 & 3));
        const int sc0_hi = (int) __byte_perm(j1 < 4 ? H0.sc_lo : H0.sc_hi, 0, 0x4440 + (j1 & 3));
        const int sc1_hi = (int) __byte_perm(j1 < 4 ? H1.sc_lo : H1.sc_hi, 0, 0x4440 + (j1 & 3));
#pragma unroll
        for (int n = 0; n < NT; n++) {
            int c_lo[4] = { 0, 0, 0, 0 }, c_hi[4] = { 0, 0, 0, 0 };
            mma_u8s8(c_lo, (int) (wx & 0x0F0F0F0Fu), (int) (wz & 0x0F0F0F0Fu), (int) (wy & 0x0F0F0F0Fu), (int) (ww & 0x0F0F0F0Fu),
                     aq[n][c].x, aq[n][c].y);
            mma_u8s8(c_hi, (int) (wx & 0xF0F0F0F0u), (int) (wz & 0xF0F0F0F0u), (int) (wy & 0xF0F0F0F0u), (int) (ww & 0xF0F0F0F0u),
                     aq[n][c].z, aq[n][c].w);
            sumi[n][0] += sc0_lo * c_lo[0]; sumh[n][0] += sc0_hi * c_hi[0];
            sumi[n][1] += sc0_lo * c_lo[1]; sumh[n][1] += sc0_hi * c_hi[1];
            sumi[n][2] += sc1_lo * c_lo[2]; sumh[n][2] += sc1_hi * c_hi[2];
            sumi[n][3] += sc1_lo * c_lo[3]; sumh[n][3] += sc1_hi * c_hi[3];
        }
    }
    const int am0 = t == 0 ? (int) H0.m_lo : t == 1 ? (int) H0.m_hi : 0;
    const int am1 = t == 0 ? (int) H1.m_lo : t == 1 ? (int) H1.m_hi : 0;
#pragma unroll
    for (int n = 0; n < NT; n++) {
        int c_mh[4] = { 0, 0, 0, 0 }, c_ml[4] = { 0, 0, 0, 0 };
        const int bmh = t < 2 ? amh[n] : 0, bml = t < 2 ? aml[n] : 0;
        mma_s8(c_mh, am0, am1, 0, 0, bmh, 0);
        mma_s8(c_ml, am0, am1, 0, 0, bml, 0);
#pragma unroll
        for (int e = 0; e < 2; e++) {
            const float dx = adx[n][e];
            acc[n][e]     = fmaf(__fmul_rn(dx, H0.d), (float) (sumi[n][e] + (sumh[n][e] >> 4)), acc[n][e]);
            acc[n][e]     = fmaf(__fmul_rn(-dx, H0.dmin), (float) (64 * c_mh[e] + c_ml[e]), acc[n][e]);
            acc[n][2 + e] = fmaf(__fmul_rn(dx, H1.d), (float) (sumi[n][2 + e] + (sumh[n][2 + e] >> 4)), acc[n][2 + e]);
            acc[n][2 + e] = fmaf(__fmul_rn(-dx, H1.dmin), (float) (64 * c_mh[2 + e] + c_ml[2 + e]), acc[n][2 + e]);
        }
    }
}

// ---- direct variant: one CTA per (row tile, k split), the four warps interleave the super-blocks, weights prefetched one
// super-block ahead into registers with streaming 128-bit loads (no shared-memory staging); same arithmetic core -------------
#ifndef MM_L2_AHEAD
#define MM_L2_AHEAD 0
#endif
template <int NT>
__global__ void __launch_bounds__(MM_THREADS, NT == 1 ? 4 : 3)
gemv_q4k_direct_kernel(const uint8_t * __restrict__ tiles, int nsb, int N, int K, const int8_t * __restrict__ aqs,
                       const float * __restrict__ ad, const int8_t * __restrict__ abs8, int b, float * __restrict__ y,
                       float * __restrict__ partial, unsigned int * __restrict__ counters, const int * __restrict__ cancel) {
    if (PIB_CANCELLED(cancel)) {
        pib_pdl_wait();
        // cancelled: no math, but the arrival protocol of the k split still runs (see the persistent kernel above)
        if (gridDim.y > 1 && threadIdx.x == 0 && atomicAdd(&counters[blockIdx.x], 1u) == gridDim.y - 1) counters[blockIdx.x] = 0;
        return;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int T = blockIdx.x;
    const int nsplit = gridDim.y, split = blockIdx.y;
    const int sb_per = (nsb + nsplit - 1) / nsplit;
    const int sb0 = split * sb_per, sb1 = min(nsb, sb0 + sb_per);
    const uint8_t * tbase = tiles + (size_t) T * nsb * MM_TILE_BYTES + lane * 16;
    const uint8_t * hbase = tiles + (size_t) T * nsb * MM_TILE_BYTES + 2048 + g * 16;

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

    int4 wq[4], hq[2];
    int  s = sb0 + warp;
    if (s < sb1) {
#pragma unroll
        for (int c = 0; c < 4; c++) wq[c] = ld_stream_v4(tbase + (size_t) s * MM_TILE_BYTES + c * 512);
        hq[0] = ld_stream_v4(hbase + (size_t) s * MM_TILE_BYTES);
        hq[1] = ld_stream_v4(hbase + (size_t) s * MM_TILE_BYTES + 128);
    }
    // the weights above are immutable, so their loads were issued before waiting for the kernel that produces the activations
    pib_pdl_wait();
    pib_pdl_trigger();
    for (; s < sb1; s += MM_WARPS) {
        int4 wn[4], hn[2];
        const int sn = s + MM_WARPS;
#if MM_L2_AHEAD > 0
        {   // keep more bytes in flight than the registers can hold: pull the record MM_L2_AHEAD iterations ahead into L2
            const int sp = s + (1 + MM_L2_AHEAD) * MM_WARPS;
            if (sp < sb1 && lane < MM_TILE_BYTES / 128)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(tiles + ((size_t) T * nsb + sp) * MM_TILE_BYTES + lane * 128));
        }
#endif
        if (sn < sb1) {
#pragma unroll
            for (int c = 0; c < 4; c++) wn[c] = ld_stream_v4(tbase + (size_t) sn * MM_TILE_BYTES + c * 512);
            hn[0] = ld_stream_v4(hbase + (size_t) sn * MM_TILE_BYTES);
            hn[1] = ld_stream_v4(hbase + (size_t) sn * MM_TILE_BYTES + 128);
        }
        q4k_record<NT>(wq, hq, s, t, a_qs, a_b8, a_dx, acc);
        if (sn < sb1) {
#pragma unroll
            for (int c = 0; c < 4; c++) wq[c] = wn[c];
            hq[0] = hn[0]; hq[1] = hn[1];
        }
    }

    // cross-warp reduction in warp order, then (k split) cross-CTA reduction in split order by the last CTA to arrive
    __shared__ float red[MM_WARPS][NT * 8][16];
    __shared__ bool  s_last;
#pragma unroll
    for (int n = 0; n < NT; n++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
            red[warp][n * 8 + 2 * t + e][g]     = acc[n][e];
            red[warp][n * 8 + 2 * t + e][g + 8] = acc[n][2 + e];
        }
    __syncthreads();
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
    __syncthreads();
    if (threadIdx.x == 0) { __threadfence(); s_last = atomicAdd(&counters[T], 1u) == (unsigned) nsplit - 1; }
    __syncthreads();
    if (!s_last) return;
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

// ---- ring variant (default): same decomposition and arithmetic as the direct kernel, but every warp streams its records through
// a private ring in shared memory filled by cp.async (16 bytes per lane and copy, L2 only): RG_STAGES - 1 records in flight per warp
// instead of one register-prefetched record.  With one wave of CTAs (attn_out, down: 512 row tiles for 592 slots) a warp's time is
// a chain of record latencies, and that chain is what the deeper prefetch shortens.  WARPS = warps per CTA (interleaving the
// super-blocks of the tile): 8 for single-wave shapes, 4 otherwise.
#define RG_STAGES 4
template <int NT, int WARPS, int STAGES>
__global__ void __launch_bounds__(WARPS * 32, NT == 1 ? (WARPS == 4 ? 5 : 2) : (WARPS == 4 ? 3 : 1))
gemv_q4k_ring_kernel(const uint8_t * __restrict__ tiles, int nsb, int N, int K, const int8_t * __restrict__ aqs,
                     const float * __restrict__ ad, const int8_t * __restrict__ abs8, int b, float * __restrict__ y,
                     float * __restrict__ partial, unsigned int * __restrict__ counters, const int * __restrict__ cancel) {
    if (PIB_CANCELLED(cancel)) {
        pib_pdl_wait();
        if (gridDim.y > 1 && threadIdx.x == 0 && atomicAdd(&counters[blockIdx.x], 1u) == gridDim.y - 1) counters[blockIdx.x] = 0;
        return;
    }
    extern __shared__ __align__(16) uint8_t rg_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int T = blockIdx.x;
    const int nsplit = gridDim.y, split = blockIdx.y;
    const int sb_per = (nsb + nsplit - 1) / nsplit;
    const int sb0 = split * sb_per, sb1 = min(nsb, sb0 + sb_per);
    const uint8_t * tbase = tiles + (size_t) T * nsb * MM_TILE_BYTES + lane * 16;
    uint8_t *       ring  = rg_smem + warp * (STAGES * MM_TILE_BYTES);
    const uint32_t  ring_a = smem_u32(ring) + lane * 16;

    // prefetch cursor: this warp's records sb0 + warp, + WARPS, ...
    int ps = sb0 + warp;
    auto issue = [&](int stage) {
        if (ps < sb1) {
            const uint8_t * src = tbase + (size_t) ps * MM_TILE_BYTES;
            const uint32_t  dst = ring_a + stage * MM_TILE_BYTES;
#pragma unroll
            for (int c = 0; c < 4; c++) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + c * 512), "l"(src + c * 512) : "memory");
            if (lane < 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 2048), "l"(src + 2048) : "memory");
            ps += WARPS;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // the weights are immutable: their copies start before the wait on the kernel that produces the activations
#pragma unroll
    for (int i = 0; i < STAGES - 1; i++) issue(i);

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
    for (int s = sb0 + warp; s < sb1; s += WARPS) {
        __syncwarp();                                           // every lane has read the stage about to be refilled
        issue((stage + STAGES - 1) % STAGES);
        asm volatile("cp.async.wait_group %0;" ::"n"(STAGES - 1) : "memory");
        __syncwarp();
        const uint8_t * rs = ring + stage * MM_TILE_BYTES;
        int4 wq[4], hq[2];
#pragma unroll
        for (int c = 0; c < 4; c++) wq[c] = *reinterpret_cast<const int4 *>(rs + c * 512 + lane * 16);
        hq[0] = *reinterpret_cast<const int4 *>(rs + 2048 + g * 16);
        hq[1] = *reinterpret_cast<const int4 *>(rs + 2048 + (g + 8) * 16);
        q4k_record<NT>(wq, hq, s, t, a_qs, a_b8, a_dx, acc);
        stage = stage + 1 == STAGES ? 0 : stage + 1;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                                            // the rings are done with: their memory becomes the reduction buffer

    // cross-warp reduction in warp order, then (k split) cross-CTA reduction in split order by the last CTA to arrive
    float (*red)[NT * 8][16] = reinterpret_cast<float (*)[NT * 8][16]>(rg_smem);
    __shared__ bool s_last;
#pragma unroll
    for (int n = 0; n < NT; n++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
            red[warp][n * 8 + 2 * t + e][g]     = acc[n][e];
            red[warp][n * 8 + 2 * t + e][g + 8] = acc[n][2 + e];
        }
    __syncthreads();
    for (int i = threadIdx.x; i < NT * 8 * 16; i += WARPS * 32) {
        const int tok = i / 16, r = i % 16, row = T * 16 + r;
        if (tok < b && row < N) {
            float v = red[0][tok][r];
#pragma unroll
            for (int w2 = 1; w2 < WARPS; w2++) v += red[w2][tok][r];
            if (nsplit == 1) y[(size_t) tok * N + row] = v;
            else partial[((size_t) split * b + tok) * N + row] = v;
        }
    }
    if (nsplit == 1) return;
    __syncthreads();
    if (threadIdx.x == 0) { __threadfence(); s_last = atomicAdd(&counters[T], 1u) == (unsigned) nsplit - 1; }
    __syncthreads();
    if (!s_last) return;
    for (int i = threadIdx.x; i < NT * 8 * 16; i += WARPS * 32) {
        const int tok = i / 16, r = i % 16, row = T * 16 + r;
        if (tok < b && row < N) {
            float v = 0.f;
            for (int sp = 0; sp < nsplit; sp++) v += __ldcg(partial + ((size_t) sp * b + tok) * N + row);
            y[(size_t) tok * N + row] = v;
        }
    }
    if (threadIdx.x == 0) counters[T] = 0;          // self-cleaning: ready for the next launch
}

// ---- prologue-fused variant (small models): the GEMV quantises its own activations ------------------------------------------------
// A model as small as the 1.1B draft is bound by the number of launches, not by bytes: every kernel costs a ~1.7 us hop however
// little it does.  Here the kernel that used to stand in front of a matmul -- residual add + RMSNorm + quantise, plain quantise, or
// SiLU(gate) * up + quantise -- runs inside the matmul: every CTA (one 16-row tile, as in the ring kernel) forms the quantised
// activation planes of all b tokens in its own shared memory (K <= 8192 values per token: a few microseconds of redundant work per
// CTA instead of a launch), then streams its records exactly as the ring kernel does -- same decomposition, same order of fp32
// additions, same bytes out.  The quantiser is the one of act_quant.cu (q8k_quant.cuh).  CTA 0 also writes the residual sum.
enum { GP_PLAIN = 0, GP_RMSNORM = 1, GP_SILU_MUL = 2 };
#define GP_WARPS 4
template <int PRO>
__global__ void __launch_bounds__(GP_WARPS * 32, 4)
gemv_q4k_pro_kernel(const uint8_t * __restrict__ tiles, int nsb, int N, int K, int b, float * __restrict__ y,
                    const float * __restrict__ x, const float * __restrict__ x2, const float * __restrict__ addend, float * __restrict__ sum_out,
                    int64_t xstride, float eps, const uint16_t * __restrict__ lut, const int * __restrict__ cancel) {
    if (PIB_CANCELLED(cancel)) { pib_pdl_wait(); return; }
    extern __shared__ __align__(16) uint8_t rg_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int T = blockIdx.x;
    constexpr int STAGES = RG_STAGES;
    const uint8_t * tbase  = tiles + (size_t) T * nsb * MM_TILE_BYTES + lane * 16;
    uint8_t *       ring   = rg_smem + warp * (STAGES * MM_TILE_BYTES);
    const uint32_t  ring_a = smem_u32(ring) + lane * 16;
    // activation planes behind the rings: qsp [b][K] int8, bs8 [b][nsb][16] int8, d [b][nsb] float
    int8_t * s_qsp = reinterpret_cast<int8_t *>(rg_smem + GP_WARPS * STAGES * MM_TILE_BYTES);
    int8_t * s_bs8 = s_qsp + (size_t) b * K;
    float *  s_d   = reinterpret_cast<float *>(s_bs8 + (size_t) b * nsb * 16);
    __shared__ double s_red[GP_WARPS][8];
    int ps = warp;
    auto issue = [&](int stage) {
        if (ps < nsb) {
            const uint8_t * src = tbase + (size_t) ps * MM_TILE_BYTES;
            const uint32_t  dst = ring_a + stage * MM_TILE_BYTES;
#pragma unroll
            for (int c = 0; c < 4; c++) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + c * 512), "l"(src + c * 512) : "memory");
            if (lane < 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 2048), "l"(src + 2048) : "memory");
            ps += GP_WARPS;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int i = 0; i < STAGES - 1; i++) issue(i);
    pib_pdl_wait();
    pib_pdl_trigger();

    // ---- prologue: the activation planes of all b tokens ---------------------------------------------------------------------
    float scale[8];
#pragma unroll
    for (int tk = 0; tk < 8; tk++) scale[tk] = 1.f;
    if (PRO == GP_RMSNORM) {
        // sum of squares per token in double (ggml.c:9028), residual add in front when there is an addend; CTA 0 stores the sums
        const int n4 = K / 4;
#pragma unroll 1
        for (int tk = 0; tk < b; tk++) {
            const float4 * x4 = reinterpret_cast<const float4 *>(x + (size_t) tk * xstride);
            const float4 * a4 = addend ? reinterpret_cast<const float4 *>(addend + (size_t) tk * K) : nullptr;
            float4 *       o4 = (sum_out && T == 0) ? reinterpret_cast<float4 *>(sum_out + (size_t) tk * K) : nullptr;
            double sm = 0.0;
#pragma unroll 4
            for (int i = threadIdx.x; i < n4; i += GP_WARPS * 32) {
                float4 v = x4[i];
                if (a4) {
                    const float4 a = a4[i];
                    v.x = __fadd_rn(v.x, a.x); v.y = __fadd_rn(v.y, a.y); v.z = __fadd_rn(v.z, a.z); v.w = __fadd_rn(v.w, a.w);
                    if (o4) o4[i] = v;
                }
                sm += (double) __fmul_rn(v.x, v.x); sm += (double) __fmul_rn(v.y, v.y); sm += (double) __fmul_rn(v.z, v.z); sm += (double) __fmul_rn(v.w, v.w);
            }
            sm = warp_sum_d(sm);
            if (lane == 0) s_red[warp][tk] = sm;
        }
        __syncthreads();
#pragma unroll
        for (int tk = 0; tk < 8; tk++)
            if (tk < b) {
                double tt = 0.0;
#pragma unroll
                for (int w = 0; w < GP_WARPS; w++) tt += s_red[w][tk];
                scale[tk] = 1.0f / sqrtf((float) (tt / (double) K) + eps);
            }
    }
    for (int task = warp; task < b * nsb; task += GP_WARPS) {
        const int tk = task / nsb, sb = task - tk * nsb;
        const int64_t i0 = (int64_t) sb * 256 + lane * 8;
        const float * xr = x + (size_t) tk * xstride + i0;
        const float4 xa = *reinterpret_cast<const float4 *>(xr), xb = *reinterpret_cast<const float4 *>(xr + 4);
        float xv[8] = { xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w };
        float v[8];
        if (PRO == GP_RMSNORM) {
            if (addend) {
                const float4 aa = *reinterpret_cast<const float4 *>(addend + (size_t) tk * K + i0), ab = *reinterpret_cast<const float4 *>(addend + (size_t) tk * K + i0 + 4);
                xv[0] = __fadd_rn(xv[0], aa.x); xv[1] = __fadd_rn(xv[1], aa.y); xv[2] = __fadd_rn(xv[2], aa.z); xv[3] = __fadd_rn(xv[3], aa.w);
                xv[4] = __fadd_rn(xv[4], ab.x); xv[5] = __fadd_rn(xv[5], ab.y); xv[6] = __fadd_rn(xv[6], ab.z); xv[7] = __fadd_rn(xv[7], ab.w);
            }
            const float4 wa = *reinterpret_cast<const float4 *>(x2 + i0), wb = *reinterpret_cast<const float4 *>(x2 + i0 + 4);
            const float  w8[8] = { wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w };
            float sc = 1.f;
#pragma unroll
            for (int q = 0; q < 8; q++) if (q == tk) sc = scale[q];
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = __fmul_rn(__fmul_rn(xv[j], sc), w8[j]);
        } else if (PRO == GP_SILU_MUL) {
            const float * ur = x2 + (size_t) tk * xstride + i0;
            const float4 ua = *reinterpret_cast<const float4 *>(ur), ub = *reinterpret_cast<const float4 *>(ur + 4);
            const float  u8[8] = { ua.x, ua.y, ua.z, ua.w, ub.x, ub.y, ub.z, ub.w };
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = __fmul_rn(pib_h2f(lut[pib_f2h(xv[j])]), u8[j]);
        } else {
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = xv[j];
        }
        q8k_quantise_sb(v, lane, nullptr, s_qsp + (size_t) tk * K + sb * 256, nullptr, s_bs8 + ((size_t) tk * nsb + sb) * 16, nullptr, nullptr,
                        s_d + (size_t) tk * nsb + sb);
    }
    __syncthreads();

    // ---- main loop: the ring kernel's, activations from shared memory ---------------------------------------------------------------
    float acc[1][4] = { { 0.f, 0.f, 0.f, 0.f } };
    const int8_t * a_qs[1];
    const int8_t * a_b8[1];
    const float *  a_dx[1][2];
    {
        const int tok = min(g, b - 1);
        a_qs[0] = s_qsp + (size_t) tok * K + t * 64;
        a_b8[0] = s_bs8 + (size_t) tok * nsb * 16 + (t & 1) * 4;
#pragma unroll
        for (int e = 0; e < 2; e++) a_dx[0][e] = s_d + (size_t) min(2 * t + e, b - 1) * nsb;
    }
    int stage = 0;
    for (int s = warp; s < nsb; s += GP_WARPS) {
        __syncwarp();
        issue((stage + STAGES - 1) % STAGES);
        asm volatile("cp.async.wait_group %0;" ::"n"(STAGES - 1) : "memory");
        __syncwarp();
        const uint8_t * rs = ring + stage * MM_TILE_BYTES;
        int4 wq[4], hq[2];
#pragma unroll
        for (int c = 0; c < 4; c++) wq[c] = *reinterpret_cast<const int4 *>(rs + c * 512 + lane * 16);
        hq[0] = *reinterpret_cast<const int4 *>(rs + 2048 + g * 16);
        hq[1] = *reinterpret_cast<const int4 *>(rs + 2048 + (g + 8) * 16);
        q4k_record<1, true>(wq, hq, s, t, a_qs, a_b8, a_dx, acc);
        stage = stage + 1 == STAGES ? 0 : stage + 1;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                                            // the rings are done with: their memory becomes the reduction buffer
    float (*red)[8][16] = reinterpret_cast<float (*)[8][16]>(rg_smem);
#pragma unroll
    for (int e = 0; e < 2; e++) {
        red[warp][2 * t + e][g]     = acc[0][e];
        red[warp][2 * t + e][g + 8] = acc[0][2 + e];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 8 * 16; i += GP_WARPS * 32) {
        const int tok = i / 16, r = i % 16, row = T * 16 + r;
        if (tok < b && row < N) {
            float v = red[0][tok][r];
#pragma unroll
            for (int w2 = 1; w2 < GP_WARPS; w2++) v += red[w2][tok][r];
            y[(size_t) tok * N + row] = v;
        }
    }
}

// ---- balanced variant: the matrix as ONE stream of records ---------------------------------------------------------------------
// At b <= 4 an SM spends ~100 cycles of issue slots per record, close to what its share of HBM delivers: the SM with the most
// records sets the kernel's time.  Row tiles are too coarse a unit for that (512 tiles over 148 SMs: 4 on some, 3 on others),
// so here the n_tiles x nsb records -- contiguous in memory in exactly this order -- are cut into W equal runs, one per warp of a
// grid that fills the machine once.  A warp streams its run through its ring, and whenever the run leaves a row tile (or ends)
// it hands the tile's partial sums on: the only piece of that tile -> written to y; otherwise stored as piece k of the tile, and
// the warp that brings the last piece adds the pieces in run order (fixed order: the result does not depend on timing).
// No block-wide barrier anywhere: the warps of a CTA share nothing but the launch.
#define SK_WARPS 4
__device__ __forceinline__ int sk_owner(long long u, int W, long long U) { return (int) (((u + 1) * W - 1) / U); }   // the warp whose run holds record u

template <int NT>
__global__ void __launch_bounds__(SK_WARPS * 32, NT == 1 ? 5 : 3)
gemv_q4k_stream_kernel(const uint8_t * __restrict__ tiles, int nsb, int n_tiles, int N, int K, const int8_t * __restrict__ aqs,
                       const float * __restrict__ ad, const int8_t * __restrict__ abs8, int b, float * __restrict__ y,
                       float * __restrict__ partial, unsigned int * __restrict__ counters, int pmax, int W, const int * __restrict__ cancel) {
    extern __shared__ __align__(16) uint8_t rg_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const long long U  = (long long) n_tiles * nsb;                        // W <= U runs: none is empty
    const int       gw = blockIdx.x * SK_WARPS + warp;
    if (gw >= W) return;
    const long long u0 = U * gw / W, u1 = U * (gw + 1) / W;
    if (PIB_CANCELLED(cancel)) {
        // a cancelled launch still has to leave the tile counters at zero: every piece arrives, the last one resets
        pib_pdl_wait();
        if (lane == 0)
            for (long long u = u0; u < u1;) {
                const int       T     = (int) (u / nsb);
                const long long tend  = (long long) (T + 1) * nsb;
                const int       first = sk_owner((long long) T * nsb, W, U), last = sk_owner(tend - 1, W, U);
                if (last > first && atomicAdd(&counters[T], 1u) == (unsigned) (last - first)) counters[T] = 0;
                u = tend;
            }
        return;
    }
    uint8_t *       ring   = rg_smem + warp * (RG_STAGES * MM_TILE_BYTES);
    const uint32_t  ring_a = smem_u32(ring) + lane * 16;
    const uint8_t * wbase  = tiles + lane * 16;
    long long pu = u0;
    auto issue = [&](int stage) {
        if (pu < u1) {
            const uint8_t * src = wbase + (size_t) pu * MM_TILE_BYTES;
            const uint32_t  dst = ring_a + stage * MM_TILE_BYTES;
#pragma unroll
            for (int c = 0; c < 4; c++) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + c * 512), "l"(src + c * 512) : "memory");
            if (lane < 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 2048), "l"(src + 2048) : "memory");
            pu++;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // the weights are immutable: their copies start before the wait on the kernel that produces the activations
#pragma unroll
    for (int i = 0; i < RG_STAGES - 1; i++) issue(i);

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
    int T = (int) (u0 / nsb), sb = (int) (u0 - (long long) T * nsb), stage = 0;
    for (long long u = u0; u < u1; u++) {
        __syncwarp();                                           // every lane has read the stage about to be refilled
        issue((stage + RG_STAGES - 1) % RG_STAGES);
        asm volatile("cp.async.wait_group %0;" ::"n"(RG_STAGES - 1) : "memory");
        __syncwarp();
        const uint8_t * rs = ring + stage * MM_TILE_BYTES;
        int4 wq[4], hq[2];
#pragma unroll
        for (int c = 0; c < 4; c++) wq[c] = *reinterpret_cast<const int4 *>(rs + c * 512 + lane * 16);
        hq[0] = *reinterpret_cast<const int4 *>(rs + 2048 + g * 16);
        hq[1] = *reinterpret_cast<const int4 *>(rs + 2048 + (g + 8) * 16);
        q4k_record<NT>(wq, hq, sb, t, a_qs, a_b8, a_dx, acc);
        stage = stage + 1 == RG_STAGES ? 0 : stage + 1;
        if (++sb < nsb && u + 1 < u1) continue;

        // ---- the run leaves tile T: hand its partial sums on -------------------------------------------------------------------
        const int first = sk_owner((long long) T * nsb, W, U), last = sk_owner((long long) (T + 1) * nsb - 1, W, U);
        const int P = last - first + 1, k = gw - first;
        bool write_y = P == 1;
        if (P > 1) {
            float4 * mine = reinterpret_cast<float4 *>(partial + ((size_t) T * pmax + k) * (NT * 128)) + lane;
#pragma unroll
            for (int n = 0; n < NT; n++) mine[n * 32] = make_float4(acc[n][0], acc[n][1], acc[n][2], acc[n][3]);
            // publish: the warp barrier orders every lane's stores before lane 0's release; the last arriver's acquire sees them all
            __syncwarp();
            unsigned int old = 0;
            if (lane == 0) asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(counters + T) : "memory");
            old = __shfl_sync(0xffffffffu, old, 0);
            if (old == (unsigned) P - 1) {
                const float4 * pc = reinterpret_cast<const float4 *>(partial + (size_t) T * pmax * (NT * 128)) + lane;
#pragma unroll
                for (int n = 0; n < NT; n++) {
                    float4 v = __ldcg(pc + n * 32);
                    for (int q = 1; q < P; q++) {
                        const float4 o = __ldcg(pc + (size_t) q * (NT * 32) + n * 32);
                        v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
                    }
                    acc[n][0] = v.x; acc[n][1] = v.y; acc[n][2] = v.z; acc[n][3] = v.w;
                }
                if (lane == 0) counters[T] = 0;          // self-cleaning: ready for the next launch
                write_y = true;
            }
        }
        if (write_y) {
#pragma unroll
            for (int n = 0; n < NT; n++)
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const int tok = n * 8 + 2 * t + e, row = T * 16 + g;
                    if (tok < b) {
                        if (row < N) y[(size_t) tok * N + row] = acc[n][e];
                        if (row + 8 < N) y[(size_t) tok * N + row + 8] = acc[n][2 + e];
                    }
                }
        }
#pragma unroll
        for (int n = 0; n < NT; n++) { acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f; }
        T++;
        sb = 0;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// ---- launcher -----------------------------------------------------------------------------------------------------------
// scratch for the k-split path: partial sums [split][b][N] + one counter per row tile.  One set per (device, stream):
// launches on one stream are serialised, so they can share it; two contexts (target and draft) running on their own
// streams at the same time must not.
struct mm_scratch { float * partial = nullptr; unsigned int * counters = nullptr; size_t partial_floats = 0; size_t n_counters = 0; };
static std::map<std::pair<int, cudaStream_t>, mm_scratch> g_scratch;
static std::mutex                                        g_scratch_mu;

// Fixed size, allocated once per (device, stream) and never again: captured CUDA graphs keep these addresses.  A launch whose
// split would not fit is given a smaller split by its caller (fit_split); one that cannot fit at all is refused.
#define MM_SCRATCH_FLOATS   ((size_t) 8 << 20)          /* 32 MiB: two splits of 16 tokens x 118784 rows */
#define MM_SCRATCH_COUNTERS ((size_t) 8192)
static int ensure_scratch(size_t floats, size_t counters, cudaStream_t st, mm_scratch ** out) {
    if (floats > MM_SCRATCH_FLOATS || counters > MM_SCRATCH_COUNTERS) return PIB_ERR_ARG;
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    mm_scratch & s = g_scratch[std::make_pair(dev, st)];
    if (!s.partial) {
        if (cudaMalloc(&s.partial, MM_SCRATCH_FLOATS * sizeof(float)) != cudaSuccess) { cudaGetLastError(); s.partial = nullptr; return PIB_ERR_NOMEM; }
        if (cudaMalloc(&s.counters, MM_SCRATCH_COUNTERS * sizeof(unsigned int)) != cudaSuccess) { cudaGetLastError(); cudaFree(s.partial); s.partial = nullptr; return PIB_ERR_NOMEM; }
        cudaMemset(s.counters, 0, MM_SCRATCH_COUNTERS * sizeof(unsigned int));
        s.partial_floats = MM_SCRATCH_FLOATS; s.n_counters = MM_SCRATCH_COUNTERS;
    }
    *out = &s;
    return PIB_OK;
}
// largest split <= want whose partial sums and counters fit the fixed scratch (1 = no scratch needed)
static int fit_split(int want, int b, int64_t N, int n_tiles) {
    while (want > 1 && ((size_t) want * b * N > MM_SCRATCH_FLOATS || (size_t) n_tiles > MM_SCRATCH_COUNTERS)) want /= 2;
    return want;
}

bool pib_gemv_mma_supported(const pib_weight * w, int b) {
    return w && w->type == PIB_TYPE_Q4_K && w->tiles != nullptr && b >= 1 && b <= PIB_MAX_BATCH;
}

// log2 of the number of k splits: items (row tiles x splits) should outnumber the resident warps a few times over for load
// balance; a split must divide the super-block count and keep at least two super-blocks per item
static int choose_lg_split(int n_tiles, int nsb) {
    const int target = 3 * 4 * MM_WARPS * pib_num_sms();
    int lg = 0;
    while ((n_tiles << lg) < target && nsb % (2 << lg) == 0 && (nsb >> (lg + 1)) >= 2 && lg < 6) lg++;
    return lg;
}

static size_t mm_smem_bytes() { return (size_t) MM_WARPS * MM_STAGES * MM_STAGE_BYTES + (size_t) MM_WARPS * MM_STAGES * 8; }

// k splits of the direct variant: enough CTAs to fill the machine when there are few row tiles
int pib_split_fill_target() {
    static int pct = -1;
    if (pct < 0) { const char * e = getenv("PIB_SPLIT_FILL_PCT"); pct = e ? atoi(e) : 80; }
    return 4 * pib_num_sms() * pct / 100;
}

static int choose_split_direct(int n_tiles, int nsb) {
    // a k split costs a second phase (partial sums through L2, an atomic, a re-read by the last CTA): worth it only while
    // the row tiles alone leave a good part of the CTA slots (4 per SM) empty.  Measured on the 70B shapes: with 512 tiles for
    // 592 slots, not splitting takes 14.8 instead of 18.8 us (attn_out) and 33.6 instead of 37.3 us (down)
    const int target = pib_split_fill_target();
    int split = 1;
    while (n_tiles * split < target && split * 2 * MM_WARPS <= nsb && split < 16) split *= 2;
    return split;
}

extern "C" int pib_gemv_direct_splits(int64_t N, int64_t K) {
    if (N <= 0 || K <= 0 || K % 256) return -1;
    return choose_split_direct((int) ((N + 15) / 16), (int) (K / 256));
}

int pib_gemv_direct_launch(const pib_weight * w, const void * act_dev, int b, float * y_dev, int split, float * partial,
                           unsigned int * counters, const int * cancel, cudaStream_t st) {
    const pib_act_layout L   = pib_act_layout_of(w->type, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    const int            nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    const dim3           grid(n_tiles, split);
    const int8_t *       aqs = (const int8_t *) (act + L.qsp_off);
    const float *        ad  = (const float *) (act + L.d_off);
    const int8_t *       bs  = (const int8_t *) (act + L.bs8_off);
    if (b <= 8) pib_launch(gemv_q4k_direct_kernel<1>, grid, dim3(MM_THREADS), 0, st, w->tiles, nsb, (int) w->N, (int) w->K, aqs, ad, bs, b, y_dev, partial, counters, cancel);
    else        pib_launch(gemv_q4k_direct_kernel<2>, grid, dim3(MM_THREADS), 0, st, w->tiles, nsb, (int) w->N, (int) w->K, aqs, ad, bs, b, y_dev, partial, counters, cancel);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

int pib_gemv_mma_launch(const pib_weight * w, const void * act_dev, int b, float * y_dev, int split, float * partial,
                        unsigned int * counters, const int * cancel, cudaStream_t st) {
    const pib_act_layout L   = pib_act_layout_of(w->type, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    const int            nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    const int8_t *       aqs = (const int8_t *) (act + L.qsp_off);      // fragment-order quants
    const float *        ad  = (const float *) (act + L.d_off);
    const int8_t *       bs  = (const int8_t *) (act + L.bs8_off);
    static bool configured = false;
    if (!configured) {
        PIB_CHECK(cudaFuncSetAttribute(gemv_q4k_mma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) mm_smem_bytes()));
        PIB_CHECK(cudaFuncSetAttribute(gemv_q4k_mma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) mm_smem_bytes()));
        configured = true;
    }
    const int items = n_tiles << split;                                            // `split` is log2 of the split count here
    const int ctas  = (items + MM_WARPS - 1) / MM_WARPS;
    const int grid  = ctas < 4 * pib_num_sms() ? ctas : 4 * pib_num_sms();         // persistent: 4 CTAs per SM
    if (b <= 8) gemv_q4k_mma_kernel<1><<<grid, MM_THREADS, mm_smem_bytes(), st>>>(w->tiles, nsb, n_tiles, split, (int) w->N, (int) w->K, aqs, ad, bs, b, y_dev, partial, counters, cancel);
    else        gemv_q4k_mma_kernel<2><<<grid, MM_THREADS, mm_smem_bytes(), st>>>(w->tiles, nsb, n_tiles, split, (int) w->N, (int) w->K, aqs, ad, bs, b, y_dev, partial, counters, cancel);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

template <int NT, int WARPS, int STAGES>
static int ring_launch(const pib_weight * w, const int8_t * aqs, const float * ad, const int8_t * bs, int b, float * y_dev, int split, float * partial,
                       unsigned int * counters, const int * cancel, cudaStream_t st) {
    const size_t smem = (size_t) WARPS * STAGES * MM_TILE_BYTES;
    static bool configured = false;
    if (!configured) { PIB_CHECK(cudaFuncSetAttribute(gemv_q4k_ring_kernel<NT, WARPS, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem)); configured = true; }
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    pib_launch(gemv_q4k_ring_kernel<NT, WARPS, STAGES>, dim3(n_tiles, split), dim3(WARPS * 32), smem, st, w->tiles, nsb, (int) w->N, (int) w->K, aqs, ad, bs, b, y_dev,
               partial, counters, cancel);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

// variant 0 / 2: the ring kernel (cp.async-staged); PIB_Q4K_FAST=direct selects the register-prefetched direct kernel instead
int pib_gemv_stream(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
static bool stream_plan(int n_tiles, int nsb, int b, int * runs, int * pmax);

int pib_gemv_ring(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!pib_gemv_mma_supported(w, b)) { pib_set_error("pib_gemv: variant 2 needs Q4_K weights with tiles"); return PIB_ERR_UNSUPP; }
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    // 9 .. 16 tokens (two column blocks per record): a record costs twice the issue slots, the SM with the most records sets the
    // time, and the balanced stream kernel wins on every 70B shape (gate|up 100 vs 108 us, down 53 vs 63, attn_out 21 vs 23);
    // at <= 8 tokens the tile-per-CTA ring is ahead (42 vs 53, 28 vs 30, 10.4 vs 12.7)
    {
        int runs = 0, pmax = 0;
        if (b > 8 && (long long) n_tiles * nsb >= 16384 && stream_plan(n_tiles, nsb, b, &runs, &pmax)) return pib_gemv_stream(w, act_dev, b, y_dev, cancel, st);
    }
    // one wave of CTAs or less (4 CTAs of 4 warps per SM): eight warps per tile halve every warp's chain of records
    static const int wide_on = getenv("PIB_RING_WIDE") ? atoi(getenv("PIB_RING_WIDE")) : 1;
    const bool wide = wide_on && n_tiles <= 2 * pib_num_sms() * 2 && nsb >= 16;          // <= 2 CTAs of 8 warps per SM
    static const int split_env = getenv("PIB_RING_SPLIT") ? atoi(getenv("PIB_RING_SPLIT")) : 0;
    const int  split = fit_split(split_env > 0 ? split_env : wide ? 1 : choose_split_direct(n_tiles, nsb), b, w->N, n_tiles);
    mm_scratch * s = nullptr;
    if (split > 1 && ensure_scratch((size_t) split * b * w->N, (size_t) n_tiles, st, &s) != PIB_OK) { pib_set_error("pib_gemv: scratch allocation failed"); return PIB_ERR_NOMEM; }
    const pib_act_layout L   = pib_act_layout_of(w->type, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    const int8_t * aqs = (const int8_t *) (act + L.qsp_off);
    const float *  ad  = (const float *) (act + L.d_off);
    const int8_t * bs  = (const int8_t *) (act + L.bs8_off);
    float * pp = s ? s->partial : nullptr;
    unsigned int * cc = s ? s->counters : nullptr;
    if (wide) return b <= 8 ? ring_launch<1, 8, 4>(w, aqs, ad, bs, b, y_dev, split, pp, cc, cancel, st) : ring_launch<2, 8, 4>(w, aqs, ad, bs, b, y_dev, split, pp, cc, cancel, st);
    return b <= 8 ? ring_launch<1, 4, 4>(w, aqs, ad, bs, b, y_dev, split, pp, cc, cancel, st) : ring_launch<2, 4, 4>(w, aqs, ad, bs, b, y_dev, split, pp, cc, cancel, st);
}

// variant 8 (and the default for shapes whose row tiles do not fill the machine evenly): the balanced stream kernel
template <int NT>
static int stream_launch(const pib_weight * w, const int8_t * aqs, const float * ad, const int8_t * bs, int b, float * y_dev, int W, int pmax,
                         float * partial, unsigned int * counters, const int * cancel, cudaStream_t st) {
    const size_t smem = (size_t) SK_WARPS * RG_STAGES * MM_TILE_BYTES;
    static bool configured = false;
    if (!configured) { PIB_CHECK(cudaFuncSetAttribute(gemv_q4k_stream_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem)); configured = true; }
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    pib_launch(gemv_q4k_stream_kernel<NT>, dim3((W + SK_WARPS - 1) / SK_WARPS), dim3(SK_WARPS * 32), smem, st, w->tiles, nsb, n_tiles, (int) w->N, (int) w->K, aqs,
               ad, bs, b, y_dev, partial, counters, pmax, W, cancel);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

// grid and piece bound of the stream kernel for a shape; false when the scratch cannot hold the pieces
static bool stream_plan(int n_tiles, int nsb, int b, int * runs, int * pmax) {
    const int       per_sm = b <= 8 ? 5 : 3;
    const long long U = (long long) n_tiles * nsb;
    long long       W = (long long) pib_num_sms() * per_sm * SK_WARPS;
    if (W > U) W = U;                                                          // tiny matrices: one record per run
    *runs = (int) W;
    *pmax = (int) (((long long) nsb * W + U - 1) / U) + 1;                     // pieces a tile can be cut into
    return (size_t) n_tiles * *pmax * (b <= 8 ? 128 : 256) <= MM_SCRATCH_FLOATS && (size_t) n_tiles <= MM_SCRATCH_COUNTERS;
}

int pib_gemv_stream(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!pib_gemv_mma_supported(w, b)) { pib_set_error("pib_gemv: variant 8 needs Q4_K weights with tiles"); return PIB_ERR_UNSUPP; }
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    int runs = 0, pmax = 0;
    if (!stream_plan(n_tiles, nsb, b, &runs, &pmax)) { pib_set_error("pib_gemv: variant 8: shape too large for the piece scratch"); return PIB_ERR_UNSUPP; }
    mm_scratch * s = nullptr;
    if (ensure_scratch((size_t) n_tiles * pmax * (b <= 8 ? 128 : 256), (size_t) n_tiles, st, &s) != PIB_OK) { pib_set_error("pib_gemv: scratch allocation failed"); return PIB_ERR_NOMEM; }
    const pib_act_layout L   = pib_act_layout_of(w->type, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    const int8_t * aqs = (const int8_t *) (act + L.qsp_off);
    const float *  ad  = (const float *) (act + L.d_off);
    const int8_t * bs  = (const int8_t *) (act + L.bs8_off);
    return b <= 8 ? stream_launch<1>(w, aqs, ad, bs, b, y_dev, runs, pmax, s->partial, s->counters, cancel, st)
                  : stream_launch<2>(w, aqs, ad, bs, b, y_dev, runs, pmax, s->partial, s->counters, cancel, st);
}

// matmul with the activation prologue inside (see gemv_q4k_pro_kernel).  pro: 0 plain (x), 1 [x + addend ->] RMSNorm * x2, 2 silu(x) * x2
bool pib_gemv_pro_supported(const pib_weight * w, int b) {
    if (!w || w->type != PIB_TYPE_Q4_K || !w->tiles || b < 1 || b > 8 || w->K > 8192 || w->K % 256) return false;
    const int nsb = (int) (w->K / 256);
    return (size_t) GP_WARPS * RG_STAGES * MM_TILE_BYTES + (size_t) b * (w->K + nsb * 20) <= 100 * 1024;
}

template <int PRO>
static int pro_launch(const pib_weight * w, int b, float * y, const float * x, const float * x2, const float * addend, float * sum_out, int64_t xstride,
                      float eps, const uint16_t * lut, const int * cancel, cudaStream_t st) {
    const int    nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    const size_t smem = (size_t) GP_WARPS * RG_STAGES * MM_TILE_BYTES + (size_t) b * (w->K + nsb * 20);
    static size_t configured = 0;
    if (smem > configured) {
        PIB_CHECK(cudaFuncSetAttribute(gemv_q4k_pro_kernel<PRO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (100 * 1024)));
        PIB_CHECK(cudaFuncSetAttribute(gemv_q4k_pro_kernel<PRO>, cudaFuncAttributePreferredSharedMemoryCarveout, (int) cudaSharedmemCarveoutMaxShared));
        configured = 100 * 1024;
    }
    pib_launch(gemv_q4k_pro_kernel<PRO>, dim3(n_tiles), dim3(GP_WARPS * 32), smem, st, w->tiles, nsb, (int) w->N, (int) w->K, b, y, x, x2, addend, sum_out,
               xstride, eps, lut, cancel);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

int pib_gemv_pro(const pib_weight * w, int pro, const float * x, const float * x2, const float * addend, float * sum_out, int64_t xstride, float eps,
                 int b, float * y, const int * cancel, cudaStream_t st) {
    if (!pib_gemv_pro_supported(w, b)) { pib_set_error("pib_gemv_pro: unsupported matrix / batch"); return PIB_ERR_UNSUPP; }
    if (xstride == 0) xstride = w->K;
    switch (pro) {
        case GP_PLAIN:    return pro_launch<GP_PLAIN>(w, b, y, x, nullptr, nullptr, nullptr, xstride, 0.f, nullptr, cancel, st);
        case GP_RMSNORM:  return pro_launch<GP_RMSNORM>(w, b, y, x, x2, addend, sum_out, xstride, eps, nullptr, cancel, st);
        case GP_SILU_MUL: return pro_launch<GP_SILU_MUL>(w, b, y, x, x2, nullptr, nullptr, xstride, 0.f, pib_get_luts().silu_f16, cancel, st);
    }
    pib_set_error("pib_gemv_pro: unknown prologue %d", pro);
    return PIB_ERR_ARG;
}

// variant 5: direct (register-prefetched) kernel
int pib_gemv_direct(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!pib_gemv_mma_supported(w, b)) { pib_set_error("pib_gemv: variant 2 needs Q4_K weights with tiles"); return PIB_ERR_UNSUPP; }
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    const int split = fit_split(choose_split_direct(n_tiles, nsb), b, w->N, n_tiles);
    mm_scratch * s = nullptr;
    if (split > 1) {
        if (ensure_scratch((size_t) split * b * w->N, (size_t) n_tiles, st, &s) != PIB_OK) { pib_set_error("pib_gemv: scratch allocation failed"); return PIB_ERR_NOMEM; }
    }
    return pib_gemv_direct_launch(w, act_dev, b, y_dev, split, s ? s->partial : nullptr, s ? s->counters : nullptr, cancel, st);
}

// variant 3: persistent TMA-staged kernel
int pib_gemv_mma(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!pib_gemv_mma_supported(w, b)) { pib_set_error("pib_gemv: variant 3 needs Q4_K weights with tiles"); return PIB_ERR_UNSUPP; }
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    int split = choose_lg_split(n_tiles, nsb);
    while (split > 0 && (((size_t) b << split) * w->N > MM_SCRATCH_FLOATS || (size_t) n_tiles > MM_SCRATCH_COUNTERS)) split--;
    mm_scratch * s = nullptr;
    if (split > 0) {
        if (ensure_scratch(((size_t) b << split) * w->N, (size_t) n_tiles, st, &s) != PIB_OK) { pib_set_error("pib_gemv: scratch allocation failed"); return PIB_ERR_NOMEM; }
    }
    return pib_gemv_mma_launch(w, act_dev, b, y_dev, split, s ? s->partial : nullptr, s ? s->counters : nullptr, cancel, st);
}

// ---- allocation-free entry point ------------------------------------------------------------------------------------
// The k-split reduction above needs global scratch (partial sums + arrival counters).  pib_gemv gets it from cudaMalloc
// on first use, which some callers cannot afford: a real-time serving loop that must not take the allocator's lock or
// its implicit device synchronisation in the middle of a decode, code running while another component has device
// memory fully committed, or a host that forbids allocation after start-up.  For them the same kernel runs with its
// scratch in a fixed array that lives in the library's own device image.
//
// `noalloc_scratch` below is, on the host side, only a handle for the symbol: its device address is obtained once per
// device with cudaGetSymbolAddress (with that device current) and cached.  That look-up happens inside the FIRST call
// on each device, so make that first call outside any stream capture.
//
// The array is shared by every call on a device and keeps whatever the previous call left in it.  The arrival counters
// must start at zero, so each call clears the ones it uses (stream-ordered, in front of the kernel); partial sums are
// always written before they are read.  Two calls in flight at once on the same device (different streams or threads)
// would use the same counters and sums and corrupt each other: serialise them.
#define PIB_NOALLOC_COUNTER_BYTES (16u * 1024u)                                    /* 4096 row tiles */
static __device__ __align__(256) unsigned char noalloc_scratch[PIB_GEMV_NOALLOC_SCRATCH_BYTES];
static unsigned char * g_noalloc_ptr[PIB_MAX_DEV_NOALLOC];

extern "C" int pib_gemv_q4k_noalloc(const pib_weight * w, const void * act_dev, int b, float * y_dev, pib_stream stream) {
    if (!w || !act_dev || !y_dev) { pib_set_error("pib_gemv_q4k_noalloc: null argument"); return PIB_ERR_ARG; }
    if (b < 1 || b > PIB_MAX_BATCH) { pib_set_error("pib_gemv_q4k_noalloc: b = %d outside 1..%d", b, PIB_MAX_BATCH); return PIB_ERR_ARG; }
    if (!pib_gemv_mma_supported(w, b)) { pib_set_error("pib_gemv_q4k_noalloc: needs a Q4_K weight handle"); return PIB_ERR_UNSUPP; }
    int dev = -1;
    PIB_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= PIB_MAX_DEV_NOALLOC) { pib_set_error("pib_gemv_q4k_noalloc: device ordinal %d not supported (max %d)", dev, PIB_MAX_DEV_NOALLOC - 1); return PIB_ERR_UNSUPP; }
    if (!g_noalloc_ptr[dev]) {                       // once per device, with that device current
        void * p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, noalloc_scratch);
        if (e != cudaSuccess || !p) { pib_set_error("pib_gemv_q4k_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e)); return PIB_ERR_CUDA; }
        g_noalloc_ptr[dev] = (unsigned char *) p;
    }
    const int    nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    const int    split = choose_split_direct(n_tiles, nsb);
    float *        partial  = nullptr;
    unsigned int * counters = nullptr;
    if (split > 1) {
        const size_t need_partial  = (size_t) split * b * w->N * sizeof(float);
        const size_t need_counters = (size_t) n_tiles * sizeof(unsigned int);
        if (need_counters > PIB_NOALLOC_COUNTER_BYTES || need_partial > PIB_GEMV_NOALLOC_SCRATCH_BYTES - PIB_NOALLOC_COUNTER_BYTES) {
            pib_set_error("pib_gemv_q4k_noalloc: N = %lld, b = %d with %d k-splits needs %zu + %zu bytes of scratch; the built-in "
                          "array holds %u + %u (use pib_gemv)", (long long) w->N, b, split, need_partial, need_counters,
                          PIB_GEMV_NOALLOC_SCRATCH_BYTES - PIB_NOALLOC_COUNTER_BYTES, PIB_NOALLOC_COUNTER_BYTES);
            return PIB_ERR_ARG;
        }
        counters = reinterpret_cast<unsigned int *>(g_noalloc_ptr[dev]);
        partial  = reinterpret_cast<float *>(g_noalloc_ptr[dev] + PIB_NOALLOC_COUNTER_BYTES);
        PIB_CHECK(cudaMemsetAsync(counters, 0, need_counters, (cudaStream_t) stream));
    }
    return pib_gemv_direct_launch(w, act_dev, b, y_dev, split, partial, counters, nullptr, (cudaStream_t) stream);
}
