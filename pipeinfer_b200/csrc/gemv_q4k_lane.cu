// gemv_q4k_lane.cu -- the Q4_K x Q8_K matmul (K1) of exact mode (PIB_EXACT=1, variant 6): tensor-core integer dot products that
// return THE SAME BITS as the reference's CPU kernel (ggml_vec_dot_q4_K_q8_K, AVX2 branch, ggml-quants.c:5318-5381), not merely the same value up to fp32
// reassociation.  That matters because everything downstream of a matmul is rounded discontinuously (int8 activations, fp16
// K/V/P, fp16 look-up tables): a 1e-7 difference now and then lands on the other side of a rounding step, and a 22-layer model
// turns a handful of those into ~5e-4 relative on the logits.  With the reference's own order there is nothing to flip.
//
// What the reference computes for one output (row r, token):   8 fp32 lane accumulators acc[l], 4 for the mins acc_m[i]
//   per super-block s, in order:   sumi[l] = sum_j sc_j * (sum of the 4 products of elements 4l .. 4l+3 of sub-block j)   (int32)
//                                  acc[l]   = fma(d_x * d_w,    float(sumi[l]), acc[l])
//                                  acc_m[i] = fma(-d_x * dmin_w, float(m_2i * bsum_2i + m_2i+1 * bsum_2i+1), acc_m[i])
//   result = ((a0+a4)+(a2+a6)) + ((a1+a5)+(a3+a7))  +  ((m0+m2)+(m1+m3))
// So the integer part is needed per LANE GROUP of four elements, and the fp32 part is a sequential chain over the super-blocks.
//
// Integer part on the tensor core: mma.m16n8k32 sums all 32 k-slots of a column, so the eight columns of the B operand are made
// the eight lane groups of ONE token: column l holds the token's quants of lane group l and zeros elsewhere.  One MMA per
// (16-row tile, 32-element sub-block, token) yields the 16 x 8 lane-group sums; b tokens cost b MMAs (the int8 pipe has ~5x the
// rate one token needs at full HBM bandwidth).  Weight records are the fragment-major tiles of gemv_mma.cu, unchanged: thread
// (g, t) holds elements 8t..8t+3 (lane group 2t) in its k-slots 4t.. and 8t+4..8t+7 (lane group 2t+1) in k-slots 16+4t.., so
// column l is fed by thread t = l/2 alone, through B register l%2.
//
// fp32 part: the four warps of a CTA interleave the super-blocks of a tile; each writes its integers ("terms") to shared memory
// (double buffered, one barrier per group of four super-blocks) and the 16 x b x 12 chains advance through them in order.
//
// Balance: when there are more row tiles than CTA slots the grid is persistent and the weight stream is cut into equal pieces
// that ignore tile boundaries.  A tile cut in two is started by the CTA that holds its head -- FIRST thing that CTA does -- which
// hands the 16 x b x 12 chain state to the next CTA through global memory; that CTA finishes the tile LAST, when the state has
// long arrived.  No piece lies strictly inside a tile (pieces are at least a tile long), so a hand-off never waits on a hand-off.
#include <map>
#include <mutex>
#include <utility>

#include "pib_common.cuh"

#define LX_WARPS   4
#define LX_THREADS (LX_WARPS * 32)
#define LX_TILE_BYTES 2304
#define LX_STAGES 3                 // records per warp ring: two in flight while one is multiplied
#define LX_MAX_SEGS 24              // a CTA never gets more than ~8 tiles (tiles / slots + 2 pieces)
#define LX_MAXB_W 8                 // largest batch of the warp-per-tile variant
#define LX_TERMS 12                 // 8 lane sums + 4 min products per (row, token, super-block)

__device__ __forceinline__ void lx_mma_u8s8(int (&c)[4], int a0, int a1, int a2, int a3, int b0, int b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

struct lx_hdr { float d, dmin; uint32_t sc_lo, sc_hi, m_lo, m_hi; };
__device__ __forceinline__ lx_hdr lx_decode(const int4 h) {
    lx_hdr r;
    const uint32_t x = (uint32_t) h.x, y = (uint32_t) h.y, z = (uint32_t) h.z, w = (uint32_t) h.w;
    r.d    = pib_h2f((uint16_t) (x & 0xffff));
    r.dmin = pib_h2f((uint16_t) (x >> 16));
    r.sc_lo = y & 0x3f3f3f3fu;                                       // get_scale_min_k4 (ggml-quants.c:1446) for all eight sub-blocks
    r.m_lo  = z & 0x3f3f3f3fu;
    r.sc_hi = (w & 0x0f0f0f0fu) | (((y >> 6) & 0x03030303u) << 4);
    r.m_hi  = ((w >> 4) & 0x0f0f0f0fu) | (((z >> 6) & 0x03030303u) << 4);
    return r;
}

// device scratch of one (device, stream): chain states handed from CTA to CTA, their ready flags, the launch epoch
struct lx_sync {
    unsigned int epoch;        // number of launches that have completed on this scratch
    unsigned int ticket;       // CTAs of the current launch that have exited
};

// One CTA = 4 warps.  Integer part: warp w takes super-block 4 it + w of the current unit.  Chains: thread (row = tid / 8,
// l = tid % 8) owns lane accumulator l of that row for every token, and, when l < 4, min accumulator l as well.
// BMAX = tokens the chain registers are sized for (1, 2, 4, 8, 16); TB = tokens multiplied per pass over a chunk's masked nibbles.

// 128-bit read-only load that returns zeros where the predicate is off (no branch)
__device__ __forceinline__ int4 lx_ldg_pred(const void * p, int pred) {
    int4 r;
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %5, 0;\n\tmov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\tmov.b32 %2, 0;\n\tmov.b32 %3, 0;\n\t"
                 "@q ld.global.nc.v4.s32 {%0, %1, %2, %3}, [%4];\n\t}"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "r"(pred));
    return r;
}

template <int BMAX, int TB>
__global__ void __launch_bounds__(LX_THREADS, BMAX <= 2 ? 5 : (BMAX <= 4 ? 4 : (BMAX <= 8 ? 3 : 2)))
gemv_q4k_lane_kernel(const uint8_t * __restrict__ tiles, int nsb, int n_tiles, int N, int K, const int2 * __restrict__ aqsl,
                     const float * __restrict__ ad, const int16_t * __restrict__ asbs, int b, float * __restrict__ y,
                     float * __restrict__ handoff, unsigned int * __restrict__ flags, lx_sync * __restrict__ sync,
                     const int * __restrict__ cancel) {
    extern __shared__ __align__(16) int lx_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int crow = tid >> 3, cl = tid & 7;                           // chain role
    constexpr int TS = 16 * LX_TERMS;                                   // terms of one (warp, token)
    // shared memory: terms [2][LX_WARPS][b][16][12] int32, headers [2][LX_WARPS][16] float2, the warps' weight rings
    int *     T  = lx_smem;
    float2 *  HD = reinterpret_cast<float2 *>(T + 2 * LX_WARPS * b * TS);
    float *   DX = reinterpret_cast<float *>(HD + 2 * LX_WARPS * 16);        // [2][LX_WARPS][16] activation scales of the unit's super-blocks
    uint8_t * W  = reinterpret_cast<uint8_t *>(DX + 2 * LX_WARPS * 16);      // [LX_WARPS][LX_STAGES] records
    // [LX_WARPS][2] activation stages: per token 512 bytes of B fragments, then per token 16 bytes of sub-block sums, then per token
    // the fp32 scale -- what a warp needs of the activations for one super-block, copied one unit ahead
    uint8_t * F  = W + LX_WARPS * LX_STAGES * LX_TILE_BYTES;
    const int fstage_bytes = b * 512 + b * 16 + ((b * 4 + 15) & ~15);

    // ---- this CTA's piece of the weight stream, in units of 4 consecutive super-blocks of a tile ----------------------------
    const int       ipt = (nsb + LX_WARPS - 1) / LX_WARPS;
    const long long U   = (long long) n_tiles * ipt;
    const int       G   = gridDim.x, c = blockIdx.x;
    const long long u0 = U * c / G, u1 = U * (c + 1) / G;
    const int X = (int) (u0 / ipt), a = (int) (u0 % ipt);               // first tile, first unit inside it
    const int Y = (int) ((u1 - 1) / ipt), e = (int) (u1 - (long long) Y * ipt);     // last tile, end unit inside it (1 .. ipt)
    const bool has_tail = a > 0;                                        // tile X was started by CTA c - 1
    const bool has_head = e < ipt;                                      // tile Y is finished by CTA c + 1 (then a == 0 if X == Y: pieces >= a tile)
    const int  F0 = has_tail ? X + 1 : X, F1 = has_head ? Y : Y + 1;    // full tiles [F0, F1)
    // segments in processing order: the head piece first (its state is handed on at once), then the full tiles, the tail piece
    // last.  One int4 {tile, first unit, end unit, kind: 0 head / 1 full / 2 tail} each, written once by thread 0.
    __shared__ int4 segs[LX_MAX_SEGS];
    __shared__ int  s_nseg;
    if (tid == 0) {
        int n = 0;
        if (has_head) segs[n++] = make_int4(Y, 0, e, 0);
        for (int f = F0; f < F1 && n < LX_MAX_SEGS - 1; f++) segs[n++] = make_int4(f, 0, ipt, 1);
        if (has_tail) segs[n++] = make_int4(X, a, ipt, 2);
        s_nseg = n;
    }
    __syncthreads();
    const int nseg = s_nseg;
    const size_t unit_bytes = (size_t) LX_WARPS * LX_TILE_BYTES;
    // prefetch cursors (segment, unit): `p` for the weights, LX_STAGES - 1 units ahead, with this lane's 16 bytes of chunk 0 of this
    // warp's record in that unit; `q` for the activations, one unit ahead
    int  pseg = 0, pit = 0, pend = 0, qseg = 0, qit = 0, qend = 0;
    const uint8_t * nrec = tiles;
    auto pseg_load = [&]() {
        if (pseg < nseg) {
            const int4 sg = segs[pseg];
            pit = sg.y; pend = sg.z;
            nrec = tiles + ((size_t) sg.x * nsb + sg.y * LX_WARPS + warp) * LX_TILE_BYTES + lane * 16;
        }
    };
    pseg_load();
    if (nseg > 0) { qit = segs[0].y; qend = segs[0].z; }
    // every warp streams its records through a private ring in shared memory filled by cp.async (16 bytes per lane and copy, L2
    // only): LX_STAGES - 1 records in flight per warp whatever the registers hold, and no CTA-wide wait on anybody's weights
    const uint32_t ring  = (uint32_t) __cvta_generic_to_shared(W + warp * (LX_STAGES * LX_TILE_BYTES)) + lane * 16;
    auto issue = [&](int stage) {                // one commit group per unit, empty when this warp has no record in it
        if (pseg < nseg) {
            if (pit * LX_WARPS + warp < nsb) {
                const uint32_t dst = ring + stage * LX_TILE_BYTES;
#pragma unroll
                for (int ch = 0; ch < 4; ch++) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + ch * 512), "l"(nrec + ch * 512) : "memory");
                if (lane < 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 2048), "l"(nrec + 2048) : "memory");
            }
            if (++pit < pend) nrec += unit_bytes; else { pseg++; pseg_load(); }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // activations of the unit the `q` cursor points at: B fragments (one 16-byte copy per lane and token), sub-block sums, scales
    const uint32_t fring = (uint32_t) __cvta_generic_to_shared(F + warp * 2 * fstage_bytes);
    const uint8_t * xsrc = reinterpret_cast<const uint8_t *>(aqsl) + lane * 16;
    auto issue_act = [&](int fstage) {
        if (qseg < nseg) {
            const int sq = qit * LX_WARPS + warp;
            if (sq < nsb) {
                const uint32_t dst = fring + fstage * fstage_bytes;
                for (int k = 0; k < b; k++)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + k * 512 + lane * 16), "l"(xsrc + ((size_t) k * nsb + sq) * 512) : "memory");
                if (lane < b) {
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + b * 512 + lane * 16), "l"(asbs + ((size_t) lane * nsb + sq) * 8) : "memory");
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + b * 528 + lane * 4), "l"(ad + (size_t) lane * nsb + sq) : "memory");
                }
            }
            if (++qit >= qend) { qseg++; if (qseg < nseg) { qit = segs[qseg].y; qend = segs[qseg].z; } }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // weights are immutable, so the first copies go out before the wait on the producer kernel; the activations are what that
    // kernel writes.  Group order from here on: act(k + 1), weights(k + LX_STAGES - 1) per unit k, so that waiting for all but the
    // last three groups means "act(k) and weights(k) have landed" without waiting for the deeper weight prefetch
#pragma unroll
    for (int i = 0; i < LX_STAGES - 1; i++) issue(i);
    pib_pdl_wait();
    pib_pdl_trigger();
    issue_act(0);
    const unsigned int ready = sync->epoch + 1;                         // flag value that means "handed off in THIS launch"
    const bool cancelled = PIB_CANCELLED(cancel);

    float acc[BMAX], accm[BMAX];
    const int b_active = (t == (g >> 1)) ? 1 : 0;                       // this thread feeds column g (lane group g) of the B operand
    // per-thread bases: B fragments (512 bytes per (token, super-block), this thread's 64 at g * 64), sub-block sums, terms
    int  buf = 0, stage = 0;
    bool first_unit = true;

    for (int seg = 0; seg < nseg && !cancelled; seg++) {
      const int4 sg   = segs[seg];
      const int  tile = sg.x, kind = sg.w;
      for (int it = sg.y; it < sg.z; it++) {
        const int  s         = it * LX_WARPS + warp;
        const bool seg_first = it == sg.y, seg_last = it == sg.z - 1;
        // start the copies of the next unit's activations and of the weights LX_STAGES - 1 units ahead (both into stages consumed in
        // the previous iteration: its barrier lies between the last read and these writes), then wait for this unit's
        issue_act(buf ^ 1);
        issue((stage + LX_STAGES - 1) % LX_STAGES);
        // (in the very first unit act(0) was committed after weights(1), so it is among the last three groups: wait for one more)
        if (first_unit) { asm volatile("cp.async.wait_group 2;" ::: "memory"); first_unit = false; }
        else asm volatile("cp.async.wait_group 3;" ::: "memory");
        __syncwarp();
        int *    Tb  = T + buf * (LX_WARPS * b * TS);
        float2 * HDb = HD + buf * (LX_WARPS * 16);
        // ---- integer part of super-block s for every token --------------------------------------------------------------------
        if (s < nsb) {
            const uint8_t * rs = W + (warp * LX_STAGES + stage) * LX_TILE_BYTES;
            int4 wq[4];
#pragma unroll
            for (int ch = 0; ch < 4; ch++) wq[ch] = *reinterpret_cast<const int4 *>(rs + ch * 512 + lane * 16);
            const lx_hdr H0 = lx_decode(*reinterpret_cast<const int4 *>(rs + 2048 + g * 16)), H1 = lx_decode(*reinterpret_cast<const int4 *>(rs + 2048 + (g + 8) * 16));
            if (t == 0) {
                HDb[warp * 16 + g]     = make_float2(H0.d, H0.dmin);
                HDb[warp * 16 + g + 8] = make_float2(H1.d, H1.dmin);
            }
            // the two mins this thread multiplies (min pair t of rows g and g + 8), as bytes 0 / 1 of a register for dp2a
            const int mw0 = (int) (((t < 2 ? H0.m_lo : H0.m_hi) >> (16 * (t & 1))) & 0xffffu);
            const int mw1 = (int) (((t < 2 ? H1.m_lo : H1.m_hi) >> (16 * (t & 1))) & 0xffffu);
            int *           Tt = Tb + warp * b * TS + g * LX_TERMS + 2 * t;
            const uint8_t * fs = F + (warp * 2 + buf) * fstage_bytes;          // this unit's activation stage
            const uint8_t * xs = fs + g * 64;                                  // this thread's 64 bytes of a token's B fragments
            const int *     ss = reinterpret_cast<const int *>(fs + b * 512) + t;
            // the chains of all four warps need this super-block's activation scales: out of the warp-private stage (which is
            // refilled while slower warps still run their chains) into a double-buffered array, like the terms
            if (lane < b) DX[(buf * LX_WARPS + warp) * 16 + lane] = *reinterpret_cast<const float *>(fs + b * 528 + lane * 4);
            for (int tok0 = 0; tok0 < b; tok0 += TB, Tt += TB * TS, xs += TB * 512, ss += TB * 4) {
                int sumi[TB][4], sumh[TB][4];
#pragma unroll
                for (int k = 0; k < TB; k++)
#pragma unroll
                    for (int i = 0; i < 4; i++) { sumi[k][i] = 0; sumh[k][i] = 0; }
                // an odd batch repeats its last token in the second slot (computed, not stored)
                const int xk1 = (TB > 1 && tok0 + 1 < b) ? 512 : 0;
#pragma unroll
                for (int ch = 0; ch < 4; ch++) {
                    const uint32_t wx = (uint32_t) wq[ch].x, wy = (uint32_t) wq[ch].y, wz = (uint32_t) wq[ch].z, ww = (uint32_t) wq[ch].w;
                    const int al0 = (int) (wx & 0x0F0F0F0Fu), al1 = (int) (wz & 0x0F0F0F0Fu), al2 = (int) (wy & 0x0F0F0F0Fu), al3 = (int) (ww & 0x0F0F0F0Fu);
                    const int ah0 = (int) (wx & 0xF0F0F0F0u), ah1 = (int) (wz & 0xF0F0F0F0u), ah2 = (int) (wy & 0xF0F0F0F0u), ah3 = (int) (ww & 0xF0F0F0F0u);
                    const int j0 = 2 * ch, j1 = 2 * ch + 1;
                    const int sc0_lo = (int) __byte_perm(j0 < 4 ? H0.sc_lo : H0.sc_hi, 0, 0x4440 + (j0 & 3));
                    const int sc1_lo = (int) __byte_perm(j0 < 4 ? H1.sc_lo : H1.sc_hi, 0, 0x4440 + (j0 & 3));
                    const int sc0_hi = (int) __byte_perm(j1 < 4 ? H0.sc_lo : H0.sc_hi, 0, 0x4440 + (j1 & 3));
                    const int sc1_hi = (int) __byte_perm(j1 < 4 ? H1.sc_lo : H1.sc_hi, 0, 0x4440 + (j1 & 3));
#pragma unroll
                    for (int k = 0; k < TB; k++) {
                        // {b0, b1} of sub-blocks 2 ch and 2 ch + 1: zeros unless this thread feeds its column
                        int4 xq = *reinterpret_cast<const int4 *>(xs + (k ? xk1 : 0) + ch * 16);
                        if (!b_active) xq = make_int4(0, 0, 0, 0);
                        int c_lo[4] = { 0, 0, 0, 0 }, c_hi[4] = { 0, 0, 0, 0 };
                        lx_mma_u8s8(c_lo, al0, al1, al2, al3, xq.x, xq.y);
                        lx_mma_u8s8(c_hi, ah0, ah1, ah2, ah3, xq.z, xq.w);
                        sumi[k][0] += sc0_lo * c_lo[0]; sumh[k][0] += sc0_hi * c_hi[0];
                        sumi[k][1] += sc0_lo * c_lo[1]; sumh[k][1] += sc0_hi * c_hi[1];
                        sumi[k][2] += sc1_lo * c_lo[2]; sumh[k][2] += sc1_hi * c_hi[2];
                        sumi[k][3] += sc1_lo * c_lo[3]; sumh[k][3] += sc1_hi * c_hi[3];
                    }
                }
#pragma unroll
                for (int k = 0; k < TB; k++) {
                    if (k == 0 || tok0 + k < b) {
                        // the high nibbles were multiplied in place (q << 4): every term of sumh is a multiple of 16
                        int * Tk = Tt + k * TS;
                        *reinterpret_cast<int2 *>(Tk)                = make_int2(sumi[k][0] + (sumh[k][0] >> 4), sumi[k][1] + (sumh[k][1] >> 4));
                        *reinterpret_cast<int2 *>(Tk + 8 * LX_TERMS) = make_int2(sumi[k][2] + (sumh[k][2] >> 4), sumi[k][3] + (sumh[k][3] >> 4));
                        // mins: m[2t] * q8s[2t] + m[2t+1] * q8s[2t+1], q8s = sub-block sums (_mm_hadd_epi16 + _mm_madd_epi16, :5336-5339)
                        const int q8 = ss[k * 4];
                        Tk[8 - t]                = __dp2a_lo(q8, mw0, 0);          // Tk = row g, term 2t: term 8 + t is 8 - t further on
                        Tk[8 * LX_TERMS + 8 - t] = __dp2a_lo(q8, mw1, 0);
                    }
                }
            }
        }
        __syncthreads();
        // ---- fp32 chains: super-blocks 4 it .. 4 it + 3 in order ------------------------------------------------------------------
        if (seg_first) {
            if (kind == 2) {
                // the chain state of this tile's head was written by CTA c - 1 at the very start of its run
                const volatile unsigned int * f = flags + (c - 1);
                while (*f != ready) { if (PIB_CANCELLED(cancel)) break; }
                __threadfence();
                const float2 * hin = reinterpret_cast<const float2 *>(handoff) + (size_t) (c - 1) * PIB_MAX_BATCH * LX_THREADS;
#pragma unroll
                for (int k = 0; k < BMAX; k++) if (k < b) { const float2 v = __ldcg(hin + k * LX_THREADS + tid); acc[k] = v.x; accm[k] = v.y; }
            } else {
#pragma unroll
                for (int k = 0; k < BMAX; k++) { acc[k] = 0.f; accm[k] = 0.f; }
            }
        }
        {
            const int     nw  = nsb - it * LX_WARPS;                  // super-blocks of this unit (>= 4 except in a ragged last unit)
            const int *   Tc  = Tb + crow * LX_TERMS + cl;
            const float * dxs = DX + buf * LX_WARPS * 16;
#pragma unroll
            for (int w = 0; w < LX_WARPS; w++) {
                if (w < nw) {
                    const float2 hd = HDb[w * 16 + crow];
#pragma unroll
                    for (int k = 0; k < BMAX; k++)
                        if (k < b) {
                            const float dx = dxs[w * 16 + k];
                            acc[k] = fmaf(__fmul_rn(dx, hd.x), (float) Tc[(w * b + k) * TS], acc[k]);
                            if (cl < 4) accm[k] = fmaf(__fmul_rn(-dx, hd.y), (float) Tc[(w * b + k) * TS + 8], accm[k]);
                        }
                }
            }
        }
        if (seg_last) {
            if (kind == 0) {
                float2 * hout = reinterpret_cast<float2 *>(handoff) + (size_t) c * PIB_MAX_BATCH * LX_THREADS;
#pragma unroll
                for (int k = 0; k < BMAX; k++) if (k < b) __stcg(hout + k * LX_THREADS + tid, make_float2(acc[k], accm[k]));
                __threadfence();
                __syncthreads();
                if (tid == 0) { __threadfence(); *(volatile unsigned int *) (flags + c) = ready; }
            } else {
                // hsum_float_8 (ggml-quants.c:67-73) over the eight lanes of the row, the 4-lane min sum the same way, then their sum
                const int grow = tile * 16 + crow;
#pragma unroll
                for (int k = 0; k < BMAX; k++)
                    if (k < b) {
                        float v = acc[k], m = accm[k];
                        v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 4));        // lanes 0..3: a[l] + a[l + 4]
                        v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 2));        // lanes 0, 1: (a0+a4)+(a2+a6), (a1+a5)+(a3+a7)
                        v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 1));
                        m = __fadd_rn(m, __shfl_xor_sync(0xffffffffu, m, 2));        // lanes 0, 1: m0+m2, m1+m3
                        m = __fadd_rn(m, __shfl_xor_sync(0xffffffffu, m, 1));
                        if (cl == 0 && grow < N) y[(size_t) k * N + grow] = __fadd_rn(v, m);
                    }
            }
        }
        buf ^= 1;
        stage = stage + 1 == LX_STAGES ? 0 : stage + 1;
      }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // ---- launch epoch: the last CTA out advances it (cancelled CTAs too, so that flags of this launch can never look current again)
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        if (atomicAdd(&sync->ticket, 1u) == (unsigned) G - 1) { sync->ticket = 0; __threadfence(); atomicAdd(&sync->epoch, 1u); }
    }
}

// ---- warp-per-tile variant (many row tiles: gate|up, lm_head, qkv) ------------------------------------------------------------
// Same arithmetic, other decomposition: a warp walks ITS tiles' super-blocks in order, so the twelve chains of an output stay in
// that warp's registers (thread (g, t): lanes 2t, 2t+1 and min t of rows g and g + 8) -- no terms through shared memory, no CTA
// barrier.  The weight stream is cut per WARP into equal pieces of records; a tile cut in two is started by the warp holding
// its head (first thing it does) and finished by the next warp (last thing it does), the chain state (6 floats per thread and
// token) going through global memory.  Needs at least as many tiles as warps in the grid, else every warp simply takes one tile.
template <int BMAX>
__global__ void __launch_bounds__(LX_THREADS, BMAX <= 2 ? 5 : (BMAX <= 4 ? 4 : 3))
gemv_q4k_lanew_kernel(const uint8_t * __restrict__ tiles, int nsb, int n_tiles, int N, int K, const int2 * __restrict__ aqsl,
                      const float * __restrict__ ad, const int16_t * __restrict__ asbs, int b, float * __restrict__ y,
                      float * __restrict__ handoff, unsigned int * __restrict__ flags, lx_sync * __restrict__ sync,
                      const int * __restrict__ cancel) {
    extern __shared__ __align__(16) int lx_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int fstage_bytes = b * 512 + b * 16 + ((b * 4 + 15) & ~15);
    uint8_t * W = reinterpret_cast<uint8_t *>(lx_smem) + warp * (LX_STAGES * LX_TILE_BYTES + 2 * fstage_bytes);   // this warp's rings
    uint8_t * F = W + LX_STAGES * LX_TILE_BYTES;

    // ---- this warp's piece of the weight stream, in records -------------------------------------------------------------------
    const int       Wn = gridDim.x * LX_WARPS, gw = blockIdx.x * LX_WARPS + warp;
    const long long U  = (long long) n_tiles * nsb;
    const long long u0 = U * gw / Wn, u1 = U * (gw + 1) / Wn;
    const int X = (int) (u0 / nsb), a = (int) (u0 % nsb);
    const int Y = u1 > u0 ? (int) ((u1 - 1) / nsb) : X, e = u1 > u0 ? (int) (u1 - (long long) Y * nsb) : nsb;
    const bool has_tail = a > 0 && u1 > u0, has_head = e < nsb && u1 > u0;
    const int  F0 = has_tail ? X + 1 : X, F1 = has_head ? Y : Y + 1;
    const int  nfull = u1 > u0 ? (F1 > F0 ? F1 - F0 : 0) : 0;
    const int  nseg = (has_head ? 1 : 0) + nfull + (has_tail ? 1 : 0);
    auto seg_of = [&](int i, int & tile, int & s0, int & s1, int & kind) {
        if (has_head) { if (i == 0) { tile = Y; s0 = 0; s1 = e; kind = 0; return; } i--; }
        if (i < nfull) { tile = F0 + i; s0 = 0; s1 = nsb; kind = 1; return; }
        tile = X; s0 = a; s1 = nsb; kind = 2;
    };
    // prefetch cursors: weights (LX_STAGES - 1 records ahead) and activations (one ahead)
    int pseg = 0, ps = 0, pend = 0, qseg = 0, qs = 0, qend = 0;
    const uint8_t * nrec = tiles;
    auto pseg_load = [&]() { if (pseg < nseg) { int tl, k; seg_of(pseg, tl, ps, pend, k); nrec = tiles + ((size_t) tl * nsb + ps) * LX_TILE_BYTES + lane * 16; } };
    auto qseg_load = [&]() { if (qseg < nseg) { int tl, k; seg_of(qseg, tl, qs, qend, k); } };
    pseg_load();
    qseg_load();
    const uint32_t ring  = (uint32_t) __cvta_generic_to_shared(W) + lane * 16;
    const uint32_t fring = (uint32_t) __cvta_generic_to_shared(F);
    auto issue = [&](int stage) {
        if (pseg < nseg) {
            const uint32_t dst = ring + stage * LX_TILE_BYTES;
#pragma unroll
            for (int ch = 0; ch < 4; ch++) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + ch * 512), "l"(nrec + ch * 512) : "memory");
            if (lane < 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 2048), "l"(nrec + 2048) : "memory");
            if (++ps < pend) nrec += LX_TILE_BYTES; else { pseg++; pseg_load(); }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    const uint8_t * xsrc = reinterpret_cast<const uint8_t *>(aqsl) + lane * 16;
    auto issue_act = [&](int fstage) {
        if (qseg < nseg) {
            const uint32_t dst = fring + fstage * fstage_bytes;
            for (int k = 0; k < b; k++)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + k * 512 + lane * 16), "l"(xsrc + ((size_t) k * nsb + qs) * 512) : "memory");
            if (lane < b) {
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + b * 512 + lane * 16), "l"(asbs + ((size_t) lane * nsb + qs) * 8) : "memory");
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + b * 528 + lane * 4), "l"(ad + (size_t) lane * nsb + qs) : "memory");
            }
            if (++qs >= qend) { qseg++; qseg_load(); }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int i = 0; i < LX_STAGES - 1; i++) issue(i);          // weights are immutable: before the wait on the producer kernel
    pib_pdl_wait();
    pib_pdl_trigger();
    issue_act(0);
    const unsigned int ready = sync->epoch + 1;
    const bool cancelled = PIB_CANCELLED(cancel);

    const int b_active = (t == (g >> 1)) ? 1 : 0;
    float acc[BMAX][4], accm[BMAX][2];
    int   fbuf = 0, stage = 0;
    bool  first_unit = true;
    float * hbase = handoff + ((size_t) gw * LX_MAXB_W * 32 + lane) * 6;          // [warp][token][lane][6]

    for (int seg = 0; seg < nseg && !cancelled; seg++) {
        int tile, s0, s1, kind;
        seg_of(seg, tile, s0, s1, kind);
        if (kind == 2) {
            const volatile unsigned int * f = flags + (gw - 1);
            while (*f != ready) { if (PIB_CANCELLED(cancel)) break; }
            __threadfence();
            const float * hin = hbase - (size_t) LX_MAXB_W * 32 * 6;
#pragma unroll
            for (int k = 0; k < BMAX; k++)
                if (k < b) {
                    const float2 v0 = __ldcg(reinterpret_cast<const float2 *>(hin + k * 192)), v1 = __ldcg(reinterpret_cast<const float2 *>(hin + k * 192 + 2)),
                                 v2 = __ldcg(reinterpret_cast<const float2 *>(hin + k * 192 + 4));
                    acc[k][0] = v0.x; acc[k][1] = v0.y; acc[k][2] = v1.x; acc[k][3] = v1.y; accm[k][0] = v2.x; accm[k][1] = v2.y;
                }
        } else {
#pragma unroll
            for (int k = 0; k < BMAX; k++) { acc[k][0] = acc[k][1] = acc[k][2] = acc[k][3] = 0.f; accm[k][0] = accm[k][1] = 0.f; }
        }
        for (int s = s0; s < s1; s++) {
            __syncwarp();                         // every lane is done with the stages about to be refilled
            issue_act(fbuf ^ 1);
            issue((stage + LX_STAGES - 1) % LX_STAGES);
            if (first_unit) { asm volatile("cp.async.wait_group 2;" ::: "memory"); first_unit = false; }
            else asm volatile("cp.async.wait_group 3;" ::: "memory");
            __syncwarp();
            const uint8_t * rs = W + stage * LX_TILE_BYTES;
            int4 wq[4];
#pragma unroll
            for (int ch = 0; ch < 4; ch++) wq[ch] = *reinterpret_cast<const int4 *>(rs + ch * 512 + lane * 16);
            const lx_hdr H0 = lx_decode(*reinterpret_cast<const int4 *>(rs + 2048 + g * 16)), H1 = lx_decode(*reinterpret_cast<const int4 *>(rs + 2048 + (g + 8) * 16));
            const int mw0 = (int) (((t < 2 ? H0.m_lo : H0.m_hi) >> (16 * (t & 1))) & 0xffffu);
            const int mw1 = (int) (((t < 2 ? H1.m_lo : H1.m_hi) >> (16 * (t & 1))) & 0xffffu);
            const uint8_t * fs = F + fbuf * fstage_bytes;
            const uint8_t * xs = fs + g * 64;
            const int *     ss = reinterpret_cast<const int *>(fs + b * 512) + t;
            const float *   dxs = reinterpret_cast<const float *>(fs + b * 528);
#pragma unroll
            for (int k0 = 0; k0 < BMAX; k0 += 2) {
                if (k0 < b) {
                    constexpr int TB = BMAX > 1 ? 2 : 1;
                    int sumi[TB][4], sumh[TB][4];
#pragma unroll
                    for (int k = 0; k < TB; k++)
#pragma unroll
                        for (int i = 0; i < 4; i++) { sumi[k][i] = 0; sumh[k][i] = 0; }
                    const int xk1 = (TB > 1 && k0 + 1 < b) ? 512 : 0;        // an odd batch repeats its last token (computed, not used)
#pragma unroll
                    for (int ch = 0; ch < 4; ch++) {
                        const uint32_t wx = (uint32_t) wq[ch].x, wy = (uint32_t) wq[ch].y, wz = (uint32_t) wq[ch].z, ww = (uint32_t) wq[ch].w;
                        const int al0 = (int) (wx & 0x0F0F0F0Fu), al1 = (int) (wz & 0x0F0F0F0Fu), al2 = (int) (wy & 0x0F0F0F0Fu), al3 = (int) (ww & 0x0F0F0F0Fu);
                        const int ah0 = (int) (wx & 0xF0F0F0F0u), ah1 = (int) (wz & 0xF0F0F0F0u), ah2 = (int) (wy & 0xF0F0F0F0u), ah3 = (int) (ww & 0xF0F0F0F0u);
                        const int j0 = 2 * ch, j1 = 2 * ch + 1;
                        const int sc0_lo = (int) __byte_perm(j0 < 4 ? H0.sc_lo : H0.sc_hi, 0, 0x4440 + (j0 & 3));
                        const int sc1_lo = (int) __byte_perm(j0 < 4 ? H1.sc_lo : H1.sc_hi, 0, 0x4440 + (j0 & 3));
                        const int sc0_hi = (int) __byte_perm(j1 < 4 ? H0.sc_lo : H0.sc_hi, 0, 0x4440 + (j1 & 3));
                        const int sc1_hi = (int) __byte_perm(j1 < 4 ? H1.sc_lo : H1.sc_hi, 0, 0x4440 + (j1 & 3));
#pragma unroll
                        for (int k = 0; k < TB; k++) {
                            int4 xq = *reinterpret_cast<const int4 *>(xs + k0 * 512 + (k ? xk1 : 0) + ch * 16);
                            if (!b_active) xq = make_int4(0, 0, 0, 0);
                            int c_lo[4] = { 0, 0, 0, 0 }, c_hi[4] = { 0, 0, 0, 0 };
                            lx_mma_u8s8(c_lo, al0, al1, al2, al3, xq.x, xq.y);
                            lx_mma_u8s8(c_hi, ah0, ah1, ah2, ah3, xq.z, xq.w);
                            sumi[k][0] += sc0_lo * c_lo[0]; sumh[k][0] += sc0_hi * c_hi[0];
                            sumi[k][1] += sc0_lo * c_lo[1]; sumh[k][1] += sc0_hi * c_hi[1];
                            sumi[k][2] += sc1_lo * c_lo[2]; sumh[k][2] += sc1_hi * c_hi[2];
                            sumi[k][3] += sc1_lo * c_lo[3]; sumh[k][3] += sc1_hi * c_hi[3];
                        }
                    }
#pragma unroll
                    for (int k = 0; k < TB; k++) {
                        if (k0 + k < BMAX && (k == 0 || k0 + k < b)) {
                            const float dx = dxs[k0 + k];
                            const float d0 = __fmul_rn(dx, H0.d), d1 = __fmul_rn(dx, H1.d);
                            acc[k0 + k][0] = fmaf(d0, (float) (sumi[k][0] + (sumh[k][0] >> 4)), acc[k0 + k][0]);
                            acc[k0 + k][1] = fmaf(d0, (float) (sumi[k][1] + (sumh[k][1] >> 4)), acc[k0 + k][1]);
                            acc[k0 + k][2] = fmaf(d1, (float) (sumi[k][2] + (sumh[k][2] >> 4)), acc[k0 + k][2]);
                            acc[k0 + k][3] = fmaf(d1, (float) (sumi[k][3] + (sumh[k][3] >> 4)), acc[k0 + k][3]);
                            const int q8 = ss[(k0 + k) * 4];
                            accm[k0 + k][0] = fmaf(__fmul_rn(-dx, H0.dmin), (float) __dp2a_lo(q8, mw0, 0), accm[k0 + k][0]);
                            accm[k0 + k][1] = fmaf(__fmul_rn(-dx, H1.dmin), (float) __dp2a_lo(q8, mw1, 0), accm[k0 + k][1]);
                        }
                    }
                }
            }
            fbuf ^= 1;
            stage = stage + 1 == LX_STAGES ? 0 : stage + 1;
        }
        if (kind == 0) {
#pragma unroll
            for (int k = 0; k < BMAX; k++)
                if (k < b) {
                    __stcg(reinterpret_cast<float2 *>(hbase + k * 192), make_float2(acc[k][0], acc[k][1]));
                    __stcg(reinterpret_cast<float2 *>(hbase + k * 192 + 2), make_float2(acc[k][2], acc[k][3]));
                    __stcg(reinterpret_cast<float2 *>(hbase + k * 192 + 4), make_float2(accm[k][0], accm[k][1]));
                }
            __threadfence();
            __syncwarp();
            if (lane == 0) { __threadfence(); *(volatile unsigned int *) (flags + gw) = ready; }
        } else {
            // hsum_float_8 over the quad: thread t holds lanes 2t, 2t+1; ((a0+a4)+(a2+a6)) + ((a1+a5)+(a3+a7)), then the mins ((m0+m2)+(m1+m3))
            const int r0 = tile * 16 + g, r1 = r0 + 8;
#pragma unroll
            for (int k = 0; k < BMAX; k++)
                if (k < b) {
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        float p = acc[k][2 * h], q = acc[k][2 * h + 1], m = accm[k][h];
                        p = __fadd_rn(p, __shfl_xor_sync(0xffffffffu, p, 2));
                        q = __fadd_rn(q, __shfl_xor_sync(0xffffffffu, q, 2));
                        p = __fadd_rn(p, __shfl_xor_sync(0xffffffffu, p, 1));
                        q = __fadd_rn(q, __shfl_xor_sync(0xffffffffu, q, 1));
                        m = __fadd_rn(m, __shfl_xor_sync(0xffffffffu, m, 2));
                        m = __fadd_rn(m, __shfl_xor_sync(0xffffffffu, m, 1));
                        const int row = h ? r1 : r0;
                        if (t == 0 && row < N) y[(size_t) k * N + row] = __fadd_rn(__fadd_rn(p, q), m);
                    }
                }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        if (atomicAdd(&sync->ticket, 1u) == gridDim.x - 1) { sync->ticket = 0; __threadfence(); atomicAdd(&sync->epoch, 1u); }
    }
}

// ---- launcher -----------------------------------------------------------------------------------------------------------
struct lx_scratch { float * handoff = nullptr; unsigned int * flags = nullptr; lx_sync * sync = nullptr; int max_ctas = 0; };
static std::map<std::pair<int, cudaStream_t>, lx_scratch> g_lx;
static std::mutex                                        g_lx_mu;

static size_t lx_smem_bytes(int b) {
    return (size_t) 2 * LX_WARPS * 16 * LX_TERMS * b * 4 + 2 * LX_WARPS * 16 * 8 + 2 * LX_WARPS * 16 * 4 + (size_t) LX_WARPS * LX_STAGES * LX_TILE_BYTES +
           (size_t) LX_WARPS * 2 * (b * 512 + b * 16 + ((b * 4 + 15) & ~15));
}

// sized once per (device, stream) for the largest batch and grid the kernel can be launched with: nothing is ever reallocated, so
// addresses held by captured CUDA graphs stay valid
static int lx_get_scratch(cudaStream_t st, lx_scratch ** out) {
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lock(g_lx_mu);
    lx_scratch & s = g_lx[std::make_pair(dev, st)];
    if (!s.sync) {
        const int max_ctas = 5 * pib_num_sms();
        // [cta][token][thread] {lane acc, min acc} for the CTA-per-tile kernel, [warp][token][lane][6] for the warp-per-tile one
        const size_t hb = (size_t) max_ctas * LX_WARPS * (PIB_MAX_BATCH * 32 * sizeof(float2) > LX_MAXB_W * 32 * 6 * sizeof(float) ? PIB_MAX_BATCH * 32 * sizeof(float2) : LX_MAXB_W * 32 * 6 * sizeof(float));
        if (cudaMalloc(&s.handoff, hb) != cudaSuccess || cudaMalloc(&s.flags, (size_t) max_ctas * LX_WARPS * 4) != cudaSuccess ||
            cudaMalloc(&s.sync, sizeof(lx_sync)) != cudaSuccess) { cudaGetLastError(); return PIB_ERR_NOMEM; }
        cudaMemset(s.flags, 0, (size_t) max_ctas * LX_WARPS * 4);
        cudaMemset(s.sync, 0, sizeof(lx_sync));
        s.max_ctas = max_ctas;
    }
    *out = &s;
    return PIB_OK;
}

template <int BMAX, int TB>
static int lx_launch(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st, lx_scratch * s) {
    const size_t smem = lx_smem_bytes(b);
    static bool configured = false;
    static int  per_sm_of[PIB_MAX_BATCH + 1];                            // CTAs per SM for each batch size (shared memory grows with b)
    if (!configured) {
        PIB_CHECK(cudaFuncSetAttribute(gemv_q4k_lane_kernel<BMAX, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) lx_smem_bytes(PIB_MAX_BATCH)));
        for (int i = 0; i <= PIB_MAX_BATCH; i++) per_sm_of[i] = 0;
        configured = true;
    }
    if (!per_sm_of[b]) PIB_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_of[b], gemv_q4k_lane_kernel<BMAX, TB>, LX_THREADS, smem));
    int per_sm = per_sm_of[b];
    if (per_sm < 1) { pib_set_error("pib_gemv: lane kernel does not fit (b = %d)", b); return PIB_ERR_UNSUPP; }
    if (per_sm > 5) per_sm = 5;
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    const int slots = per_sm * pib_num_sms();
    // one tile per CTA while they all fit; else a persistent grid with equal pieces (a CTA's segment list holds LX_MAX_SEGS: beyond
    // ~20 tiles per CTA fall back to one tile per CTA in several waves)
    const int grid  = n_tiles <= slots ? n_tiles : ((long long) n_tiles <= (long long) (LX_MAX_SEGS - 4) * slots ? slots : n_tiles);
    const pib_act_layout L   = pib_act_layout_of(w->type, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    pib_launch(gemv_q4k_lane_kernel<BMAX, TB>, dim3(grid), dim3(LX_THREADS), smem, st, w->tiles, nsb, n_tiles, (int) w->N, (int) w->K,
               (const int2 *) (act + L.qsl_off), (const float *) (act + L.d_off), (const int16_t *) (act + L.sbs_off), b, y_dev,
               s->handoff, s->flags, s->sync, cancel);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

static size_t lxw_smem_bytes(int b) { return (size_t) LX_WARPS * (LX_STAGES * LX_TILE_BYTES + 2 * (b * 512 + b * 16 + ((b * 4 + 15) & ~15))); }

template <int BMAX>
static int lxw_launch(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st, lx_scratch * s) {
    const size_t smem = lxw_smem_bytes(b);
    static bool configured = false;
    static int  per_sm_of[PIB_MAX_BATCH + 1];
    if (!configured) {
        PIB_CHECK(cudaFuncSetAttribute(gemv_q4k_lanew_kernel<BMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) lxw_smem_bytes(LX_MAXB_W)));
        for (int i = 0; i <= PIB_MAX_BATCH; i++) per_sm_of[i] = 0;
        configured = true;
    }
    if (!per_sm_of[b]) PIB_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_of[b], gemv_q4k_lanew_kernel<BMAX>, LX_THREADS, smem));
    int per_sm = per_sm_of[b];
    if (per_sm < 1) { pib_set_error("pib_gemv: warp-per-tile kernel does not fit (b = %d)", b); return PIB_ERR_UNSUPP; }
    if (per_sm > 5) per_sm = 5;
    const int nsb = (int) (w->K / 256), n_tiles = (int) ((w->N + 15) / 16);
    const int slots = per_sm * pib_num_sms();                           // CTAs; 4 warps each
    // balanced pieces need a tile or more per warp; with fewer tiles every warp takes one tile
    const int grid = n_tiles >= slots * LX_WARPS ? slots : n_tiles / LX_WARPS;        // never more warps than tiles: a piece is a tile or more
    const pib_act_layout L   = pib_act_layout_of(w->type, w->K, b);
    const uint8_t *      act = (const uint8_t *) act_dev;
    pib_launch(gemv_q4k_lanew_kernel<BMAX>, dim3(grid), dim3(LX_THREADS), smem, st, w->tiles, nsb, n_tiles, (int) w->N, (int) w->K,
               (const int2 *) (act + L.qsl_off), (const float *) (act + L.d_off), (const int16_t *) (act + L.sbs_off), b, y_dev,
               s->handoff, s->flags, s->sync, cancel);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

bool pib_gemv_mma_supported(const pib_weight * w, int b);
static int lx_mode() { static int m = -1; if (m < 0) { const char * e = getenv("PIB_Q4K_MODE"); m = e ? atoi(e) : 0; } return m; }     // 1: CTA-per-tile always, 2: warp-per-tile whenever b <= 8

// variant 2 / default for Q4_K
int pib_gemv_q4k_lane(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    if (!pib_gemv_mma_supported(w, b)) { pib_set_error("pib_gemv: the Q4_K kernel needs Q4_K weights with tiles"); return PIB_ERR_UNSUPP; }
    lx_scratch * s = nullptr;
    if (lx_get_scratch(st, &s) != PIB_OK) { pib_set_error("pib_gemv: scratch allocation failed"); return PIB_ERR_NOMEM; }
    const int n_tiles = (int) ((w->N + 15) / 16);
    if (b <= LX_MAXB_W && lx_mode() != 1 && n_tiles >= LX_WARPS && (n_tiles >= 600 || lx_mode() == 2)) {
        if (b <= 1) return lxw_launch<1>(w, act_dev, b, y_dev, cancel, st, s);
        if (b <= 2) return lxw_launch<2>(w, act_dev, b, y_dev, cancel, st, s);
        if (b <= 4) return lxw_launch<4>(w, act_dev, b, y_dev, cancel, st, s);
        return lxw_launch<8>(w, act_dev, b, y_dev, cancel, st, s);
    }
    if (b <= 1) return lx_launch<1, 1>(w, act_dev, b, y_dev, cancel, st, s);
    if (b <= 2) return lx_launch<2, 2>(w, act_dev, b, y_dev, cancel, st, s);
    if (b <= 4) return lx_launch<4, 2>(w, act_dev, b, y_dev, cancel, st, s);
    if (b <= 8) return lx_launch<8, 2>(w, act_dev, b, y_dev, cancel, st, s);
    return lx_launch<16, 2>(w, act_dev, b, y_dev, cancel, st, s);
}
