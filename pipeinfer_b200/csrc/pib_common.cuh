// pib_common.cuh -- shared device/host helpers for libpipeinfer_b200 (sm_100a only).
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/pipeinfer_b200.h"

#define PIB_CHECK(expr)                                                                                      \
    do {                                                                                                     \
        cudaError_t e__ = (expr);                                                                            \
        if (e__ != cudaSuccess) {                                                                            \
            pib_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__));            \
            return PIB_ERR_CUDA;                                                                             \
        }                                                                                                    \
    } while (0)

#define PIB_CHECK_PTR(expr)                                                                                  \
    do {                                                                                                     \
        cudaError_t e__ = (expr);                                                                            \
        if (e__ != cudaSuccess) {                                                                            \
            pib_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__));            \
            return nullptr;                                                                                  \
        }                                                                                                    \
    } while (0)

void pib_set_error(const char * fmt, ...);

// block geometry of the three weight formats on the path (reference: ggml-quants.h:10-15, 42-47, 116-122, 161-166)
#define PIB_QK   32
#define PIB_QK_K 256

// fp16 look-up tables (exp / silu / gelu), uploaded by pib_init from host-computed values so that they are the very
// tables the reference builds with the host libm (ggml.c:2185-2203)
struct pib_luts {
    const uint16_t * exp_f16;
    const uint16_t * silu_f16;
    const uint16_t * gelu_f16;
};
const pib_luts & pib_get_luts();   // asserts pib_init was called on the current device
int pib_num_sms();

// ---- device helpers ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float pib_h2f(uint16_t h) { return __half2float(__ushort_as_half(h)); }
__device__ __forceinline__ uint16_t pib_f2h(float f) { return __half_as_ushort(__float2half_rn(f)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Programmatic dependent launch (PDL): a kernel launched with pib_launch may become resident while its predecessor in
// the stream is still running; it must call pib_pdl_wait() before it touches anything the predecessor wrote (or before it
// exits: a cancelled kernel that returned without waiting would let ITS successor overtake the predecessor).  Only
// immutable data (weights, tables) may be read before the wait.  pib_pdl_trigger() lets the successor start its own
// prologue; it is issued after the wait, so at most two kernels overlap.
__device__ __forceinline__ void pib_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pib_pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Cancellation block of a context (device): int cur = run slot whose kernels are executing on the context's stream (set in
// stream order in front of every run), then one flag per run slot.  Kernels get the block's address, which never
// changes, so a captured CUDA graph serves every run; a null pointer means "not cancellable".
#define PIB_CANCELLED(cancel) ((cancel) != nullptr && (cancel)[1 + (cancel)[0]] != 0)

// 128-bit read-only streaming load that does not allocate in L1 (weights are read exactly once per launch)
__device__ __forceinline__ int4 ld_stream_v4(const void * p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ int2 ld_stream_v2(const void * p) {
    int2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.s32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}

// launch with the PDL attribute (env PIB_PDL=0 turns it off: plain stream order)
bool pib_pdl_enabled();
void pib_pdl_suspend(bool suspend);   // event-timed launches must not overlap their predecessor
template <typename... KArgs, typename... Args>
static inline cudaError_t pib_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr;
    memset(&attr, 0, sizeof(attr));
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr; cfg.numAttrs = pib_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// device view of a repacked weight matrix (see weights.cu for the layouts)
struct pib_weight {
    int      type;      // PIB_TYPE_*
    int64_t  N, K;      // rows (outputs), columns (inputs)
    void *   base;      // one allocation
    size_t   bytes;     // algorithmic bytes == bytes of the reference block layout (N * K/blk * blk_bytes)
    // v1 planes (row-major SoA)
    uint8_t * qs;       // Q8_0: int8 [N][K];  Q4_0: [N][K/2] nibbles in block order;  Q4_K: [N][K/2] nibbles
    uint16_t * d;       // Q8_0/Q4_0: fp16 [N][K/32];  Q4_K: 16-byte headers {d, dmin, scales[12]} as [N][K/256][8] halves
    // v2 tiles (fragment-major, see gemv_mma.cu)
    uint8_t * tiles;
    size_t    tile_bytes;
    int       owns;
};

// activation format a weight type multiplies with, named by the weight type that owns the layout below: the K-quants share
// Q8_K (reference type_traits[].vec_dot_type, ggml.c:455-560)
static inline int pib_act_family(int wtype) { return (wtype == PIB_TYPE_Q6_K || wtype == PIB_TYPE_Q5_K || wtype == PIB_TYPE_Q3_K) ? PIB_TYPE_Q4_K : wtype; }

// byte layout of a quantised-activation buffer for `b` tokens (device, produced by act_quant.cu)
struct pib_act_layout {
    size_t qs_off;     // int8  [b][K]
    size_t d_off;      // float [b][K/32] (Q8_0 family) or [b][K/256] (Q8_K family)
    size_t bs_off;     // Q8_0 family: int32 [b][K/32] block sums (unused by the reference, used for the -8 offset);
                       // Q8_K family: int16 [b][K/16] bsums exactly as block_q8_K
    size_t bs8_off;    // Q8_K family: int8 [b][K/256][16] = {hi_0..hi_7, lo_0..lo_7}: sub-block sums bsums[2j] + bsums[2j+1]
                       //   split as 64 * hi + lo (hi in [-64, 63], lo in [0, 63]) so that the min term of Q4_K can be a
                       //   tensor-core dot product; redundant with bs_off (same numbers)
    size_t qsp_off;    // Q8_K family: int8 [b][K], the same quants as qs_off in tensor-core fragment order: inside each
                       //   super-block, element e = 64c + 32h + 8t + i (chunk c, low/high-nibble half h, lane quarter t) is
                       //   stored at 64t + 16c + 8h + i, so a lane's B fragments for all four chunks are 64 contiguous bytes
    size_t qsl_off;    // Q8_K family: int2 [b][K/256][8 lane groups g][8 sub-blocks j]: {x if g even else 0, x if g odd else 0} with x =
                       //   the four quants 4g .. 4g+3 of sub-block j -- the two B registers of the lane-exact kernel (gemv_q4k_lane.cu)
    size_t sbs_off;    // Q8_K family: int16 [b][K/32] sub-block sums bsums[2j] + bsums[2j+1]
    size_t q6p_off;    // Q8_K family: int8 [b][K], the quants in the B-fragment order of mma.m16n8k16 over the 16-element groups of
                       //   Q6_K: inside each super-block element e = 16j + 4t + i (group j, lane quarter t) sits at 64t + 4j + i
    size_t b16_off;    // Q8_K family: int8 [b][K/256][32] = {hi_0..hi_15, lo_0..lo_15}: the 16 bsums split as 64 * hi + lo, so that the
                       //   -32 offset of Q6_K (sum_j scale_j * bsum_j) is two more tensor-core dot products (gemv_q6k.cu)
    size_t total;
};
__host__ __device__ inline pib_act_layout pib_act_layout_of(int wtype, int64_t K, int b) {
    pib_act_layout L;
    L.qs_off = 0;
    size_t qs = (size_t) b * K;
    L.d_off  = (qs + 255) & ~(size_t) 255;
    size_t nd = (wtype == PIB_TYPE_Q4_K) ? (size_t) b * (K / PIB_QK_K) : (size_t) b * (K / PIB_QK);
    L.bs_off = (L.d_off + nd * 4 + 255) & ~(size_t) 255;
    size_t nbs = (wtype == PIB_TYPE_Q4_K) ? (size_t) b * (K / 16) * 2 : (size_t) b * (K / PIB_QK) * 4;
    L.bs8_off = (L.bs_off + nbs + 255) & ~(size_t) 255;
    // 32-element block formats: bs8 holds the "dxp" plane (fp16 block scales [b][ngr * 8], ngr = groups of 8 blocks, zero
    // padded) and qsp the fragment-order quants [b][ngr * 256] of gemv_blk32.cu
    const size_t ngr = (size_t) ((K + 255) / 256);
    size_t nbs8 = (wtype == PIB_TYPE_Q4_K) ? (size_t) b * (K / PIB_QK_K) * 16 : (size_t) b * ngr * 8 * 2;
    L.qsp_off = (L.bs8_off + nbs8 + 255) & ~(size_t) 255;
    L.qsl_off = (L.qsp_off + ((wtype == PIB_TYPE_Q4_K) ? qs : (size_t) b * ngr * 256) + 255) & ~(size_t) 255;
    L.sbs_off = (L.qsl_off + ((wtype == PIB_TYPE_Q4_K) ? 2 * qs : 0) + 255) & ~(size_t) 255;
    L.q6p_off = (L.sbs_off + ((wtype == PIB_TYPE_Q4_K) ? (size_t) b * (K / 32) * 2 : 0) + 255) & ~(size_t) 255;
    L.b16_off = (L.q6p_off + ((wtype == PIB_TYPE_Q4_K) ? qs : 0) + 255) & ~(size_t) 255;
    L.total   = (L.b16_off + ((wtype == PIB_TYPE_Q4_K) ? (size_t) b * (K / PIB_QK_K) * 32 : 0) + 255) & ~(size_t) 255;
    return L;
}
