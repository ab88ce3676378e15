// weights.cu -- upload and repack of quantised weight matrices.
//
// The reference keeps weights as arrays of blocks (block_q4_0 18 B, block_q8_0 34 B, block_q4_K 144 B; reference
// ggml-quants.h:10-15, 42-47, 116-122).  18- and 34-byte blocks cannot be read with aligned 128-bit loads or moved by
// TMA bulk copies (16-byte granularity), so at load time each matrix is split into planes.  Values are untouched:
//
//   Q8_0:  qs  int8  [N][K]          d   fp16 [N][K/32]
//   Q4_0:  qs  uint8 [N][K/2]        d   fp16 [N][K/32]      (byte j of a block = element j low nibble, j+16 high)
//   Q4_K:  qs  uint8 [N][K/2]        hdr 16 B [N][K/256]     {fp16 d, fp16 dmin, uint8 scales[12]}
//   Q5_K:  qs  160 B [N][K/256]      hdr 16 B [N][K/256]     {qs[128], qh[32]} and {fp16 d, fp16 dmin, uint8 scales[12]}
//   Q3_K:  qs   96 B [N][K/256]      hdr 16 B [N][K/256]     {qs[64] low bit pairs, hmask[32]} and {scales[12], fp16 d, pad}
//   Q6_K:  qs  192 B [N][K/256]      hdr 32 B [N][K/256]     {ql[128], qh[64]} and {int8 scales[16], fp16 d, pad}
//
// The tensor-core GEMV additionally builds fragment-major tiles from these planes (gemv_mma.cu).
#include "pib_common.cuh"

#include <string.h>

static __global__ void repack_q8_0(const uint8_t * __restrict__ src, int8_t * __restrict__ qs, uint16_t * __restrict__ d,
                                   int64_t nblk) {
    int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblk) return;
    const uint8_t * b = src + i * 34;
    d[i]              = (uint16_t) (b[0] | (b[1] << 8));
    int8_t * o        = qs + i * 32;
#pragma unroll
    for (int j = 0; j < 32; j++) o[j] = (int8_t) b[2 + j];
}

static __global__ void repack_q4_0(const uint8_t * __restrict__ src, uint8_t * __restrict__ qs, uint16_t * __restrict__ d,
                                   int64_t nblk) {
    int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblk) return;
    const uint8_t * b = src + i * 18;
    d[i]              = (uint16_t) (b[0] | (b[1] << 8));
    uint8_t * o       = qs + i * 16;
#pragma unroll
    for (int j = 0; j < 16; j++) o[j] = b[2 + j];
}

static __global__ void repack_q4_K(const uint8_t * __restrict__ src, uint8_t * __restrict__ qs, uint8_t * __restrict__ hdr,
                                   int64_t nblk) {
    int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblk * 9) return;                 // 9 x 16-byte pieces per 144-byte super-block
    int64_t     blk = i / 9;
    int         p   = (int) (i % 9);
    const int4  v   = *reinterpret_cast<const int4 *>(src + blk * 144 + p * 16);
    if (p == 0) *reinterpret_cast<int4 *>(hdr + blk * 16) = v;
    else        *reinterpret_cast<int4 *>(qs + blk * 128 + (p - 1) * 16) = v;
}

// block_q5_K (ggml-quants.h:138-145): d, dmin, scales[12], qh[32], qs[128] -- 176 bytes = 11 x 16
static __global__ void repack_q5_K(const uint8_t * __restrict__ src, uint8_t * __restrict__ qs, uint8_t * __restrict__ hdr, int64_t nblk) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblk * 11) return;
    const int64_t blk = i / 11;
    const int     p   = (int) (i % 11);
    const int4    v   = *reinterpret_cast<const int4 *>(src + blk * 176 + p * 16);
    if (p == 0)     *reinterpret_cast<int4 *>(hdr + blk * 16) = v;
    else if (p < 3) *reinterpret_cast<int4 *>(qs + blk * 160 + 128 + (p - 1) * 16) = v;
    else            *reinterpret_cast<int4 *>(qs + blk * 160 + (p - 3) * 16) = v;
}

// block_q6_K (ggml-quants.h:152-157): ql[128], qh[64], scales[16], d -- 210 bytes, not even 4-byte aligned in a row
static __global__ void repack_q6_K(const uint8_t * __restrict__ src, uint8_t * __restrict__ qs, uint8_t * __restrict__ hdr, int64_t nblk) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblk * 16) return;                // 16 threads per super-block: 12 x 16 quant bytes, 1 x scales, 1 x d, 2 idle
    const int64_t blk = i / 16;
    const int     p   = (int) (i % 16);
    const uint8_t * b = src + blk * 210;
    if (p < 12) { for (int k = 0; k < 16; k++) qs[blk * 192 + p * 16 + k] = b[p * 16 + k]; }
    else if (p == 12) { for (int k = 0; k < 16; k++) hdr[blk * 32 + k] = b[192 + k]; }
    else if (p == 13) { hdr[blk * 32 + 16] = b[208]; hdr[blk * 32 + 17] = b[209]; for (int k = 18; k < 32; k++) hdr[blk * 32 + k] = 0; }
}

// block_q3_K (ggml-quants.h:95-101): hmask[32], qs[64], scales[12], d -- 110 bytes
static __global__ void repack_q3_K(const uint8_t * __restrict__ src, uint8_t * __restrict__ qs, uint8_t * __restrict__ hdr, int64_t nblk) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblk * 8) return;                 // 8 threads per super-block: 4 x 16 bytes of qs, 2 x 16 of hmask, 1 x header, 1 idle
    const int64_t blk = i / 8;
    const int     p   = (int) (i % 8);
    const uint8_t * b = src + blk * 110;
    if (p < 4)       { for (int k = 0; k < 16; k++) qs[blk * 96 + p * 16 + k] = b[32 + p * 16 + k]; }
    else if (p < 6)  { for (int k = 0; k < 16; k++) qs[blk * 96 + 64 + (p - 4) * 16 + k] = b[(p - 4) * 16 + k]; }
    else if (p == 6) { for (int k = 0; k < 14; k++) hdr[blk * 16 + k] = b[96 + k]; hdr[blk * 16 + 14] = 0; hdr[blk * 16 + 15] = 0; }
}

static int block_geom(int type, int * elems, int * bytes) {
    switch (type) {
        case PIB_TYPE_Q4_0: *elems = 32;  *bytes = 18;  return 0;
        case PIB_TYPE_Q8_0: *elems = 32;  *bytes = 34;  return 0;
        case PIB_TYPE_Q4_K: *elems = 256; *bytes = 144; return 0;
        case PIB_TYPE_Q3_K: *elems = 256; *bytes = 110; return 0;
        case PIB_TYPE_Q5_K: *elems = 256; *bytes = 176; return 0;
        case PIB_TYPE_Q6_K: *elems = 256; *bytes = 210; return 0;
    }
    return -1;
}

static pib_weight * alloc_weight(int type, int64_t N, int64_t K) {
    int be, bb;
    if (block_geom(type, &be, &bb) != 0 || N <= 0 || K <= 0 || K % be != 0) {
        pib_set_error("pib_weight: unsupported type %d or shape %lld x %lld", type, (long long) N, (long long) K);
        return nullptr;
    }
    pib_weight * w = (pib_weight *) calloc(1, sizeof(pib_weight));
    w->type  = type;
    w->N     = N;
    w->K     = K;
    w->bytes = (size_t) N * (K / be) * bb;
    size_t qs_bytes = (type == PIB_TYPE_Q8_0) ? (size_t) N * K : (type == PIB_TYPE_Q6_K) ? (size_t) N * (K / 256) * 192 : (type == PIB_TYPE_Q5_K) ? (size_t) N * (K / 256) * 160
                    : (type == PIB_TYPE_Q3_K) ? (size_t) N * (K / 256) * 96 : (size_t) N * K / 2;
    size_t d_bytes  = (type == PIB_TYPE_Q4_K || type == PIB_TYPE_Q5_K || type == PIB_TYPE_Q3_K) ? (size_t) N * (K / 256) * 16 : (type == PIB_TYPE_Q6_K) ? (size_t) N * (K / 256) * 32
                                                                            : (size_t) N * (K / 32) * 2;
    size_t d_off    = (qs_bytes + 255) & ~(size_t) 255;
    if (cudaMalloc(&w->base, d_off + d_bytes + 256) != cudaSuccess) {
        pib_set_error("pib_weight: cudaMalloc of %zu bytes failed", d_off + d_bytes);
        cudaGetLastError();
        free(w);
        return nullptr;
    }
    w->qs   = (uint8_t *) w->base;
    w->d    = (uint16_t *) ((uint8_t *) w->base + d_off);
    w->owns = 1;
    return w;
}

static int repack_from_device(pib_weight * w, const uint8_t * src, int64_t row0, int64_t rows, cudaStream_t st) {
    int be, bb;
    block_geom(w->type, &be, &bb);
    const int64_t bpr  = w->K / be;
    const int64_t nblk = rows * bpr;
    const int64_t blk0 = row0 * bpr;
    const int     T    = 256;
    switch (w->type) {
        case PIB_TYPE_Q8_0:
            repack_q8_0<<<(unsigned) ((nblk + T - 1) / T), T, 0, st>>>(src, (int8_t *) w->qs + blk0 * 32, w->d + blk0, nblk);
            break;
        case PIB_TYPE_Q4_0:
            repack_q4_0<<<(unsigned) ((nblk + T - 1) / T), T, 0, st>>>(src, w->qs + blk0 * 16, w->d + blk0, nblk);
            break;
        case PIB_TYPE_Q5_K:
            repack_q5_K<<<(unsigned) ((nblk * 11 + T - 1) / T), T, 0, st>>>(src, w->qs + blk0 * 160, (uint8_t *) w->d + blk0 * 16, nblk);
            break;
        case PIB_TYPE_Q3_K:
            repack_q3_K<<<(unsigned) ((nblk * 8 + T - 1) / T), T, 0, st>>>(src, w->qs + blk0 * 96, (uint8_t *) w->d + blk0 * 16, nblk);
            break;
        case PIB_TYPE_Q6_K:
            repack_q6_K<<<(unsigned) ((nblk * 16 + T - 1) / T), T, 0, st>>>(src, w->qs + blk0 * 192, (uint8_t *) w->d + blk0 * 32, nblk);
            break;
        case PIB_TYPE_Q4_K:
            repack_q4_K<<<(unsigned) ((nblk * 9 + T - 1) / T), T, 0, st>>>(src, w->qs + blk0 * 128,
                                                                            (uint8_t *) w->d + blk0 * 16, nblk);
            break;
    }
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

int pib_weight_build_tiles(pib_weight * w);   // gemv_mma.cu

pib_weight * pib_weight_alloc(int type, int64_t N, int64_t K) { return alloc_weight(type, N, K); }

extern "C" pib_weight * pib_weight_upload(int type, const void * host_blocks, int64_t N, int64_t K) {
    pib_weight * w = alloc_weight(type, N, K);
    if (!w) return nullptr;
    int be, bb;
    block_geom(type, &be, &bb);
    const size_t row_bytes = (size_t) (K / be) * bb;
    // staged in <= 256 MiB pieces of whole rows so that a 100 GB model never needs a second full-size buffer
    int64_t rows_per = (int64_t) ((256u << 20) / row_bytes);
    if (rows_per < 1) rows_per = 1;
    if (rows_per > N) rows_per = N;
    uint8_t * stage = nullptr;
    if (cudaMalloc(&stage, (size_t) rows_per * row_bytes + 16) != cudaSuccess) {
        pib_set_error("pib_weight_upload: staging cudaMalloc failed");
        cudaGetLastError();
        pib_weight_free(w);
        return nullptr;
    }
    for (int64_t r0 = 0; r0 < N; r0 += rows_per) {
        int64_t rows = N - r0 < rows_per ? N - r0 : rows_per;
        if (cudaMemcpy(stage, (const uint8_t *) host_blocks + (size_t) r0 * row_bytes, (size_t) rows * row_bytes,
                       cudaMemcpyHostToDevice) != cudaSuccess ||
            repack_from_device(w, stage, r0, rows, 0) != PIB_OK || cudaDeviceSynchronize() != cudaSuccess) {
            pib_set_error("pib_weight_upload: copy/repack failed: %s", cudaGetErrorString(cudaGetLastError()));
            cudaFree(stage);
            pib_weight_free(w);
            return nullptr;
        }
    }
    cudaFree(stage);
    if (pib_weight_build_tiles(w) != PIB_OK || cudaDeviceSynchronize() != cudaSuccess) { pib_weight_free(w); return nullptr; }
    return w;
}

extern "C" pib_weight * pib_weight_from_device(int type, const void * dev_blocks, int64_t N, int64_t K) {
    pib_weight * w = alloc_weight(type, N, K);
    if (!w) return nullptr;
    if (repack_from_device(w, (const uint8_t *) dev_blocks, 0, N, 0) != PIB_OK || cudaDeviceSynchronize() != cudaSuccess) {
        pib_set_error("pib_weight_from_device: repack failed: %s", cudaGetErrorString(cudaGetLastError()));
        pib_weight_free(w);
        return nullptr;
    }
    if (pib_weight_build_tiles(w) != PIB_OK || cudaDeviceSynchronize() != cudaSuccess) { pib_weight_free(w); return nullptr; }
    return w;
}

extern "C" void pib_weight_free(pib_weight * w) {
    if (!w) return;
    if (w->owns && w->base) cudaFree(w->base);
    if (w->tiles) cudaFree(w->tiles);
    free(w);
}

// The tensor-core kernels read only the fragment-major tiles; the plane form (needed by the dp4a variant, get_rows and the
// block round trip) is a second full copy.  The engine drops it for the matrices it only ever multiplies with, which
// halves the resident bytes of a model (70B Q4_K: 39 GB instead of 78 GB).
void pib_weight_drop_planes(pib_weight * w) {
    if (!w || !w->tiles || !w->owns || !w->base) return;
    cudaFree(w->base);
    w->base = nullptr; w->qs = nullptr; w->d = nullptr;
}

extern "C" size_t pib_weight_bytes(const pib_weight * w) { return w ? w->bytes : 0; }

extern "C" int pib_weight_row_blocks(const pib_weight * w, int64_t row, void * out) {
    if (!w || row < 0 || row >= w->N) { pib_set_error("pib_weight_row_blocks: bad row"); return PIB_ERR_ARG; }
    if (!w->qs) { pib_set_error("pib_weight_row_blocks: this matrix keeps only its tensor-core tiles"); return PIB_ERR_UNSUPP; }
    uint8_t * o = (uint8_t *) out;
    if (w->type == PIB_TYPE_Q5_K) {
        const int64_t nsb = w->K / 256;
        for (int64_t s = 0; s < nsb; s++) {
            PIB_CHECK(cudaMemcpy(o + s * 176, (uint8_t *) w->d + (row * nsb + s) * 16, 16, cudaMemcpyDeviceToHost));
            PIB_CHECK(cudaMemcpy(o + s * 176 + 16, w->qs + (row * nsb + s) * 160 + 128, 32, cudaMemcpyDeviceToHost));
            PIB_CHECK(cudaMemcpy(o + s * 176 + 48, w->qs + (row * nsb + s) * 160, 128, cudaMemcpyDeviceToHost));
        }
        return PIB_OK;
    }
    if (w->type == PIB_TYPE_Q3_K) {
        const int64_t nsb = w->K / 256;
        for (int64_t s = 0; s < nsb; s++) {
            PIB_CHECK(cudaMemcpy(o + s * 110, w->qs + (row * nsb + s) * 96 + 64, 32, cudaMemcpyDeviceToHost));
            PIB_CHECK(cudaMemcpy(o + s * 110 + 32, w->qs + (row * nsb + s) * 96, 64, cudaMemcpyDeviceToHost));
            PIB_CHECK(cudaMemcpy(o + s * 110 + 96, (uint8_t *) w->d + (row * nsb + s) * 16, 14, cudaMemcpyDeviceToHost));
        }
        return PIB_OK;
    }
    if (w->type == PIB_TYPE_Q6_K) {
        const int64_t nsb = w->K / 256;
        for (int64_t s = 0; s < nsb; s++) {
            PIB_CHECK(cudaMemcpy(o + s * 210, w->qs + (row * nsb + s) * 192, 192, cudaMemcpyDeviceToHost));
            PIB_CHECK(cudaMemcpy(o + s * 210 + 192, (uint8_t *) w->d + (row * nsb + s) * 32, 18, cudaMemcpyDeviceToHost));
        }
        return PIB_OK;
    }
    if (w->type == PIB_TYPE_Q4_K) {
        const int64_t nsb = w->K / 256;
        for (int64_t s = 0; s < nsb; s++) {
            PIB_CHECK(cudaMemcpy(o + s * 144, (uint8_t *) w->d + (row * nsb + s) * 16, 16, cudaMemcpyDeviceToHost));
            PIB_CHECK(cudaMemcpy(o + s * 144 + 16, w->qs + (row * nsb + s) * 128, 128, cudaMemcpyDeviceToHost));
        }
        return PIB_OK;
    }
    const int64_t nb  = w->K / 32;
    const int     qb  = w->type == PIB_TYPE_Q8_0 ? 32 : 16;
    uint8_t *     qs  = (uint8_t *) malloc((size_t) nb * qb);
    uint16_t *    d   = (uint16_t *) malloc((size_t) nb * 2);
    cudaError_t   e1  = cudaMemcpy(qs, w->qs + (size_t) row * nb * qb, (size_t) nb * qb, cudaMemcpyDeviceToHost);
    cudaError_t   e2  = cudaMemcpy(d, w->d + (size_t) row * nb, (size_t) nb * 2, cudaMemcpyDeviceToHost);
    if (e1 == cudaSuccess && e2 == cudaSuccess) {
        for (int64_t b = 0; b < nb; b++) {
            memcpy(o + b * (2 + qb), d + b, 2);
            memcpy(o + b * (2 + qb) + 2, qs + b * qb, (size_t) qb);
        }
    }
    free(qs);
    free(d);
    PIB_CHECK(e1);
    PIB_CHECK(e2);
    return PIB_OK;
}
