// lib.cu -- library state (device, look-up tables, error string), small C-ABI helpers, GEMV dispatch, embedding rows.
#include "pib_common.cuh"

#include <math.h>
#include <stdarg.h>
#include <string.h>

#include <mutex>

static thread_local char g_err[512] = "";

void pib_set_error(const char * fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char * pib_last_error(void) { return g_err; }

#define PIB_MAX_DEV 16
static pib_luts   g_luts[PIB_MAX_DEV];
static int        g_sms[PIB_MAX_DEV];
static bool       g_ready[PIB_MAX_DEV];
static std::mutex g_mu;

static inline uint16_t host_f2h(float f) { const __half h = __float2half_rn(f); uint16_t u; memcpy(&u, &h, 2); return u; }
static inline float    host_h2f(uint16_t u) { __half h; memcpy(&h, &u, 2); return __half2float(h); }

extern "C" int pib_init(int device) {
    std::lock_guard<std::mutex> lk(g_mu);
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        pib_set_error("pib_init: no CUDA device (libpipeinfer_b200 has no CPU path)");
        return PIB_ERR_CUDA;
    }
    if (device < 0 || device >= n || device >= PIB_MAX_DEV) { pib_set_error("pib_init: bad device %d", device); return PIB_ERR_ARG; }
    PIB_CHECK(cudaSetDevice(device));
    if (g_ready[device]) return PIB_OK;
    cudaDeviceProp prop;
    PIB_CHECK(cudaGetDeviceProperties(&prop, device));
    g_sms[device] = prop.multiProcessorCount;
    // the reference fills these tables in ggml_init with the host libm (ggml.c:2190-2200): value = fp16(f(fp32(fp16 i)))
    const int  n16 = 1 << 16;
    uint16_t * h   = (uint16_t *) malloc(3 * n16 * sizeof(uint16_t));
    for (int i = 0; i < n16; i++) {
        const float f = host_h2f((uint16_t) i);
        h[i]           = host_f2h(expf(f));
        h[n16 + i]     = host_f2h(f / (1.0f + expf(-f)));
        h[2 * n16 + i] = host_f2h(0.5f * f * (1.0f + tanhf(0.79788456080286535587989211986876f * f * (1.0f + 0.044715f * f * f))));
    }
    uint16_t * d = nullptr;
    PIB_CHECK(cudaMalloc(&d, 3 * n16 * sizeof(uint16_t)));
    PIB_CHECK(cudaMemcpy(d, h, 3 * n16 * sizeof(uint16_t), cudaMemcpyHostToDevice));
    free(h);
    g_luts[device].exp_f16  = d;
    g_luts[device].silu_f16 = d + n16;
    g_luts[device].gelu_f16 = d + 2 * n16;
    g_ready[device]         = true;
    return PIB_OK;
}

const pib_luts & pib_get_luts() {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= PIB_MAX_DEV || !g_ready[dev]) {
        fprintf(stderr, "libpipeinfer_b200: pib_init() was not called for device %d\n", dev);
        abort();
    }
    return g_luts[dev];
}

static int g_pdl_suspended = 0;
void pib_pdl_suspend(bool suspend) { g_pdl_suspended = suspend ? 1 : 0; }
bool pib_pdl_enabled() {
    static int on = -1;
    if (on < 0) { const char * e = getenv("PIB_PDL"); on = (!e || atoi(e) != 0) ? 1 : 0; }
    return on != 0 && !g_pdl_suspended;
}

int pib_num_sms() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < PIB_MAX_DEV && g_ready[dev]) ? g_sms[dev] : 148;
}

extern "C" int pib_device_sm_count(void) { return pib_num_sms(); }

extern "C" void * pib_dev_alloc(size_t bytes) {
    void * p = nullptr;
    if (cudaMalloc(&p, bytes ? bytes : 1) != cudaSuccess) {
        pib_set_error("pib_dev_alloc(%zu): %s", bytes, cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return p;
}
extern "C" void pib_dev_free(void * p) { if (p) cudaFree(p); }
extern "C" int  pib_memcpy_h2d(void * dev, const void * host, size_t bytes, pib_stream s) {
    PIB_CHECK(cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, (cudaStream_t) s));
    return PIB_OK;
}
extern "C" int pib_memcpy_d2h(void * host, const void * dev, size_t bytes, pib_stream s) {
    PIB_CHECK(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, (cudaStream_t) s));
    return PIB_OK;
}
extern "C" int pib_stream_sync(pib_stream s) {
    PIB_CHECK(cudaStreamSynchronize((cudaStream_t) s));
    return PIB_OK;
}

// ---- K1 dispatch ----------------------------------------------------------------------------------------------------
int pib_gemv_dp4a(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
int pib_gemv_mma(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
bool pib_gemv_mma_supported(const pib_weight * w, int b);
bool pib_gemv_blk32_supported(const pib_weight * w, int b);
int  pib_gemv_q5k(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
int  pib_gemv_q3k(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
int  pib_get_rows_q3k(const pib_weight * w, const int32_t * ids_dev, int n, float * out_dev, cudaStream_t st);
int  pib_get_rows_q5k(const pib_weight * w, const int32_t * ids_dev, int n, float * out_dev, cudaStream_t st);
int  pib_get_rows_q6k(const pib_weight * w, const int32_t * ids_dev, int n, float * out_dev, cudaStream_t st);
int  pib_gemv_q6k(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
int  pib_gemv_blk32(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
int  pib_gemv_ring(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
int  pib_gemv_direct(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
int  pib_gemv_stream(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
int  pib_gemv_q4k_lane(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);
int  pib_gemv_exact_ref(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st);

int pib_gemv_internal(const pib_weight * w, const void * act_dev, int b, float * y_dev, int variant, const int * cancel,
                      cudaStream_t st) {
    // the other K-quants have one fast kernel each; exact mode (7) and variant 4 take the reference-order kernel when the planes are there
    if ((variant == 7 || variant == 4) && (w->type == PIB_TYPE_Q6_K || w->type == PIB_TYPE_Q5_K || w->type == PIB_TYPE_Q3_K) && w->qs)
        return pib_gemv_exact_ref(w, act_dev, b, y_dev, cancel, st);
    if (w->type == PIB_TYPE_Q6_K) return pib_gemv_q6k(w, act_dev, b, y_dev, cancel, st);
    if (w->type == PIB_TYPE_Q5_K) return pib_gemv_q5k(w, act_dev, b, y_dev, cancel, st);
    if (w->type == PIB_TYPE_Q3_K) return pib_gemv_q3k(w, act_dev, b, y_dev, cancel, st);
    // variants: 0 / 2 the fast tensor-core kernels (fp32 block terms added in the kernel's own order), 1 dp4a, 3 TMA-staged persistent
    // (Q4_K), 4 the reference's summation order restated literally (one thread per output; bit-identical to the CPU, slow), 6 the
    // tensor-core kernel that keeps the reference's order (Q4_K; bit-identical to the CPU).  7 = "exact mode": 6 where it exists, else 4.
    if (variant == 7) variant = (w->type == PIB_TYPE_Q4_K && pib_gemv_mma_supported(w, b)) ? 6 : 4;
    if (variant == 4) return pib_gemv_exact_ref(w, act_dev, b, y_dev, cancel, st);
    const bool tc = pib_gemv_mma_supported(w, b);
    if (variant == 6 && tc) return pib_gemv_q4k_lane(w, act_dev, b, y_dev, cancel, st);
    if (variant == 3 && tc) return pib_gemv_mma(w, act_dev, b, y_dev, cancel, st);
    if ((variant == 2 || variant == 0) && tc) return pib_gemv_ring(w, act_dev, b, y_dev, cancel, st);
    if (variant == 5 && tc) return pib_gemv_direct(w, act_dev, b, y_dev, cancel, st);
    if (variant == 8 && tc) return pib_gemv_stream(w, act_dev, b, y_dev, cancel, st);
    if ((variant == 2 || variant == 0) && pib_gemv_blk32_supported(w, b)) return pib_gemv_blk32(w, act_dev, b, y_dev, cancel, st);
    return pib_gemv_dp4a(w, act_dev, b, y_dev, cancel, st);
}

extern "C" int pib_gemv(const pib_weight * w, const void * act_dev, int b, float * y_dev, int variant, pib_stream s) {
    if (!w || !act_dev || !y_dev || b < 1 || b > PIB_MAX_BATCH) {
        pib_set_error("pib_gemv: bad arguments (b = %d, limit %d)", b, PIB_MAX_BATCH);
        return PIB_ERR_ARG;
    }
    if ((w->type == PIB_TYPE_Q6_K || w->type == PIB_TYPE_Q5_K || w->type == PIB_TYPE_Q3_K) && variant != 4 && variant != 7) variant = 0;   // one fast kernel for these types
    if (((variant == 3 || variant == 6) && !pib_gemv_mma_supported(w, b)) || (variant == 2 && !pib_gemv_mma_supported(w, b) && !pib_gemv_blk32_supported(w, b))) {
        pib_set_error("pib_gemv: variant 2 needs tiles (Q4_K, Q4_0, Q8_0), variants 3 and 6 Q4_K");
        return PIB_ERR_UNSUPP;
    }
    return pib_gemv_internal(w, act_dev, b, y_dev, variant, nullptr, (cudaStream_t) s);
}

int pib_gemv_pro(const pib_weight * w, int pro, const float * x, const float * x2, const float * addend, float * sum_out, int64_t xstride, float eps,
                 int b, float * y, const int * cancel, cudaStream_t st);
extern "C" int pib_gemv_fused(const pib_weight * w, int prologue, const float * x_dev, const float * x2_dev, const float * addend_dev, float * sum_out_dev,
                              int64_t xstride, float eps, int b, float * y_dev, pib_stream s) {
    if (!w || !x_dev || !y_dev || prologue < 0 || prologue > 2 || (prologue > 0 && !x2_dev) || (sum_out_dev && !addend_dev)) {
        pib_set_error("pib_gemv_fused: bad arguments");
        return PIB_ERR_ARG;
    }
    return pib_gemv_pro(w, prologue, x_dev, x2_dev, addend_dev, sum_out_dev, xstride, eps, b, y_dev, nullptr, (cudaStream_t) s);
}

extern "C" int pib_mul_mat_host(const pib_weight * w, const float * x_host, int b, float * y_host, int variant) {
    if (!w || b < 1 || b > PIB_MAX_BATCH) { pib_set_error("pib_mul_mat_host: bad arguments"); return PIB_ERR_ARG; }
    const size_t xb = (size_t) b * w->K * 4, yb = (size_t) b * w->N * 4, ab = pib_act_bytes(pib_act_family(w->type), w->K, b);
    uint8_t *    buf = nullptr;
    PIB_CHECK(cudaMalloc(&buf, xb + yb + ab + 512));
    float * x   = (float *) buf;
    float * y   = (float *) (buf + ((xb + 255) & ~(size_t) 255));
    void *  act = buf + ((xb + 255) & ~(size_t) 255) + ((yb + 255) & ~(size_t) 255);
    int     rc  = PIB_OK;
    if (cudaMemcpy(x, x_host, xb, cudaMemcpyHostToDevice) != cudaSuccess) rc = PIB_ERR_CUDA;
    if (rc == PIB_OK) rc = pib_quantize_act(pib_act_family(w->type), x, b, w->K, act, nullptr);
    if (rc == PIB_OK) rc = pib_gemv(w, act, b, y, variant, nullptr);
    if (rc == PIB_OK && cudaMemcpy(y_host, y, yb, cudaMemcpyDeviceToHost) != cudaSuccess) rc = PIB_ERR_CUDA;
    if (rc == PIB_ERR_CUDA) pib_set_error("pib_mul_mat_host: %s", cudaGetErrorString(cudaGetLastError()));
    cudaFree(buf);
    return rc;
}

// ---- K3: embedding rows ---------------------------------------------------------------------------------------------
// out[n][K] = dequantised row ids[n] (reference dequantize_row_q4_0 / q8_0 / q4_K: ggml-quants.c:1074, 1168, 1917;
// every element is one fp32 multiply (and one subtract for Q4_K), so results are bit-exact)
__global__ void get_rows_kernel(int type, const uint8_t * __restrict__ qs, const uint16_t * __restrict__ d, int64_t K,
                                const int32_t * __restrict__ ids, float * __restrict__ out) {
    pib_pdl_wait();
    pib_pdl_trigger();
    const int64_t row = ids[blockIdx.x];
    float *       o   = out + (size_t) blockIdx.x * K;
    if (type == PIB_TYPE_Q8_0) {
        const int8_t * q = (const int8_t *) qs + row * K;
        for (int64_t i = threadIdx.x; i < K; i += blockDim.x) o[i] = (float) q[i] * pib_h2f(d[row * (K / 32) + i / 32]);
    } else if (type == PIB_TYPE_Q4_0) {
        const uint8_t * q = qs + row * (K / 2);
        for (int64_t i = threadIdx.x; i < K; i += blockDim.x) {
            const int64_t blk = i / 32;
            const int     j   = (int) (i % 32);
            const uint8_t by  = q[blk * 16 + (j & 15)];
            const int     v   = (j < 16 ? (by & 0x0F) : (by >> 4)) - 8;
            o[i]              = (float) v * pib_h2f(d[row * (K / 32) + blk]);
        }
    } else {
        const uint8_t * q   = qs + row * (K / 2);
        const uint8_t * hdr = (const uint8_t *) d + row * (K / 256) * 16;
        for (int64_t i = threadIdx.x; i < K; i += blockDim.x) {
            const int64_t   sb = i / 256;
            const int       e  = (int) (i % 256);
            const int       c  = e / 64, r = e % 64;        // chunk of 64: low nibbles = first 32 elements
            const uint8_t * h  = hdr + sb * 16;
            const float     dd = pib_h2f((uint16_t) (h[0] | (h[1] << 8)));
            const float     mn = pib_h2f((uint16_t) (h[2] | (h[3] << 8)));
            const int       j  = 2 * c + (r >= 32);
            const uint8_t * sc = h + 4;
            int s, m;
            if (j < 4) { s = sc[j] & 63; m = sc[j + 4] & 63; }
            else       { s = (sc[j + 4] & 0xF) | ((sc[j - 4] >> 6) << 4); m = (sc[j + 4] >> 4) | ((sc[j] >> 6) << 4); }
            const uint8_t by = q[sb * 128 + c * 32 + (r & 31)];
            const int     v  = r < 32 ? (by & 0xF) : (by >> 4);
            const float   d1 = __fmul_rn(dd, (float) s), m1 = __fmul_rn(mn, (float) m);
            o[i]             = __fsub_rn(__fmul_rn(d1, (float) v), m1);
        }
    }
}

extern "C" int pib_get_rows(const pib_weight * w, const int32_t * ids_dev, int n, float * out_dev, pib_stream s) {
    if (!w || !ids_dev || !out_dev || n < 1) { pib_set_error("pib_get_rows: bad arguments"); return PIB_ERR_ARG; }
    if (!w->qs) { pib_set_error("pib_get_rows: this matrix keeps only its tensor-core tiles"); return PIB_ERR_UNSUPP; }
    if (w->type == PIB_TYPE_Q3_K) return pib_get_rows_q3k(w, ids_dev, n, out_dev, (cudaStream_t) s);
    if (w->type == PIB_TYPE_Q5_K) return pib_get_rows_q5k(w, ids_dev, n, out_dev, (cudaStream_t) s);
    if (w->type == PIB_TYPE_Q6_K) return pib_get_rows_q6k(w, ids_dev, n, out_dev, (cudaStream_t) s);
    pib_launch(get_rows_kernel, dim3(n), dim3(256), 0, (cudaStream_t) s, w->type, w->qs, w->d, w->K, ids_dev, out_dev);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

// ---- profiler range control (ncu --profile-from-start off) ------------------------------------------------------------
#include <cuda_profiler_api.h>
extern "C" void pib_profiler_start(void) { cudaProfilerStart(); }
extern "C" void pib_profiler_stop(void) { cudaDeviceSynchronize(); cudaProfilerStop(); }
