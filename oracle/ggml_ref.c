/*
 * oracle/ggml_ref.c -- TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * CPU restatement of the arithmetic the reference's hot path performs.  The reference
 * (graz-dev/llms-on-kubernetes) contains no compute code: its ramalama chart launches
 * `llama-server` from the image quay.io/ramalama/ramalama:latest
 * (ramalama-models/helm-chart/templates/model-deployments.yaml:26-35, image pin
 * ramalama-models/helm-chart/values.yaml:21-24 -- a floating tag).  The algorithm therefore
 * lives in a third-party dependency (llama.cpp / ggml, version unpinned) that is absent from
 * /root/reference and from this machine.  This file restates ggml's published CPU algorithm:
 *
 *   - block formats and dequantisation: checked bit-for-bit against gguf-py 0.19.0
 *     (gguf/quants.py:395-401 Q8_0, :475-522 Q4_K, :525-549 Q5_K, :552-572 Q6_K), which is
 *     llama.cpp's own Python statement of the same formats -> PINNED (tests/test_oracle_blocks.py,
 *     tests/golden/blocks_*.npz);
 *   - the fp32 graph (RMSNorm, RoPE, attention, SwiGLU) : checked against HF transformers'
 *     LlamaForCausalLM / Qwen2ForCausalLM fed by transformers.modeling_gguf_pytorch_utils
 *     (tests/golden/graph_*.npz) -> PINNED to a stated tolerance;
 *   - ggml's CPU mat-vec arithmetic (activations quantised to Q8_K / Q8_0, integer dot
 *     products, fp32 scaling: quantize_row_q8_K, ggml_vec_dot_q4_K_q8_K and friends) is
 *     restated from upstream knowledge and CANNOT be checked against upstream here
 *     -> PARITY UNPINNED for that part (tier T3).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (llms-on-kubernetes_b200/csrc) never links or calls it.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define QK_K 256
#define QK8_0 32

enum { T_F32 = 0, T_F16 = 1, T_Q8_0 = 8, T_Q4_K = 12, T_Q5_K = 13, T_Q6_K = 14, T_BF16 = 30 };

/* ---- fp16 <-> fp32 (IEEE binary16, round-to-nearest-even; ggml: GGML_FP16_TO_FP32 / GGML_FP32_TO_FP16) ---- */
static inline float h2f(uint16_t h) {
    _Float16 v;
    memcpy(&v, &h, 2);
    return (float)v;
}
static inline uint16_t f2h(float f) {
    _Float16 v = (_Float16)f;
    uint16_t h;
    memcpy(&h, &v, 2);
    return h;
}
static inline float bf2f(uint16_t h) {
    uint32_t u = (uint32_t)h << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

/* ---- block layouts (ggml-common.h; sizes match gguf/constants.py:4231-4267) ---- */
#pragma pack(push, 1)
typedef struct { uint16_t d; int8_t qs[32]; } block_q8_0;                                    /* 34  */
typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qs[128]; } block_q4_K;        /* 144 */
typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qh[32]; uint8_t qs[128]; } block_q5_K; /* 176 */
typedef struct { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; uint16_t d; } block_q6_K;        /* 210 */
typedef struct { float d; int8_t qs[256]; int16_t bsums[16]; } block_q8_K;                   /* 292 */
#pragma pack(pop)

int64_t ref_row_bytes(int type, int64_t k) {
    switch (type) {
        case T_F32: return k * 4;
        case T_F16: case T_BF16: return k * 2;
        case T_Q8_0: return k / 32 * 34;
        case T_Q4_K: return k / 256 * 144;
        case T_Q5_K: return k / 256 * 176;
        case T_Q6_K: return k / 256 * 210;
    }
    return -1;
}

/* 6-bit scale/min unpack shared by Q4_K and Q5_K (ggml get_scale_min_k4; gguf/quants.py:479-501) */
static inline void get_scale_min_k4(int j, const uint8_t* q, uint8_t* d, uint8_t* m) {
    if (j < 4) {
        *d = q[j] & 63;
        *m = q[j + 4] & 63;
    } else {
        *d = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        *m = (q[j + 4] >> 4) | ((q[j - 0] >> 6) << 4);
    }
}

/* ---- dequantisation: y = (d*sc)*q - (dmin*m), association as gguf/quants.py:516-522 ---- */
static void dequant_q8_0(const block_q8_0* x, float* y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const float d = h2f(x[i].d);
        for (int j = 0; j < 32; j++) y[i * 32 + j] = x[i].qs[j] * d;
    }
}
static void dequant_q4_K(const block_q4_K* x, float* y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = h2f(x[i].d), dmin = h2f(x[i].dmin);
        const uint8_t* q = x[i].qs;
        int is = 0;
        for (int j = 0; j < QK_K; j += 64) {
            uint8_t sc, m;
            get_scale_min_k4(is + 0, x[i].scales, &sc, &m);
            const float d1 = d * sc, m1 = dmin * m;
            get_scale_min_k4(is + 1, x[i].scales, &sc, &m);
            const float d2 = d * sc, m2 = dmin * m;
            for (int l = 0; l < 32; l++) *y++ = d1 * (q[l] & 0xF) - m1;
            for (int l = 0; l < 32; l++) *y++ = d2 * (q[l] >> 4) - m2;
            q += 32;
            is += 2;
        }
    }
}
static void dequant_q5_K(const block_q5_K* x, float* y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = h2f(x[i].d), dmin = h2f(x[i].dmin);
        const uint8_t* ql = x[i].qs;
        const uint8_t* qh = x[i].qh;
        int is = 0;
        uint8_t u1 = 1, u2 = 2;
        for (int j = 0; j < QK_K; j += 64) {
            uint8_t sc, m;
            get_scale_min_k4(is + 0, x[i].scales, &sc, &m);
            const float d1 = d * sc, m1 = dmin * m;
            get_scale_min_k4(is + 1, x[i].scales, &sc, &m);
            const float d2 = d * sc, m2 = dmin * m;
            for (int l = 0; l < 32; l++) *y++ = d1 * ((ql[l] & 0xF) + (qh[l] & u1 ? 16 : 0)) - m1;
            for (int l = 0; l < 32; l++) *y++ = d2 * ((ql[l] >> 4) + (qh[l] & u2 ? 16 : 0)) - m2;
            ql += 32;
            is += 2;
            u1 <<= 2;
            u2 <<= 2;
        }
    }
}
static void dequant_q6_K(const block_q6_K* x, float* y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = h2f(x[i].d);
        const uint8_t* ql = x[i].ql;
        const uint8_t* qh = x[i].qh;
        const int8_t* sc = x[i].scales;
        for (int n = 0; n < QK_K; n += 128) {
            for (int l = 0; l < 32; l++) {
                const int is = l / 16;
                const int8_t q1 = (int8_t)((ql[l + 0] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                const int8_t q2 = (int8_t)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                const int8_t q3 = (int8_t)((ql[l + 0] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                const int8_t q4 = (int8_t)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                y[l + 0] = (d * sc[is + 0]) * q1;
                y[l + 32] = (d * sc[is + 2]) * q2;
                y[l + 64] = (d * sc[is + 4]) * q3;
                y[l + 96] = (d * sc[is + 6]) * q4;
            }
            y += 128;
            ql += 64;
            qh += 32;
            sc += 8;
        }
    }
}

int ref_dequantize(int type, const void* src, float* dst, int64_t n) {
    switch (type) {
        case T_F32: memcpy(dst, src, (size_t)n * 4); return 0;
        case T_F16: for (int64_t i = 0; i < n; i++) dst[i] = h2f(((const uint16_t*)src)[i]); return 0;
        case T_BF16: for (int64_t i = 0; i < n; i++) dst[i] = bf2f(((const uint16_t*)src)[i]); return 0;
        case T_Q8_0: dequant_q8_0((const block_q8_0*)src, dst, n); return 0;
        case T_Q4_K: dequant_q4_K((const block_q4_K*)src, dst, n); return 0;
        case T_Q5_K: dequant_q5_K((const block_q5_K*)src, dst, n); return 0;
        case T_Q6_K: dequant_q6_K((const block_q6_K*)src, dst, n); return 0;
    }
    return -1;
}

/* ---- activation quantisation, ggml quantize_row_q8_0_ref / quantize_row_q8_K_ref [UPSTREAM-MEMORY] ---- */
void ref_quantize_q8_0(const float* x, void* vy, int64_t k) {
    block_q8_0* y = (block_q8_0*)vy;
    for (int64_t i = 0; i < k / 32; i++) {
        float amax = 0.0f;
        for (int j = 0; j < 32; j++) {
            const float v = fabsf(x[i * 32 + j]);
            if (v > amax) amax = v;
        }
        const float d = amax / 127.0f;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = f2h(d);
        for (int j = 0; j < 32; j++) y[i].qs[j] = (int8_t)roundf(x[i * 32 + j] * id);
    }
}
static inline int nearest_int(float fval) {
    /* ggml's nearest_int: add 1.5*2^23, read the mantissa -> round-half-to-even */
    float val = fval + 12582912.f;
    int i;
    memcpy(&i, &val, sizeof(int));
    return (i & 0x007fffff) - 0x00400000;
}
void ref_quantize_q8_K(const float* x, void* vy, int64_t k) {
    block_q8_K* y = (block_q8_K*)vy;
    for (int64_t i = 0; i < k / QK_K; i++) {
        float max = 0, amax = 0;
        for (int j = 0; j < QK_K; j++) {
            const float ax = fabsf(x[j]);
            if (ax > amax) { amax = ax; max = x[j]; }
        }
        if (!amax) {
            y[i].d = 0;
            memset(y[i].qs, 0, QK_K);
            memset(y[i].bsums, 0, sizeof(y[i].bsums));
            x += QK_K;
            continue;
        }
        const float iscale = -127.f / max;
        for (int j = 0; j < QK_K; j++) {
            int v = nearest_int(iscale * x[j]);
            y[i].qs[j] = (int8_t)(v < 127 ? v : 127);
        }
        for (int j = 0; j < QK_K / 16; j++) {
            int sum = 0;
            for (int ii = 0; ii < 16; ii++) sum += y[i].qs[j * 16 + ii];
            y[i].bsums[j] = (int16_t)sum;
        }
        y[i].d = 1 / iscale;
        x += QK_K;
    }
}

/* ---- dot products, ggml_vec_dot_*_generic [UPSTREAM-MEMORY] ---- */
static float vec_dot_q8_0_q8_0(int64_t n, const block_q8_0* x, const block_q8_0* y) {
    float sumf = 0;
    for (int64_t ib = 0; ib < n / 32; ib++) {
        int sumi = 0;
        for (int j = 0; j < 32; j++) sumi += x[ib].qs[j] * y[ib].qs[j];
        sumf += sumi * (h2f(x[ib].d) * h2f(y[ib].d));
    }
    return sumf;
}

static inline void unpack_scales_k4(const uint8_t* packed, uint8_t* scales, uint8_t* mins) {
    for (int j = 0; j < 8; j++) get_scale_min_k4(j, packed, &scales[j], &mins[j]);
}

static float vec_dot_q4_K_q8_K(int64_t n, const block_q4_K* x, const block_q8_K* y) {
    int8_t aux8[QK_K];
    float sums[8] = {0};
    float sumf = 0;
    for (int64_t i = 0; i < n / QK_K; i++) {
        const uint8_t* q4 = x[i].qs;
        const int8_t* q8 = y[i].qs;
        int32_t aux32[8] = {0};
        int8_t* a = aux8;
        for (int j = 0; j < QK_K / 64; j++) {
            for (int l = 0; l < 32; l++) a[l] = (int8_t)(q4[l] & 0xF);
            a += 32;
            for (int l = 0; l < 32; l++) a[l] = (int8_t)(q4[l] >> 4);
            a += 32;
            q4 += 32;
        }
        uint8_t scales[8], mins[8];
        unpack_scales_k4(x[i].scales, scales, mins);
        int sumi = 0;
        for (int j = 0; j < QK_K / 16; j++) sumi += y[i].bsums[j] * mins[j / 2];
        a = aux8;
        for (int j = 0; j < QK_K / 32; j++) {
            const int32_t scale = scales[j];
            for (int g = 0; g < 4; g++) {
                for (int l = 0; l < 8; l++) aux32[l] += scale * ((int16_t)q8[l] * a[l]);
                q8 += 8;
                a += 8;
            }
        }
        const float d = h2f(x[i].d) * y[i].d;
        for (int l = 0; l < 8; l++) sums[l] += d * aux32[l];
        const float dmin = h2f(x[i].dmin) * y[i].d;
        sumf -= dmin * sumi;
    }
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}

static float vec_dot_q5_K_q8_K(int64_t n, const block_q5_K* x, const block_q8_K* y) {
    int8_t aux8[QK_K];
    float sums[8] = {0};
    float sumf = 0;
    for (int64_t i = 0; i < n / QK_K; i++) {
        const uint8_t* q4 = x[i].qs;
        const uint8_t* hm = x[i].qh;
        const int8_t* q8 = y[i].qs;
        int32_t aux32[8] = {0};
        int8_t* a = aux8;
        uint8_t m = 1;
        for (int j = 0; j < QK_K / 64; j++) {
            for (int l = 0; l < 32; l++) a[l] = (int8_t)((q4[l] & 0xF) + (hm[l] & m ? 16 : 0));
            a += 32;
            m <<= 1;
            for (int l = 0; l < 32; l++) a[l] = (int8_t)((q4[l] >> 4) + (hm[l] & m ? 16 : 0));
            a += 32;
            m <<= 1;
            q4 += 32;
        }
        uint8_t scales[8], mins[8];
        unpack_scales_k4(x[i].scales, scales, mins);
        int sumi = 0;
        for (int j = 0; j < QK_K / 16; j++) sumi += y[i].bsums[j] * mins[j / 2];
        a = aux8;
        for (int j = 0; j < QK_K / 32; j++) {
            const int32_t scale = scales[j];
            for (int g = 0; g < 4; g++) {
                for (int l = 0; l < 8; l++) aux32[l] += scale * ((int16_t)q8[l] * a[l]);
                q8 += 8;
                a += 8;
            }
        }
        const float d = h2f(x[i].d) * y[i].d;
        for (int l = 0; l < 8; l++) sums[l] += d * aux32[l];
        const float dmin = h2f(x[i].dmin) * y[i].d;
        sumf -= dmin * sumi;
    }
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}

static float vec_dot_q6_K_q8_K(int64_t n, const block_q6_K* x, const block_q8_K* y) {
    int8_t aux8[QK_K];
    float sums[8] = {0};
    float sumf = 0;
    for (int64_t i = 0; i < n / QK_K; i++) {
        const uint8_t* q4 = x[i].ql;
        const uint8_t* qh = x[i].qh;
        const int8_t* q8 = y[i].qs;
        int32_t aux32[8] = {0};
        int8_t* a = aux8;
        for (int j = 0; j < QK_K; j += 128) {
            for (int l = 0; l < 32; l++) {
                a[l + 0] = (int8_t)((q4[l + 0] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                a[l + 32] = (int8_t)((q4[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                a[l + 64] = (int8_t)((q4[l + 0] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                a[l + 96] = (int8_t)((q4[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
            }
            a += 128;
            q4 += 64;
            qh += 32;
        }
        a = aux8;
        for (int j = 0; j < QK_K / 16; j++) {
            const int scale = x[i].scales[j];
            for (int g = 0; g < 2; g++) {
                for (int l = 0; l < 8; l++) aux32[l] += scale * ((int16_t)q8[l] * a[l]);
                q8 += 8;
                a += 8;
            }
        }
        const float d = h2f(x[i].d) * y[i].d;
        for (int l = 0; l < 8; l++) sums[l] += d * aux32[l];
    }
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}

/* F16/BF16/F32 weights: ggml converts the activation row to the weight's vec_dot_type (F16 for F16,
 * BF16 for BF16, F32 for F32) and accumulates in fp32. */
static float vec_dot_f16(int64_t n, const uint16_t* x, const uint16_t* y) {
    float s = 0;
    for (int64_t i = 0; i < n; i++) s += h2f(x[i]) * h2f(y[i]);
    return s;
}

/* activation-side format a weight type needs (ggml type_traits_cpu[].vec_dot_type) */
static int act_type(int wtype) {
    switch (wtype) {
        case T_Q8_0: return T_Q8_0;
        case T_Q4_K: case T_Q5_K: case T_Q6_K: return 15; /* Q8_K */
        case T_F16: return T_F16;
        default: return T_F32;
    }
}
static int64_t act_bytes(int wtype, int64_t k) {
    switch (act_type(wtype)) {
        case T_Q8_0: return k / 32 * 34;
        case 15: return k / 256 * 292;
        case T_F16: return k * 2;
        default: return k * 4;
    }
}
static void quantize_act(int wtype, const float* x, void* out, int64_t k) {
    switch (act_type(wtype)) {
        case T_Q8_0: ref_quantize_q8_0(x, out, k); break;
        case 15: ref_quantize_q8_K(x, out, k); break;
        case T_F16: for (int64_t i = 0; i < k; i++) ((uint16_t*)out)[i] = f2h(x[i]); break;
        default: memcpy(out, x, (size_t)k * 4);
    }
}
static float vec_dot_row(int wtype, int64_t k, const void* wrow, const void* act) {
    switch (wtype) {
        case T_Q8_0: return vec_dot_q8_0_q8_0(k, (const block_q8_0*)wrow, (const block_q8_0*)act);
        case T_Q4_K: return vec_dot_q4_K_q8_K(k, (const block_q4_K*)wrow, (const block_q8_K*)act);
        case T_Q5_K: return vec_dot_q5_K_q8_K(k, (const block_q5_K*)wrow, (const block_q8_K*)act);
        case T_Q6_K: return vec_dot_q6_K_q8_K(k, (const block_q6_K*)wrow, (const block_q8_K*)act);
        case T_F16: return vec_dot_f16(k, (const uint16_t*)wrow, (const uint16_t*)act);
        case T_BF16: {
            float s = 0;
            for (int64_t i = 0; i < k; i++) s += bf2f(((const uint16_t*)wrow)[i]) * ((const float*)act)[i];
            return s;
        }
        case T_F32: {
            float s = 0;
            for (int64_t i = 0; i < k; i++) s += ((const float*)wrow)[i] * ((const float*)act)[i];
            return s;
        }
    }
    return NAN;
}

/* y[N] = W[N x K] . x[K] the way ggml_compute_forward_mul_mat does it on the CPU:
 * quantise the activation row once, then one vec_dot per weight row. */
int ref_gemv(int wtype, const void* W, const float* x, float* y, int64_t N, int64_t K) {
    const int64_t rb = ref_row_bytes(wtype, K);
    if (rb < 0) return -1;
    void* act = malloc((size_t)act_bytes(wtype, K) + 64);
    quantize_act(wtype, x, act, K);
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < N; r++) y[r] = vec_dot_row(wtype, K, (const char*)W + r * rb, act);
    free(act);
    return 0;
}

/* ---- elementwise ops (ggml_compute_forward_rms_norm_f32, ggml_compute_forward_rope_f32,
 *      ggml_compute_forward_soft_max_f32) [UPSTREAM-MEMORY] ---- */
void ref_rmsnorm(const float* x, const float* w, float* y, int64_t n, float eps) {
    double sum = 0.0;
    for (int64_t i = 0; i < n; i++) sum += (double)(x[i] * x[i]);
    const float mean = (float)(sum / (double)n);
    const float scale = 1.0f / sqrtf(mean + eps);
    for (int64_t i = 0; i < n; i++) y[i] = (x[i] * scale) * w[i];
}

/* cos/sin cache for one position: theta_i = pos * base^(-2i/n_dims) built by repeated multiplication */
void ref_rope_cache(int32_t pos, int n_dims, float base, const float* freq_factors, float* cache /* [n_dims] cos,sin pairs */) {
    const float theta_scale = powf(base, -2.0f / n_dims);
    float theta = (float)pos;
    for (int i0 = 0; i0 < n_dims; i0 += 2) {
        const float ff = freq_factors ? freq_factors[i0 / 2] : 1.0f;
        const float t = theta / ff;
        cache[i0 + 0] = cosf(t);
        cache[i0 + 1] = sinf(t);
        theta *= theta_scale;
    }
}
/* mode 0 = NORM (adjacent pairs, `llama`), 2 = NEOX (i, i + n_dims/2, `qwen2`) */
void ref_rope(float* x, int n_heads, int head_dim, int n_dims, int mode, const float* cache) {
    for (int h = 0; h < n_heads; h++) {
        float* v = x + (int64_t)h * head_dim;
        if (mode == 0) {
            for (int i0 = 0; i0 < n_dims; i0 += 2) {
                const float c = cache[i0], s = cache[i0 + 1];
                const float x0 = v[i0], x1 = v[i0 + 1];
                v[i0] = x0 * c - x1 * s;
                v[i0 + 1] = x0 * s + x1 * c;
            }
        } else {
            for (int i0 = 0; i0 < n_dims; i0 += 2) {
                const float c = cache[i0], s = cache[i0 + 1];
                const int ic = i0 / 2;
                const float x0 = v[ic], x1 = v[ic + n_dims / 2];
                v[ic] = x0 * c - x1 * s;
                v[ic + n_dims / 2] = x0 * s + x1 * c;
            }
        }
    }
}

static inline float silu_f(float x) { return x / (1.0f + expf(-x)); }

/* ---- model ---- */
typedef struct { const void* data; int32_t type; int32_t _pad; int64_t rows, cols; } ref_tensor;
typedef struct {
    ref_tensor attn_norm, wq, wk, wv, wo, ffn_norm, w_gate, w_up, w_down;
    const float *bq, *bk, *bv;
} ref_layer;
typedef struct {
    int32_t arch; /* 0 llama, 1 qwen2 */
    int32_t n_layer, d_model, d_ff, n_head, n_kv, head_dim, vocab, n_ctx, rope_dim;
    float eps, rope_base;
    ref_tensor tok_embd, output_norm, output;
    const float* rope_freqs;
    const ref_layer* layers;
} ref_model;

typedef struct {
    const ref_model* m;
    uint16_t* k_cache; /* [layer][pos][n_kv*head_dim] fp16 */
    uint16_t* v_cache;
    float *x, *xn, *q, *k, *v, *att, *attn_out, *proj, *gate, *up, *h, *rope;
    float* scores;
} ref_ctx;

ref_ctx* ref_ctx_new(const ref_model* m) {
    ref_ctx* c = (ref_ctx*)calloc(1, sizeof(ref_ctx));
    c->m = m;
    const int64_t kvd = (int64_t)m->n_kv * m->head_dim;
    c->k_cache = (uint16_t*)calloc((size_t)m->n_layer * m->n_ctx * kvd, 2);
    c->v_cache = (uint16_t*)calloc((size_t)m->n_layer * m->n_ctx * kvd, 2);
    const int64_t qd = (int64_t)m->n_head * m->head_dim;
    c->x = (float*)malloc(sizeof(float) * m->d_model);
    c->xn = (float*)malloc(sizeof(float) * m->d_model);
    c->q = (float*)malloc(sizeof(float) * qd);
    c->k = (float*)malloc(sizeof(float) * kvd);
    c->v = (float*)malloc(sizeof(float) * kvd);
    c->attn_out = (float*)malloc(sizeof(float) * qd);
    c->proj = (float*)malloc(sizeof(float) * m->d_model);
    c->gate = (float*)malloc(sizeof(float) * m->d_ff);
    c->up = (float*)malloc(sizeof(float) * m->d_ff);
    c->h = (float*)malloc(sizeof(float) * m->d_ff);
    c->rope = (float*)malloc(sizeof(float) * m->head_dim);
    c->scores = (float*)malloc(sizeof(float) * (size_t)m->n_ctx * m->n_head);
    return c;
}
void ref_ctx_free(ref_ctx* c) {
    if (!c) return;
    free(c->k_cache); free(c->v_cache); free(c->x); free(c->xn); free(c->q); free(c->k); free(c->v);
    free(c->attn_out); free(c->proj); free(c->gate); free(c->up); free(c->h); free(c->rope); free(c->scores);
    free(c);
}

/* attention for one new token at position pos over the cache [0..pos]; classic (non-flash) ggml path:
 * q and the soft-max probabilities are rounded to fp16 because K and V are stored as F16 and
 * mul_mat converts src1 to the vec_dot_type of src0. */
static void ref_attention(ref_ctx* c, int il, int pos) {
    const ref_model* m = c->m;
    const int hd = m->head_dim, H = m->n_head, KV = m->n_kv, g = H / KV;
    const int64_t kvd = (int64_t)KV * hd;
    const uint16_t* Kc = c->k_cache + (int64_t)il * m->n_ctx * kvd;
    const uint16_t* Vc = c->v_cache + (int64_t)il * m->n_ctx * kvd;
    const float scale = 1.0f / sqrtf((float)hd);
    const int T = pos + 1;
#pragma omp parallel for schedule(static)
    for (int h = 0; h < H; h++) {
        const int kvh = h / g;
        float* s = c->scores + (int64_t)h * m->n_ctx;
        uint16_t qh[512];
        for (int d = 0; d < hd; d++) qh[d] = f2h(c->q[(int64_t)h * hd + d]);
        float mx = -INFINITY;
        for (int t = 0; t < T; t++) {
            const uint16_t* kr = Kc + (int64_t)t * kvd + (int64_t)kvh * hd;
            float acc = 0;
            for (int d = 0; d < hd; d++) acc += h2f(kr[d]) * h2f(qh[d]);
            s[t] = acc * scale;
            if (s[t] > mx) mx = s[t];
        }
        double sum = 0;
        for (int t = 0; t < T; t++) {
            const float e = expf(s[t] - mx);
            s[t] = e;
            sum += (double)e;
        }
        const float inv = (float)(1.0 / sum);
        for (int t = 0; t < T; t++) s[t] *= inv;
        float* o = c->attn_out + (int64_t)h * hd;
        for (int d = 0; d < hd; d++) o[d] = 0;
        for (int t = 0; t < T; t++) {
            const float p = h2f(f2h(s[t]));
            const uint16_t* vr = Vc + (int64_t)t * kvd + (int64_t)kvh * hd;
            for (int d = 0; d < hd; d++) o[d] += h2f(vr[d]) * p;
        }
    }
}

/* One token through the decoder (llama.cpp build_llama / build_qwen2 graph for n_tokens = 1).
 * logits_out may be NULL.  Returns the greedy token (argmax of logits, lowest index on ties). */
/* Sensitivity probe (tests only): with eps > 0 every residual-stream update and every SwiGLU output is multiplied by
 * (1 + eps * r), r uniform in [-0.5, 0.5) from a hash -- the size of change a different but equally valid fp32
 * summation order makes.  Comparing a perturbed run with the plain one measures how far two CORRECT implementations of
 * this model can drift apart (int8 re-quantisation of activations turns a last-bit change into a rounding flip that
 * the following layers amplify); the parity tests state their tolerance relative to that measured floor. */
static float g_perturb_eps = 0.0f;
static uint32_t g_perturb_seed = 0;
void ref_set_perturb(float eps, uint32_t seed) { g_perturb_eps = eps; g_perturb_seed = seed; }
static void perturb(float* v, int n, uint32_t salt) {
    if (g_perturb_eps <= 0.0f) return;
    for (int i = 0; i < n; i++) {
        uint32_t h = ((uint32_t)i * 2654435761u) ^ (salt * 40503u) ^ g_perturb_seed;
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        v[i] *= 1.0f + g_perturb_eps * ((float)(h & 0xffff) / 65536.0f - 0.5f);
    }
}

int32_t ref_decode(ref_ctx* c, int32_t token, int32_t pos, float* logits_out) {
    const ref_model* m = c->m;
    const int d = m->d_model, hd = m->head_dim;
    const int64_t qd = (int64_t)m->n_head * hd, kvd = (int64_t)m->n_kv * hd;
    if (token < 0 || token >= m->vocab || pos < 0 || pos >= m->n_ctx) return -1;
    /* get_rows on token_embd */
    ref_dequantize(m->tok_embd.type, (const char*)m->tok_embd.data + (int64_t)token * ref_row_bytes(m->tok_embd.type, d), c->x, d);
    ref_rope_cache(pos, m->rope_dim, m->rope_base, m->rope_freqs, c->rope);
    for (int il = 0; il < m->n_layer; il++) {
        const ref_layer* L = &m->layers[il];
        ref_rmsnorm(c->x, (const float*)L->attn_norm.data, c->xn, d, m->eps);
        ref_gemv(L->wq.type, L->wq.data, c->xn, c->q, qd, d);
        ref_gemv(L->wk.type, L->wk.data, c->xn, c->k, kvd, d);
        ref_gemv(L->wv.type, L->wv.data, c->xn, c->v, kvd, d);
        if (L->bq) for (int64_t i = 0; i < qd; i++) c->q[i] += L->bq[i];
        if (L->bk) for (int64_t i = 0; i < kvd; i++) c->k[i] += L->bk[i];
        if (L->bv) for (int64_t i = 0; i < kvd; i++) c->v[i] += L->bv[i];
        const int mode = m->arch == 1 ? 2 : 0;
        ref_rope(c->q, m->n_head, hd, m->rope_dim, mode, c->rope);
        ref_rope(c->k, m->n_kv, hd, m->rope_dim, mode, c->rope);
        uint16_t* kdst = c->k_cache + ((int64_t)il * m->n_ctx + pos) * kvd;
        uint16_t* vdst = c->v_cache + ((int64_t)il * m->n_ctx + pos) * kvd;
        for (int64_t i = 0; i < kvd; i++) { kdst[i] = f2h(c->k[i]); vdst[i] = f2h(c->v[i]); }
        ref_attention(c, il, pos);
        ref_gemv(L->wo.type, L->wo.data, c->attn_out, c->proj, d, qd);
        for (int i = 0; i < d; i++) c->x[i] += c->proj[i];
        perturb(c->x, d, 4u * (uint32_t)il + 1u);
        ref_rmsnorm(c->x, (const float*)L->ffn_norm.data, c->xn, d, m->eps);
        ref_gemv(L->w_gate.type, L->w_gate.data, c->xn, c->gate, m->d_ff, d);
        ref_gemv(L->w_up.type, L->w_up.data, c->xn, c->up, m->d_ff, d);
        for (int i = 0; i < m->d_ff; i++) c->h[i] = silu_f(c->gate[i]) * c->up[i];
        perturb(c->h, m->d_ff, 4u * (uint32_t)il + 2u);
        ref_gemv(L->w_down.type, L->w_down.data, c->h, c->proj, d, m->d_ff);
        for (int i = 0; i < d; i++) c->x[i] += c->proj[i];
    }
    ref_rmsnorm(c->x, (const float*)m->output_norm.data, c->xn, d, m->eps);
    float* logits = logits_out ? logits_out : (float*)malloc(sizeof(float) * m->vocab);
    ref_gemv(m->output.type, m->output.data, c->xn, logits, m->vocab, d);
    int32_t best = 0;
    for (int32_t i = 1; i < m->vocab; i++) if (logits[i] > logits[best]) best = i;
    if (!logits_out) free(logits);
    return best;
}

/* expose the hidden state after the last layer (before output_norm) and the last layer's intermediates */
const float* ref_ctx_hidden(const ref_ctx* c) { return c->x; }
const float* ref_ctx_buffer(const ref_ctx* c, int which) {
    switch (which) {
        case 0: return c->x;
        case 1: return c->q;
        case 2: return c->attn_out;
        case 3: return c->h;
    }
    return 0;
}
/* fp16 K / V rows [n_kv*head_dim] of a layer at a position */
const uint16_t* ref_ctx_k(const ref_ctx* c, int layer, int pos) {
    return c->k_cache + ((int64_t)layer * c->m->n_ctx + pos) * c->m->n_kv * c->m->head_dim;
}
const uint16_t* ref_ctx_v(const ref_ctx* c, int layer, int pos) {
    return c->v_cache + ((int64_t)layer * c->m->n_ctx + pos) * c->m->n_kv * c->m->head_dim;
}

/* dot of one weight row with an fp32 activation row (quantises the activation first) */
float ref_vec_dot(int wtype, const void* wrow, const float* x, int64_t K) {
    void* act = malloc((size_t)act_bytes(wtype, K) + 64);
    quantize_act(wtype, x, act, K);
    const float r = vec_dot_row(wtype, K, wrow, act);
    free(act);
    return r;
}

int ref_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* bench.py's CPU arm under torch.distributed.run inherits OMP_NUM_THREADS=1: it sets the team size itself */
void ref_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
