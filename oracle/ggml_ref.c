/*
 * oracle/ggml_ref.c -- CPU restatement of the reference's numeric path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file.  The product (localai_b200/) never links, imports or calls it.
 *
 * PARITY STATUS: the arithmetic of the reference's llama-cpp backend lives in llama.cpp @ d5cb868
 * (/root/reference/Makefile:11), which is fetched at build time and is NOT under /root/reference
 * (backend/cpp/llama/Makefile:48-55).  This file restates that library's *published scalar
 * reference algorithms* (ggml-quants.c `*_ref` / generic `ggml_vec_dot_*` paths, ggml.c
 * compute_forward_{rms_norm,rope,soft_max,mul_mat}, llama.cpp `llm_build_llama`) as called from the
 * reference's own call sites:
 *   llama_decode            grpc-server.cpp:1910        -> ref_eval()
 *   gpt_sampler_sample      grpc-server.cpp:1946 (temp<=0 greedy) -> ref_argmax()
 *   llama_get_embeddings    grpc-server.cpp:1313        -> ref_eval(..., embd_out)
 * PINNED: dequantisation of 13 formats and Q8_0 quantisation bit-exactly against gguf-py 0.19.0 (llama.cpp's own
 * Python package; tests/golden/dequant_*.npz); the forward pass (RoPE pairing, GQA, norm placement, SwiGLU, head)
 * against HuggingFace transformers' LlamaForCausalLM loaded from the same GGUF (tools/crosscheck_transformers.py,
 * tests/golden/transformers_crosscheck.json, tests/test_oracle.py::test_oracle_matches_transformers) -- the
 * reference's own tests hold no logits or token-id vectors for this path.
 * The scaled RoPE variants (linear, YaRN, the rope_freqs.weight divisors of Llama-3.1; ref_rope_cache_ext below), the
 * qwen2 graph (QKV biases, NEOX pairs) and the phi3 graph (fused tensors) are pinned the same way: against transformers'
 * "linear" / "yarn" / "llama3" rotary scaling, Qwen2ForCausalLM and Phi3ForCausalLM on the same GGUFs (same script and
 * golden file; 2-3e-3 sigma on F16 files, where leaving the scaling out costs more than one sigma).
 *
 * Build: see oracle/Makefile (gcc -O3 -fopenmp -ffp-contract=off).  Contraction is disabled so
 * that `d*q - m` is two roundings, as in gguf-py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define QK8_0 32
#define QK_K 256
#define K_SCALE_SIZE 12

enum { T_F32 = 0, T_F16 = 1, T_Q4_0 = 2, T_Q4_1 = 3, T_Q5_0 = 6, T_Q5_1 = 7, T_Q8_0 = 8, T_Q8_1 = 9, T_Q2_K = 10, T_Q3_K = 11,
       T_Q4_K = 12, T_Q5_K = 13,
       T_Q6_K = 14, T_Q8_K = 15, T_IQ4_NL = 20, T_IQ4_XS = 23 };

typedef uint16_t fp16_t;

static inline float fp16_to_fp32(fp16_t h) {
    _Float16 f;
    memcpy(&f, &h, 2);
    return (float)f;
}
static inline fp16_t fp32_to_fp16(float x) {
    _Float16 f = (_Float16)x;
    fp16_t h;
    memcpy(&h, &f, 2);
    return h;
}

#pragma pack(push, 1)
typedef struct { fp16_t d; int8_t qs[QK8_0]; } block_q8_0;                                   /* 34 B  */
typedef struct { fp16_t d, s; int8_t qs[QK8_0]; } block_q8_1;                                /* 36 B: s = d * sum(qs) */
typedef struct { fp16_t d; uint8_t qs[16]; } block_q4_0;                                     /* 18 B  */
typedef struct { fp16_t d, m; uint8_t qs[16]; } block_q4_1;                                  /* 20 B  */
typedef struct { fp16_t d; uint8_t qh[4]; uint8_t qs[16]; } block_q5_0;                      /* 22 B  */
typedef struct { fp16_t d, m; uint8_t qh[4]; uint8_t qs[16]; } block_q5_1;                   /* 24 B  */
typedef struct { uint8_t scales[QK_K / 16]; uint8_t qs[QK_K / 4]; fp16_t d, dmin; } block_q2_K;     /* 84 B  */
typedef struct { uint8_t hmask[QK_K / 8]; uint8_t qs[QK_K / 4]; uint8_t scales[12]; fp16_t d; } block_q3_K; /* 110 B */
typedef struct { fp16_t d, dmin; uint8_t scales[K_SCALE_SIZE]; uint8_t qs[QK_K / 2]; } block_q4_K; /* 144 B */
typedef struct { fp16_t d, dmin; uint8_t scales[K_SCALE_SIZE]; uint8_t qh[QK_K / 8]; uint8_t qs[QK_K / 2]; } block_q5_K; /* 176 B */
typedef struct { uint8_t ql[QK_K / 2]; uint8_t qh[QK_K / 4]; int8_t scales[QK_K / 16]; fp16_t d; } block_q6_K; /* 210 B */
typedef struct { float d; int8_t qs[QK_K]; int16_t bsums[QK_K / 16]; } block_q8_K;             /* 292 B */
typedef struct { fp16_t d; uint8_t qs[16]; } block_iq4_nl;                                   /* 18 B: 4-bit indices into a fixed non-linear table */
typedef struct { fp16_t d; uint16_t scales_h; uint8_t scales_l[QK_K / 64]; uint8_t qs[QK_K / 2]; } block_iq4_xs; /* 136 B */
#pragma pack(pop)

static int64_t type_block(int t) {
    if (t == T_F32 || t == T_F16) return 1;
    return (t == T_Q8_0 || t == T_Q8_1 || t == T_Q4_0 || t == T_Q4_1 || t == T_Q5_0 || t == T_Q5_1 || t == T_IQ4_NL) ? 32 : 256;
}
static int64_t type_bytes(int t) {
    switch (t) {
        case T_Q2_K: return 84; case T_Q3_K: return 110; case T_IQ4_NL: return 18; case T_IQ4_XS: return 136;
        case T_Q4_0: return 18; case T_Q4_1: return 20; case T_Q5_0: return 22; case T_Q5_1: return 24; case T_Q8_1: return 36;
        case T_F32: return 4; case T_F16: return 2; case T_Q8_0: return 34; case T_Q4_K: return 144;
        case T_Q5_K: return 176; case T_Q6_K: return 210; case T_Q8_K: return 292;
    }
    return 0;
}
int64_t ref_row_bytes(int type, int64_t cols) { return cols / type_block(type) * type_bytes(type); }

/* ------------------------------------------------------------------ dequantisation (ggml-quants.c dequantize_row_*) */
static inline void get_scale_min_k4(int j, const uint8_t *q, uint8_t *d, uint8_t *m) {
    if (j < 4) {
        *d = q[j] & 63;
        *m = q[j + 4] & 63;
    } else {
        *d = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        *m = (q[j + 4] >> 4) | ((q[j - 0] >> 6) << 4);
    }
}

static void dequantize_row_q8_0(const block_q8_0 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK8_0; i++) {
        const float d = fp16_to_fp32(x[i].d);
        for (int j = 0; j < QK8_0; ++j) y[i * QK8_0 + j] = x[i].qs[j] * d;
    }
}

/* legacy 32-element blocks (ggml-quants.c dequantize_row_q4_0 / q4_1 / q5_0 / q5_1) */
static void dequantize_row_q4_0(const block_q4_0 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const float d = fp16_to_fp32(x[i].d);
        for (int j = 0; j < 16; ++j) {
            y[i * 32 + j] = ((x[i].qs[j] & 0x0F) - 8) * d;
            y[i * 32 + j + 16] = ((x[i].qs[j] >> 4) - 8) * d;
        }
    }
}
static void dequantize_row_q4_1(const block_q4_1 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const float d = fp16_to_fp32(x[i].d), m = fp16_to_fp32(x[i].m);
        for (int j = 0; j < 16; ++j) {
            y[i * 32 + j] = (x[i].qs[j] & 0x0F) * d + m;
            y[i * 32 + j + 16] = (x[i].qs[j] >> 4) * d + m;
        }
    }
}
static void dequantize_row_q5_0(const block_q5_0 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const float d = fp16_to_fp32(x[i].d);
        uint32_t qh;
        memcpy(&qh, x[i].qh, sizeof(qh));
        for (int j = 0; j < 16; ++j) {
            const uint8_t xh_0 = ((qh >> (j + 0)) << 4) & 0x10;
            const uint8_t xh_1 = ((qh >> (j + 12))) & 0x10;
            y[i * 32 + j] = (((x[i].qs[j] & 0x0F) | xh_0) - 16) * d;
            y[i * 32 + j + 16] = (((x[i].qs[j] >> 4) | xh_1) - 16) * d;
        }
    }
}
static void dequantize_row_q5_1(const block_q5_1 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const float d = fp16_to_fp32(x[i].d), m = fp16_to_fp32(x[i].m);
        uint32_t qh;
        memcpy(&qh, x[i].qh, sizeof(qh));
        for (int j = 0; j < 16; ++j) {
            const uint8_t xh_0 = ((qh >> (j + 0)) << 4) & 0x10;
            const uint8_t xh_1 = ((qh >> (j + 12))) & 0x10;
            y[i * 32 + j] = ((x[i].qs[j] & 0x0F) | xh_0) * d + m;
            y[i * 32 + j + 16] = ((x[i].qs[j] >> 4) | xh_1) * d + m;
        }
    }
}

/* ggml-quants.c dequantize_row_q2_K / dequantize_row_q3_K */
static void dequantize_row_q2_K(const block_q2_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = fp16_to_fp32(x[i].d), min = fp16_to_fp32(x[i].dmin);
        const uint8_t *q = x[i].qs;
        int is = 0;
        for (int n = 0; n < QK_K; n += 128) {
            int shift = 0;
            for (int j = 0; j < 4; ++j) {
                uint8_t sc = x[i].scales[is++];
                float dl = d * (sc & 0xF), ml = min * (sc >> 4);
                for (int l = 0; l < 16; ++l) *y++ = dl * ((int8_t)((q[l] >> shift) & 3)) - ml;
                sc = x[i].scales[is++];
                dl = d * (sc & 0xF);
                ml = min * (sc >> 4);
                for (int l = 0; l < 16; ++l) *y++ = dl * ((int8_t)((q[l + 16] >> shift) & 3)) - ml;
                shift += 2;
            }
            q += 32;
        }
    }
}
static void q3_K_scales(const uint8_t *packed, int8_t scales[16]) {
    const uint32_t km1 = 0x03030303, km2 = 0x0f0f0f0f;
    uint32_t aux[4];
    memcpy(aux, packed, 12);
    const uint32_t tmp = aux[2];
    aux[2] = ((aux[0] >> 4) & km2) | (((tmp >> 4) & km1) << 4);
    aux[3] = ((aux[1] >> 4) & km2) | (((tmp >> 6) & km1) << 4);
    aux[0] = (aux[0] & km2) | (((tmp >> 0) & km1) << 4);
    aux[1] = (aux[1] & km2) | (((tmp >> 2) & km1) << 4);
    memcpy(scales, aux, 16);
}
static void dequantize_row_q3_K(const block_q3_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d_all = fp16_to_fp32(x[i].d);
        const uint8_t *q = x[i].qs, *hm = x[i].hmask;
        uint8_t m = 1;
        int8_t scales[16];
        q3_K_scales(x[i].scales, scales);
        int is = 0;
        for (int n = 0; n < QK_K; n += 128) {
            int shift = 0;
            for (int j = 0; j < 4; ++j) {
                float dl = d_all * (scales[is++] - 32);
                for (int l = 0; l < 16; ++l) *y++ = dl * ((int8_t)((q[l + 0] >> shift) & 3) - ((hm[l + 0] & m) ? 0 : 4));
                dl = d_all * (scales[is++] - 32);
                for (int l = 0; l < 16; ++l) *y++ = dl * ((int8_t)((q[l + 16] >> shift) & 3) - ((hm[l + 16] & m) ? 0 : 4));
                shift += 2;
                m <<= 1;
            }
            q += 32;
        }
    }
}

static void dequantize_row_q4_K(const block_q4_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const uint8_t *q = x[i].qs;
        const float d = fp16_to_fp32(x[i].d), min = fp16_to_fp32(x[i].dmin);
        int is = 0;
        uint8_t sc, m;
        for (int j = 0; j < QK_K; j += 64) {
            get_scale_min_k4(is + 0, x[i].scales, &sc, &m);
            const float d1 = d * sc, m1 = min * m;
            get_scale_min_k4(is + 1, x[i].scales, &sc, &m);
            const float d2 = d * sc, m2 = min * m;
            for (int l = 0; l < 32; ++l) *y++ = d1 * (q[l] & 0xF) - m1;
            for (int l = 0; l < 32; ++l) *y++ = d2 * (q[l] >> 4) - m2;
            q += 32;
            is += 2;
        }
    }
}

static void dequantize_row_q5_K(const block_q5_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const uint8_t *ql = x[i].qs, *qh = x[i].qh;
        const float d = fp16_to_fp32(x[i].d), min = fp16_to_fp32(x[i].dmin);
        int is = 0;
        uint8_t sc, m, u1 = 1, u2 = 2;
        for (int j = 0; j < QK_K; j += 64) {
            get_scale_min_k4(is + 0, x[i].scales, &sc, &m);
            const float d1 = d * sc, m1 = min * m;
            get_scale_min_k4(is + 1, x[i].scales, &sc, &m);
            const float d2 = d * sc, m2 = min * m;
            for (int l = 0; l < 32; ++l) *y++ = d1 * ((ql[l] & 0xF) + (qh[l] & u1 ? 16 : 0)) - m1;
            for (int l = 0; l < 32; ++l) *y++ = d2 * ((ql[l] >> 4) + (qh[l] & u2 ? 16 : 0)) - m2;
            ql += 32;
            is += 2;
            u1 <<= 2;
            u2 <<= 2;
        }
    }
}

static void dequantize_row_q6_K(const block_q6_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = fp16_to_fp32(x[i].d);
        const uint8_t *ql = x[i].ql, *qh = x[i].qh;
        const int8_t *sc = x[i].scales;
        for (int n = 0; n < QK_K; n += 128) {
            for (int l = 0; l < 32; ++l) {
                int is = l / 16;
                const int8_t q1 = (int8_t)((ql[l + 0] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                const int8_t q2 = (int8_t)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                const int8_t q3 = (int8_t)((ql[l + 0] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                const int8_t q4 = (int8_t)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                /* gguf-py groups (d*scale) first, then *q: same here */
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

/* ggml-quants.c kvalues_iq4nl, dequantize_row_iq4_nl / dequantize_row_iq4_xs */
static const int8_t kvalues_iq4nl[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};

static void dequantize_row_iq4_nl(const block_iq4_nl *x, float *y, int64_t k) {
    const int64_t nb = k / 32;
    for (int64_t i = 0; i < nb; i++) {
        const float d = fp16_to_fp32(x[i].d);
        for (int j = 0; j < 16; ++j) {
            y[j] = d * kvalues_iq4nl[x[i].qs[j] & 0xf];
            y[j + 16] = d * kvalues_iq4nl[x[i].qs[j] >> 4];
        }
        y += 32;
    }
}

static void dequantize_row_iq4_xs(const block_iq4_xs *x, float *y, int64_t k) {
    const int64_t nb = k / QK_K;
    for (int64_t i = 0; i < nb; i++) {
        const uint8_t *qs = x[i].qs;
        const float d = fp16_to_fp32(x[i].d);
        for (int ib = 0; ib < QK_K / 32; ++ib) {
            const int ls = ((x[i].scales_l[ib / 2] >> 4 * (ib % 2)) & 0xf) | (((x[i].scales_h >> 2 * ib) & 3) << 4);
            const float dl = d * (ls - 32);
            for (int j = 0; j < 16; ++j) {
                y[j] = dl * kvalues_iq4nl[qs[j] & 0xf];
                y[j + 16] = dl * kvalues_iq4nl[qs[j] >> 4];
            }
            y += 32;
            qs += 16;
        }
    }
}

/* one row of `k` elements of type `type` -> f32 */
void ref_dequantize_row(int type, const void *x, float *y, int64_t k) {
    switch (type) {
        case T_F32: memcpy(y, x, k * sizeof(float)); break;
        case T_F16: for (int64_t i = 0; i < k; i++) y[i] = fp16_to_fp32(((const fp16_t *)x)[i]); break;
        case T_Q8_0: dequantize_row_q8_0((const block_q8_0 *)x, y, k); break;
        case T_Q4_0: dequantize_row_q4_0((const block_q4_0 *)x, y, k); break;
        case T_Q4_1: dequantize_row_q4_1((const block_q4_1 *)x, y, k); break;
        case T_Q5_0: dequantize_row_q5_0((const block_q5_0 *)x, y, k); break;
        case T_Q5_1: dequantize_row_q5_1((const block_q5_1 *)x, y, k); break;
        case T_Q2_K: dequantize_row_q2_K((const block_q2_K *)x, y, k); break;
        case T_Q3_K: dequantize_row_q3_K((const block_q3_K *)x, y, k); break;
        case T_Q4_K: dequantize_row_q4_K((const block_q4_K *)x, y, k); break;
        case T_Q5_K: dequantize_row_q5_K((const block_q5_K *)x, y, k); break;
        case T_Q6_K: dequantize_row_q6_K((const block_q6_K *)x, y, k); break;
        case T_IQ4_NL: dequantize_row_iq4_nl((const block_iq4_nl *)x, y, k); break;
        case T_IQ4_XS: dequantize_row_iq4_xs((const block_iq4_xs *)x, y, k); break;
        default: break;
    }
}

/* ------------------------------------------------------------------ activation quantisation (quantize_row_q8_0_ref / q8_K_ref) */
static inline int nearest_int(float fval) {
    float val = fval + 12582912.f;
    int i;
    memcpy(&i, &val, sizeof(int));
    return (i & 0x007fffff) - 0x00400000;
}

void ref_quantize_row_q8_0(const float *x, void *vy, int64_t k) {
    block_q8_0 *y = (block_q8_0 *)vy;
    for (int64_t i = 0; i < k / QK8_0; i++) {
        float amax = 0.0f;
        for (int j = 0; j < QK8_0; j++) {
            const float v = fabsf(x[i * QK8_0 + j]);
            if (v > amax) amax = v;
        }
        const float d = amax / ((1 << 7) - 1);
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = fp32_to_fp16(d);
        for (int j = 0; j < QK8_0; ++j) y[i].qs[j] = (int8_t)roundf(x[i * QK8_0 + j] * id);
    }
}

/* quantize_row_q8_1_ref: like q8_0 plus s = d * sum(qs), both stored as fp16 */
void ref_quantize_row_q8_1(const float *x, void *vy, int64_t k) {
    block_q8_1 *y = (block_q8_1 *)vy;
    for (int64_t i = 0; i < k / 32; i++) {
        float amax = 0.0f;
        for (int j = 0; j < 32; j++) {
            const float v = fabsf(x[i * 32 + j]);
            if (v > amax) amax = v;
        }
        const float d = amax / ((1 << 7) - 1);
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = fp32_to_fp16(d);
        int sum = 0;
        for (int j = 0; j < 16; ++j) {
            const float v0 = x[i * 32 + j] * id, v1 = x[i * 32 + 16 + j] * id;
            y[i].qs[j] = (int8_t)roundf(v0);
            y[i].qs[16 + j] = (int8_t)roundf(v1);
            sum += y[i].qs[j] + y[i].qs[16 + j];
        }
        y[i].s = fp32_to_fp16(sum * d);
    }
}

void ref_quantize_row_q8_K(const float *x, void *vy, int64_t k) {
    block_q8_K *y = (block_q8_K *)vy;
    for (int64_t i = 0; i < k / QK_K; i++) {
        float max = 0, amax = 0;
        for (int j = 0; j < QK_K; ++j) {
            float ax = fabsf(x[j]);
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
        for (int j = 0; j < QK_K; ++j) {
            int v = nearest_int(iscale * x[j]);
            y[i].qs[j] = (int8_t)(v < 127 ? v : 127);
        }
        for (int j = 0; j < QK_K / 16; ++j) {
            int sum = 0;
            for (int ii = 0; ii < 16; ++ii) sum += y[i].qs[j * 16 + ii];
            y[i].bsums[j] = (int16_t)sum;
        }
        y[i].d = 1 / iscale;
        x += QK_K;
    }
}

/* ------------------------------------------------------------------ block dot products (generic ggml_vec_dot_* paths) */
static float vec_dot_q8_0_q8_0(int n, const block_q8_0 *x, const block_q8_0 *y) {
    float sumf = 0;
    for (int i = 0; i < n / QK8_0; i++) {
        int sumi = 0;
        for (int j = 0; j < QK8_0; j++) sumi += x[i].qs[j] * y[i].qs[j];
        sumf += sumi * (fp16_to_fp32(x[i].d) * fp16_to_fp32(y[i].d));
    }
    return sumf;
}

/* ggml_vec_dot_q4_0_q8_0 / q4_1_q8_1 / q5_0_q8_0 / q5_1_q8_1, generic paths */
static float vec_dot_q4_0_q8_0(int n, const block_q4_0 *x, const block_q8_0 *y) {
    float sumf = 0;
    for (int i = 0; i < n / 32; i++) {
        int sumi = 0;
        for (int j = 0; j < 16; ++j) {
            const int v0 = (x[i].qs[j] & 0x0F) - 8, v1 = (x[i].qs[j] >> 4) - 8;
            sumi += (v0 * y[i].qs[j]) + (v1 * y[i].qs[j + 16]);
        }
        sumf += sumi * fp16_to_fp32(x[i].d) * fp16_to_fp32(y[i].d);
    }
    return sumf;
}
static float vec_dot_q4_1_q8_1(int n, const block_q4_1 *x, const block_q8_1 *y) {
    float sumf = 0;
    for (int i = 0; i < n / 32; i++) {
        int sumi = 0;
        for (int j = 0; j < 16; ++j) {
            const int v0 = (x[i].qs[j] & 0x0F), v1 = (x[i].qs[j] >> 4);
            sumi += (v0 * y[i].qs[j]) + (v1 * y[i].qs[j + 16]);
        }
        sumf += (fp16_to_fp32(x[i].d) * fp16_to_fp32(y[i].d)) * sumi + fp16_to_fp32(x[i].m) * fp16_to_fp32(y[i].s);
    }
    return sumf;
}
static float vec_dot_q5_0_q8_0(int n, const block_q5_0 *x, const block_q8_0 *y) {
    float sumf = 0;
    for (int i = 0; i < n / 32; i++) {
        uint32_t qh;
        memcpy(&qh, x[i].qh, sizeof(qh));
        int sumi = 0;
        for (int j = 0; j < 16; ++j) {
            const uint8_t xh_0 = ((qh & (1u << (j + 0))) >> (j + 0)) << 4;
            const uint8_t xh_1 = ((qh & (1u << (j + 16))) >> (j + 12));
            const int32_t x0 = (int8_t)(((x[i].qs[j] & 0x0F) | xh_0) - 16);
            const int32_t x1 = (int8_t)(((x[i].qs[j] >> 4) | xh_1) - 16);
            sumi += (x0 * y[i].qs[j]) + (x1 * y[i].qs[j + 16]);
        }
        sumf += (fp16_to_fp32(x[i].d) * fp16_to_fp32(y[i].d)) * sumi;
    }
    return sumf;
}
static float vec_dot_q5_1_q8_1(int n, const block_q5_1 *x, const block_q8_1 *y) {
    float sumf = 0;
    for (int i = 0; i < n / 32; i++) {
        uint32_t qh;
        memcpy(&qh, x[i].qh, sizeof(qh));
        int sumi = 0;
        for (int j = 0; j < 16; ++j) {
            const uint8_t xh_0 = ((qh >> (j + 0)) << 4) & 0x10;
            const uint8_t xh_1 = ((qh >> (j + 12))) & 0x10;
            const int32_t x0 = (x[i].qs[j] & 0xF) | xh_0;
            const int32_t x1 = (x[i].qs[j] >> 4) | xh_1;
            sumi += (x0 * y[i].qs[j]) + (x1 * y[i].qs[j + 16]);
        }
        sumf += (fp16_to_fp32(x[i].d) * fp16_to_fp32(y[i].d)) * sumi + fp16_to_fp32(x[i].m) * fp16_to_fp32(y[i].s);
    }
    return sumf;
}

static const uint32_t kmask1 = 0x3f3f3f3f, kmask2 = 0x0f0f0f0f, kmask3 = 0x03030303;

static inline void unpack_k4_scales(const uint8_t *packed, uint32_t utmp[4]) {
    memcpy(utmp, packed, 12);
    utmp[3] = ((utmp[2] >> 4) & kmask2) | (((utmp[1] >> 6) & kmask3) << 4);
    const uint32_t uaux = utmp[1] & kmask1;
    utmp[1] = (utmp[2] & kmask2) | (((utmp[0] >> 6) & kmask3) << 4);
    utmp[2] = uaux;
    utmp[0] &= kmask1;
}

static float k4_accumulate(const int8_t *aux8, const uint8_t *packed_scales, const block_q8_K *y, float xd, float xdmin,
                           float sums[8], float *sumf) {
    uint32_t utmp[4];
    int32_t aux32[8];
    int16_t aux16[8];
    unpack_k4_scales(packed_scales, utmp);
    const uint8_t *scales = (const uint8_t *)&utmp[0];
    const uint8_t *mins = (const uint8_t *)&utmp[2];
    memset(aux32, 0, sizeof(aux32));
    int sumi = 0;
    for (int j = 0; j < QK_K / 16; ++j) sumi += y->bsums[j] * mins[j / 2];
    const int8_t *a = aux8;
    const int8_t *q8 = y->qs;
    int is = 0;
    for (int j = 0; j < QK_K / 32; ++j) {
        int32_t scale = scales[is++];
        for (int r = 0; r < 4; ++r) {
            for (int l = 0; l < 8; ++l) aux16[l] = (int16_t)(q8[l] * a[l]);
            for (int l = 0; l < 8; ++l) aux32[l] += scale * aux16[l];
            q8 += 8;
            a += 8;
        }
    }
    const float d = xd * y->d;
    for (int l = 0; l < 8; ++l) sums[l] += d * aux32[l];
    const float dmin = xdmin * y->d;
    *sumf -= dmin * sumi;
    return 0;
}

/* ggml_vec_dot_q2_K_q8_K (generic path) */
static float vec_dot_q2_K_q8_K(int n, const block_q2_K *x, const block_q8_K *y) {
    float sumf = 0;
    for (int i = 0; i < n / QK_K; ++i) {
        const uint8_t *q2 = x[i].qs, *sc = x[i].scales;
        const int8_t *q8 = y[i].qs;
        int summs = 0;
        for (int j = 0; j < 16; ++j) summs += y[i].bsums[j] * (sc[j] >> 4);
        const float dall = y[i].d * fp16_to_fp32(x[i].d), dmin = y[i].d * fp16_to_fp32(x[i].dmin);
        int isum = 0, is = 0;
        for (int k = 0; k < QK_K / 128; ++k) {
            int shift = 0;
            for (int j = 0; j < 4; ++j) {
                int d = sc[is++] & 0xF, isuml = 0;
                for (int l = 0; l < 16; ++l) isuml += q8[l] * ((q2[l] >> shift) & 3);
                isum += d * isuml;
                d = sc[is++] & 0xF;
                isuml = 0;
                for (int l = 16; l < 32; ++l) isuml += q8[l] * ((q2[l] >> shift) & 3);
                isum += d * isuml;
                shift += 2;
                q8 += 32;
            }
            q2 += 32;
        }
        sumf += dall * isum - dmin * summs;
    }
    return sumf;
}
/* ggml_vec_dot_q3_K_q8_K: the same integer sums as the generic path; the generic path accumulates them in eight fp32
 * lanes, here one fp32 product per super-block (this type runs on the bf16 tensor-core path on the device, whose
 * tolerance is far wider than that re-association) */
static float vec_dot_q3_K_q8_K(int n, const block_q3_K *x, const block_q8_K *y) {
    float sumf = 0;
    for (int i = 0; i < n / QK_K; ++i) {
        const uint8_t *q = x[i].qs, *hm = x[i].hmask;
        const int8_t *q8 = y[i].qs;
        int8_t scales[16];
        q3_K_scales(x[i].scales, scales);
        uint8_t m = 1;
        int isum = 0, is = 0;
        for (int k = 0; k < QK_K / 128; ++k) {
            int shift = 0;
            for (int j = 0; j < 4; ++j) {
                int s0 = 0, s1 = 0;
                for (int l = 0; l < 16; ++l) s0 += q8[l] * ((int)((q[l] >> shift) & 3) - ((hm[l] & m) ? 0 : 4));
                for (int l = 16; l < 32; ++l) s1 += q8[l] * ((int)((q[l] >> shift) & 3) - ((hm[l] & m) ? 0 : 4));
                isum += (scales[is] - 32) * s0 + (scales[is + 1] - 32) * s1;
                is += 2;
                shift += 2;
                m <<= 1;
                q8 += 32;
            }
            q += 32;
        }
        sumf += (fp16_to_fp32(x[i].d) * y[i].d) * isum;
    }
    return sumf;
}

static float vec_dot_q4_K_q8_K(int n, const block_q4_K *x, const block_q8_K *y) {
    int8_t aux8[QK_K];
    float sums[8] = {0};
    float sumf = 0;
    for (int i = 0; i < n / QK_K; ++i) {
        const uint8_t *q4 = x[i].qs;
        int8_t *a = aux8;
        for (int j = 0; j < QK_K / 64; ++j) {
            for (int l = 0; l < 32; ++l) a[l] = (int8_t)(q4[l] & 0xF);
            a += 32;
            for (int l = 0; l < 32; ++l) a[l] = (int8_t)(q4[l] >> 4);
            a += 32;
            q4 += 32;
        }
        k4_accumulate(aux8, x[i].scales, &y[i], fp16_to_fp32(x[i].d), fp16_to_fp32(x[i].dmin), sums, &sumf);
    }
    for (int l = 0; l < 8; ++l) sumf += sums[l];
    return sumf;
}

static float vec_dot_q5_K_q8_K(int n, const block_q5_K *x, const block_q8_K *y) {
    int8_t aux8[QK_K];
    float sums[8] = {0};
    float sumf = 0;
    for (int i = 0; i < n / QK_K; ++i) {
        const uint8_t *q4 = x[i].qs, *hm = x[i].qh;
        int8_t *a = aux8;
        uint8_t m = 1;
        for (int j = 0; j < QK_K / 64; ++j) {
            for (int l = 0; l < 32; ++l) a[l] = (int8_t)(q4[l] & 0xF);
            for (int l = 0; l < 32; ++l) a[l] += (hm[l] & m ? 16 : 0);
            a += 32;
            m <<= 1;
            for (int l = 0; l < 32; ++l) a[l] = (int8_t)(q4[l] >> 4);
            for (int l = 0; l < 32; ++l) a[l] += (hm[l] & m ? 16 : 0);
            a += 32;
            m <<= 1;
            q4 += 32;
        }
        k4_accumulate(aux8, x[i].scales, &y[i], fp16_to_fp32(x[i].d), fp16_to_fp32(x[i].dmin), sums, &sumf);
    }
    for (int l = 0; l < 8; ++l) sumf += sums[l];
    return sumf;
}

static float vec_dot_q6_K_q8_K(int n, const block_q6_K *x, const block_q8_K *y) {
    int8_t aux8[QK_K];
    int16_t aux16[8];
    float sums[8] = {0};
    int32_t aux32[8];
    float sumf = 0;
    for (int i = 0; i < n / QK_K; ++i) {
        const uint8_t *q4 = x[i].ql, *qh = x[i].qh;
        const int8_t *q8 = y[i].qs;
        memset(aux32, 0, sizeof(aux32));
        int8_t *a = aux8;
        for (int j = 0; j < QK_K; j += 128) {
            for (int l = 0; l < 32; ++l) {
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
        int is = 0;
        for (int j = 0; j < QK_K / 16; ++j) {
            int scale = x[i].scales[is++];
            for (int r = 0; r < 2; ++r) {
                for (int l = 0; l < 8; ++l) aux16[l] = (int16_t)(q8[l] * a[l]);
                for (int l = 0; l < 8; ++l) aux32[l] += scale * aux16[l];
                q8 += 8;
                a += 8;
            }
        }
        const float d = fp16_to_fp32(x[i].d) * y[i].d;
        for (int l = 0; l < 8; ++l) sums[l] += d * aux32[l];
    }
    for (int l = 0; l < 8; ++l) sumf += sums[l];
    return sumf;
}

/* ggml_vec_dot_iq4_nl_q8_0 (generic path): one float product per 32-element block */
static float vec_dot_iq4_nl_q8_0(int n, const block_iq4_nl *x, const block_q8_0 *y) {
    const int nb = n / 32;
    float sumf = 0;
    for (int ib = 0; ib < nb; ++ib) {
        const float d = fp16_to_fp32(y[ib].d) * fp16_to_fp32(x[ib].d);
        int sumi1 = 0, sumi2 = 0;
        for (int j = 0; j < 16; ++j) {
            sumi1 += y[ib].qs[j] * kvalues_iq4nl[x[ib].qs[j] & 0xf];
            sumi2 += y[ib].qs[j + 16] * kvalues_iq4nl[x[ib].qs[j] >> 4];
        }
        sumf += d * (sumi1 + sumi2);
    }
    return sumf;
}

/* ggml_vec_dot_iq4_xs_q8_K (generic path): sub-block scales applied in float, two sub-blocks per turn */
static float vec_dot_iq4_xs_q8_K(int n, const block_iq4_xs *x, const block_q8_K *y) {
    const int nb = n / QK_K;
    float sumf = 0;
    for (int ibl = 0; ibl < nb; ++ibl) {
        const float d4d8 = fp16_to_fp32(x[ibl].d) * y[ibl].d;
        uint16_t h = x[ibl].scales_h;
        const uint8_t *qs = x[ibl].qs;
        const int8_t *q8 = y[ibl].qs;
        for (int ib = 0; ib < QK_K / 32; ib += 2) {
            const uint8_t ls1 = (uint8_t)((x[ibl].scales_l[ib / 2] & 0xf) | ((h << 4) & 0x30));
            const uint8_t ls2 = (uint8_t)((x[ibl].scales_l[ib / 2] >> 4) | ((h << 2) & 0x30));
            h >>= 4;
            const float d1 = d4d8 * (ls1 - 32);
            const float d2 = d4d8 * (ls2 - 32);
            int sumi1 = 0, sumi2 = 0;
            for (int j = 0; j < 16; ++j) {
                sumi1 += q8[j + 0] * kvalues_iq4nl[qs[j] & 0xf];
                sumi2 += q8[j + 16] * kvalues_iq4nl[qs[j] >> 4];
            }
            sumf += d1 * (sumi1 + sumi2);
            qs += 16;
            q8 += 32;
            sumi1 = sumi2 = 0;
            for (int j = 0; j < 16; ++j) {
                sumi1 += q8[j + 0] * kvalues_iq4nl[qs[j] & 0xf];
                sumi2 += q8[j + 16] * kvalues_iq4nl[qs[j] >> 4];
            }
            sumf += d2 * (sumi1 + sumi2);
            qs += 16;
            q8 += 32;
        }
    }
    return sumf;
}

/* generic ggml_vec_dot_f16 / ggml_vec_dot_f32: double accumulator over f32 products */
static float vec_dot_f16(int n, const fp16_t *x, const fp16_t *y) {
    double sumf = 0.0;
    for (int i = 0; i < n; ++i) sumf += (double)(fp16_to_fp32(x[i]) * fp16_to_fp32(y[i]));
    return (float)sumf;
}
static float vec_dot_f32(int n, const float *x, const float *y) {
    double sumf = 0.0;
    for (int i = 0; i < n; ++i) sumf += (double)(x[i] * y[i]);
    return (float)sumf;
}

/* the type activations are converted to before the dot (ggml type_traits.vec_dot_type) */
int ref_vec_dot_type(int wtype) {
    switch (wtype) {
        case T_F32: return T_F32;
        case T_F16: return T_F16;
        case T_Q8_0: case T_Q4_0: case T_Q5_0: case T_IQ4_NL: return T_Q8_0;
        case T_Q4_1: case T_Q5_1: return T_Q8_1;
        default: return T_Q8_K;
    }
}

void ref_quantize_activations(int wtype, const float *x, void *y, int64_t k) {
    switch (ref_vec_dot_type(wtype)) {
        case T_F32: memcpy(y, x, k * 4); break;
        case T_F16: for (int64_t i = 0; i < k; i++) ((fp16_t *)y)[i] = fp32_to_fp16(x[i]); break;
        case T_Q8_0: ref_quantize_row_q8_0(x, y, k); break;
        case T_Q8_1: ref_quantize_row_q8_1(x, y, k); break;
        default: ref_quantize_row_q8_K(x, y, k); break;
    }
}

float ref_vec_dot(int wtype, int n, const void *w, const void *yq) {
    switch (wtype) {
        case T_F32: return vec_dot_f32(n, (const float *)w, (const float *)yq);
        case T_F16: return vec_dot_f16(n, (const fp16_t *)w, (const fp16_t *)yq);
        case T_Q8_0: return vec_dot_q8_0_q8_0(n, (const block_q8_0 *)w, (const block_q8_0 *)yq);
        case T_Q4_0: return vec_dot_q4_0_q8_0(n, (const block_q4_0 *)w, (const block_q8_0 *)yq);
        case T_Q4_1: return vec_dot_q4_1_q8_1(n, (const block_q4_1 *)w, (const block_q8_1 *)yq);
        case T_Q5_0: return vec_dot_q5_0_q8_0(n, (const block_q5_0 *)w, (const block_q8_0 *)yq);
        case T_Q5_1: return vec_dot_q5_1_q8_1(n, (const block_q5_1 *)w, (const block_q8_1 *)yq);
        case T_Q2_K: return vec_dot_q2_K_q8_K(n, (const block_q2_K *)w, (const block_q8_K *)yq);
        case T_Q3_K: return vec_dot_q3_K_q8_K(n, (const block_q3_K *)w, (const block_q8_K *)yq);
        case T_Q4_K: return vec_dot_q4_K_q8_K(n, (const block_q4_K *)w, (const block_q8_K *)yq);
        case T_Q5_K: return vec_dot_q5_K_q8_K(n, (const block_q5_K *)w, (const block_q8_K *)yq);
        case T_Q6_K: return vec_dot_q6_K_q8_K(n, (const block_q6_K *)w, (const block_q8_K *)yq);
        case T_IQ4_NL: return vec_dot_iq4_nl_q8_0(n, (const block_iq4_nl *)w, (const block_q8_0 *)yq);
        case T_IQ4_XS: return vec_dot_iq4_xs_q8_K(n, (const block_iq4_xs *)w, (const block_q8_K *)yq);
    }
    return NAN;
}

/* ggml_compute_forward_mul_mat: out[t][r] = dot(W[r,:], quantise(x[t,:])) */
void ref_mul_mat(int wtype, const void *W, int64_t rows, int64_t cols, const float *x, int n_tok, float *out) {
    const int vt = ref_vec_dot_type(wtype);
    const int64_t wrb = ref_row_bytes(wtype, cols);
    const int64_t xrb = ref_row_bytes(vt, cols);
    uint8_t *xq = (uint8_t *)malloc((size_t)(xrb * n_tok));
#pragma omp parallel for schedule(static)
    for (int t = 0; t < n_tok; t++) ref_quantize_activations(wtype, x + (int64_t)t * cols, xq + (int64_t)t * xrb, cols);
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < rows; r++) {
        const uint8_t *w = (const uint8_t *)W + r * wrb;
        for (int t = 0; t < n_tok; t++) out[(int64_t)t * rows + r] = ref_vec_dot(wtype, (int)cols, w, xq + (int64_t)t * xrb);
    }
    free(xq);
}

/* ------------------------------------------------------------------ the llama graph (llm_build_llama) */
typedef struct { const void *data; int type; int64_t rows, cols; } ref_tensor;

enum { TK_TOK_EMBD = 0, TK_OUTPUT_NORM, TK_OUTPUT, TK_ATTN_NORM, TK_WQ, TK_WK, TK_WV, TK_WO, TK_FFN_NORM, TK_GATE, TK_UP,
       TK_DOWN, TK_BQ, TK_BK, TK_BV, TK_COUNT }; /* TK_BQ/BK/BV: attn_q/k/v.bias (f32 rows; qwen2) */

typedef struct {
    int n_layer, n_embd, n_head, n_head_kv, head_dim, n_ff, n_vocab, n_ctx;
    float rms_eps, rope_theta, rope_freq_scale;
    /* RoPE scaling (llama.cpp cparams / hparams): ext_factor 1 for YaRN else 0, attn_factor, the YaRN correction range,
     * optional per-dimension frequency divisors (rope_freqs.weight, the Llama-3.1 scheme) */
    float rope_ext_factor, rope_attn_factor, rope_corr_dims[2];
    const float *rope_freq_factors;
    int rope_neox; /* LLAMA_ROPE_TYPE_NEOX (qwen2): the rotated pairs are (i, i + n_dims / 2) instead of (2i, 2i + 1) */
    ref_tensor global[3];
    ref_tensor *layer[TK_COUNT]; /* indices >= TK_ATTN_NORM used */
    fp16_t *k_cache, *v_cache;   /* [n_layer][n_ctx][n_head_kv*head_dim] */
    float *dbg_hidden;           /* optional [n_layer+1][n_embd]: residual stream of the LAST token after each layer */
} ref_ctx;

ref_ctx *ref_ctx_new(int n_layer, int n_embd, int n_head, int n_head_kv, int n_ff, int n_vocab, int n_ctx, float rms_eps,
                     float rope_theta, float rope_freq_scale) {
    ref_ctx *c = (ref_ctx *)calloc(1, sizeof(ref_ctx));
    c->n_layer = n_layer; c->n_embd = n_embd; c->n_head = n_head; c->n_head_kv = n_head_kv;
    c->head_dim = n_embd / n_head; c->n_ff = n_ff; c->n_vocab = n_vocab; c->n_ctx = n_ctx;
    c->rms_eps = rms_eps; c->rope_theta = rope_theta; c->rope_freq_scale = rope_freq_scale;
    c->rope_ext_factor = 0.0f; c->rope_attn_factor = 1.0f; c->rope_corr_dims[0] = c->rope_corr_dims[1] = 0.0f; c->rope_freq_factors = NULL;
    for (int k = TK_ATTN_NORM; k < TK_COUNT; k++) c->layer[k] = (ref_tensor *)calloc(n_layer, sizeof(ref_tensor));
    size_t kv = (size_t)n_layer * n_ctx * n_head_kv * c->head_dim;
    c->k_cache = (fp16_t *)calloc(kv, sizeof(fp16_t));
    c->v_cache = (fp16_t *)calloc(kv, sizeof(fp16_t));
    return c;
}

void ref_rope_yarn_corr_dims(int n_dims, int n_ctx_orig, float freq_base, float beta_fast, float beta_slow, float dims[2]);
/* llama.cpp: freq_scale = 1 / rope.scaling.factor (linear and YaRN), ext_factor = 1 for YaRN, beta_fast 32, beta_slow 1 */
void ref_ctx_set_rope_scaling(ref_ctx *c, float freq_scale, float ext_factor, float attn_factor, int n_ctx_orig, float beta_fast,
                              float beta_slow, const float *freq_factors) {
    c->rope_freq_scale = freq_scale; c->rope_ext_factor = ext_factor; c->rope_attn_factor = attn_factor;
    c->rope_freq_factors = freq_factors;
    if (ext_factor != 0.0f) ref_rope_yarn_corr_dims(c->head_dim, n_ctx_orig, c->rope_theta, beta_fast, beta_slow, c->rope_corr_dims);
}

void ref_ctx_set_rope_neox(ref_ctx *c, int neox) { c->rope_neox = neox; }

void ref_ctx_free(ref_ctx *c) {
    if (!c) return;
    for (int k = TK_ATTN_NORM; k < TK_COUNT; k++) free(c->layer[k]);
    free(c->k_cache); free(c->v_cache); free(c);
}

void ref_ctx_set_tensor(ref_ctx *c, int kind, int layer, const void *data, int type, int64_t rows, int64_t cols) {
    ref_tensor t = {data, type, rows, cols};
    if (kind < TK_ATTN_NORM) c->global[kind] = t; else c->layer[kind][layer] = t;
}

void ref_ctx_set_debug(ref_ctx *c, float *buf) { c->dbg_hidden = buf; }

/* ggml_compute_forward_rms_norm_f32 followed by ggml_mul with the norm weight */
static void rms_norm_mul(const float *x, const float *w, float *y, int n, float eps) {
    double sum = 0.0;
    for (int i = 0; i < n; i++) sum += (double)(x[i] * x[i]);
    const float mean = (float)(sum / n);
    const float scale = 1.0f / sqrtf(mean + eps);
    for (int i = 0; i < n; i++) y[i] = x[i] * scale;
    for (int i = 0; i < n; i++) y[i] = y[i] * w[i];
}

/* ggml_rope_cache_init (ext_factor = 0, attn_factor = 1, no freq_factors) */
void ref_rope_cache(float theta_base, float freq_scale, float freq_base, int n_dims, float *cache) {
    const float theta_scale = powf(freq_base, -2.0f / n_dims);
    float theta = theta_base;
    for (int i0 = 0; i0 < n_dims; i0 += 2) {
        const float t = freq_scale * theta;
        cache[i0 + 0] = cosf(t);
        cache[i0 + 1] = sinf(t);
        theta *= theta_scale;
    }
}

/* The scaled variants, restated from ggml's published algorithm (ggml.c: rope_yarn, rope_yarn_ramp,
 * ggml_rope_yarn_corr_dims, ggml_rope_cache_init with freq_factors; YaRN: arXiv 2309.00071).  llama.cpp is an
 * un-vendored submodule of the reference, so this part of the oracle is pinned the way the unscaled path is: against
 * transformers' implementations of the same three schemes (tools/crosscheck_transformers.py: rope_linear / rope_yarn /
 * rope_l31 entries of tests/golden/transformers_crosscheck.json). */
static float rope_yarn_ramp(float low, float high, int i0) {
    const float y = (i0 / 2 - low) / fmaxf(0.001f, high - low);
    return 1.0f - fminf(1.0f, fmaxf(0.0f, y));
}
static float rope_yarn_corr_dim(int n_dims, int n_ctx_orig, float n_rot, float base) {
    return n_dims * logf(n_ctx_orig / (n_rot * 2 * (float)M_PI)) / (2 * logf(base));
}
void ref_rope_yarn_corr_dims(int n_dims, int n_ctx_orig, float freq_base, float beta_fast, float beta_slow, float dims[2]) {
    const float start = floorf(rope_yarn_corr_dim(n_dims, n_ctx_orig, beta_fast, freq_base));
    const float end = ceilf(rope_yarn_corr_dim(n_dims, n_ctx_orig, beta_slow, freq_base));
    dims[0] = fmaxf(0, start);
    dims[1] = fminf(n_dims - 1, end);
}
void ref_rope_cache_ext(float theta_base, float freq_scale, float freq_base, int n_dims, const float *freq_factors, float ext_factor,
                        float attn_factor, const float corr_dims[2], float *cache) {
    const float theta_scale = powf(freq_base, -2.0f / n_dims);
    float theta = theta_base;
    for (int i0 = 0; i0 < n_dims; i0 += 2) {
        const float ff = freq_factors ? freq_factors[i0 / 2] : 1.0f;
        const float theta_extrap = theta / ff;
        const float theta_interp = freq_scale * theta_extrap;
        float t = theta_interp, mscale = attn_factor;
        if (ext_factor != 0.0f) {
            const float ramp_mix = rope_yarn_ramp(corr_dims[0], corr_dims[1], i0) * ext_factor;
            t = theta_interp * (1 - ramp_mix) + theta_extrap * ramp_mix;
            mscale *= 1.0f + 0.1f * logf(1.0f / freq_scale);
        }
        cache[i0 + 0] = cosf(t) * mscale;
        cache[i0 + 1] = sinf(t) * mscale;
        theta *= theta_scale;
    }
}

/* mode 2 ("NEOX", ggml.c rope_neox): pair i = (x[i], x[i + n_dims / 2]), rotated by cache entry i */
static void rope_neox(float *x, const float *cache, int n_dims) {
    for (int i0 = 0; i0 < n_dims; i0 += 2) {
        const int ic = i0 / 2;
        const float c = cache[i0], s = cache[i0 + 1];
        const float x0 = x[ic], x1 = x[ic + n_dims / 2];
        x[ic] = x0 * c - x1 * s;
        x[ic + n_dims / 2] = x0 * s + x1 * c;
    }
}

/* mode 0 ("NORM"): adjacent pairs */
static void rope_norm(float *x, const float *cache, int n_dims) {
    for (int i0 = 0; i0 < n_dims; i0 += 2) {
        const float c = cache[i0], s = cache[i0 + 1];
        const float x0 = x[i0], x1 = x[i0 + 1];
        x[i0] = x0 * c - x1 * s;
        x[i0 + 1] = x0 * s + x1 * c;
    }
}

static void mat(const ref_tensor *t, const float *x, int n_tok, float *out) {
    ref_mul_mat(t->type, t->data, t->rows, t->cols, x, n_tok, out);
}

/*
 * One llama_decode() over `n_tokens` tokens of ONE sequence occupying cache positions
 * [n_past, n_past+n_tokens).  logits: NULL, or [n_tokens][n_vocab] if logits_all else [n_vocab] (last token).
 * embd_out: NULL or [n_embd] = "result_norm" of the last token (what llama_get_embeddings returns with pooling NONE).
 * Returns 0, or -1 when the context is exhausted.
 */
int ref_eval(ref_ctx *c, const int32_t *tokens, int n_tokens, int n_past, float *logits, int logits_all, float *embd_out) {
    const int d = c->n_embd, hd = c->head_dim, nh = c->n_head, nkv = c->n_head_kv, ff = c->n_ff;
    const int kvd = nkv * hd, gqa = nh / nkv;
    if (n_past + n_tokens > c->n_ctx) return -1;
    const int64_t T = n_tokens;
    float *inpL = (float *)malloc(sizeof(float) * T * d);
    float *cur = (float *)malloc(sizeof(float) * T * d);
    float *q = (float *)malloc(sizeof(float) * T * d);
    float *k = (float *)malloc(sizeof(float) * T * kvd);
    float *v = (float *)malloc(sizeof(float) * T * kvd);
    float *att = (float *)malloc(sizeof(float) * T * d);
    float *g = (float *)malloc(sizeof(float) * T * ff);
    float *u = (float *)malloc(sizeof(float) * T * ff);
    float *ropec = (float *)malloc(sizeof(float) * T * hd);
    const float kq_scale = 1.0f / sqrtf((float)hd);

    /* get_rows(tok_embd) */
    const ref_tensor *te = &c->global[TK_TOK_EMBD];
    const int64_t erb = ref_row_bytes(te->type, d);
    for (int t = 0; t < T; t++) ref_dequantize_row(te->type, (const uint8_t *)te->data + (int64_t)tokens[t] * erb, inpL + t * d, d);
    for (int t = 0; t < T; t++) {
        if (c->rope_ext_factor == 0.0f && c->rope_attn_factor == 1.0f && !c->rope_freq_factors)
            ref_rope_cache((float)(n_past + t), c->rope_freq_scale, c->rope_theta, hd, ropec + t * hd);
        else
            ref_rope_cache_ext((float)(n_past + t), c->rope_freq_scale, c->rope_theta, hd, c->rope_freq_factors, c->rope_ext_factor,
                               c->rope_attn_factor, c->rope_corr_dims, ropec + t * hd);
    }

    for (int il = 0; il < c->n_layer; il++) {
        fp16_t *kc = c->k_cache + (size_t)il * c->n_ctx * kvd;
        fp16_t *vc = c->v_cache + (size_t)il * c->n_ctx * kvd;
        const float *wn = (const float *)c->layer[TK_ATTN_NORM][il].data;
        for (int t = 0; t < T; t++) rms_norm_mul(inpL + t * d, wn, cur + t * d, d, c->rms_eps);
        mat(&c->layer[TK_WQ][il], cur, (int)T, q);
        mat(&c->layer[TK_WK][il], cur, (int)T, k);
        mat(&c->layer[TK_WV][il], cur, (int)T, v);
        /* qwen2 (llm_build_qwen2): Qcur = ggml_add(Qcur, bq) etc. before the rotation */
        const float *bq = (const float *)c->layer[TK_BQ][il].data, *bk = (const float *)c->layer[TK_BK][il].data,
                    *bv = (const float *)c->layer[TK_BV][il].data;
        for (int t = 0; t < T; t++) {
            if (bq) for (int i = 0; i < d; i++) q[t * d + i] += bq[i];
            if (bk) for (int i = 0; i < kvd; i++) k[t * kvd + i] += bk[i];
            if (bv) for (int i = 0; i < kvd; i++) v[t * kvd + i] += bv[i];
        }
        for (int t = 0; t < T; t++) {
            for (int h = 0; h < nh; h++) (c->rope_neox ? rope_neox : rope_norm)(q + t * d + h * hd, ropec + t * hd, hd);
            for (int h = 0; h < nkv; h++) (c->rope_neox ? rope_neox : rope_norm)(k + t * kvd + h * hd, ropec + t * hd, hd);
            for (int i = 0; i < kvd; i++) {
                kc[(size_t)(n_past + t) * kvd + i] = fp32_to_fp16(k[t * kvd + i]);
                vc[(size_t)(n_past + t) * kvd + i] = fp32_to_fp16(v[t * kvd + i]);
            }
        }
        /* kq = mul_mat(k_f16, q) [q -> f16] ; soft_max_ext(scale, causal mask) ; kqv = mul_mat(v_f16, kq) [kq -> f16] */
#pragma omp parallel for collapse(2) schedule(static)
        for (int t = 0; t < (int)T; t++) {
            for (int h = 0; h < nh; h++) {
                const int n_kv = n_past + t + 1;
                const int hk = h / gqa;
                fp16_t qh[512];
                float *p = (float *)malloc(sizeof(float) * n_kv);
                fp16_t *ph = (fp16_t *)malloc(sizeof(fp16_t) * n_kv);
                for (int i = 0; i < hd; i++) qh[i] = fp32_to_fp16(q[t * d + h * hd + i]);
                float mx = -INFINITY;
                for (int j = 0; j < n_kv; j++) {
                    p[j] = vec_dot_f16(hd, kc + (size_t)j * kvd + hk * hd, qh) * kq_scale;
                    if (p[j] > mx) mx = p[j];
                }
                double sum = 0.0;
                for (int j = 0; j < n_kv; j++) {
                    const float e = expf(p[j] - mx);
                    sum += (double)e;
                    p[j] = e;
                }
                const float inv = (float)(1.0 / sum);
                for (int j = 0; j < n_kv; j++) ph[j] = fp32_to_fp16(p[j] * inv);
                for (int i = 0; i < hd; i++) {
                    double acc = 0.0; /* vec_dot_f16 over the kv axis (v cache is stored transposed upstream) */
                    for (int j = 0; j < n_kv; j++)
                        acc += (double)(fp16_to_fp32(vc[(size_t)j * kvd + hk * hd + i]) * fp16_to_fp32(ph[j]));
                    att[t * d + h * hd + i] = (float)acc;
                }
                free(p);
                free(ph);
            }
        }
        mat(&c->layer[TK_WO][il], att, (int)T, cur);
        for (int64_t i = 0; i < T * d; i++) inpL[i] = cur[i] + inpL[i]; /* ffn_inp = add(cur, inpSA) */
        const float *fn = (const float *)c->layer[TK_FFN_NORM][il].data;
        for (int t = 0; t < T; t++) rms_norm_mul(inpL + t * d, fn, cur + t * d, d, c->rms_eps);
        mat(&c->layer[TK_UP][il], cur, (int)T, u);
        mat(&c->layer[TK_GATE][il], cur, (int)T, g);
        for (int64_t i = 0; i < T * ff; i++) {
            const float s = g[i] / (1.0f + expf(-g[i])); /* ggml_silu_f32 */
            g[i] = s * u[i];
        }
        mat(&c->layer[TK_DOWN][il], g, (int)T, cur);
        for (int64_t i = 0; i < T * d; i++) inpL[i] = cur[i] + inpL[i];
        if (c->dbg_hidden) memcpy(c->dbg_hidden + (size_t)il * d, inpL + (T - 1) * d, sizeof(float) * d);
    }
    const float *on = (const float *)c->global[TK_OUTPUT_NORM].data;
    const int t0 = logits_all ? 0 : (int)T - 1;
    for (int t = t0; t < T; t++) rms_norm_mul(inpL + t * d, on, cur + t * d, d, c->rms_eps);
    if (c->dbg_hidden) memcpy(c->dbg_hidden + (size_t)c->n_layer * d, cur + (T - 1) * d, sizeof(float) * d);
    if (embd_out) memcpy(embd_out, cur + (T - 1) * d, sizeof(float) * d);
    if (logits) {
        const ref_tensor *out = c->global[TK_OUTPUT].data ? &c->global[TK_OUTPUT] : &c->global[TK_TOK_EMBD];
        mat(out, cur + (int64_t)t0 * d, (int)T - t0, logits);
    }
    free(inpL); free(cur); free(q); free(k); free(v); free(att); free(g); free(u); free(ropec);
    return 0;
}

/* greedy sampler: first maximal element (std::max_element semantics) */
int ref_argmax(const float *x, int n) {
    int best = 0;
    for (int i = 1; i < n; i++) if (x[i] > x[best]) best = i;
    return best;
}

/* prompt + greedy generation; returns number of generated tokens written to out */
int ref_generate_greedy(ref_ctx *c, const int32_t *prompt, int n_prompt, int n_gen, int32_t *out, float *last_logits,
                        float *margins) {
    float *logits = (float *)malloc(sizeof(float) * c->n_vocab);
    if (ref_eval(c, prompt, n_prompt, 0, logits, 0, NULL) != 0) { free(logits); return 0; }
    int n_past = n_prompt, n = 0;
    while (n < n_gen) {
        int tok = ref_argmax(logits, c->n_vocab);
        if (margins) { /* top-1 minus top-2 logit: how robust the greedy choice is */
            float second = -INFINITY;
            for (int i = 0; i < c->n_vocab; i++) if (i != tok && logits[i] > second) second = logits[i];
            margins[n] = logits[tok] - second;
        }
        out[n++] = tok;
        if (n == n_gen || n_past >= c->n_ctx) break;
        if (ref_eval(c, &tok, 1, n_past, logits, 0, NULL) != 0) break;
        n_past++;
    }
    if (last_logits) memcpy(last_logits, logits, sizeof(float) * c->n_vocab);
    free(logits);
    return n;
}

/* launchers such as torchrun export OMP_NUM_THREADS=1; the CPU baseline is meant to use every host core */
void ref_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int ref_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
