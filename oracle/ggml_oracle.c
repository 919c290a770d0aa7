/*
 * oracle/ggml_oracle.c — CPU restatement of the reference's quantized MUL_MAT / MUL_MAT_ID path.
 *
 * TEST INFRASTRUCTURE ONLY (see ggml_oracle.h).  Parity status: PINNED against the reference compiled
 * from /root/reference (oracle/_ref) and against tests/golden/ by tests/test_oracle.py.
 *
 * Plain scalar C, one function per reference function, each citing the reference file:line it
 * follows (paths relative to /root/reference).  Compiled with -ffp-contract=off so that no FMA
 * contraction changes a rounding: the reference's generic build evaluates `x*id` and `sum*d` as
 * separate IEEE operations and so do we.
 */
#include "ggml_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------------
 * block layouts (ggml/src/ggml-common.h:184-189, 241-259, 317-328, 352-366); all little-endian,
 * no padding, fields only 2-byte aligned -> accessed through memcpy / byte pointers.
 * ---------------------------------------------------------------------------------------------- */
#define QK   32   /* elements per simple block   */
#define QKK  256  /* elements per k-quant superblock */

#define SZ_Q4_0 18   /* u16 d; u8 qs[16]                                  */
#define SZ_Q8_0 34   /* u16 d; i8 qs[32]                                  */
#define SZ_Q8_1 36   /* u16 d; u16 s; i8 qs[32]                           */
#define SZ_Q4_K 144  /* u16 d; u16 dmin; u8 scales[12]; u8 qs[128]        */
#define SZ_Q6_K 210  /* u8 ql[128]; u8 qh[64]; i8 scales[16]; u16 d       */
#define SZ_Q8_K 292  /* f32 d; i8 qs[256]; i16 bsums[16]                  */

int orc_blck_size(int type) {
    switch (type) {
        case ORC_TYPE_F32:  return 1;
        case ORC_TYPE_Q4_0: case ORC_TYPE_Q8_0: case ORC_TYPE_Q8_1: return QK;
        case ORC_TYPE_Q4_K: case ORC_TYPE_Q6_K: case ORC_TYPE_Q8_K: return QKK;
        default: return 0;
    }
}

size_t orc_type_size(int type) {
    switch (type) {
        case ORC_TYPE_F32:  return 4;
        case ORC_TYPE_Q4_0: return SZ_Q4_0;
        case ORC_TYPE_Q8_0: return SZ_Q8_0;
        case ORC_TYPE_Q8_1: return SZ_Q8_1;
        case ORC_TYPE_Q4_K: return SZ_Q4_K;
        case ORC_TYPE_Q6_K: return SZ_Q6_K;
        case ORC_TYPE_Q8_K: return SZ_Q8_K;
        default: return 0;
    }
}

size_t orc_row_size(int type, int64_t k) {
    const int bs = orc_blck_size(type);
    return bs ? orc_type_size(type) * (size_t) (k / bs) : 0;
}

static inline uint16_t ld_u16(const uint8_t * p) { uint16_t v; memcpy(&v, p, 2); return v; }
static inline void     st_u16(uint8_t * p, uint16_t v) { memcpy(p, &v, 2); }

/* ------------------------------------------------------------------------------------------------
 * binary16 conversions.  The reference's software versions (ggml/src/ggml-impl.h:384-431) and the
 * F16C instructions its x86 build uses (ggml-cpu/simd-mappings.h:56-63) are both exact IEEE
 * conversions with round-to-nearest-even; this is an independent formulation of the same mapping.
 * ---------------------------------------------------------------------------------------------- */
float orc_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = (uint32_t) (h & 0x8000u) << 16;
    const uint32_t exp  = (h >> 10) & 0x1fu;
    uint32_t       man  = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else { /* subnormal: renormalise */
            int e = -1;
            do { man <<= 1; e++; } while ((man & 0x400u) == 0);
            bits = sign | ((uint32_t) (127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7f800000u | (man << 13);
    } else {
        bits = sign | ((exp + (127 - 15)) << 23) | (man << 13);
    }
    float f; memcpy(&f, &bits, 4); return f;
}

uint16_t orc_fp32_to_fp16(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    const uint16_t sign = (uint16_t) ((x >> 16) & 0x8000u);
    const uint32_t ax   = x & 0x7fffffffu;
    if (ax > 0x7f800000u) return (uint16_t) (sign | 0x7e00u);         /* NaN (reference: 0x7E00)  */
    if (ax >= 0x47800000u) return (uint16_t) (sign | 0x7c00u);        /* >= 65536 (or inf) -> inf; 65520..65536 handled below */
    const int32_t  e = (int32_t) (ax >> 23) - 127;                    /* unbiased exponent        */
    if (e < -25) return sign;                                         /* < 2^-25 -> 0             */
    uint32_t man = (ax & 0x7fffffu) | 0x800000u;                      /* 24-bit significand       */
    int shift;                                                        /* bits dropped             */
    uint32_t he;                                                      /* half exponent field      */
    if (e < -14) { shift = 13 + (-14 - e); he = 0; }                  /* subnormal half           */
    else         { shift = 13;             he = (uint32_t) (e + 15); }
    const uint32_t keep  = man >> shift;
    const uint32_t rem   = man & ((1u << shift) - 1u);
    const uint32_t halfw = 1u << (shift - 1);
    uint32_t h = (he << 10) + (he ? (keep & 0x3ffu) : keep);
    if (rem > halfw || (rem == halfw && (keep & 1u))) h += 1;         /* RNE; carries propagate into the exponent */
    return (uint16_t) (sign | h);
}

/* ------------------------------------------------------------------------------------------------
 * activation quantizers
 * ---------------------------------------------------------------------------------------------- */

/* quantize_row_q8_1_ref, ggml/src/ggml-quants.c:260-293 (== quantize_row_q8_1_generic,
 * ggml-cpu/quants.c:49-51, == the fork's s390x VXE quantize_row_q8_1, ggml-cpu/arch/s390/quants.c:94-144) */
void orc_quantize_row_q8_1(const float * x, void * vy, int64_t k) {
    uint8_t * y = (uint8_t *) vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++, x += QK, y += SZ_Q8_1) {
        float amax = 0.0f;
        for (int j = 0; j < QK; j++) { const float a = fabsf(x[j]); if (a > amax) amax = a; }
        const float d  = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        st_u16(y, orc_fp32_to_fp16(d));
        int sum = 0;
        int8_t * qs = (int8_t *) (y + 4);
        for (int j = 0; j < QK; j++) {
            const float v = x[j] * id;
            qs[j] = (int8_t) roundf(v);      /* round half away from zero */
            sum  += qs[j];
        }
        st_u16(y + 2, orc_fp32_to_fp16((float) sum * d));  /* s uses the fp32 d, not the rounded half */
    }
}

/* quantize_row_q8_0_ref, ggml/src/ggml-quants.c:234-257 */
void orc_quantize_row_q8_0(const float * x, void * vy, int64_t k) {
    uint8_t * y = (uint8_t *) vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++, x += QK, y += SZ_Q8_0) {
        float amax = 0.0f;
        for (int j = 0; j < QK; j++) { const float a = fabsf(x[j]); if (a > amax) amax = a; }
        const float d  = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        st_u16(y, orc_fp32_to_fp16(d));
        int8_t * qs = (int8_t *) (y + 2);
        for (int j = 0; j < QK; j++) qs[j] = (int8_t) roundf(x[j] * id);
    }
}

/* nearest_int, ggml/src/ggml-quants.c:559-564 (round-to-nearest-even through the 1.5*2^23 trick) */
static inline int nearest_int(float fval) {
    float val = fval + 12582912.f;
    int i; memcpy(&i, &val, sizeof(int));
    return (i & 0x007fffff) - 0x00400000;
}

/* quantize_row_q8_K_ref, ggml/src/ggml-quants.c:2692-2729 */
void orc_quantize_row_q8_K(const float * x, void * vy, int64_t k) {
    uint8_t * y = (uint8_t *) vy;
    const int64_t nb = k / QKK;
    for (int64_t i = 0; i < nb; i++, x += QKK, y += SZ_Q8_K) {
        float max = 0, amax = 0;
        for (int j = 0; j < QKK; j++) {
            const float ax = fabsf(x[j]);
            if (ax > amax) { amax = ax; max = x[j]; }
        }
        int8_t * qs = (int8_t *) (y + 4);
        if (!amax) {
            /* the reference zeroes d and qs only; bsums of an all-zero block are left untouched there
             * (uninitialised work buffer) and are multiplied by d == 0 later.  We zero them. */
            memset(y, 0, SZ_Q8_K);
            continue;
        }
        const float iscale = -127.f / max;
        for (int j = 0; j < QKK; j++) {
            const int v = nearest_int(iscale * x[j]);
            qs[j] = (int8_t) (v < 127 ? v : 127);
        }
        for (int j = 0; j < QKK / 16; j++) {
            int sum = 0;
            for (int ii = 0; ii < 16; ii++) sum += qs[j * 16 + ii];
            const int16_t s16 = (int16_t) sum;
            memcpy(y + 4 + QKK + 2 * j, &s16, 2);
        }
        const float d = 1 / iscale;
        memcpy(y, &d, 4);
    }
}

/* quantize_row_q4_0_ref, ggml/src/ggml-quants.c:71-106 */
void orc_quantize_row_q4_0(const float * x, void * vy, int64_t k) {
    uint8_t * y = (uint8_t *) vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++, x += QK, y += SZ_Q4_0) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < QK; j++) {
            const float v = x[j];
            if (amax < fabsf(v)) { amax = fabsf(v); max = v; }
        }
        const float d  = max / -8;
        const float id = d ? 1.0f / d : 0.0f;
        st_u16(y, orc_fp32_to_fp16(d));
        for (int j = 0; j < QK / 2; j++) {
            const float x0 = x[j] * id;
            const float x1 = x[QK / 2 + j] * id;
            const int8_t  t0 = (int8_t) (x0 + 8.5f);
            const int8_t  t1 = (int8_t) (x1 + 8.5f);
            const uint8_t q0 = (uint8_t) (t0 < 15 ? t0 : 15);
            const uint8_t q1 = (uint8_t) (t1 < 15 ? t1 : 15);
            y[2 + j] = (uint8_t) (q0 | (q1 << 4));
        }
    }
}

/* ------------------------------------------------------------------------------------------------
 * dequantization
 * ---------------------------------------------------------------------------------------------- */

/* get_scale_min_k4, ggml/src/ggml-quants.c:818-825 : 8 six-bit (scale, min) pairs in 12 bytes */
static inline void scale_min_k4(int j, const uint8_t * q, uint8_t * sc, uint8_t * mn) {
    if (j < 4) {
        *sc = q[j] & 63;
        *mn = q[j + 4] & 63;
    } else {
        *sc = (uint8_t) ((q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4));
        *mn = (uint8_t) ((q[j + 4] >> 4) | ((q[j] >> 6) << 4));
    }
}

/* one q6_K element's signed 6-bit value; element e of the superblock.
 * layout per dequantize_row_q6_K, ggml/src/ggml-quants.c:1877-1906 */
static inline int q6k_value(const uint8_t * blk, int e) {
    const int half = e >> 7;             /* which 128-element half          */
    const int r    = e & 127;
    const int quad = r >> 5;             /* 0..3 : which 32-run in the half  */
    const int l    = r & 31;
    const uint8_t * ql = blk + half * 64;
    const uint8_t * qh = blk + 128 + half * 32;
    const uint8_t lo = (quad & 1) ? ql[l + 32] : ql[l];
    const int nib  = (quad & 2) ? (lo >> 4) : (lo & 0xF);
    const int hi   = (qh[l] >> (2 * quad)) & 3;
    return (nib | (hi << 4)) - 32;
}

int orc_dequantize_row(int type, const void * vx, float * y, int64_t k) {
    const uint8_t * x = (const uint8_t *) vx;
    switch (type) {
        case ORC_TYPE_Q4_0: {   /* dequantize_row_q4_0, ggml-quants.c:397-415 */
            for (int64_t i = 0; i < k / QK; i++, x += SZ_Q4_0, y += QK) {
                const float d = orc_fp16_to_fp32(ld_u16(x));
                for (int j = 0; j < QK / 2; j++) {
                    y[j]          = (float) ((x[2 + j] & 0x0F) - 8) * d;
                    y[j + QK / 2] = (float) ((x[2 + j] >> 4) - 8) * d;
                }
            }
            return 0;
        }
        case ORC_TYPE_Q8_0: {   /* dequantize_row_q8_0, ggml-quants.c:491-505 */
            for (int64_t i = 0; i < k / QK; i++, x += SZ_Q8_0, y += QK) {
                const float d = orc_fp16_to_fp32(ld_u16(x));
                const int8_t * qs = (const int8_t *) (x + 2);
                for (int j = 0; j < QK; j++) y[j] = (float) qs[j] * d;
            }
            return 0;
        }
        case ORC_TYPE_Q8_1: {   /* d * q (block_q8_1, ggml-common.h:248-259) */
            for (int64_t i = 0; i < k / QK; i++, x += SZ_Q8_1, y += QK) {
                const float d = orc_fp16_to_fp32(ld_u16(x));
                const int8_t * qs = (const int8_t *) (x + 4);
                for (int j = 0; j < QK; j++) y[j] = (float) qs[j] * d;
            }
            return 0;
        }
        case ORC_TYPE_Q4_K: {   /* dequantize_row_q4_K, ggml-quants.c:1467-1489 */
            for (int64_t i = 0; i < k / QKK; i++, x += SZ_Q4_K) {
                const float d   = orc_fp16_to_fp32(ld_u16(x));
                const float min = orc_fp16_to_fp32(ld_u16(x + 2));
                const uint8_t * scales = x + 4;
                const uint8_t * q      = x + 16;
                int is = 0;
                for (int j = 0; j < QKK; j += 64, q += 32, is += 2) {
                    uint8_t sc, m;
                    scale_min_k4(is + 0, scales, &sc, &m);
                    const float d1 = d * sc, m1 = min * m;
                    scale_min_k4(is + 1, scales, &sc, &m);
                    const float d2 = d * sc, m2 = min * m;
                    for (int l = 0; l < 32; l++) *y++ = d1 * (float) (q[l] & 0xF) - m1;
                    for (int l = 0; l < 32; l++) *y++ = d2 * (float) (q[l] >> 4) - m2;
                }
            }
            return 0;
        }
        case ORC_TYPE_Q6_K: {   /* dequantize_row_q6_K, ggml-quants.c:1877-1906 */
            for (int64_t i = 0; i < k / QKK; i++, x += SZ_Q6_K, y += QKK) {
                const float d = orc_fp16_to_fp32(ld_u16(x + 208));
                const int8_t * sc = (const int8_t *) (x + 192);
                for (int e = 0; e < QKK; e++) {
                    /* reference evaluates d * sc * q left to right in float */
                    y[e] = d * (float) sc[e >> 4] * (float) q6k_value(x, e);
                }
            }
            return 0;
        }
        default: return -1;
    }
}

/* ------------------------------------------------------------------------------------------------
 * vec_dot (generic C versions the reference falls back to; the SIMD variants compute the same
 * integer sums, differing only in float summation order)
 * ---------------------------------------------------------------------------------------------- */

int orc_vec_dot_type(int wtype) {
    switch (wtype) {   /* type_traits_cpu, ggml-cpu/ggml-cpu.c:226-235, 258-267, 297-322 */
        case ORC_TYPE_Q4_0: case ORC_TYPE_Q8_0: return ORC_TYPE_Q8_0;
        case ORC_TYPE_Q4_K: case ORC_TYPE_Q6_K: return ORC_TYPE_Q8_K;
        default: return -1;
    }
}

/* ggml_vec_dot_q4_0_q8_0_generic, ggml-cpu/quants.c:174-208 */
static float vec_dot_q4_0_q8_0(int64_t k, const uint8_t * x, const uint8_t * y) {
    float sumf = 0;
    for (int64_t ib = 0; ib < k / QK; ib++, x += SZ_Q4_0, y += SZ_Q8_0) {
        const int8_t * q8 = (const int8_t *) (y + 2);
        int sumi0 = 0, sumi1 = 0;
        for (int j = 0; j < QK / 2; j++) {
            sumi0 += ((x[2 + j] & 0x0F) - 8) * q8[j];
            sumi1 += ((x[2 + j] >> 4) - 8) * q8[j + QK / 2];
        }
        const int sumi = sumi0 + sumi1;
        sumf += (float) sumi * orc_fp16_to_fp32(ld_u16(x)) * orc_fp16_to_fp32(ld_u16(y));
    }
    return sumf;
}

/* ggml_vec_dot_q8_0_q8_0_generic, ggml-cpu/quants.c:400-428 */
static float vec_dot_q8_0_q8_0(int64_t k, const uint8_t * x, const uint8_t * y) {
    float sumf = 0;
    for (int64_t ib = 0; ib < k / QK; ib++, x += SZ_Q8_0, y += SZ_Q8_0) {
        const int8_t * a = (const int8_t *) (x + 2);
        const int8_t * b = (const int8_t *) (y + 2);
        int sumi = 0;
        for (int j = 0; j < QK; j++) sumi += a[j] * b[j];
        sumf += (float) sumi * (orc_fp16_to_fp32(ld_u16(x)) * orc_fp16_to_fp32(ld_u16(y)));
    }
    return sumf;
}

/* ggml_vec_dot_q4_K_q8_K_generic, ggml-cpu/quants.c:645-718.  The reference keeps 8 float lanes
 * (sums[l], l = element index mod 8) and adds them at the end; reproduced so float rounding matches. */
static float vec_dot_q4_K_q8_K(int64_t k, const uint8_t * x, const uint8_t * y) {
    float sums[8] = {0};
    float sumf = 0;
    for (int64_t i = 0; i < k / QKK; i++, x += SZ_Q4_K, y += SZ_Q8_K) {
        float yd; memcpy(&yd, y, 4);
        const int8_t * q8 = (const int8_t *) (y + 4);
        int16_t bsums[16]; memcpy(bsums, y + 4 + QKK, 32);
        const uint8_t * q4 = x + 16;
        int32_t aux32[8] = {0};
        int sumi = 0;
        for (int j = 0; j < 8; j++) {               /* 8 sub-blocks of 32 */
            uint8_t sc, mn;
            scale_min_k4(j, x + 4, &sc, &mn);
            sumi += (bsums[2 * j] + bsums[2 * j + 1]) * mn;
            const uint8_t * qb = q4 + (j >> 1) * 32;
            for (int l = 0; l < 32; l++) {
                const int a = (j & 1) ? (qb[l] >> 4) : (qb[l] & 0xF);
                aux32[l & 7] += (int32_t) sc * (int16_t) (q8[j * 32 + l] * a);
            }
        }
        const float d = orc_fp16_to_fp32(ld_u16(x)) * yd;
        for (int l = 0; l < 8; l++) sums[l] += d * (float) aux32[l];
        const float dmin = orc_fp16_to_fp32(ld_u16(x + 2)) * yd;
        sumf -= dmin * (float) sumi;
    }
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}

/* ggml_vec_dot_q6_K_q8_K_generic, ggml-cpu/quants.c:800-853 */
static float vec_dot_q6_K_q8_K(int64_t k, const uint8_t * x, const uint8_t * y) {
    float sums[8] = {0};
    float sumf = 0;
    for (int64_t i = 0; i < k / QKK; i++, x += SZ_Q6_K, y += SZ_Q8_K) {
        float yd; memcpy(&yd, y, 4);
        const int8_t * q8 = (const int8_t *) (y + 4);
        const int8_t * sc = (const int8_t *) (x + 192);
        int32_t aux32[8] = {0};
        for (int e = 0; e < QKK; e++) {
            aux32[e & 7] += (int32_t) sc[e >> 4] * (int16_t) (q8[e] * q6k_value(x, e));
        }
        const float d = orc_fp16_to_fp32(ld_u16(x + 208)) * yd;
        for (int l = 0; l < 8; l++) sums[l] += d * (float) aux32[l];
    }
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}

float orc_vec_dot(int wtype, int64_t k, const void * w_row, const void * act) {
    const uint8_t * x = (const uint8_t *) w_row;
    const uint8_t * y = (const uint8_t *) act;
    switch (wtype) {
        case ORC_TYPE_Q4_0: return vec_dot_q4_0_q8_0(k, x, y);
        case ORC_TYPE_Q8_0: return vec_dot_q8_0_q8_0(k, x, y);
        case ORC_TYPE_Q4_K: return vec_dot_q4_K_q8_K(k, x, y);
        case ORC_TYPE_Q6_K: return vec_dot_q6_K_q8_K(k, x, y);
        default: return NAN;
    }
}

/* ------------------------------------------------------------------------------------------------
 * a tiny static-partition thread helper (the reference uses a work-stealing chunk queue,
 * ggml-cpu.c:1378-1438; results do not depend on the partition because every output element is
 * computed by exactly one vec_dot call)
 * ---------------------------------------------------------------------------------------------- */
typedef void (*orc_job_fn)(void * ctx, int64_t begin, int64_t end);
struct orc_job { orc_job_fn fn; void * ctx; int64_t begin, end; };
static void * orc_job_trampoline(void * p) {
    struct orc_job * j = (struct orc_job *) p;
    j->fn(j->ctx, j->begin, j->end);
    return NULL;
}
static void orc_parallel_for(int64_t n, int n_threads, orc_job_fn fn, void * ctx) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    if ((int64_t) n_threads > n) n_threads = (int) (n > 0 ? n : 1);
    if (n_threads == 1) { fn(ctx, 0, n); return; }
    pthread_t      th[256];
    struct orc_job jobs[256];
    for (int t = 0; t < n_threads; t++) {
        jobs[t].fn = fn; jobs[t].ctx = ctx;
        jobs[t].begin = n * t / n_threads;
        jobs[t].end   = n * (t + 1) / n_threads;
        pthread_create(&th[t], NULL, orc_job_trampoline, &jobs[t]);
    }
    for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
}

static void quantize_act(int atype, const float * x, void * y, int64_t k) {
    if (atype == ORC_TYPE_Q8_0) orc_quantize_row_q8_0(x, y, k);
    else                        orc_quantize_row_q8_K(x, y, k);
}

/* ------------------------------------------------------------------------------------------------
 * MUL_MAT, ggml_compute_forward_mul_mat (ggml-cpu/ggml-cpu.c:1241-1439) + _one_chunk (:1151-1239)
 * ---------------------------------------------------------------------------------------------- */
struct mm_ctx {
    int wtype; const uint8_t * W; const uint8_t * Xq; float * Y;
    int64_t m, n, k, ne02, ne03, ne12, ne13;
    size_t w_row, a_row;
};

static void mm_rows(void * vctx, int64_t begin, int64_t end) {
    const struct mm_ctx * c = (const struct mm_ctx *) vctx;
    const int64_t r2 = c->ne12 / c->ne02, r3 = c->ne13 / c->ne03;
    for (int64_t i13 = 0; i13 < c->ne13; i13++)
    for (int64_t i12 = 0; i12 < c->ne12; i12++) {
        const int64_t i03 = i13 / r3, i02 = i12 / r2;     /* broadcast src0 (ggml-cpu.c:1205-1207) */
        const uint8_t * w = c->W + (size_t) (i02 + i03 * c->ne02) * c->m * c->w_row;
        for (int64_t i11 = 0; i11 < c->n; i11++) {
            const uint8_t * a = c->Xq + (size_t) (i11 + i12 * c->n + i13 * c->ne12 * c->n) * c->a_row;
            float * y = c->Y + (size_t) (i11 + i12 * c->n + i13 * c->ne12 * c->n) * c->m;
            for (int64_t i0 = begin; i0 < end; i0++) {
                y[i0] = orc_vec_dot(c->wtype, c->k, w + (size_t) i0 * c->w_row, a);
            }
        }
    }
}

int orc_mul_mat(int wtype, const void * W, const float * X, float * Y,
                int64_t m, int64_t n, int64_t k,
                int64_t ne02, int64_t ne03, int64_t ne12, int64_t ne13, int n_threads) {
    const int atype = orc_vec_dot_type(wtype);
    if (atype < 0 || k <= 0 || k % orc_blck_size(wtype) != 0) return -1;
    if (ne02 <= 0 || ne03 <= 0 || ne12 % ne02 != 0 || ne13 % ne03 != 0) return -2;
    const size_t a_row = orc_row_size(atype, k);
    const int64_t ncols = n * ne12 * ne13;
    uint8_t * Xq = (uint8_t *) malloc(a_row * (size_t) (ncols > 0 ? ncols : 1));
    if (!Xq) return -3;
    /* phase 1: quantize every src1 row into wdata (ggml-cpu.c:1309-1344) */
    for (int64_t c = 0; c < ncols; c++) quantize_act(atype, X + (size_t) c * k, Xq + (size_t) c * a_row, k);
    /* phase 2: dot products (ggml-cpu.c:1378-1438) */
    struct mm_ctx ctx = { wtype, (const uint8_t *) W, Xq, Y, m, n, k, ne02, ne03, ne12, ne13,
                          orc_row_size(wtype, k), a_row };
    orc_parallel_for(m, n_threads, mm_rows, &ctx);
    free(Xq);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * MUL_MAT_ID, ggml_compute_forward_mul_mat_id (ggml-cpu/ggml-cpu.c:1521-1694) + _one_chunk
 * (:1443-1511).  The reference buckets (slot, token) pairs per expert and runs a per-expert matmul;
 * every output column is still one expert's rows dotted with one quantized activation row, so the
 * restatement walks (token, slot) directly.
 * ---------------------------------------------------------------------------------------------- */
struct mmid_ctx {
    int wtype; const uint8_t * as; const uint8_t * bq; const int32_t * ids; float * c;
    int64_t m, k, n_expert, n_used, n_tokens, b_ne1, ids_stride;
    size_t w_row, a_row;
};

static void mmid_rows(void * vctx, int64_t begin, int64_t end) {
    const struct mmid_ctx * c = (const struct mmid_ctx *) vctx;
    for (int64_t t = 0; t < c->n_tokens; t++)
    for (int64_t e = 0; e < c->n_used; e++) {
        const int32_t ex = c->ids[t * c->ids_stride + e];
        if (ex < 0 || ex >= c->n_expert) continue;          /* reference asserts (ggml-cpu.c:1616) */
        const uint8_t * w = c->as + (size_t) ex * c->m * c->w_row;
        const int64_t i11 = e % c->b_ne1;                   /* slot broadcast (ggml-cpu.c:1486) */
        const uint8_t * a = c->bq + (size_t) (i11 + t * c->b_ne1) * c->a_row;
        float * y = c->c + (size_t) (e + t * c->n_used) * c->m;
        for (int64_t i0 = begin; i0 < end; i0++) {
            y[i0] = orc_vec_dot(c->wtype, c->k, w + (size_t) i0 * c->w_row, a);
        }
    }
}

int orc_mul_mat_id(int wtype, const void * as, const float * b, const int32_t * ids, float * cdst,
                   int64_t m, int64_t k, int64_t n_expert, int64_t n_used, int64_t n_tokens,
                   int64_t b_ne1, int64_t ids_stride, int n_threads) {
    const int atype = orc_vec_dot_type(wtype);
    if (atype < 0 || k <= 0 || k % orc_blck_size(wtype) != 0) return -1;
    if (b_ne1 != 1 && b_ne1 != n_used) return -2;
    const size_t a_row = orc_row_size(atype, k);
    const int64_t ncols = b_ne1 * n_tokens;
    uint8_t * bq = (uint8_t *) malloc(a_row * (size_t) (ncols > 0 ? ncols : 1));
    if (!bq) return -3;
    for (int64_t col = 0; col < ncols; col++) quantize_act(atype, b + (size_t) col * k, bq + (size_t) col * a_row, k);
    struct mmid_ctx ctx = { wtype, (const uint8_t *) as, bq, ids, cdst, m, k, n_expert, n_used, n_tokens,
                            b_ne1, ids_stride, orc_row_size(wtype, k), a_row };
    orc_parallel_for(m, n_threads, mmid_rows, &ctx);
    free(bq);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * f64 ground truth and NMSE
 * ---------------------------------------------------------------------------------------------- */
struct f64_ctx { int wtype; const uint8_t * W; const float * X; double * Y; int64_t m, n, k; size_t w_row; };

static void f64_rows(void * vctx, int64_t begin, int64_t end) {
    const struct f64_ctx * c = (const struct f64_ctx *) vctx;
    float * wrow = (float *) malloc(sizeof(float) * (size_t) c->k);
    for (int64_t i0 = begin; i0 < end; i0++) {
        orc_dequantize_row(c->wtype, c->W + (size_t) i0 * c->w_row, wrow, c->k);
        for (int64_t j = 0; j < c->n; j++) {
            const float * x = c->X + (size_t) j * c->k;
            double acc = 0.0;
            for (int64_t kk = 0; kk < c->k; kk++) acc += (double) wrow[kk] * (double) x[kk];
            c->Y[(size_t) j * c->m + i0] = acc;
        }
    }
    free(wrow);
}

int orc_mul_mat_f64(int wtype, const void * W, const float * X, double * Y,
                    int64_t m, int64_t n, int64_t k, int n_threads) {
    if (orc_vec_dot_type(wtype) < 0 || k % orc_blck_size(wtype) != 0) return -1;
    struct f64_ctx ctx = { wtype, (const uint8_t *) W, X, Y, m, n, k, orc_row_size(wtype, k) };
    orc_parallel_for(m, n_threads, f64_rows, &ctx);
    return 0;
}

/* nmse, tests/test-backend-ops.cpp:278-291 (float differences, double accumulation) */
double orc_nmse(const float * a, const float * b, size_t n) {
    double mse_a_b = 0.0, mse_a_0 = 0.0;
    for (size_t i = 0; i < n; i++) {
        const float ai = a[i], bi = b[i];
        mse_a_b += (ai - bi) * (ai - bi);
        mse_a_0 += ai * ai;
    }
    return mse_a_b / mse_a_0;
}
