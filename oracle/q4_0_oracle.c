/*
 * q4_0_oracle.c -- CPU restatement of ggml's Q4_0 x F32 mul_mat path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and there only as the checker (or the timed CPU arm), never as the thing shipped.
 *
 * Parity status: PINNED.  Every function below is checked (tests/test_oracle.py) against
 *   (1) the reference itself, compiled unmodified into oracle/_ref/libggml_ref.so by
 *       oracle/Makefile (when /root/reference is present), and
 *   (2) the known answers SURVEY.md 8(c) recorded from the reference (FNV hashes of the
 *       test-quantize-fns generator, first-block bytes, the 4470.44141 dot, sum S_b).
 *
 * Each function cites the reference lines it restates (paths under /root/reference).
 * Plain scalar C, built with -ffp-contract=off so no multiply-add is fused that the
 * reference does not fuse.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <stdlib.h>

#define QK 32

/* ggml.c:870-875 (block_q4_0) and ggml.c:902-907 (block_q8_0); packed, 18 and 34 bytes */
#pragma pack(push, 1)
typedef struct { uint16_t d; uint8_t qs[QK / 2]; } oq4_0;
typedef struct { uint16_t d; int8_t  qs[QK];     } oq8_0;
#pragma pack(pop)

_Static_assert(sizeof(oq4_0) == 18, "q4_0 block is 18 bytes");
_Static_assert(sizeof(oq8_0) == 34, "q8_0 block is 34 bytes");

/* ---- fp16 <-> fp32 -------------------------------------------------------------------
 * ggml.c:284-292: on x86 GGML_FP32_TO_FP16 is _cvtss_sh(x, 0) (round-to-nearest-even) and
 * GGML_FP16_TO_FP32 is a table of _cvtsh_ss; both are the IEEE binary16 conversions, restated
 * here with integer arithmetic so the oracle does not depend on F16C. */
float oracle_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    const uint32_t exp  = (h >> 10) & 0x1Fu;
    uint32_t man        = h & 0x3FFu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else { /* subnormal half -> normal float */
            int e = -1;
            do { man <<= 1; e++; } while ((man & 0x400u) == 0);
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3FFu) << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7F800000u | (man << 13);
    } else {
        bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
    }
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

uint16_t oracle_fp32_to_fp16(float f) {
    uint32_t x;
    memcpy(&x, &f, 4);
    const uint16_t sign = (uint16_t)((x >> 16) & 0x8000u);
    x &= 0x7FFFFFFFu;
    if (x >= 0x7F800000u) { /* inf / nan (F16C keeps the top mantissa bits and quiets) */
        return (uint16_t)(sign | 0x7C00u | (x > 0x7F800000u ? (0x200u | ((x >> 13) & 0x3FFu)) : 0));
    }
    if (x >= 0x477FF000u) { /* rounds to >= 65520 -> inf */
        return (uint16_t)(sign | 0x7C00u);
    }
    if (x < 0x33000001u) { /* <= 2^-25 -> rounds to zero (ties-to-even at exactly 2^-25) */
        return sign;
    }
    uint32_t e = x >> 23;
    uint32_t m = (x & 0x7FFFFFu) | 0x800000u;
    uint32_t shift;
    uint32_t he;
    if (e < 113) { /* subnormal half: value = m * 2^(e-150); want units of 2^-24 */
        shift = 126 - e; /* 14..24 */
        he = 0;
    } else {
        shift = 13;
        he = e - 112;
    }
    uint32_t q   = m >> shift;
    uint32_t rem = m & ((1u << shift) - 1u);
    uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1u))) q++;
    /* he==0: q is the subnormal mantissa (may carry into the exponent, which is correct);
       he>0 : q has the implicit bit at 0x400, so add (he-1)<<10 */
    uint32_t h = (he == 0) ? q : (((he - 1) << 10) + q);
    return (uint16_t)(sign | h);
}

/* ---- weight quantizer: ggml.c:918-953 quantize_row_q4_0_reference ------------------------ */
void oracle_quantize_row_q4_0(const float *x, void *vy, int64_t k) {
    oq4_0 *y = (oq4_0 *)vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < QK; j++) {
            const float v = x[i * QK + j];
            if (amax < fabsf(v)) { amax = fabsf(v); max = v; }
        }
        const float d  = max / -8;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = oracle_fp32_to_fp16(d);
        for (int j = 0; j < QK / 2; ++j) {
            const float x0 = x[i * QK + 0 + j] * id;
            const float x1 = x[i * QK + QK / 2 + j] * id;
            const int8_t t0 = (int8_t)(x0 + 8.5f);
            const int8_t t1 = (int8_t)(x1 + 8.5f);
            const uint8_t xi0 = (uint8_t)(t0 < 15 ? t0 : 15);
            const uint8_t xi1 = (uint8_t)(t1 < 15 ? t1 : 15);
            y[i].qs[j] = (uint8_t)(xi0 | (xi1 << 4));
        }
    }
}

/* ggml.c:1500-1518 dequantize_row_q4_0 (used only by the round-trip test, test-quantize-fns.cpp:129) */
void oracle_dequantize_row_q4_0(const void *vx, float *y, int64_t k) {
    const oq4_0 *x = (const oq4_0 *)vx;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++) {
        const float d = oracle_fp16_to_fp32(x[i].d);
        for (int j = 0; j < QK / 2; ++j) {
            const int x0 = (x[i].qs[j] & 0x0F) - 8;
            const int x1 = (x[i].qs[j] >> 4) - 8;
            y[i * QK + j]          = x0 * d;
            y[i * QK + j + QK / 2] = x1 * d;
        }
    }
}

/* ---- activation quantizer -------------------------------------------------------------
 * The contract on x86 is the AVX2 branch, ggml.c:1192-1250, NOT the scalar
 * quantize_row_q8_0_reference (ggml.c:1097-1120): d = amax/127, id = 127/amax (a second
 * division, not 1/d), round-to-nearest-EVEN (_mm256_round_ps NEAREST), cvtps_epi32, then
 * saturating packs to int8. */
static inline int8_t sat_i8_from_rounded(float r) {
    /* _mm256_cvtps_epi32 gives INT_MIN for NaN/out-of-range; packs_epi32/packs_epi16 saturate */
    int32_t i;
    if (!(r >= -2147483648.0f && r < 2147483648.0f)) i = INT32_MIN; else i = (int32_t)r;
    if (i > 32767) i = 32767;
    if (i < -32768) i = -32768;
    if (i > 127) i = 127;
    if (i < -128) i = -128;
    return (int8_t)i;
}

void oracle_quantize_row_q8_0(const float *x, void *vy, int64_t k) {
    oq8_0 *y = (oq8_0 *)vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++) {
        float amax = 0.0f;
        for (int j = 0; j < QK; j++) {
            const float a = fabsf(x[i * QK + j]);
            if (a > amax) amax = a; /* maxps on finite values */
        }
        const float d  = amax / 127.f;
        y[i].d = oracle_fp32_to_fp16(d);
        const float id = (amax != 0.0f) ? 127.f / amax : 0.0f;
        for (int j = 0; j < QK; j++) {
            const float v = x[i * QK + j] * id;
            y[i].qs[j] = sat_i8_from_rounded(nearbyintf(v)); /* default mode = ties-to-even */
        }
    }
}

/* the scalar variant the reference keeps for model-file determinism (ggml.c:1097-1120);
 * exposed so a test can show it is NOT what the op calls on x86 (ties away, id = 1/d) */
void oracle_quantize_row_q8_0_scalar_reference(const float *x, void *vy, int64_t k) {
    oq8_0 *y = (oq8_0 *)vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++) {
        float amax = 0.0f;
        for (int j = 0; j < QK; j++) {
            const float a = fabsf(x[i * QK + j]);
            amax = amax > a ? amax : a;
        }
        const float d  = amax / ((1 << 7) - 1);
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = oracle_fp32_to_fp16(d);
        for (int j = 0; j < QK; ++j) {
            y[i].qs[j] = (int8_t)roundf(x[i * QK + j] * id);
        }
    }
}

/* ---- per-block integer dots: ggml.c:2593-2600 (the scalar definition of S_b) ----------- */
void oracle_block_dots_q4_0_q8_0(int64_t k, const void *vx, const void *vy, int32_t *S) {
    const oq4_0 *x = (const oq4_0 *)vx;
    const oq8_0 *y = (const oq8_0 *)vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; i++) {
        int sumi = 0;
        for (int j = 0; j < QK / 2; ++j) {
            const int v0 = (x[i].qs[j] & 0x0F) - 8;
            const int v1 = (x[i].qs[j] >> 4) - 8;
            sumi += (v0 * y[i].qs[j]) + (v1 * y[i].qs[j + QK / 2]);
        }
        S[i] = sumi;
    }
}

/* ---- dot kernel, scalar order: ggml.c:2588-2606 ---------------------------------------- */
float oracle_vec_dot_q4_0_q8_0_scalar(int64_t k, const void *vx, const void *vy) {
    const oq4_0 *x = (const oq4_0 *)vx;
    const oq8_0 *y = (const oq8_0 *)vy;
    const int64_t nb = k / QK;
    float sumf = 0.0f;
    for (int64_t i = 0; i < nb; i++) {
        int sumi = 0;
        for (int j = 0; j < QK / 2; ++j) {
            const int v0 = (x[i].qs[j] & 0x0F) - 8;
            const int v1 = (x[i].qs[j] >> 4) - 8;
            sumi += (v0 * y[i].qs[j]) + (v1 * y[i].qs[j + QK / 2]);
        }
        sumf += sumi * oracle_fp16_to_fp32(x[i].d) * oracle_fp16_to_fp32(y[i].d);
    }
    return sumf;
}

/* ---- dot kernel, AVX2 order: ggml.c:2412-2435 with helpers :632-672 and hsum :591-597 ----
 * The AVX2 code keeps 8 fp32 lanes.  After bytes_from_nibbles_32 the 32 weights of a block sit
 * in natural element order (bytes 0..15 = low nibbles = elements 0..15, bytes 16..31 = high
 * nibbles = elements 16..31); dpbusd / maddubs+madd sums 4 consecutive byte products per 32-bit
 * lane, so lane l holds sum_{e=4l..4l+3} (q_e-8)*y_e.  Each lane is converted to float (exact),
 * then acc[l] = fma(d, q[l], acc[l]).  hsum_float_8: r[i] = acc[i+4] + acc[i]; r[0]+=r[2],
 * r[1]+=r[3]; r[0]+=r[1].  Restated lane by lane with fmaf => bit-identical to the reference on
 * AVX2 hosts (checked in tests/test_oracle.py against oracle/_ref). */
float oracle_vec_dot_q4_0_q8_0_avx2_order(int64_t k, const void *vx, const void *vy) {
    const oq4_0 *x = (const oq4_0 *)vx;
    const oq8_0 *y = (const oq8_0 *)vy;
    const int64_t nb = k / QK;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int64_t i = 0; i < nb; i++) {
        const float d = oracle_fp16_to_fp32(x[i].d) * oracle_fp16_to_fp32(y[i].d);
        for (int l = 0; l < 8; l++) {
            int s = 0;
            for (int e = 4 * l; e < 4 * l + 4; e++) {
                const int q = (e < 16) ? (x[i].qs[e] & 0x0F) : (x[i].qs[e - 16] >> 4);
                s += (q - 8) * y[i].qs[e];
            }
            acc[l] = fmaf(d, (float)s, acc[l]);
        }
    }
    float r0 = acc[4] + acc[0], r1 = acc[5] + acc[1], r2 = acc[6] + acc[2], r3 = acc[7] + acc[3];
    r0 = r0 + r2;
    r1 = r1 + r3;
    return r0 + r1;
}

/* ---- the operator: ggml.c:11226-11424 (INIT :11353-11367, COMPUTE :11373-11411) -----------
 * W: M rows of K/32 q4_0 blocks, row stride nb01 bytes.  X: n rows of K floats, row stride
 * nb11 bytes.  Y: n rows of M floats, row stride nb1 bytes (dst[ic*ne0 + i0]).  wdata: scratch
 * of n*K/32 q8_0 blocks.  order: 0 = scalar dot, 1 = AVX2-order dot.  Single-threaded: the row
 * partition over threads (:11376-11383) does not change any value. */
void oracle_mul_mat_q4_0_f32(int64_t M, int64_t K, int64_t n,
                             const void *W, size_t nb01,
                             const float *X, size_t nb11,
                             float *Y, size_t nb1,
                             void *wdata, int order) {
    const size_t row_size = (size_t)(K / QK) * sizeof(oq8_0);
    for (int64_t i11 = 0; i11 < n; ++i11) {
        oracle_quantize_row_q8_0((const float *)((const char *)X + i11 * nb11),
                                 (char *)wdata + i11 * row_size, K);
    }
    for (int64_t ir = 0; ir < M; ++ir) {
        const void *src0_row = (const char *)W + ir * nb01;
        for (int64_t ic = 0; ic < n; ++ic) {
            const void *col = (const char *)wdata + ic * row_size;
            float *out = (float *)((char *)Y + ic * nb1) + ir;
            *out = order ? oracle_vec_dot_q4_0_q8_0_avx2_order(K, src0_row, col)
                         : oracle_vec_dot_q4_0_q8_0_scalar(K, src0_row, col);
        }
    }
}

/* ---- deterministic synthetic inputs: integer hash only, identical everywhere --
 * counter-based hash -> 24-bit uniform -> sum of 4 (approx. normal, unit variance), * scale;
 * every `outlier_every`-th value (by hash) is multiplied by 8 when outlier_every > 0. */
static inline uint32_t mix32(uint32_t z) {
    z ^= z >> 16; z *= 0x7feb352dU; z ^= z >> 15; z *= 0x846ca68bU; z ^= z >> 16;
    return z;
}

void oracle_fill_synthetic(float *dst, int64_t count, uint32_t seed, float scale, int outlier_every) {
    for (int64_t i = 0; i < count; i++) {
        const uint32_t base = (uint32_t)(i * 4) ^ (uint32_t)((uint64_t)i >> 30);
        int32_t acc = 0;
        for (uint32_t t = 0; t < 4; t++) {
            const uint32_t h = mix32(mix32(base + t + seed) ^ seed);
            acc += (int32_t)(h >> 8) - (1 << 23); /* uniform in [-2^23, 2^23) */
        }
        /* each term has std 2^24/sqrt(12); sum of 4 -> std 2^24/sqrt(3) */
        float g = (float)acc * (1.7320508f / 16777216.0f);
        if (outlier_every > 0 && (mix32((uint32_t)i ^ (seed * 0x9E3779B9u)) % (uint32_t)outlier_every) == 0) g *= 8.0f;
        dst[i] = g * scale;
    }
}

/* FNV-1a 64 over bytes: the hash SURVEY 8(c) used for its known answers */
uint64_t oracle_fnv1a64(const void *p, size_t nbytes) {
    const uint8_t *b = (const uint8_t *)p;
    uint64_t h = 0xcbf29ce484222325ULL;
    for (size_t i = 0; i < nbytes; i++) { h ^= b[i]; h *= 0x100000001b3ULL; }
    return h;
}

/* the reference test generator, tests/test-quantize-fns.cpp:26-30: dst[i] = 0.1 + 2*cosf(i + offset)
 * (size_t + float -> float; 2*cosf -> float; 0.1 is a double, so the sum is rounded once to float).
 * libm's cosf may differ between C libraries in the last bit, so the outputs recorded from THIS
 * container are committed under tests/golden/ and tests compare against those, not a re-run. */
void oracle_fill_cos(float *dst, int64_t count, float offset) {
    for (int64_t i = 0; i < count; i++) {
        dst[i] = 0.1 + 2 * cosf((size_t)i + offset);
    }
}
