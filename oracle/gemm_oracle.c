/*
 * gemm_oracle.c — CPU restatement of the reference's dense-contraction path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked into, imported by or called from
 * the product (tilelang_mesh_b200/, libtlb200.so).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may use it, and only as the checker.
 *
 * What it restates (Tilelang-Mesh snapshot, TileLang 0.1.7 lineage):
 *   - T.gemm semantics: C[M,N] (+)= op(A) * op(B), K_A == K_B,
 *     transpose flags as tilelang/language/gemm.py:78-81,105-116.
 *   - accumulation: fp32 accumulator cleared once (clear_accum at k == 0:
 *     tilelang/intrinsics/tcgen05_macro_generator.py:313, src/tl_templates/cuda/gemm_sm100.h:391-403),
 *     products of 16-/8-bit operands are exact in fp32 (tensor-core semantics), summed in
 *     ascending k (the hardware's intra-MMA order is unspecified, hence tolerance parity on GPU).
 *   - output cast: one round-to-nearest-even from fp32 (generated epilogue uses
 *     __float22bfloat162_rn / __float22half2_rn; fp8 via cute::float_e4m3_t(float),
 *     src/tl_templates/cuda/common.h:505-523 = saturating RNE).
 *   - product_mode 1 restates the reference's CPU ("c") target instead: its codegen emits
 *     `C_local = C_local + (float)(A_local * B_local)` with the product evaluated in the
 *     16-bit type (half.hpp, HALF_ROUND_STYLE 1), see testing/python/cpu/test_tilelang_cpu_gemm.py:66-98.
 *     This mode exists so the encode/decode/loop code can be pinned BIT-EXACTLY against the
 *     golden vector produced by that reference program (tests/golden/cpu_c_target_matmul.npz).
 *
 * Parity status: decode/encode + loop structure pinned bit-exactly by that golden vector
 * (fp16); tensor-core mode pinned by tolerance tests only (the reference stores no golden
 * vectors for its GPU path — SURVEY.md §8(c)).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

enum { DT_F16 = 0, DT_BF16 = 1, DT_E4M3 = 2, DT_E5M2 = 3, DT_F32 = 4 };

static float bits_f32(uint32_t u) {
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static uint32_t f32_bits(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
}

/* ---- decode ------------------------------------------------------------------------- */
/* Generic small-float decode: 1 sign, `eb` exponent bits, `mb` mantissa bits, bias `bias`. */
static float decode_small(uint32_t v, int eb, int mb, int bias, int ieee_inf) {
  const uint32_t sign = (v >> (eb + mb)) & 1u;
  const uint32_t e = (v >> mb) & ((1u << eb) - 1u);
  const uint32_t m = v & ((1u << mb) - 1u);
  float r;
  if (e == 0) {
    r = ldexpf((float)m, 1 - bias - mb);
  } else if (ieee_inf && e == ((1u << eb) - 1u)) {
    r = m ? NAN : INFINITY;
  } else if (!ieee_inf && e == ((1u << eb) - 1u) && m == ((1u << mb) - 1u)) {
    r = NAN; /* e4m3fn: S.1111.111 is NaN, no infinities */
  } else {
    r = ldexpf((float)(m | (1u << mb)), (int)e - bias - mb);
  }
  return sign ? -r : r;
}

float oracle_decode(int dt, const void* p, long long i) {
  switch (dt) {
    case DT_F16: return decode_small(((const uint16_t*)p)[i], 5, 10, 15, 1);
    case DT_BF16: return bits_f32((uint32_t)((const uint16_t*)p)[i] << 16);
    case DT_E4M3: return decode_small(((const uint8_t*)p)[i], 4, 3, 7, 0);
    case DT_E5M2: return decode_small(((const uint8_t*)p)[i], 5, 2, 15, 1);
    default: return ((const float*)p)[i];
  }
}

/* ---- encode (round to nearest, ties to even) ---------------------------------------- */
/* Round |x| to a float with `mb` mantissa bits and minimum normal exponent `emin`
 * (gradual underflow below it); returns the rounded magnitude as a float. */
static float round_to_format(float ax, int mb, int emin) {
  if (ax == 0.f || isinf(ax) || isnan(ax)) return ax;
  int e;
  (void)frexpf(ax, &e); /* ax = f * 2^e, f in [0.5,1) -> exponent of leading bit = e-1 */
  int lead = e - 1;
  if (lead < emin) lead = emin;
  const float ulp = ldexpf(1.0f, lead - mb);
  /* nearbyintf honours the current rounding mode (default: ties-to-even). */
  return nearbyintf(ax / ulp) * ulp;
}

static uint32_t encode_small(float x, int eb, int mb, int bias, int ieee_inf, float max_finite, int saturate) {
  const uint32_t sign = (f32_bits(x) >> 31) << (eb + mb);
  if (isnan(x)) return sign | (ieee_inf ? ((((1u << eb) - 1u) << mb) | (1u << (mb - 1))) : ((1u << (eb + mb)) - 1u));
  float ax = fabsf(x);
  float r = isinf(ax) ? ax : round_to_format(ax, mb, 1 - bias);
  if (r > max_finite) {
    if (saturate) r = max_finite;
    else return sign | (((1u << eb) - 1u) << mb); /* +-inf */
  }
  if (r == 0.f) return sign;
  int e;
  (void)frexpf(r, &e);
  int lead = e - 1;
  if (lead < 1 - bias) { /* subnormal */
    const uint32_t m = (uint32_t)ldexpf(r, bias - 1 + mb);
    return sign | m;
  }
  const uint32_t m = (uint32_t)ldexpf(r, mb - lead) - (1u << mb);
  return sign | ((uint32_t)(lead + bias) << mb) | m;
}

void oracle_encode(int dt, void* p, long long i, float x) {
  switch (dt) {
    case DT_F16: ((uint16_t*)p)[i] = (uint16_t)encode_small(x, 5, 10, 15, 1, 65504.f, 0); break;
    case DT_BF16: {
      uint32_t u = f32_bits(x);
      if (isnan(x)) { ((uint16_t*)p)[i] = (uint16_t)((u >> 16) | 0x40u); break; }
      u += 0x7fffu + ((u >> 16) & 1u); /* RNE on the discarded 16 bits */
      ((uint16_t*)p)[i] = (uint16_t)(u >> 16);
      break;
    }
    case DT_E4M3: ((uint8_t*)p)[i] = (uint8_t)encode_small(x, 4, 3, 7, 0, 448.f, 1); break;
    case DT_E5M2: ((uint8_t*)p)[i] = (uint8_t)encode_small(x, 5, 2, 15, 1, 57344.f, 1); break;
    default: ((float*)p)[i] = x;
  }
}

/* Round a float to the operand type and back (product_mode 1). */
static float requantize(int dt, float x) {
  uint32_t buf = 0;
  oracle_encode(dt, &buf, 0, x);
  return oracle_decode(dt, &buf, 0);
}

/*
 * C = cast(relu?(alpha * sum_k op(A)[m,k] * op(B)[k,n]))
 *   trans_a == 0: A[m*lda + k]   trans_a == 1: A[k*lda + m]
 *   trans_b == 1: B[n*ldb + k]   trans_b == 0: B[k*ldb + n]
 * Returns 0, or -1 on bad arguments.
 */
int oracle_gemm(int m, int n, int k, int ab_dtype, int c_dtype, int trans_a, int trans_b, const void* A,
                long long lda, const void* B, long long ldb, void* C, long long ldc, float alpha, int relu,
                int product_mode) {
  if (m < 0 || n < 0 || k < 0 || ab_dtype < 0 || ab_dtype > 3 || c_dtype < 0 || c_dtype > 4) return -1;
  for (int i = 0; i < m; ++i) {
    for (int j = 0; j < n; ++j) {
      float acc = 0.f; /* T.clear / clear_accum */
      for (int kk = 0; kk < k; ++kk) {
        const float a = oracle_decode(ab_dtype, A, trans_a ? (long long)kk * lda + i : (long long)i * lda + kk);
        const float b = oracle_decode(ab_dtype, B, trans_b ? (long long)j * ldb + kk : (long long)kk * ldb + j);
        float prod = a * b; /* exact in fp32 for <=11-bit significands */
        if (product_mode == 1) prod = requantize(ab_dtype, prod);
        acc = acc + prod;
      }
      float v = (alpha == 1.0f) ? acc : acc * alpha;
      if (relu && v < 0.f) v = 0.f;
      oracle_encode(c_dtype, C, (long long)i * ldc + j, v);
    }
  }
  return 0;
}

/*
 * oracle_gemm with epilogue vectors (bias / per-row scales; the per-row form is scales_a[m, kb] of
 * examples/deepseek_deepgemm/example_deepgemm_fp8_2xAcc.py:34,71-73 collapsed to one K block):
 *   C = cast(relu?(((alpha * acc) * row_scale[m]) * col_scale[n] + bias)),   every step rounded to fp32, in this order;
 *   bias = bias[n], or bias[m] when bias_per_row.  Null vectors are skipped.
 */
int oracle_gemm_fused(int m, int n, int k, int ab_dtype, int c_dtype, int trans_a, int trans_b, const void* A,
                      long long lda, const void* B, long long ldb, void* C, long long ldc, float alpha, int relu,
                      const float* row_scale, const float* col_scale, const float* bias, int bias_per_row) {
  if (m < 0 || n < 0 || k < 0 || ab_dtype < 0 || ab_dtype > 3 || c_dtype < 0 || c_dtype > 4) return -1;
  for (int i = 0; i < m; ++i) {
    for (int j = 0; j < n; ++j) {
      float acc = 0.f;
      for (int kk = 0; kk < k; ++kk) {
        const float a = oracle_decode(ab_dtype, A, trans_a ? (long long)kk * lda + i : (long long)i * lda + kk);
        const float b = oracle_decode(ab_dtype, B, trans_b ? (long long)j * ldb + kk : (long long)kk * ldb + j);
        acc = acc + a * b;
      }
      volatile float v = acc * alpha; /* volatile: no contraction of the multiply chain into fmas */
      if (row_scale) v = v * row_scale[i];
      if (col_scale) v = v * col_scale[j];
      if (bias) v = v + bias[bias_per_row ? i : j];
      float out = v;
      if (relu && out < 0.f) out = 0.f;
      oracle_encode(c_dtype, C, (long long)i * ldc + j, out);
    }
  }
  return 0;
}

/* fp8 contraction with fine-grained scales, the tile program of
 * examples/deepseek_deepgemm/example_deepgemm_fp8_2xAcc.py:43-83 restated per output element:
 *   for every 128-wide K block kb:  partial = sum_k A[i,k]*B[j,k]        (T.gemm into a cleared C_local, :75,79)
 *                                   scale   = scales_a[i,kb] * scales_b[j/128,kb]   (Scale_C_shared, :71-73)
 *                                   acc    += partial * scale            (C_local_accum, :77-78; nvcc contracts
 *                                                                         the multiply-add into one fmaf)
 * A is [m,k], B is [n,k] (transpose_B=True, :75); scales are fp32 with element strides. */
#include <math.h>
int oracle_gemm_blockscaled(int m, int n, int k, int ab_dtype, int c_dtype, const void* A, long long lda, const void* B,
                            long long ldb, void* C, long long ldc, const float* sa, long long sa_sm, long long sa_sk,
                            const float* sb, long long sb_sn, long long sb_sk) {
  if (m < 0 || n < 0 || k < 0 || (ab_dtype != 2 && ab_dtype != 3) || c_dtype < 0 || c_dtype > 4) return -1;
  const int nkb = (k + 127) / 128;
  for (int i = 0; i < m; ++i) {
    for (int j = 0; j < n; ++j) {
      float acc = 0.f;
      for (int kb = 0; kb < nkb; ++kb) {
        float partial = 0.f;
        const int k1 = (kb + 1) * 128 < k ? (kb + 1) * 128 : k;
        for (int kk = kb * 128; kk < k1; ++kk)
          partial += oracle_decode(ab_dtype, A, (long long)i * lda + kk) * oracle_decode(ab_dtype, B, (long long)j * ldb + kk);
        const float scale = sa[i * sa_sm + kb * sa_sk] * sb[(j / 128) * sb_sn + kb * sb_sk];
        acc = fmaf(partial, scale, acc);
      }
      oracle_encode(c_dtype, C, (long long)i * ldc + j, acc);
    }
  }
  return 0;
}
