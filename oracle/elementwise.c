/* oracle/elementwise.c -- TEST INFRASTRUCTURE (CPU restatement; never on the product path).
 * Elementwise ops of the reference's root package: math.go:133-169 Add / Multiply, :230-272 ReLU / Sigmoid,
 * activations.go:9-27 SigmoidFloat32, and fused.go:18-216 FusedKernel.Execute (one read, the op list applied to the value in
 * order, one write) with the ops LinearReLU (fused.go:252) and GELU (:261) are built from.  Every Go statement on a float32
 * variable rounds to float32; on the default GOAMD64=v1 build `val += alpha*other` is an unfused multiply then add. */
#include <stddef.h>
#include <stdlib.h>
#include "guda_oracle.h"

float orc_sigmoid_f32(float x) {                    /* activations.go:9-27 */
  if (x < -10.0f) return 0.0f;
  if (x > 10.0f) return 1.0f;
  if (x >= 0) {
    float e = orc_exp_f32(-x);
    return 1.0f / (1.0f + e);
  }
  float e = orc_exp_f32(x);                         /* also the NaN route (every comparison false) */
  return e / (1.0f + e);
}

void orc_ew_add(const float* a, const float* b, float* c, size_t n) { for (size_t i = 0; i < n; i++) c[i] = a[i] + b[i]; }
void orc_ew_mul(const float* a, const float* b, float* c, size_t n) { for (size_t i = 0; i < n; i++) c[i] = a[i] * b[i]; }
void orc_ew_relu(float* x, size_t n) { for (size_t i = 0; i < n; i++) if (x[i] < 0) x[i] = 0; }          /* NaN and -0 stay */
void orc_ew_sigmoid(float* x, size_t n) { for (size_t i = 0; i < n; i++) x[i] = orc_sigmoid_f32(x[i]); }

/* fused.go:123-216.  ops[i].type: 0 Add (val += alpha * other), 1 Mul (val *= other), 2 AddScalar, 3 MulScalar, 4 ReLU,
 * 5 Sigmoid, 6 Custom(geluFloat32).  A binary op consumes the next `others` pointer (skipped when none is left, :148,:170);
 * bsize > 0: other[(idx / bsize) % bdim] (the broadcast index of :160-161), else other[idx]. */
void orc_fused_execute(float* x, const float* const* others, int n_others, size_t n, const orc_fused_op_t* ops, int n_ops) {
  for (size_t idx = 0; idx < n; idx++) {
    float val = x[idx];
    int oi = 0;
    for (int o = 0; o < n_ops; o++) {
      const orc_fused_op_t* op = &ops[o];
      switch (op->type) {
        case 0:
        case 1:
          if (oi < n_others) {
            float ov = op->bsize > 0 ? others[oi][(idx / (size_t)op->bsize) % (size_t)op->bdim] : others[oi][idx];
            if (op->type == 0) { float t = op->alpha * ov; val = val + t; }
            else val = val * ov;
            oi++;
          }
          break;
        case 2: val = val + op->alpha; break;
        case 3: val = val * op->alpha; break;
        case 4: if (val < 0) val = 0; break;
        case 5: val = orc_sigmoid_f32(val); break;
        case 6: val = orc_gelu_f32(val); break;
        default: break;
      }
    }
    x[idx] = val;
  }
}

/* reduce.go:105-178 ReduceSum / ReduceMax over a row-major m x n matrix.  axis 0: out[j] = chain over rows; axis 1: out[i] =
 * chain over columns.  Sum starts from float32(0) and adds in index order; Max starts from the first element and replaces it
 * only when `val > max` (so a leading NaN stays, later NaNs are skipped, and of equal zeros the first one wins). */
int orc_reduce_axis(const float* x, int m, int n, int axis, int is_max, float* out) {
  if (axis != 0 && axis != 1) return -1;
  const int outer = axis == 0 ? n : m, inner = axis == 0 ? m : n;
  for (int o = 0; o < outer; o++) {
    float acc = is_max ? (axis == 0 ? x[o] : x[(size_t)o * n]) : 0.0f;
    for (int k = is_max ? 1 : 0; k < inner; k++) {
      const float v = axis == 0 ? x[(size_t)k * n + o] : x[(size_t)o * n + k];
      if (is_max) { if (v > acc) acc = v; }
      else acc = acc + v;
    }
    out[o] = acc;
  }
  return 0;
}

/* reduce.go:247-291 TopK: selection sort of (value, index) pairs, k rounds, swaps included.  values / indices hold min(k, n). */
int orc_topk(const float* x, long long n, int k, float* values, long long* indices) {
  if (k > n) k = (int)n;
  if (k <= 0) return -1;
  float* v = (float*)malloc((size_t)n * sizeof(float));
  long long* id = (long long*)malloc((size_t)n * sizeof(long long));
  for (long long i = 0; i < n; i++) { v[i] = x[i]; id[i] = i; }
  for (int i = 0; i < k; i++) {
    long long mx = i;
    for (long long j = i + 1; j < n; j++) if (v[j] > v[mx]) mx = j;
    float tv = v[i]; v[i] = v[mx]; v[mx] = tv;
    long long ti = id[i]; id[i] = id[mx]; id[mx] = ti;
  }
  for (int i = 0; i < k; i++) { values[i] = v[i]; indices[i] = id[i]; }
  free(v); free(id);
  return k;
}
