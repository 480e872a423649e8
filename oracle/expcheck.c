/* oracle/expcheck.c -- TEST INFRASTRUCTURE.  Exhaustive proof obligations for the device restatement of ExpFloat32
 * (activations.go:60-89) in guda_b200/csrc/gemm_simt.cuh, which avoids fp64 division and ldexp:
 *   (1) floor(double(x) / ln2) == floor(double(x) * (1/ln2))   for EVERY float32 x in [-87.3f, 88.7f] (2,237,661,186 values)
 *   (2) float(ldexp(double(e), k)) == e * 2^k in float32, k in [-126, 127]
 * Prints the counts; exit status 0 only if both mismatch counts are 0.  Built and run by tests/test_oracle_elementwise.py. */
#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <math.h>
#include <omp.h>
int main() {
  const double ln2 = 0.6931471805599453094;
  const double inv = 1.0 / ln2;
  long long bad = 0, badscale = 0, total = 0;
  #pragma omp parallel for reduction(+:bad,badscale,total) schedule(static)
  for (long long bits = 0; bits < (1ll << 32); bits++) {
    uint32_t u = (uint32_t)bits; float x; memcpy(&x, &u, 4);
    if (!(x <= 88.7f && x >= -87.3f)) continue;       // the range ExpFloat32 reaches the reduction in (NaN excluded)
    total++;
    int k1 = (int)floor((double)x / ln2);
    int k2 = (int)floor((double)x * inv);
    if (k1 != k2) bad++;
    // scale check on a synthetic e in [0.5, 3): (float)ldexp((double)e, k) vs e * 2^k in float
    float e = 1.0f + (float)(u & 0xffff) * (1.7f / 65536.0f);
    float s1 = (float)scalbn((double)e, k1);
    uint32_t pb = (uint32_t)(k1 + 127) << 23; float p; memcpy(&p, &pb, 4);
    float s2 = e * p;
    if (memcmp(&s1, &s2, 4) != 0) badscale++;
    if (k1 < -126 || k1 > 127) badscale++;
  }
  printf("inputs %lld  floor mismatches %lld  scale mismatches %lld\n", total, bad, badscale);
  return (bad || badscale || total != 2237661186ll) ? 1 : 0;
}
