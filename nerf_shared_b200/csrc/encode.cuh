// Positional encoding in registers, shared by the forward kernel (mlp_tc.cu) and the input-gradient
// kernel (mlp_ingrad.cu), so both evaluate the very same sin/cos values.
#pragma once
#include <cuda_runtime.h>

namespace nb {

// ------------------------------------------------------------------------------------------
// positional encoding in registers (nerf.py:28-41)
// ------------------------------------------------------------------------------------------
// sin/cos with a 3-term Cody-Waite reduction by pi/2 (|x| up to a few thousand) and degree-7/8
// minimax polynomials: max abs error 9e-8 (checked against float64 over [-400, 400]).
__device__ __forceinline__ void sincos_cw(float x, float& s, float& c) {
  float k = rintf(x * 0.6366197466850281f);
  float r = fmaf(k, -1.5707963705062866f, x);
  r = fmaf(k, 4.371138828673793e-08f, r);
  r = fmaf(k, 1.7151245100058819e-15f, r);
  int q = (int)k;
  float r2 = r * r;
  float sp = fmaf(r * r2, fmaf(r2, fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f), -1.6666654611e-1f), r);
  float cp = fmaf(r2 * r2, fmaf(r2, fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f), 4.166664568298827e-2f),
                  fmaf(-0.5f, r2, 1.0f));
  float ss = (q & 1) ? cp : sp;
  float cc = (q & 1) ? sp : cp;
  s = (q & 2) ? -ss : ss;
  c = ((q + 1) & 2) ? -cc : cc;
}

// e[0..2] = v, e[3+6l+c] = sin(2^l v_c), e[3+6l+3+c] = cos(2^l v_c); octaves 0 and 5 are evaluated
// directly, the others by angle doubling (abs error < 2e-6, two orders below a bf16 ulp).
template <int L, int N>
__device__ __forceinline__ void encode3(float v0, float v1, float v2, float (&e)[N]) {
  static_assert(3 + 6 * L <= N, "encoding does not fit the operand slab");
#pragma unroll
  for (int i = 0; i < N; ++i) e[i] = 0.f;
  float v[3] = {v0, v1, v2};
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    e[c] = v[c];
    float s = 0.f, co = 1.f;
#pragma unroll
    for (int l = 0; l < L; ++l) {
      if (l % 5 == 0) {
        sincos_cw(v[c] * (float)(1 << l), s, co);
      } else {
        float s2 = 2.0f * s * co;
        float c2 = (co - s) * (co + s);
        s = s2;
        co = c2;
      }
      e[3 + 6 * l + c] = s;
      e[3 + 6 * l + 3 + c] = co;
    }
  }
}

}  // namespace nb
