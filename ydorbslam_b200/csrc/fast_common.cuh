// Device helpers shared by the two FAST kernels (fast.cu: warp per cell; fast2.cu: CTA per block of cells).
// Packed s16x2 arithmetic on pixel pairs, see the header of fast.cu.
#pragma once
#include "kernels.cuh"

namespace ydorb {

// circular 9-arc extremum: r[k] = op3 over D[k..k+8]
#define YD_ARC_TREE(OP3, D, OUT)                                                   \
  {                                                                                \
    uint32_t w3[16];                                                               \
    _Pragma("unroll") for (int k = 0; k < 16; k++) w3[k] = OP3(D[k], D[(k + 1) & 15], D[(k + 2) & 15]); \
    _Pragma("unroll") for (int k = 0; k < 16; k++) OUT[k] = OP3(w3[k], w3[(k + 3) & 15], w3[(k + 6) & 15]); \
  }

__device__ __forceinline__ uint32_t max16_s16x2(const uint32_t (&v)[16]) {
  uint32_t a = __vimax3_s16x2(v[0], v[1], v[2]), b = __vimax3_s16x2(v[3], v[4], v[5]);
  uint32_t c = __vimax3_s16x2(v[6], v[7], v[8]), d = __vimax3_s16x2(v[9], v[10], v[11]);
  uint32_t e = __vimax3_s16x2(v[12], v[13], v[14]);
  a = __vimax3_s16x2(a, b, c);
  d = __vimax3_s16x2(d, e, v[15]);
  return __vmaxs2(a, d);
}
__device__ __forceinline__ uint32_t min16_s16x2(const uint32_t (&v)[16]) {
  uint32_t a = __vimin3_s16x2(v[0], v[1], v[2]), b = __vimin3_s16x2(v[3], v[4], v[5]);
  uint32_t c = __vimin3_s16x2(v[6], v[7], v[8]), d = __vimin3_s16x2(v[9], v[10], v[11]);
  uint32_t e = __vimin3_s16x2(v[12], v[13], v[14]);
  a = __vimin3_s16x2(a, b, c);
  d = __vimin3_s16x2(d, e, v[15]);
  return __vmins2(a, d);
}

__device__ __forceinline__ uint32_t fsr16(uint32_t lo, uint32_t hi) { return __funnelshift_r(lo, hi, 16); }

// Word j of a tile row holds tile columns (2j, 2j+1).  For the pixel pair at `row` (= &tile[r][j]): an even dx reads
// word j + dx/2, an odd dx the funnel shift of words j+(dx-1)/2 and j+(dx+1)/2.

// Both tests work on the raw circle pixels P_k instead of the differences d_k = I(p) - P_k: a minimum of differences
// over a set is I(p) minus the maximum of the pixels over it, so the 16 (8) subtractions collapse into two.
//
// Cheap necessary condition for "one of the two pixels is a corner at threshold t" (see the header of fast.cu):
//   min_k max(d_k, d_k+8) > t   <=>   I(p) - max_k min(P_k, P_k+8) > t
//   max_k min(d_k, d_k+8) < -t  <=>   min_k max(P_k, P_k+8) - I(p) > t          on k = 0, 2, 4, 6.
// k1 = (0x7FFF - (t + 256)) * 0x10001: a half h in [1, 511] is > t + 256  <=>  bit 15 of (h + k1); no carries between
// halves.
__device__ __forceinline__ bool fast_quick_pair(const uint32_t* __restrict__ row, int wpr, uint32_t k1) {
  const uint32_t cb = row[0] + 0x01000100u;  // centre + 256 in both halves
  const uint32_t p0 = row[3 * wpr], p8 = row[-3 * wpr];                      // (0,3) (0,-3)
  const uint32_t p4 = fsr16(row[1], row[2]), p12 = fsr16(row[-2], row[-1]);  // (3,0) (-3,0)
  const uint32_t p2 = row[2 * wpr + 1], p10 = row[-2 * wpr - 1];             // (2,2) (-2,-2)
  const uint32_t p6 = row[-2 * wpr + 1], p14 = row[2 * wpr - 1];             // (2,-2) (-2,2)
  const uint32_t hi = __vimax3_s16x2(__vmins2(p0, p8), __vmins2(p4, p12), __vmaxs2(__vmins2(p2, p10), __vmins2(p6, p14)));
  const uint32_t lo = __vimin3_s16x2(__vmaxs2(p0, p8), __vmaxs2(p4, p12), __vmins2(__vmaxs2(p2, p10), __vmaxs2(p6, p14)));
  // (cb - hi) and (lo + 512 - cb) have halves in [1, 511]
  return ((((cb - hi) + k1) | ((lo + 0x02000200u - cb) + k1)) & 0x80008000u) != 0u;
}

// Exact FAST-9/16 score of both pixels of the pair, returned biased: each half = m + 256, m in [-255, 255], with
//   m = max( I(p) - min_arcs max_arc P ,  max_arcs min_arc P - I(p) ).
__device__ __forceinline__ uint32_t fast_score_pair(const uint32_t* __restrict__ row, int wpr) {
  uint32_t P[16];
  {
    const uint32_t* r = row + 3 * wpr;  // dy = +3: k0 (0,3) k1 (1,3) k15 (-1,3)
    const uint32_t a = r[-1], b = r[0], c = r[1];
    P[0] = b; P[1] = fsr16(b, c); P[15] = fsr16(a, b);
  }
  {
    const uint32_t* r = row - 3 * wpr;  // dy = -3: k8 (0,-3) k7 (1,-3) k9 (-1,-3)
    const uint32_t a = r[-1], b = r[0], c = r[1];
    P[8] = b; P[7] = fsr16(b, c); P[9] = fsr16(a, b);
  }
  P[2] = row[2 * wpr + 1];    // (2,2)
  P[14] = row[2 * wpr - 1];   // (-2,2)
  P[6] = row[-2 * wpr + 1];   // (2,-2)
  P[10] = row[-2 * wpr - 1];  // (-2,-2)
  P[3] = fsr16(row[wpr + 1], row[wpr + 2]);      // (3,1)
  P[13] = fsr16(row[wpr - 2], row[wpr - 1]);     // (-3,1)
  P[5] = fsr16(row[-wpr + 1], row[-wpr + 2]);    // (3,-1)
  P[11] = fsr16(row[-wpr - 2], row[-wpr - 1]);   // (-3,-1)
  P[4] = fsr16(row[1], row[2]);                  // (3,0)
  P[12] = fsr16(row[-2], row[-1]);               // (-3,0)
  const uint32_t cb = row[0] + 0x01000100u;      // centre + 256 in both halves
  uint32_t lo9[16], hi9[16];
  YD_ARC_TREE(__vimin3_s16x2, P, lo9)
  YD_ARC_TREE(__vimax3_s16x2, P, hi9)
  const uint32_t arc_min = max16_s16x2(lo9);   // max over arcs of the arc's darkest pixel
  const uint32_t arc_max = min16_s16x2(hi9);   // min over arcs of the arc's brightest pixel
  // halves of both terms in [1, 511]: no borrow between halves
  return __vmaxs2(cb - arc_max, arc_min + 0x02000200u - cb);
}

}  // namespace ydorb
