// Impurity / split-cost arithmetic with the reference's float64 rounding.
//
// The reference evaluates (mpitree/tree/decision_tree.py:49-51, :79-91)
//     H(x)  = -np.sum(p * np.log2(p)),  p = counts_present / n
//     cost  = np.dot([nL/n, nR/n], [H(yL), H(yR)])
//     gain  = H(y) - min(cost)
// To make first-arg-min / first-arg-max ties resolve the same way on the GPU,
// every operation below is a separately rounded IEEE double operation
// (__dmul_rn / __dadd_rn stop nvcc from contracting into FMAs), the sum follows
// numpy's pairwise order (plain loop below 8 terms, 8 running accumulators
// otherwise) and the 2-term dot follows the host BLAS's rounding, probed at
// start-up (dot_mode: 0 plain, 1 fma(b,d,a*c), 2 fma(a,c,b*d)).
#pragma once

#include <math.h>
#include <stdint.h>

namespace mtb {

// Exact uint32 -> double without a conversion instruction: the bit pattern of 2^52 + k minus 2^52.
// (I2F.F64 issues on the XU pipe at ~2 per clock per SM on B200: ncu showed that pipe, not fp64,
// binding the gain phase when every k log2 k started with a conversion.)
__device__ __forceinline__ double u2d(uint32_t k) {
  return __dsub_rn(__hiloint2double(0x43300000, (int)k), 4503599627370496.0);
}

// Sums `terms` produced on the fly in numpy.add.reduce order for a contiguous
// float64 vector of length k (k <= 128).  `next()` yields the terms in order.
template <typename Next>
__device__ __forceinline__ double numpy_sum(int k, Next next) {
  if (k < 8) {
    double res = 0.0;
    for (int i = 0; i < k; ++i) res = __dadd_rn(res, next());
    return res;
  }
  double r[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) r[j] = next();
  int body = k - (k % 8);
  for (int i = 8; i < body; i += 8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], next());
  }
  double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                         __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
  for (int i = body; i < k; ++i) res = __dadd_rn(res, next());
  return res;
}

// Impurity of a class-count vector read through `cnt(c)`, total n > 0.
// criterion 0: entropy (base 2), 1: gini.
// `plogp` (may be null) is the host's table of (k/n)*log2(k/n) for n <= plogp_n.
template <typename Cnt>
__device__ __forceinline__ double impurity(int C, Cnt cnt, uint32_t n, int criterion, const double* plogp = nullptr,
                                           uint32_t plogp_n = 0) {
  int k = 0;
  for (int c = 0; c < C; ++c) k += (cnt(c) != 0u);
  const double dn = u2d(n);
  const double* row = (criterion == 0 && plogp != nullptr && n <= plogp_n)
                          ? plogp + ((size_t)(n - 1) * n) / 2 - 1
                          : nullptr;
  int c = 0;
  auto next = [&]() -> double {
    uint32_t v;
    do {
      v = cnt(c++);
    } while (v == 0u);
    if (row) return row[v];
    double p = __ddiv_rn(u2d(v), dn);
    return criterion == 0 ? __dmul_rn(p, log2(p)) : __dmul_rn(p, p);
  };
  double s = numpy_sum(k, next);
  return criterion == 0 ? -s : __dsub_rn(1.0, s);
}

__device__ __forceinline__ double split_cost(uint32_t n_left, uint32_t n_right, double h_left, double h_right,
                                             int dot_mode) {
  const double n = u2d(n_left + n_right);
  const double wl = __ddiv_rn(u2d(n_left), n);
  const double wr = __ddiv_rn(u2d(n_right), n);
  if (dot_mode == 1) return __fma_rn(wr, h_right, __dmul_rn(wl, h_left));
  if (dot_mode == 2) return __fma_rn(wl, h_left, __dmul_rn(wr, h_right));
  return __dadd_rn(__dmul_rn(wl, h_left), __dmul_rn(wr, h_right));
}

// ---------------------------------------------------------------------------------------
// Large nodes (n > plogp_n).  Rounding-level ties between *different* count configurations
// do not occur in practice there, so the evaluation only has to be a deterministic function
// of the counts (identical configurations still tie exactly and resolve by position).  The
// algebraically equal form below needs one g() and one add per class instead of a
// division, a log2 and a multiply:
//     (nL/n) H(L) + (nR/n) H(R) = ( g(nL) + g(nR) - sum_c g(kL_c) - sum_c g(kR_c) ) / n
// with g(k) = k log2 k, g(0) = g(1) = 0.
//
// g() is evaluated arithmetically on the fp64 pipe: the exponent of k gives e = floor(log2 k), the
// top 7 mantissa bits select an interval centre m0 = 1 + (i + 1/2)/128, r = (m - m0) * fl(1/m0) lies
// within +-2^-8 and log2(1 + r) is a degree-6 Horner polynomial in fused multiply-adds (truncation
// 2^-58); g = k * (fl(e + log2 m0) + r * poly).  Ten fp64 instructions and two 8-byte shared-memory
// loads per value.  Worst error found against 120-bit arithmetic: 2.5 ulp, the same class as libm's
// log2.  (History: a global-memory table of g costs one 32-byte sector per lookup and was bound by
// the L1 tag stage at ~32 cycles per warp-wide gather; int -> double conversion instructions issue
// at ~2 per clock per SM, hence u2d(); B200's fp64 pipe sustains about a quarter of the fp32 rate,
// hence the fused multiply-adds and the folded e + log2 m0 table.)
//
// Table layout (ctx.klogk, built by the host, f64): inv[128] = fl(1/m0) | el[32][128] = fl(e + log2 m0)
// | c1..c6.
constexpr int kKlogkIdxBits = 7;
constexpr int kKlogkEntries = 1 << kKlogkIdxBits;
constexpr int kKlogkTabDoubles = kKlogkEntries + 32 * kKlogkEntries;  // the part worth keeping in shared memory
constexpr int kKlogkDoubles = kKlogkTabDoubles + 6;

struct KlogK {
  const double* inv;  // [128], shared (or global) memory
  const double* el;   // [32][128]
  double c1, c2, c3, c4, c5, c6;
  __device__ __forceinline__ void bind(const double* table, const double* coefs_from) {
    inv = table;
    el = table + kKlogkEntries;
    const double* c = coefs_from + kKlogkTabDoubles;
    c1 = c[0]; c2 = c[1]; c3 = c[2]; c4 = c[3]; c5 = c[4]; c6 = c[5];
  }
  __device__ __forceinline__ double operator()(uint32_t k) const {
    const double kd = u2d(k);
    const int hi = __double2hiint(kd);
    // e * 128 + idx; masked so that k = 0 (whose exponent field is 0) stays inside the table: its
    // value is discarded below
    const int top = ((hi >> (20 - kKlogkIdxBits)) - (1023 << kKlogkIdxBits)) & (32 * kKlogkEntries - 1);
    const int idx = top & (kKlogkEntries - 1);
    const double m = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(kd));
    const double m0 = __hiloint2double(0x3ff00000 | (idx << (20 - kKlogkIdxBits)) | (1 << (19 - kKlogkIdxBits)), 0);
    const double r = __dmul_rn(__dsub_rn(m, m0), inv[idx]);
    double p = __fma_rn(c6, r, c5);
    p = __fma_rn(p, r, c4);
    p = __fma_rn(p, r, c3);
    p = __fma_rn(p, r, c2);
    p = __fma_rn(p, r, c1);
    const double lg = __fma_rn(p, r, el[top]);
    return k <= 1u ? 0.0 : __dmul_rn(kd, lg);
  }
};

// Summation order: s = g(nL) + g(nR); then for classes in ascending order s -= (g(kL_c) + g(kR_c));
// cost = s / n.  Every operation is rounded separately (no contraction), so the result is a fixed
// function of the counts on every rank.
template <typename L, typename R>
__device__ __forceinline__ double fast_split_cost(int C, L left, R right, uint32_t nl, uint32_t nr, int criterion,
                                                  const KlogK& g) {
  const double n = u2d(nl + nr);
  if (criterion == 0) {
    double s = __dadd_rn(g(nl), g(nr));
    int c = 0;
    for (; c + 2 <= C; c += 2) {  // two classes per step: four independent evaluations in flight
      const double a0 = g(left(c)), b0 = g(right(c)), a1 = g(left(c + 1)), b1 = g(right(c + 1));
      s = __dsub_rn(s, __dadd_rn(a0, b0));
      s = __dsub_rn(s, __dadd_rn(a1, b1));
    }
    if (c < C) s = __dsub_rn(s, __dadd_rn(g(left(c)), g(right(c))));
    return __ddiv_rn(s, n);
  }
  // gini: n*cost = (nL - sum kL^2 / nL) + (nR - sum kR^2 / nR); the sums of squares are exact integers
  unsigned long long sl = 0, sr = 0;
  for (int c = 0; c < C; ++c) {
    const unsigned long long a = left(c), b = right(c);
    sl += a * a;
    sr += b * b;
  }
  const double dl = __dsub_rn((double)nl, __ddiv_rn((double)sl, (double)nl));
  const double dr = __dsub_rn((double)nr, __ddiv_rn((double)sr, (double)nr));
  return __ddiv_rn(__dadd_rn(dl, dr), n);
}

template <typename Cnt>
__device__ __forceinline__ double fast_impurity(int C, Cnt cnt, uint32_t n, int criterion, const KlogK& g) {
  const double dn = u2d(n);
  if (criterion == 0) {
    double s = g(n);
    for (int c = 0; c < C; ++c) s = __dsub_rn(s, g(cnt(c)));
    return __ddiv_rn(s, dn);
  }
  unsigned long long sq = 0;
  for (int c = 0; c < C; ++c) {
    const unsigned long long a = cnt(c);
    sq += a * a;
  }
  return __dsub_rn(1.0, __ddiv_rn((double)sq, __dmul_rn(dn, dn)));
}

}  // namespace mtb
