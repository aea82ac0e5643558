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
  const double dn = (double)n;
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
    double p = __ddiv_rn((double)v, dn);
    return criterion == 0 ? __dmul_rn(p, log2(p)) : __dmul_rn(p, p);
  };
  double s = numpy_sum(k, next);
  return criterion == 0 ? -s : __dsub_rn(1.0, s);
}

__device__ __forceinline__ double split_cost(uint32_t n_left, uint32_t n_right, double h_left, double h_right,
                                             int dot_mode) {
  const double n = (double)(n_left + n_right);
  const double wl = __ddiv_rn((double)n_left, n);
  const double wr = __ddiv_rn((double)n_right, n);
  if (dot_mode == 1) return __fma_rn(wr, h_right, __dmul_rn(wl, h_left));
  if (dot_mode == 2) return __fma_rn(wl, h_left, __dmul_rn(wr, h_right));
  return __dadd_rn(__dmul_rn(wl, h_left), __dmul_rn(wr, h_right));
}

// ---------------------------------------------------------------------------------------
// Large nodes (n > plogp_n).  Rounding-level ties between *different* count configurations
// do not occur in practice there, so the evaluation only has to be a deterministic function
// of the counts (identical configurations still tie exactly and resolve by position).  The
// algebraically equal form below needs one table lookup and one add per class instead of a
// division, a log2 and a multiply:
//     (nL/n) H(L) + (nR/n) H(R) = ( g(nL) + g(nR) - sum_c g(kL_c) - sum_c g(kR_c) ) / n
// with g(k) = k log2 k, g(0) = 0.
struct KlogK {
  const double* tab;
  uint32_t n;
  __device__ __forceinline__ double operator()(uint32_t k) const {
    if (k < n) return tab[k];
    const double d = (double)k;
    return d * log2(d);
  }
};

template <typename L, typename R>
__device__ __forceinline__ double fast_split_cost(int C, L left, R right, uint32_t nl, uint32_t nr, int criterion,
                                                  const KlogK& g) {
  const double n = (double)nl + (double)nr;
  if (criterion == 0) {
    double s = g(nl) + g(nr);
    int c = 0;
    for (; c + 4 <= C; c += 4) {  // issue the eight table loads together, then accumulate
      const double a0 = g(left(c)), a1 = g(left(c + 1)), a2 = g(left(c + 2)), a3 = g(left(c + 3));
      const double b0 = g(right(c)), b1 = g(right(c + 1)), b2 = g(right(c + 2)), b3 = g(right(c + 3));
      s -= ((a0 + b0) + (a1 + b1)) + ((a2 + b2) + (a3 + b3));
    }
    for (; c < C; ++c) s -= g(left(c)) + g(right(c));
    return s / n;
  }
  double sl = 0.0, sr = 0.0;
  for (int c = 0; c < C; ++c) {
    const double a = (double)left(c), b = (double)right(c);
    sl += a * a;
    sr += b * b;
  }
  return (((double)nl - sl / (double)nl) + ((double)nr - sr / (double)nr)) / n;
}

template <typename Cnt>
__device__ __forceinline__ double fast_impurity(int C, Cnt cnt, uint32_t n, int criterion, const KlogK& g) {
  const double dn = (double)n;
  if (criterion == 0) {
    double s = g(n);
    for (int c = 0; c < C; ++c) s -= g(cnt(c));
    return s / dn;
  }
  double sq = 0.0;
  for (int c = 0; c < C; ++c) {
    const double a = (double)cnt(c);
    sq += a * a;
  }
  return 1.0 - sq / (dn * dn);
}

}  // namespace mtb
