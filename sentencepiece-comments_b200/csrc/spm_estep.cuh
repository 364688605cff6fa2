// spm_estep.cuh -- the E-step of the unigram trainer for one word:
// Lattice::PopulateMarginal (src/unigram_model.cc:272-328) + the Viterbi token
// count of Trainer::RunEStep (src/unigram_model_trainer.cc:256-267), without
// building a lattice.
//
// The reference keeps alpha / beta per NODE, but alpha[r] only depends on the
// position r begins at and beta[l] on the position l ends at:
//   A[e] = LSE over nodes (b -> e), in end_nodes_[e] order (ascending b), of score + A[b]
//   B[b] = LSE over nodes (b -> e), in begin_nodes_[b] order (ascending length,
//          the UNK node last), of score + B[e]
//   Z = A[len];  expected[id] += freq * exp(A[b] + score + B[e] - Z)
// with LogSumExp as src/unigram_model.cc:51-63 ("init on the first element";
// vmax when the gap exceeds 50; else vmax + log(exp(vmin - vmax) + 1.0) in double,
// rounded to float -- the log(exp() + 1) from a polynomial table, estep_softplus()).  An accumulator that starts at -inf reproduces "init on the first
// element" (the gap test returns the first element unchanged).
// Positions are BYTE offsets of character boundaries (pieces and text are valid UTF-8).
// The lattice of a TrainerModel: every piece is NORMAL, unk_id_ = 0, UNK score =
// min_score_ - 10 (src/unigram_model_trainer.cc:97-119).
// `__host__ __device__`: tests/host_core.cc runs the same code against ref_estep.
#pragma once
#include <math.h>

#include "spm_core.cuh"

namespace spmb {

// Leaf word .x of the E-step trie: piece id in the low bits; when the vocabulary has at most 2^20
// pieces, bits 20..31 carry the piece's slot in the per-CTA table of HOT pieces (1..kEstepHot-1; 0:
// not hot) -- the pieces with the highest scores collect most of the expected counts, and one
// global atomicAdd per node on a handful of addresses is what bounded the first kernel.
constexpr int kEstepHot = 512;
struct EstepView {
  const double *lse;       // estep_softplus()'s table: kLseN x kLseCoef
  const uint32_t *units_hot;   // copy of the first k_units units (build_xda places nodes by descending weight: the
  uint32_t k_units;            // top of the trie, where nearly every step of a walk lands); the kernel: shared memory
  const uint32_t *units;   // XDA of all pieces (spm_model.cc build_xda)
  const uint2 *leaf;       // node -> (piece id | hot slot << 20, score bits)
  float unk_score;         // min score - 10
  uint32_t unk_id;         // 0 (| its hot slot << 20)
  uint32_t id_mask;        // 0xFFFFF with hot slots packed, else 0x0FFFFFFF
};

// log(exp(d) + 1) for d in [-kLseLo, 0] as a table of polynomials (build_lse_table() in spm_model.cc): kLseN
// intervals of kLseLo / kLseN, degree 7 around the interval's centre, interpolated at Chebyshev nodes from long
// double values -- absolute error < 1e-14, against the 1e-16 of libm's exp + log; below -kLseLo the function is
// exp(d) to 1e-22.  The float the sum is rounded to differs from the libm result in about one LogSumExp of
// 10^8 (by one ulp); exp + log in double cost ~300 instructions per node of the lattice.  8 KB: the kernel keeps
// the table in shared memory.
constexpr int kLseN = 128, kLseCoef = 8;
constexpr double kLseLo = 25.0;
SPM_HD double estep_softplus(const double *tab, double d) {
  if (d < -kLseLo) return exp(d);
  int k = (int)(d * (-(double)kLseN / kLseLo));
  k = k < 0 ? 0 : (k > kLseN - 1 ? kLseN - 1 : k);       // (a NaN lands in interval 0 and stays a NaN)
  const double t = d + ((double)k + 0.5) * (kLseLo / (double)kLseN);
  const double *c = tab + k * kLseCoef;
#if defined(__CUDA_ARCH__)
  const double2 c01 = reinterpret_cast<const double2 *>(c)[0], c23 = reinterpret_cast<const double2 *>(c)[1],
                c45 = reinterpret_cast<const double2 *>(c)[2], c67 = reinterpret_cast<const double2 *>(c)[3];
  double p = c67.y;
  p = fma(p, t, c67.x); p = fma(p, t, c45.y); p = fma(p, t, c45.x); p = fma(p, t, c23.y);
  p = fma(p, t, c23.x); p = fma(p, t, c01.y); p = fma(p, t, c01.x);
  return p;
#else
  double p = c[7];
  for (int j = 6; j >= 0; --j) p = fma(p, t, c[j]);
  return p;
#endif
}
// exp() of a node's log-posterior (<= 0 up to rounding): float precision is 100 x finer than the 1e-5 the
// expected counts are pinned at; far down, where a float would flush to zero, the double routine.
SPM_HD double estep_exp(float x) {
  return x > -80.0f ? (double)expf(x) : exp((double)x);
}
// LogSumExp of src/unigram_model.cc:51-63
SPM_HD float estep_lse(const double *tab, float x, float y) {
  const float vmin = x < y ? x : y, vmax = x < y ? y : x;
  if (vmax > vmin + 50.0f) return vmax;
  return (float)((double)vmax + estep_softplus(tab, (double)(vmin - vmax)));
}

// Walks the pieces that start at byte b of the word wb(0) .. wb(n - 1).  f(e, id, score) is called for
// every node in begin_nodes_ order (ascending length; the one-character UNK node last when no
// piece covers exactly the first character, src/unigram_model.cc:582-602).
template <class WB, class F>
SPM_HD void estep_nodes_from(const EstepView &ev, WB wb, int n, int b, F f) {
  int cl = one_char_len((uint8_t)wb(b));
  if (cl > n - b) cl = n - b;
  uint32_t u = ev.units[0];
  bool single = false;
  for (int j = b; j < n; ++j) {
    const uint32_t c = wb(j);
    if (!(u & 0x200u)) break;
    const uint32_t child = (u >> 10) ^ c;
    const uint32_t v = child < ev.k_units ? ev.units_hot[child] : ev.units[child];
    if ((v & 0xFFu) != c || (v & 0x300u) == 0) break;
    u = v;
    if (v & 0x100u) {
      const uint2 lf = ev.leaf[child];
      if (j + 1 - b == cl) single = true;
      f(j + 1, lf.x, __uint_as_float_hd(lf.y));
    }
  }
  if (!single) f(b + cl, ev.unk_id, ev.unk_score);
}

// One word of n bytes, read through wb(j) (the kernel keeps a copy in shared memory: a word is read
// n / 2 times over, and with one word per thread the lines of a CTA's words do not fit what is left of L1).
// A, BV: float tables of n + 1 entries each (BV holds the Viterbi scores during the
// forward pass and beta during the backward pass: the two are never alive together), C: Viterbi
// token counts (n + 1 entries of type CT); entry k of a table lives at [k * stride] (stride 1:
// private arrays; stride = block size: [position][thread] layout in shared memory).  expected[]
// receives freq * marginal through add(x, value) with x = the leaf word (id | hot slot << 20).
// Returns Z (log-likelihood of the word, float as in the reference); *ntokens = length of the
// Viterbi path.
template <class CT, class WB, class Add>
SPM_HDN float estep_word(const EstepView &ev, WB wb, int n, float freq, float *A, float *BV, CT *C,
                         int stride, Add add, int32_t *ntokens) {
  const float NINF = -INFINITY;
  for (int k = 0; k <= n; ++k) { A[k * stride] = NINF; BV[k * stride] = NINF; C[k * stride] = 0; }
  A[0] = 0.0f; BV[0] = 0.0f;
  // forward: alpha and the Viterbi count (first strictly greater wins, src/unigram_model.cc:229-246)
  for (int b = 0; b < n;) {
    const float ab = A[b * stride], vb = BV[b * stride];
    const int32_t cb = (int32_t)C[b * stride];
    estep_nodes_from(ev, wb, n, b, [&](int e, uint32_t, float sc) {
      A[e * stride] = estep_lse(ev.lse, A[e * stride], sc + ab);
      const float cand = vb + sc;
      if (cand > BV[e * stride]) { BV[e * stride] = cand; C[e * stride] = (CT)(cb + 1); }
    });
    int cl = one_char_len((uint8_t)wb(b));
    b += cl > n - b ? n - b : cl;
  }
  const float Z = A[n * stride];
  *ntokens = (int32_t)C[n * stride];
  // backward: beta, and the expected counts of the nodes that begin at b
  BV[n * stride] = 0.0f;
  for (int b = n - 1; b >= 0; --b) {
    if (b > 0 && (wb(b) & 0xC0u) == 0x80u) continue;   // not a character boundary
    float acc = NINF;
    const float ab = A[b * stride];
    estep_nodes_from(ev, wb, n, b, [&](int e, uint32_t x, float sc) {
      const float be = BV[e * stride];
      acc = estep_lse(ev.lse, acc, sc + be);
      add(x, (double)freq * estep_exp(ab + sc + be - Z));
    });
    BV[b * stride] = acc;
  }
  return Z;
}

}  // namespace spmb
