// spm_fast.cuh -- the fast tier of the unigram encoder: one thread per
// sentence, all per-position state in shared memory / registers.
//
// Same results as unigram_encode<1> of spm_core.cuh (and therefore as
// src/unigram_model.cc:535-604 + :222-261 of the reference) for every sentence
// it does not flag; a flagged sentence (near-tied Viterbi candidates, status
// ST_CAND_OVERFLOW) is re-run by the exact tier.  Differences in *mechanics*:
//   * the trie is the "fast view" of the piece XDA: has_leaf is only set for
//     pieces PopulateNodes would insert (NORMAL / USER_DEFINED), and a parallel
//     float array carries the node score (USER_DEFINED bonus pre-computed with
//     the reference's arithmetic at load time), so a hit costs one 4-byte load;
//     the first k_units / k_scores entries of both arrays (the hot-first layout
//     puts ~95 % of all accesses there) are read from shared memory;
//   * the text is read through a 24-byte register window (three aligned 8-byte
//     loads, refilled every 8 bytes); the walk stops on the 0 byte that the
//     normaliser guarantees after every sentence (a NUL in the text can never
//     match in the reference either: darts keys are NUL-terminated);
//   * the DP keeps, per pending end position (a ring of max-piece-chars + 1
//     character slots), the best backtrace score and its (node, chars) record;
//     the record of a position is final when the scan arrives there and is
//     stored to a back-pointer row that the 32 lanes of a warp write together
//     (lanes advance one character per iteration in lock step).
// Everything is `__host__ __device__` so tests/host_core.cc can run the very
// same code on the CPU against the oracle.
#pragma once
#include "spm_core.cuh"

namespace spmb {

constexpr uint32_t kFastUnk = 0x00FFFFFFu;   // node field of an UNK back record
constexpr int kFastMaxChars = 1024;          // longest sentence (characters) of the fast tier
constexpr int kNormPad = 40;                 // zero bytes guaranteed after a normalized sentence

struct FastView {
  const uint32_t *units;      // global copy of the fast units
  const float *score;         // global: node score (valid where has_leaf)
  const uint32_t *leaf_id;    // global: node -> piece id (valid where has_leaf)
  const uint32_t *units_hot;  // first k_units entries (shared memory on the device)
  const float *score_hot;     // first k_scores entries
  uint32_t k_units, k_scores, n_units;
  int32_t ring;               // R = max piece characters + 1
  float unk_score, smax;
  int32_t unk_id;
};

SPM_HD uint64_t fast_load8(const uint8_t *p) {  // p is 8-byte aligned
  return *reinterpret_cast<const uint64_t *>(p);
}
// bytes [r, r+8) of the 16-byte little-endian pair (lo, hi), r in [0, 8)
SPM_HD uint64_t fast_window(uint64_t lo, uint64_t hi, int r) {
  const int sh = r * 8;
  return (lo >> sh) | ((hi << 1) << (63 - sh));
}

// One sentence.  s: normalized text (8-byte aligned, followed by >= kNormPad zero
// bytes), n bytes.  ring_bt / ring_bk: R slots, element k at [k * rstride].
// back: rows of back records, row c at [c * bstride]; rows 0..nchars are written.
// out: ids in REVERSE order (at most nchars); returns their count.
SPM_HDN int32_t fast_unigram(const FastView &fv, const uint8_t *s, int32_t n, float *ring_bt,
                             uint32_t *ring_bk, int rstride, uint32_t *back, int64_t bstride,
                             int32_t *out, uint32_t *status) {
  if (n <= 0) return 0;
  const int R = fv.ring;
  const float NEG = -INFINITY;
  for (int k = 0; k < R; ++k) ring_bt[k * rstride] = NEG;
  ring_bt[0] = 0.0f;
  ring_bk[0] = 0;
  const uint32_t root = fv.units_hot[0];
  uint32_t flag = 0;
  int32_t p = 0, ci = 0, ridx = 0, q = 0;
  uint64_t a0 = fast_load8(s), a1 = fast_load8(s + 8), a2 = fast_load8(s + 16);
  while (p < n) {
    const float bt_p = ring_bt[ridx * rstride];
    back[(int64_t)ci * bstride] = ring_bk[ridx * rstride];
    ring_bt[ridx * rstride] = NEG;
    const int r = p - q;
    uint64_t cur = fast_window(a0, a1, r);
    int cl = one_char_len((uint8_t)cur);
    if (cl > n - p) cl = n - p;
    int nb = cl, k = 1, d = 0;
    bool single = false;
    uint32_t u = root;
    for (;;) {
      const uint32_t c = (uint32_t)cur & 0xFFu;
      if (!(u & 0x200u) || c == 0) break;
      const uint32_t child = (u >> 10) ^ c;
      u = child < fv.k_units ? fv.units_hot[child] : SPM_LDG(fv.units + child);
      if ((u & 0xFFu) != c || !(u & 0x300u)) break;
      if (d == nb) { nb = d + one_char_len((uint8_t)c); ++k; }
      ++d;
      if (u & 0x100u) {
        const float sc = child < fv.k_scores ? fv.score_hot[child] : SPM_LDG(fv.score + child);
        const float bt = bt_p + sc;
        int slot = ridx + k;
        if (slot >= R) slot -= R;
        const float M = ring_bt[slot * rstride];
        if (bt > M) {
          if (M != NEG && bt - M <= (fmaxf(fabsf(M), fabsf(bt)) + fv.smax) * 9.5367431640625e-07f) flag = ST_CAND_OVERFLOW;
          ring_bt[slot * rstride] = bt;
          ring_bk[slot * rstride] = child | ((uint32_t)k << 24);
        } else if (bt < M) {
          if (M - bt <= (fmaxf(fabsf(M), fabsf(bt)) + fv.smax) * 9.5367431640625e-07f) flag = ST_CAND_OVERFLOW;
        }
        if (k == 1) single = true;
      }
      cur >>= 8;
      if ((d & 7) == 0) {  // crossed into the next 8 bytes of text
        if (d == 8) cur = fast_window(a1, a2, r);
        else {             // pieces longer than 16 bytes: read the text directly
          cur = 0;
          for (int j = 7; j >= 0; --j) cur = (cur << 8) | (p + d + j < n ? s[p + d + j] : 0);
        }
      }
    }
    if (!single) {  // UNK node of one character (unigram_model.cc:598-602)
      const float bt = bt_p + fv.unk_score;
      int slot = ridx + 1;
      if (slot >= R) slot -= R;
      const float M = ring_bt[slot * rstride];
      if (bt > M) {
        if (M != NEG && bt - M <= (fmaxf(fabsf(M), fabsf(bt)) + fv.smax) * 9.5367431640625e-07f) flag = ST_CAND_OVERFLOW;
        ring_bt[slot * rstride] = bt;
        ring_bk[slot * rstride] = kFastUnk | (1u << 24);
      } else if (bt < M) {
        if (M - bt <= (fmaxf(fabsf(M), fabsf(bt)) + fv.smax) * 9.5367431640625e-07f) flag = ST_CAND_OVERFLOW;
      }
    }
    p += cl;
    ++ci;
    if (++ridx == R) ridx = 0;
    while (p - q >= 8) { a0 = a1; a1 = a2; a2 = fast_load8(s + q + 24); q += 8; }
  }
  back[(int64_t)ci * bstride] = ring_bk[ridx * rstride];  // EOS picks the best node ending at n
  // backtrace; runs of unknown pieces collapse to one id (sentencepiece_processor.cc:525-529)
  int32_t cnt = 0;
  bool prev_unk = false;
  while (ci > 0) {
    const uint32_t rec = back[(int64_t)ci * bstride];
    const uint32_t node = rec & 0x00FFFFFFu;
    const bool is_unk = node == kFastUnk;
    if (!(is_unk && prev_unk)) out[cnt++] = is_unk ? fv.unk_id : (int32_t)SPM_LDG(fv.leaf_id + node);
    prev_unk = is_unk;
    ci -= (int32_t)(rec >> 24);
  }
  *status |= flag;
  return cnt;
}

}  // namespace spmb
