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
//     the record of a position is final when the scan arrives there and goes to
//     the thread's own back-pointer array, four records per 16-byte store;
//   * the loop is flattened (one trie step OR one character advance per
//     iteration), so the lanes of a warp never wait for the deepest walk.
// Everything is `__host__ __device__` so tests/host_core.cc can run the very
// same code on the CPU against the oracle.
#pragma once
#include <string.h>

#include "spm_core.cuh"

namespace spmb {

constexpr uint32_t kFastUnk = 0x00FFFFFFu;   // node field of an UNK back record
constexpr int kFastMaxRaw = 4096;            // fast tier: raw sentence length < this (bytes); also the sort key range
constexpr int kMidMaxRaw = 8192;             // lines of kFastMaxRaw <= bytes < this: normalized in pieces, then the fast tier
constexpr int kFastExtraChars = 16;          // fast tier: characters <= raw bytes + this
constexpr int kLongMaxRaw = 65536;           // long lines (kFastMaxRaw <= bytes < this): same kernels, few threads
constexpr int kNormPad = 40;                 // zero bytes guaranteed after a normalized sentence

struct FastView {
  const uint32_t *units;      // global copy of the fast units
  const float *score;         // global: node score (valid where has_leaf)
  const uint32_t *leaf_id;    // global: node -> piece id (valid where has_leaf)
  const uint32_t *units_hot;  // first k_units entries (shared memory on the device)
  const float *score_hot;     // first k_scores entries
  uint32_t units_hot_saddr, score_hot_saddr;  // device: 32-bit shared addresses of the hot copies
  uint32_t k_units, k_scores, n_units;
  int32_t ring;               // R = max piece characters + 1
  float unk_score, smax;
  int32_t unk_id;
};

SPM_HD int ffs64(uint64_t x) {   // 1-based index of the lowest set bit, 0 for x == 0
#if defined(__CUDA_ARCH__)
  return __ffsll((long long)x);
#else
  return x ? __builtin_ctzll(x) + 1 : 0;
#endif
}
SPM_HD uint64_t fast_load8(const uint8_t *p) {  // p is 8-byte aligned
  return *reinterpret_cast<const uint64_t *>(p);
}
// bytes [r, r+8) of the 16-byte little-endian pair (lo, hi), r in [0, 8)
SPM_HD uint64_t fast_window(uint64_t lo, uint64_t hi, int r) {
  const int sh = r * 8;
  return (lo >> sh) | ((hi << 1) << (63 - sh));
}
// bytes [r, r+4) of the 8-byte little-endian pair (lo, hi), r in [0, 4): one PRMT
SPM_HD uint32_t fast_window4(uint32_t lo, uint32_t hi, uint32_t sel) {
#if defined(__CUDA_ARCH__)
  return __byte_perm(lo, hi, sel);
#else
  const int sh = (int)(sel & 7) * 8;   // sel = 0x3210 + 0x1111 * r
  return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#endif
}
// table[idx]: the first `k` entries live in the hot copy (shared memory on the
// device, addressed by its 32-bit shared address `hot_saddr`), the rest in global.
// Two predicated loads, no branch.
template <class T>
SPM_HD T fast_ld_tiered(const T *hot, uint32_t hot_saddr, const T *cold, uint32_t idx, uint32_t k) {
#if defined(__CUDA_ARCH__)
  uint32_t v;
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.lt.u32 p, %1, %2;\n\t"
      "@p ld.shared.u32 %0, [%3];\n\t@!p ld.global.nc.u32 %0, [%4];\n\t}"
      : "=r"(v)
      : "r"(idx), "r"(k), "r"(hot_saddr + idx * 4u), "l"(cold + idx));
  T out;
  memcpy(&out, &v, 4);
  return out;
#else
  (void)hot_saddr;
  return idx < k ? hot[idx] : cold[idx];
#endif
}

// Text prefetch: the next word of the sentence travels global -> shared with cp.async
// (no register is tied up while it is in flight); the host build just loads it.
SPM_HD void fast_prefetch_word(uint32_t *slot, uint32_t slot_saddr, const uint32_t *g) {
#if defined(__CUDA_ARCH__)
  (void)slot;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(slot_saddr), "l"(g) : "memory");
#else
  (void)slot_saddr;
  *slot = *g;
#endif
}
SPM_HD void fast_prefetch_wait() {
#if defined(__CUDA_ARCH__)
  asm volatile("cp.async.wait_all;" ::: "memory");
#endif
}

// Near ties, in line.  A position whose best candidates are within the tolerance of each other keeps up to
// kFastCapc of them (unigram_encode()'s rule: within eps of the running maximum, exact duplicates dropped, in
// begin order), the first in the ring as always, the others in a per-thread SIDE area in global memory -- a
// near tie is met a few times per 10^4 characters, so the area is touched rarely and costs the common path one
// register (`multi`: bit k = ring slot k holds more than one candidate; bit 31 = the current position does)
// and two tests per trie hit.  Round 1 flagged such a sentence and had another kernel run it again: with the
// tolerance growing with the running score, that was nearly every line beyond 2 KB of Japanese text.
// Side area (32-bit words): [0, R) candidate count of the slot, [R, 2R) its running maximum, [2R, 5R) scores and
// [5R, 8R) records of candidates 1..3; [8R, 8R + 4) the current position's count and scores 1..3; then the
// back records of candidates 1..3 of the positions that had them: [8R + 4] = entries, pairs (character index
// << 2 | candidate, record).  Needs R <= 31 and pieces of at most 63 characters (record: trie node, 24 bits |
// characters, 6 bits | the predecessor's candidate, 2 bits); else, or with side == nullptr, the sentence is
// flagged at the first near tie as before.
constexpr int kFastCapc = 4;
constexpr int kFastSideEntries = 56;
constexpr int kFastSideWords = 8 * 31 + 4 + 1 + 2 * kFastSideEntries + 3;   // 368

// One sentence.  s: normalized VALID UTF-8 text (8-byte aligned, followed by >= kNormPad
// zero bytes), n bytes.  ring_bt / ring_bk: R slots, element k at [k * rstride].
// back: this thread's own back-record array (nchars + 2 entries).  side: kFastSideWords words, or nullptr.
// out: ids in REVERSE order (at most nchars); returns their count.
//
// The loop is flattened: every iteration is either one trie step of the walk that
// starts at the current character, or -- when that walk has ended -- the move to
// the next character (all nodes ending there are known by then, so its backtrace
// score is final).  Lanes of a warp therefore never wait for the deepest walk.
// txt: two words of per-thread staging for the text prefetch, element j at [j * rstride].
SPM_HDN int32_t fast_unigram(const FastView &fv, const uint8_t *s, int32_t n, float *ring_bt,
                             uint32_t *ring_bk, uint32_t *txt, int rstride, uint32_t *back, int32_t *out,
                             uint32_t *status, uint32_t *side = nullptr) {
  if (n <= 0) return 0;
  const int R = fv.ring;
  if (R > 31) side = nullptr;
  const float NEG = -3.0e38f;     // "no node ends here yet" (finite, so the tie tolerance stays finite)
  const float TOL = 2.384185791015625e-07f;  // 2^-22, see unigram_encode() in spm_core.cuh
  for (int k = 0; k < R; ++k) ring_bt[k * rstride] = NEG;
  const uint32_t root = fv.units_hot[0];
  const uint32_t *sw = reinterpret_cast<const uint32_t *>(s);
  uint32_t flag = 0;
  uint32_t multi = 0;             // see above
  if (side) side[8 * R + 4] = 0;
  int32_t p = 0, ci = 0, ridx = 0;
  // text window: words w0..w3 = bytes [4*qw, 4*qw + 16), w4 prefetched; r = p - 4*qw in [0, 4)
  int32_t qw = 0;
  uint32_t w0 = sw[0], w1 = sw[1], w2 = sw[2], w3 = sw[3], w4 = sw[4];
  uint32_t r = 0, sel = 0x3210u;
#if defined(__CUDA_ARCH__)
  const uint32_t txt_saddr = (uint32_t)__cvta_generic_to_shared(txt);
#else
  const uint32_t txt_saddr = 0;
#endif
  uint32_t tslot = 0;             // staging slot that holds word qw + 5
  fast_prefetch_word(txt, txt_saddr, sw + 5);
  float bt_p = 0.0f;              // BOS
  // Near-tie tolerance of the walk that starts at the current character.  A push is a near tie
  // when |bt - M| <= 2^-22 (max(|M|,|bt|) + smax) and bt != M (see unigram_encode()); since
  // |bt| <= |bt_p| + smax and a near-tied M is within the tolerance of bt, the per-character
  // bound 2^-22 (|bt_p| + 2 smax + 1) catches a superset -- cheaper; the side path then applies the
  // rule itself.
  float tol_p = (2.0f * fv.smax + 1.0f) * TOL;
  uint32_t cur = w0;
  int cl = one_char_len((uint8_t)cur);
  uint32_t u = root;
  int d = 0, k = 0;
  bool single = false;
  uint32_t *bp = back;            // back record of character ci goes to bp[0]
  // One node (current character -> slot, score sc, record body rec = node | chars << 24).
  auto push = [&](int slot, float sc, uint32_t rec) {
    float bt = bt_p + sc;
    if (multi >> 31) {            // several candidates end at the current character: first strictly greater, in order
      const uint32_t nc = side[8 * R];
      for (uint32_t c = 1; c < nc; ++c) {
        const float v = __uint_as_float_hd(side[8 * R + c]) + sc;
        if (v > bt) { bt = v; rec = (rec & 0x3FFFFFFFu) | (c << 30); }
      }
    }
    const float M = ring_bt[slot * rstride];
    const float diff = bt - M;
    const bool near = fabsf(diff) <= tol_p && diff != 0.0f;
    if (near || ((multi >> slot) & 1u)) {
      if (side == nullptr) {
        flag = ST_CAND_OVERFLOW;
        if (diff > 0.0f) { ring_bt[slot * rstride] = bt; ring_bk[slot * rstride] = rec; }
        return;
      }
      // unigram_encode()'s push on the slot's candidate list
      const bool many = ((multi >> slot) & 1u) != 0;
      const uint32_t cnt = many ? side[slot] : 1u;
      const float Mrun = many ? __uint_as_float_hd(side[R + slot]) : M;
      const float am = fabsf(Mrun), ab = fabsf(bt);
      const float eps = ((am > ab ? am : ab) + fv.smax) * TOL;
      if (bt - Mrun > eps) {
        ring_bt[slot * rstride] = bt; ring_bk[slot * rstride] = rec;
        multi &= ~(1u << slot);
      } else if (!(Mrun - bt > eps)) {
        bool dup = M == bt;       // an equal score loses the first-wins tie for every right score
        for (uint32_t c = 1; c < cnt; ++c)
          if (__uint_as_float_hd(side[2 * R + 3 * slot + (c - 1)]) == bt) dup = true;
        if (!dup) {
          if (cnt == (uint32_t)kFastCapc) {
            flag = ST_CAND_OVERFLOW;
          } else {
            side[2 * R + 3 * slot + (cnt - 1)] = __float_as_uint_hd(bt);
            side[5 * R + 3 * slot + (cnt - 1)] = rec;
            side[slot] = cnt + 1;
            side[R + slot] = __float_as_uint_hd(bt > Mrun ? bt : Mrun);
            multi |= 1u << slot;
          }
        }
      }
      return;
    }
    if (diff > 0.0f) {            // strict '>' keeps the first (longest) of equals
      ring_bt[slot * rstride] = bt;
      ring_bk[slot * rstride] = rec;
    }
  };
  for (;;) {
    const uint32_t c = cur & 0xFFu;
    bool ok = (u & 0x200u) && c != 0;
    uint32_t child = 0, v = 0;
    if (ok) {
      child = (u >> 10) ^ c;
      v = fast_ld_tiered(fv.units_hot, fv.units_hot_saddr, fv.units, child, fv.k_units);
      ok = (v & 0xFFu) == c && (v & 0x300u) != 0;
    }
    if (ok) {
      // ---- one step down the trie
      u = v;
      ++d;
      if ((c & 0xC0u) != 0x80u) ++k;       // characters started so far (text is valid UTF-8)
      cur >>= 8;
      if (v & 0x100u) {                    // a piece ends here: node (ci -> ci + k)
        const float sc = fast_ld_tiered(fv.score_hot, fv.score_hot_saddr, fv.score, child, fv.k_scores);
        int slot = ridx + k;
        if (slot >= R) slot -= R;
        if (k > 63) flag = ST_CAND_OVERFLOW;   // (does not fit the record; such models stay with the exact tier)
        push(slot, sc, child | ((uint32_t)k << 24));
        if (k == 1) single = true;
      }
      if ((d & 3) == 0) {                  // crossed into the next 4 bytes of text
        if (d == 4) cur = fast_window4(w1, w2, sel);
        else if (d == 8) cur = fast_window4(w2, w3, sel);
        else if (d == 12) cur = fast_window4(w3, w4, sel);
        else {                             // pieces longer than 16 bytes: read the text directly
          cur = 0;
          for (int j = 3; j >= 0; --j) cur = (cur << 8) | (p + d + j < n ? s[p + d + j] : 0);
        }
      }
    } else {
      // ---- the walk from character ci is over: move to character ci + 1
      if (!single) {  // UNK node of one character (unigram_model.cc:598-602)
        int slot = ridx + 1;
        if (slot >= R) slot -= R;
        push(slot, fv.unk_score, kFastUnk | (1u << 24));
      }
      p += cl;
      r += cl;
      ++ci;
      ++bp;
      if (++ridx == R) ridx = 0;
      bt_p = ring_bt[ridx * rstride];
      *bp = ring_bk[ridx * rstride];
      ring_bt[ridx * rstride] = NEG;
      multi &= 0x7FFFFFFFu;
      if ((multi >> ridx) & 1u) {          // the candidates of the new position: scores for its pushes, records to the side list
        multi = (multi & ~(1u << ridx)) | 0x80000000u;
        const uint32_t nc = side[ridx];
        side[8 * R] = nc;
        uint32_t ne = side[8 * R + 4];
        for (uint32_t c2 = 1; c2 < nc; ++c2) {
          side[8 * R + c2] = side[2 * R + 3 * ridx + (c2 - 1)];
          if (ne < (uint32_t)kFastSideEntries) {
            side[8 * R + 5 + 2 * ne] = ((uint32_t)ci << 2) | c2;
            side[8 * R + 6 + 2 * ne] = side[5 * R + 3 * ridx + (c2 - 1)];
            ++ne;
          } else {
            flag = ST_CAND_OVERFLOW;
          }
        }
        side[8 * R + 4] = ne;
      }
      if (p >= n) break;
      tol_p = (fabsf(bt_p) + 2.0f * fv.smax + 1.0f) * TOL;
      if (r >= 4) {                        // slide the window by one word; word qw + 5 was requested one slide ago
        r -= 4; ++qw;
        fast_prefetch_wait();
        w0 = w1; w1 = w2; w2 = w3; w3 = w4;
        w4 = txt[tslot * rstride];
        tslot ^= 1;
        fast_prefetch_word(txt + tslot * rstride, txt_saddr + tslot * rstride * 4, sw + qw + 5);
      }
      sel = 0x3210u + 0x1111u * r;
      cur = fast_window4(w0, w1, sel);
      cl = one_char_len((uint8_t)cur);
      u = root; d = 0; k = 0; single = false;
    }
  }
  // EOS (score 0) picks among the candidates that end at the last character (first strictly greater)
  uint32_t pick = 0;
  if (multi >> 31) {
    float best = bt_p + 0.0f;
    const uint32_t nc = side[8 * R];
    for (uint32_t c2 = 1; c2 < nc; ++c2) {
      const float v2 = __uint_as_float_hd(side[8 * R + c2]) + 0.0f;
      if (v2 > best) { best = v2; pick = c2; }
    }
  }
  // backtrace; runs of unknown pieces collapse to one id (sentencepiece_processor.cc:525-529)
  int32_t cnt = 0;
  bool prev_unk = false;
  int32_t se = side ? (int32_t)side[8 * R + 4] : 0;   // side entries are in ascending character order: search from the end
  while (ci > 0) {
    uint32_t rec = back[ci];
    if (pick != 0) {
      const uint32_t key = ((uint32_t)ci << 2) | pick;
      while (se > 0 && side[8 * R + 5 + 2 * (se - 1)] > key) --se;
      rec = (se > 0 && side[8 * R + 5 + 2 * (se - 1)] == key) ? side[8 * R + 6 + 2 * (se - 1)] : rec;
    }
    const uint32_t node = rec & 0x00FFFFFFu;
    const bool is_unk = node == kFastUnk;
    if (!(is_unk && prev_unk)) out[cnt++] = is_unk ? fv.unk_id : (int32_t)SPM_LDG(fv.leaf_id + node);
    prev_unk = is_unk;
    ci -= (int32_t)((rec >> 24) & 0x3Fu);
    pick = rec >> 30;
  }
  *status |= flag;
  return cnt;
}


// ------------------------------------------------------------------ BPE fast tier
constexpr int kBpeMaxSyms = 20;    // symbols (characters) per whitespace-delimited word
constexpr int kBpeMaxBytes = 48;   // bytes per word
constexpr int kBpeWordsPerThread = kBpeMaxSyms * 2 + kBpeMaxBytes / 4 + (kBpeMaxSyms + 4) / 4;  // smem words

// Per-thread scratch of one word, element j of every array at [j * stride].
struct BpeScratch {
  uint32_t *node;   // [kBpeMaxSyms] trie node of symbol i (kFastUnk: not a trie path)
  float *ps;        // [kBpeMaxSyms] score of the pair (symbol i, next alive symbol) or NEG
  uint32_t *text;   // [kBpeMaxBytes / 4] bytes of the word
  uint32_t *cb;     // [(kBpeMaxSyms + 4) / 4] byte offset of symbol i inside the word (u8 each)
  int stride;
};
SPM_HD uint32_t bpe_get_u8(const uint32_t *a, int stride, int j) {
  return (a[(j >> 2) * stride] >> ((j & 3) * 8)) & 0xFFu;
}
SPM_HD void bpe_set_u8(uint32_t *a, int stride, int j, uint32_t v) {
  uint32_t &w = a[(j >> 2) * stride];
  const int sh = (j & 3) * 8;
  w = (w & ~(0xFFu << sh)) | (v << sh);
}

// Walks bytes [b, e) of the word from trie node `nd` (unit word `u`).  Returns false when
// the path leaves the trie.
SPM_HD bool bpe_walk(const FastView &fv, const BpeScratch &sc, uint32_t *nd, uint32_t *u, int b, int e) {
  for (int j = b; j < e; ++j) {
    const uint32_t c = bpe_get_u8(sc.text, sc.stride, j);
    if (!(*u & 0x200u)) return false;
    const uint32_t child = (*u >> 10) ^ c;
    const uint32_t v = fast_ld_tiered(fv.units_hot, fv.units_hot_saddr, fv.units, child, fv.k_units);
    if ((v & 0xFFu) != c || !(v & 0x300u)) return false;
    *u = v; *nd = child;
  }
  return true;
}
// score of the piece (symbol at node nd) + bytes [b, e), or NEG when it is not a piece
SPM_HD float bpe_pair_score(const FastView &fv, const BpeScratch &sc, uint32_t nd, int b, int e, float NEG) {
  if (nd == kFastUnk) return NEG;
  uint32_t u = fast_ld_tiered(fv.units_hot, fv.units_hot_saddr, fv.units, nd, fv.k_units);
  if (!bpe_walk(fv, sc, &nd, &u, b, e) || !(u & 0x100u)) return NEG;
  return fast_ld_tiered(fv.score_hot, fv.score_hot_saddr, fv.score, nd, fv.k_scores);
}

// One sentence with a BPE model whose pieces never carry U+2581 at a non-initial position
// (so no merge can cross a word boundary) and that has no USER_DEFINED / UNUSED pieces:
// bpe::Model::Encode (src/bpe_model.cc:37-199) word by word.  Inside a word the agenda
// semantics are restated exactly as "merge the adjacent pair that forms a piece with the
// highest score, leftmost on ties, until none is left" (the agenda's live entries are
// exactly the current adjacent pairs; stale ones are skipped, :148-152).
// s: normalized valid UTF-8 text (escaped whitespace), padded as for fast_unigram.
// out: ids in FORWARD order.  A word longer than the scratch sets ST_TOO_LONG (the
// caller re-runs the sentence on the general tier).
// txt: two words of per-thread staging for the text prefetch (as in fast_unigram).
SPM_HDN int32_t fast_bpe(const FastView &fv, const uint8_t *s, int32_t n, const BpeScratch &sc, uint32_t *txt,
                         int32_t *out, uint32_t *status) {
  const float NEG = -3.0e38f;
  const uint32_t root = fv.units_hot[0];
  const uint32_t *sw = reinterpret_cast<const uint32_t *>(s);
  // text window (see fast_unigram): words w0..w4 = bytes [4*qw, 4*qw + 20), r = q - 4*qw
  int32_t qw = 0;
  uint32_t w0 = sw[0], w1 = sw[1], w2 = sw[2], w3 = sw[3], w4 = sw[4];
  uint32_t r = 0;
#if defined(__CUDA_ARCH__)
  const uint32_t txt_saddr = (uint32_t)__cvta_generic_to_shared(txt);
#else
  const uint32_t txt_saddr = 0;
#endif
  uint32_t tslot = 0;
  fast_prefetch_word(txt, txt_saddr, sw + 5);
  int32_t cnt = 0, p = 0;
  bool prev_unk = false;
  while (p < n) {
    // ---- split the word into characters; single-character nodes and initial pair scores
    int nsym = 0, wlen = 0;
    uint32_t prev_nd = kFastUnk, prev_u = 0;
    for (;;) {
      if (p + wlen >= n) break;
      const uint32_t cur = fast_window4(w0, w1, 0x3210u + 0x1111u * r);   // 4 bytes at p + wlen
      if (nsym > 0 && (cur & 0x00FFFFFFu) == 0x008196E2u) break;          // U+2581: the next word
      const int cl = one_char_len((uint8_t)cur);
      if (nsym == kBpeMaxSyms || wlen + cl > kBpeMaxBytes) { *status |= ST_TOO_LONG; return 0; }
      for (int j = 0; j < cl; ++j) bpe_set_u8(sc.text, sc.stride, wlen + j, (cur >> (8 * j)) & 0xFFu);
      bpe_set_u8(sc.cb, sc.stride, nsym, (uint32_t)wlen);
      // node of the character alone, and of (previous symbol + this character)
      uint32_t nd = 0, u = root, pn = prev_nd, pu = prev_u;
      bool ok1 = true, ok2 = prev_nd != kFastUnk;
      for (int j = 0; j < cl; ++j) {
        const uint32_t c = (cur >> (8 * j)) & 0xFFu;
        if (ok1) {
          ok1 = (u & 0x200u) != 0;
          if (ok1) {
            const uint32_t child = (u >> 10) ^ c;
            const uint32_t v = fast_ld_tiered(fv.units_hot, fv.units_hot_saddr, fv.units, child, fv.k_units);
            ok1 = (v & 0xFFu) == c && (v & 0x300u) != 0;
            u = v; nd = child;
          }
        }
        if (ok2) {
          ok2 = (pu & 0x200u) != 0;
          if (ok2) {
            const uint32_t child = (pu >> 10) ^ c;
            const uint32_t v = fast_ld_tiered(fv.units_hot, fv.units_hot_saddr, fv.units, child, fv.k_units);
            ok2 = (v & 0xFFu) == c && (v & 0x300u) != 0;
            pu = v; pn = child;
          }
        }
      }
      if (!ok1) nd = kFastUnk;
      sc.node[nsym * sc.stride] = nd;
      if (nsym > 0) {
        float v = NEG;
        if (ok2 && (pu & 0x100u)) v = fast_ld_tiered(fv.score_hot, fv.score_hot_saddr, fv.score, pn, fv.k_scores);
        sc.ps[(nsym - 1) * sc.stride] = v;
      }
      prev_nd = nd; prev_u = u;
      ++nsym;
      wlen += cl;
      r += cl;
      if (r >= 4) {
        r -= 4; ++qw;
        fast_prefetch_wait();
        w0 = w1; w1 = w2; w2 = w3; w3 = w4;
        w4 = txt[tslot * sc.stride];
        tslot ^= 1;
        fast_prefetch_word(txt + tslot * sc.stride, txt_saddr + tslot * sc.stride * 4, sw + qw + 5);
      }
    }
    sc.ps[(nsym - 1) * sc.stride] = NEG;
    // ---- merges
    uint32_t alive = nsym >= 32 ? 0xFFFFFFFFu : ((1u << nsym) - 1u);
    for (;;) {
      float best = NEG;
      int bi = -1;
      for (uint32_t m = alive; m;) {
        int i = 0;
        while (!((m >> i) & 1u)) ++i;
        m &= m - 1;
        if (!m) break;                       // the last alive symbol has no right neighbour
        const float v = sc.ps[i * sc.stride];
        if (v > best) { best = v; bi = i; }  // strict '>' keeps the leftmost of equals (:55-61)
      }
      if (bi < 0) break;
      int r = bi + 1;
      while (!((alive >> r) & 1u)) ++r;
      int nr = r + 1;                        // symbol after r (nsym: none)
      while (nr < nsym && !((alive >> nr) & 1u)) ++nr;
      const int rb = (int)bpe_get_u8(sc.cb, sc.stride, r);
      const int re = nr < nsym ? (int)bpe_get_u8(sc.cb, sc.stride, nr) : wlen;
      uint32_t nd = sc.node[bi * sc.stride];
      uint32_t u = fast_ld_tiered(fv.units_hot, fv.units_hot_saddr, fv.units, nd, fv.k_units);
      bpe_walk(fv, sc, &nd, &u, rb, re);     // succeeds: the pair was scored
      sc.node[bi * sc.stride] = nd;
      alive &= ~(1u << r);
      // the two new pairs (:167-168)
      float v = NEG;
      if (nr < nsym) {
        int ne = nr + 1;
        while (ne < nsym && !((alive >> ne) & 1u)) ++ne;
        const int nre = ne < nsym ? (int)bpe_get_u8(sc.cb, sc.stride, ne) : wlen;
        uint32_t n2 = nd, u2 = u;
        if (bpe_walk(fv, sc, &n2, &u2, re, nre) && (u2 & 0x100u))
          v = fast_ld_tiered(fv.score_hot, fv.score_hot_saddr, fv.score, n2, fv.k_scores);
      }
      sc.ps[bi * sc.stride] = v;
      int pl = bi - 1;
      while (pl >= 0 && !((alive >> pl) & 1u)) --pl;
      if (pl >= 0)
        sc.ps[pl * sc.stride] = bpe_pair_score(fv, sc, sc.node[pl * sc.stride],
                                               (int)bpe_get_u8(sc.cb, sc.stride, bi), re, NEG);
    }
    // ---- emit (PieceToId, src/model_interface.cc:87-97; unk runs merge, processor.cc:525-529)
    for (int i = 0; i < nsym; ++i) {
      if (!((alive >> i) & 1u)) continue;
      const uint32_t nd = sc.node[i * sc.stride];
      bool is_unk = nd == kFastUnk;
      if (!is_unk) {
        const uint32_t u = fast_ld_tiered(fv.units_hot, fv.units_hot_saddr, fv.units, nd, fv.k_units);
        is_unk = !(u & 0x100u);
      }
      if (!(is_unk && prev_unk)) out[cnt++] = is_unk ? fv.unk_id : (int32_t)SPM_LDG(fv.leaf_id + nd);
      prev_unk = is_unk;
    }
    p += wlen;
  }
  return cnt;
}

// ------------------------------------------------------------------ normaliser
// 8 bytes at the 8-byte aligned address a, never touching memory outside [lo, hi)
// (the caller's text buffer): whole-word load when it fits, guarded bytes at the edges.
SPM_HD uint64_t fast_load8_guard(const uint8_t *a, const uint8_t *lo, const uint8_t *hi) {
  if (a >= lo && a + 8 <= hi) return *reinterpret_cast<const uint64_t *>(a);
  uint64_t v = 0;
  for (int j = 7; j >= 0; --j) v = (v << 8) | ((a + j >= lo && a + j < hi) ? a[j] : 0);
  return v;
}

// Output byte stream: accumulates into a 64-bit register, stores aligned 8-byte words.
struct FastOut {
  uint8_t *base;     // 8-byte aligned slot
  uint64_t acc;
  int32_t len;       // bytes appended so far
  int32_t cap;       // capacity of the slot
  SPM_HD bool room(int64_t more) const { return (int64_t)len + more <= cap; }   // can `more` bytes still follow?
  SPM_HD bool done() const { return false; }
  static constexpr bool kRuns = true;            // runs of plain bytes are taken from the window in one piece
  static constexpr bool kFirst3 = true;          // consult the per-character rule filter (text sinks: CJK lines)
  // `run` (1..8) printable-ASCII bytes from the low end of w, appended as they are (until run v77 this sink took
  // them one by one through the general step: ~40 instructions per byte on a lone thread, 0.17 ms for the 6 k
  // sentences the word memo leaves over per 10 M)
  SPM_HD int take(uint64_t w, int run) {
    const int f = len & 7;
    const uint64_t v = run == 8 ? w : (w & ((1ull << (8 * run)) - 1));
    acc |= v << (8 * f);
    if (f + run >= 8) {
      *reinterpret_cast<uint64_t *>(base + (len & ~7)) = acc;
      acc = f ? v >> (8 * (8 - f)) : 0;
    }
    len += run;
    return run;
  }
  SPM_HD void at(int32_t) {}                     // (input position of the step: only AlignedOut records it)
  SPM_HD void put(uint32_t v, int k) {  // k in 1..4 bytes, little-endian in v
    const int f = len & 7;
    acc |= (uint64_t)v << (8 * f);
    if (f + k >= 8) {
      *reinterpret_cast<uint64_t *>(base + (len & ~7)) = acc;
      acc = f + k > 8 ? (uint64_t)v >> (8 * (8 - f)) : 0;
    }
    len += k;
  }
  SPM_HD void flush() {
    if (len & 7) *reinterpret_cast<uint64_t *>(base + (len & ~7)) = acc;
  }
};

// FastOut that also records norm_to_orig (src/normalizer.cc:132-176: every normalized byte
// maps to `consumed`, the input offset at which its NormalizePrefix step started).
struct AlignedOut : FastOut {
  int32_t *n2o;
  int32_t cur;
  static constexpr bool kRuns = true;   // runs of plain bytes are taken from the window in one piece
  static constexpr bool kFirst3 = true;
  SPM_HD void at(int32_t p) { cur = p; }
  SPM_HD void put(uint32_t v, int k) {
    for (int j = 0; j < k; ++j) n2o[len + j] = cur;
    FastOut::put(v, k);
  }
  // `run` (1..8) printable-ASCII bytes from the low end of w: each is a NormalizePrefix step of
  // its own (one byte in, the same byte out), so byte j maps to input offset cur + j.
  SPM_HD int take(uint64_t w, int run) {
    for (int j = 0; j < run; ++j) n2o[len + j] = cur + j;
    return FastOut::take(w, run);
  }
};

// Whole-sentence normalisation (src/normalizer.cc:88-211), same results as
// normalize_sentence() of spm_core.cuh, restructured for one thread per sentence:
// the raw text is read through a register window, the common case (no charsmap
// rule / user symbol can start at this byte pair: 64 Kbit filter `first2`) is
// resolved from the window alone, and output goes out as aligned 8-byte words.
// in[0..n) lies inside the buffer [lo, hi).  out: 8-byte aligned slot of capacity
// `cap` bytes + kNormPad + 8.  Returns the normalized length (the kNormPad bytes
// after it are zeroed), or -1 when the text does not fit (caller: general tier).
// *nchars_ub: upper bound of the character count.
// The scan itself, writing every normalized character to `o` (FastOut: the padded text
// slot; MemoSink of spm_memo.cuh: straight into word keys).  Returns false when the sink
// ran out of room; *nchars_ub counts the characters written.
template <class Sink>
SPM_HDN bool fast_normalize_scan(const ModelView &m, const uint32_t *first2, const uint8_t *in, int64_t n,
                                 const uint8_t *lo, const uint8_t *hi, Sink &o, int32_t *nchars_ub,
                                 bool *content,     // *content: something other than leading spaces was seen
                                 bool mid = false) {  // in[] continues a line whose content has started (norm_find_cut)
  const bool esc = m.escape_ws, rws = m.remove_extra_ws;
  const uint32_t WS = esc ? 0x8196E2u : 0x20u;  // U+2581 little-endian, or ' '
  const int WSL = esc ? 3 : 1;
  const uint8_t *q = (const uint8_t *)((uintptr_t)in & ~(uintptr_t)7);
  int r = (int)(in - q);
  uint64_t a0 = fast_load8_guard(q, lo, hi), a1 = fast_load8_guard(q + 8, lo, hi);
  int32_t nch = 0;
  bool lead = rws && !mid;        // still stripping leading spaces (:103-112)
  bool started = mid;             // dummy prefix emitted
  bool prev_space = rws && !mid;  // is_prev_space
  bool overflow = false;
  int64_t p = 0;
  while (p < n) {
    uint64_t w = fast_window(a0, a1, r);
    if (Sink::kRuns && m.ascii_plain && started && !(w & 0x80u)) {   // (a non-ASCII byte first: no run to look for)
      // A run of printable ASCII (0x21..0x7E) inside a word: no rule, user symbol or whitespace
      // handling applies to a byte whose successor is ASCII as well (HostModel::ascii_plain),
      // so the run goes to the sink in one piece.  SWAR: 0x80 marks bytes < 0x21 or > 0x7E
      // (borrows / carries only disturb bytes above the first marked one).
      const uint64_t np = (((w - 0x2121212121212121ull) & ~w) | ((w + 0x0101010101010101ull) | w)) & 0x8080808080808080ull;
      int run = np ? (ffs64(np) - 1) >> 3 : 8;
      if (run > n - p) run = (int)(n - p);
      // the successor of the last byte must be known to be ASCII (or the end of the sentence)
      if (run == 8 || (run < n - p && ((w >> (8 * run)) & 0x80u))) --run;
      if (run > 0) {
        o.at((int32_t)p);
        run = o.take(w, run);
        if (run > 0) {
          lead = false; prev_space = false;
          nch += run; p += run; r += run;
          while (r >= 8) { a0 = a1; q += 8; a1 = fast_load8_guard(q + 8, lo, hi); r -= 8; }
          if (p >= n) break;
          w = fast_window(a0, a1, r);
        }
      }
    }
    const uint32_t b0 = (uint32_t)w & 0xFF, b1 = ((uint32_t)w >> 8) & 0xFF;
    int consumed;
    // ---- what does NormalizePrefix return here?
    bool slow = m.user.n != 0 && ((m.user_first[b0 >> 5] >> (b0 & 31)) & 1);
    if (!slow && m.cm_n != 0) {
      const uint32_t key = (b0 << 8) | b1;
      slow = n - p < 2 || ((first2[key >> 5] >> (key & 31)) & 1);  // first2 may live in shared memory
      // a 3-byte character (kana, CJK, ...): first2 passes its whole block of 64 code points; first3 knows the character
      if (Sink::kFirst3 && slow && m.cm_first3 != nullptr && n - p >= 3 && (b0 & 0xF0u) == 0xE0u && (b1 & 0xC0u) == 0x80u &&
          ((uint32_t)(w >> 16) & 0xC0u) == 0x80u) {
        const uint32_t cp = ((b0 & 15u) << 12) | ((b1 & 63u) << 6) | ((uint32_t)(w >> 16) & 63u);
        slow = (SPM_LDG(m.cm_first3 + (cp >> 5)) >> (cp & 31)) & 1;
      }
    }
    if (!slow) {
      // one character straight from the window (src/normalizer.cc:271-291)
      uint32_t v; int k;
      if (b0 < 0x80) { v = b0; k = 1; }
      else {
        int mb;
        uint8_t tmp[4] = {(uint8_t)b0, (uint8_t)b1, (uint8_t)(w >> 16), (uint8_t)(w >> 24)};
        if (valid_decode_utf8(tmp, n - p, &mb)) { k = mb; v = (uint32_t)w & (mb == 4 ? 0xFFFFFFFFu : ((1u << (8 * mb)) - 1)); }
        else { k = 3; v = 0xBDBFEFu; mb = 1; }   // U+FFFD, one byte consumed
        consumed = mb;
      }
      if (b0 < 0x80) consumed = 1;
      const bool is_space = b0 == 0x20;
      if (lead && is_space) { p += 1; r += 1; goto advance; }   // np == " ": keep stripping
      lead = false;
      o.at((int32_t)p);
      if (!started) {  // first real content: dummy prefix (:132-153)
        started = true;
        if (!m.ws_suffix && m.add_dummy_prefix) { o.put(WS, WSL); ++nch; }
      }
      if (is_space) {
        if (!prev_space) { o.put(WS, WSL); ++nch; prev_space = true; }
      } else {
        o.put(v, k); ++nch;
        prev_space = false;
      }
      if (!rws) prev_space = false;
      p += consumed; r += consumed;
    } else {
      // a rule or a user-defined symbol may start here: the general step
      const uint8_t *rep; int rl;
      consumed = normalize_prefix(m, in + p, n - p, &rep, &rl);
      const bool one_space = rep != nullptr && rl == 1 && rep[0] == ' ';
      if (lead && one_space) { p += consumed; r += consumed; goto advance; }
      lead = false;
      o.at((int32_t)p);
      if (!started) {
        started = true;
        if (!m.ws_suffix && m.add_dummy_prefix) { o.put(WS, WSL); ++nch; }
      }
      if (rep == nullptr) {
        o.put(0xBDBFEFu, 3); ++nch;
        prev_space = false;
      } else {
        int k = 0;
        while (prev_space && k < rl && rep[k] == ' ') ++k;
        if (k < rl) {
          // keep room for the worst case of everything still to come (3 bytes per raw byte + suffix)
          if (!o.room(3 * (rl - k) + 3 * (n - p - consumed) + 3)) { overflow = true; break; }
          for (; k < rl; ++k) {
            const uint8_t b = rep[k];
            if (b == ' ') o.put(WS, WSL); else o.put(b, 1);
            ++nch;
          }
          prev_space = rep[rl - 1] == ' ';
        }
      }
      if (!rws) prev_space = false;
      p += consumed; r += consumed;
    }
  advance:
    if (o.done()) break;   // the sink gave up on the sentence
    while (r >= 8) { a0 = a1; q += 8; a1 = fast_load8_guard(q + 8, lo, hi); r -= 8; }
  }
  *nchars_ub = nch;
  *content = started;
  return !overflow;
}

// pc_plain_length() (spm_core.cuh) eight bytes at a time: aligned words of the batch's text [lo, hi), bytes
// outside the sentence replaced by 'a'.  A byte outside 0x20..0x7E marks its bit 7 (or one further up: only
// "any" is asked); spaces are counted with a popcount, a repeated space is two neighbouring marks.
SPM_HDN int32_t pc_plain_length_swar(const ModelView &m, const uint8_t *s, int64_t n, const uint8_t *lo, const uint8_t *hi) {
  if (n <= 0 || n > (1 << 24)) return -1;
  if (s[0] == 0x20u || s[n - 1] == 0x20u) return -1;
  const uint64_t H = 0x8080808080808080ull, L7 = 0x7F7F7F7F7F7F7F7Full, A = 0x6161616161616161ull;
  const uint8_t *q = (const uint8_t *)((uintptr_t)s & ~(uintptr_t)7);
  const uint8_t *e = s + n;
  int32_t spaces = 0;
  uint64_t prev_sp = 0;   // bit 63: the byte before this word was a space
  for (; q < e; q += 8) {
    uint64_t w = fast_load8_guard(q, lo, hi);
    if (q < s) { const int k = (int)(s - q); w = (w & (~0ull << (8 * k))) | (A & ~(~0ull << (8 * k))); }
    if (q + 8 > e) { const int k = (int)(e - q); w = (w & ~(~0ull << (8 * k))) | (A & (~0ull << (8 * k))); }
    const uint64_t bad = (((w - 0x2020202020202020ull) & ~w) | ((w + 0x0101010101010101ull) | w)) & H;
    if (bad) return -1;
    const uint64_t x = w ^ 0x2020202020202020ull;
    const uint64_t sp = ~(((x & L7) + L7) | x) & H;          // bit 7 of every space
    if (sp & ((sp << 8) | (prev_sp >> 56))) return -1;       // a space after a space
    prev_sp = sp;
    spaces += (int32_t)
#if defined(__CUDA_ARCH__)
        __popcll(sp);
#else
        __builtin_popcountll(sp);
#endif
  }
  return (int32_t)n + 2 * spaces + (m.add_dummy_prefix ? 3 : 0);
}

// Long lines are normalized in pieces (k_normalize_long).  A piece may start at `p` when the
// scan is guaranteed to arrive there with a fresh match and a state that does not matter:
//   * in[p] starts a character that never sits inside a rule key / user symbol (cm_cont2),
//     so no match that began earlier can reach over p;
//   * no rule / user symbol starts at it (first2, user_first) and it is not a space, so it
//     is copied as it is whatever the whitespace state was, and leaves prev_space = false.
// Returns the first such position >= from, or n.
SPM_HD int64_t norm_find_cut(const ModelView &m, const uint32_t *first2, const uint8_t *in, int64_t from, int64_t n) {
  for (int64_t p = from; p < n; ++p) {
    const uint32_t b0 = in[p];
    if ((b0 & 0xC0u) == 0x80u || b0 == 0x20u) continue;
    if (p + 1 >= n) continue;                       // the scan treats a last lone byte as "slow"
    const uint32_t key = (b0 << 8) | in[p + 1];
    if (m.cm_cont2 != nullptr && ((m.cm_cont2[key >> 5] >> (key & 31)) & 1)) continue;
    if (m.cm_n != 0 && ((first2[key >> 5] >> (key & 31)) & 1)) continue;
    if (m.user.n != 0 && ((m.user_first[b0 >> 5] >> (b0 & 31)) & 1)) continue;
    return p;
  }
  return n;
}

// Piece `idx` = in[a..b) of a line cut by norm_find_cut: normalized into `slot` (8-byte
// aligned, 3 * (b - a) + 72 + 8 bytes).  *content is only meaningful for piece 0.
#ifndef SPM_NORM_CHUNK
#define SPM_NORM_CHUNK 128
#endif
constexpr int kNormChunk = SPM_NORM_CHUNK;   // raw bytes per piece (before the cut is moved to a safe position)
SPM_HD int64_t norm_piece_slot(int64_t a, int idx) { return (3 * a + 88 * (int64_t)idx + 7) & ~(int64_t)7; }
SPM_HDN bool norm_piece(const ModelView &m, const uint32_t *first2, const uint8_t *in, int64_t a, int64_t b,
                        const uint8_t *lo, const uint8_t *hi, uint8_t *slot, int idx, int32_t *out_len,
                        int32_t *nchars, bool *content) {
  FastOut o;
  o.base = slot; o.acc = 0; o.len = 0; o.cap = (int32_t)(3 * (b - a) + 72);
  const bool ok = fast_normalize_scan(m, first2, in + a, b - a, lo, hi, o, nchars, content, idx > 0);
  o.flush();
  *out_len = o.len;
  return ok;
}
// After the pieces are concatenated in out[0..len): what fast_normalize() does behind its scan.
// raw = raw bytes of the line (slot capacity 3 * raw + 72).  Returns the final length, -1 when
// the line does not fit the fast tier's bounds.  The caller zeroes out[len .. align8(len) + kNormPad).
SPM_HD int32_t norm_finish(const ModelView &m, uint8_t *out, int32_t len, int32_t nch, bool started, int64_t raw) {
  if (len + 3 > 3 * raw + 72) return -1;
  if (m.remove_extra_ws) {
    if (m.escape_ws) { while (len >= 3 && out[len - 3] == 0xE2 && out[len - 2] == 0x96 && out[len - 1] == 0x81) { len -= 3; --nch; } }
    else { while (len >= 1 && out[len - 1] == ' ') { len -= 1; --nch; } }
  }
  if (m.ws_suffix && m.add_dummy_prefix && started) {
    if (m.escape_ws) { out[len] = 0xE2; out[len + 1] = 0x96; out[len + 2] = 0x81; len += 3; } else { out[len++] = ' '; }
    ++nch;
  }
  if (nch > raw + kFastExtraChars) return -1;
  return len;
}

SPM_HDN int32_t fast_normalize(const ModelView &m, const uint32_t *first2, const uint8_t *in, int64_t n,
                               const uint8_t *lo, const uint8_t *hi, uint8_t *out, int32_t cap,
                               int32_t *nchars_ub) {
  *nchars_ub = 0;
  if (n == 0) return 0;
  const bool esc = m.escape_ws, rws = m.remove_extra_ws;
  const uint32_t WS = esc ? 0x8196E2u : 0x20u;
  const int WSL = esc ? 3 : 1;
  FastOut o;
  o.base = out; o.acc = 0; o.len = 0; o.cap = cap;
  int32_t nch = 0;
  bool started = false;
  if (!fast_normalize_scan(m, first2, in, n, lo, hi, o, &nch, &started)) return -1;
  o.flush();
  int32_t len = o.len;
  if (rws) {  // strip trailing whitespace symbols (:191-201): re-read our own tail
    if (esc) { while (len >= 3 && out[len - 3] == 0xE2 && out[len - 2] == 0x96 && out[len - 1] == 0x81) { len -= 3; --nch; } }
    else { while (len >= 1 && out[len - 1] == ' ') { len -= 1; --nch; } }
  }
  if (m.ws_suffix && m.add_dummy_prefix && started) {
    o.len = len;
    o.acc = (len & 7) ? *reinterpret_cast<uint64_t *>(out + (len & ~7)) & ((1ull << (8 * (len & 7))) - 1) : 0;
    o.put(WS, WSL); ++nch;
    o.flush();
    len = o.len;
  }
  // zero padding: finish the current word, then kNormPad more bytes
  {
    const int f = len & 7;
    int32_t z = len;
    if (f) {
      uint64_t *wp = reinterpret_cast<uint64_t *>(out + (len & ~7));
      *wp &= (1ull << (8 * f)) - 1;
      z = (len & ~7) + 8;
    }
    for (int32_t k = 0; k < kNormPad; k += 8) *reinterpret_cast<uint64_t *>(out + z + k) = 0;
  }
  *nchars_ub = nch;
  return len;
}


// fast_normalize() that also fills norm_to_orig: n2o[0..len] (one entry per normalized byte and
// the final one, src/normalizer.cc:191-206).  Same contract otherwise; -1: caller falls back to
// normalize_sentence_t<true>.
SPM_HDN int32_t fast_normalize_aligned(const ModelView &m, const uint32_t *first2, const uint8_t *in, int64_t n,
                                       const uint8_t *lo, const uint8_t *hi, uint8_t *out, int32_t cap, int32_t *n2o) {
  if (n == 0) return 0;
  const bool esc = m.escape_ws, rws = m.remove_extra_ws;
  const uint32_t WS = esc ? 0x8196E2u : 0x20u;
  const int WSL = esc ? 3 : 1;
  AlignedOut o;
  o.base = out; o.acc = 0; o.len = 0; o.cap = cap; o.n2o = n2o; o.cur = 0;
  int32_t nch = 0;
  bool started = false;
  if (!fast_normalize_scan(m, first2, in, n, lo, hi, o, &nch, &started)) return -1;
  o.flush();
  int32_t len = o.len;
  int32_t consumed = (int32_t)n;
  if (rws) {  // strip trailing whitespace symbols; consumed = norm_to_orig[new length] (:191-201)
    if (esc) { while (len >= 3 && out[len - 3] == 0xE2 && out[len - 2] == 0x96 && out[len - 1] == 0x81) { len -= 3; consumed = n2o[len]; } }
    else { while (len >= 1 && out[len - 1] == ' ') { len -= 1; consumed = n2o[len]; } }
  }
  if (m.ws_suffix && m.add_dummy_prefix && started) {
    o.len = len;
    o.acc = (len & 7) ? *reinterpret_cast<uint64_t *>(out + (len & ~7)) & ((1ull << (8 * (len & 7))) - 1) : 0;
    o.at(consumed);
    o.put(WS, WSL);
    o.flush();
    len = o.len;
  }
  if (started || len > 0) n2o[len] = consumed;
  return len;
}

}  // namespace spmb
