// spm_core.cuh -- per-sentence algorithms of the B200 batch encoder.
//
// Everything here is `__host__ __device__` so that the very same arithmetic
// that runs inside the CUDA kernels (spm_kernels.cu) can be compiled for the
// host by the *test suite* (tests/host_core.cc) and compared with the oracle
// in the GPU-less build container.  The shipped library never runs these on
// the host.
//
// Reformulations relative to the reference (behaviour is bit-identical):
//   * piece matching walks our own flat double array ("XDA", built by
//     spm_model.cc) whose units carry has_leaf/has_child bits and whose leaf
//     table carries (id, type, score) -- replaces Darts commonPrefixSearch +
//     pieces(id) lookups (third_party/darts_clone/darts.h:467-513,
//     src/unigram_model.cc:572-595);
//   * the unigram lattice is never materialised: nodes are pushed into a ring
//     of pending end positions as the trie walk finds them and Viterbi is
//     evaluated on the fly with *pruned candidate lists* that provably select
//     the same predecessor as the reference's strict-'>' scan over
//     end_nodes_[pos] (src/unigram_model.cc:222-261); see unigram_encode().
//   * BPE keeps the reference's agenda semantics (max score, then smallest
//     left index; lazy invalidation by size; src/bpe_model.cc:55-61,143-169)
//     on a flat binary heap.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define SPM_HD __host__ __device__ __forceinline__
#define SPM_HDN __host__ __device__
#else
#define SPM_HD inline
#define SPM_HDN inline
#endif

#if !defined(__CUDACC__)
struct uint2 { uint32_t x, y; };
#endif

namespace spmb {

// piece types (src/sentencepiece_model.proto:243-250)
enum : uint32_t { T_NORMAL = 1, T_UNKNOWN = 2, T_CONTROL = 3, T_USER_DEFINED = 4, T_UNUSED = 5 };
enum : int32_t { M_UNIGRAM = 1, M_BPE = 2, M_WORD = 3, M_CHAR = 4 };

// per-sentence status bits written by the kernels
enum : uint32_t {
  ST_OK = 0,
  ST_NORM_OVERFLOW = 1,   // normalized text did not fit the tier's capacity
  ST_IDS_OVERFLOW = 2,    // ids did not fit the tier's staging capacity
  ST_CAND_OVERFLOW = 4,   // more near-tied Viterbi candidates than the tier keeps
  ST_TOO_LONG = 8,        // sentence exceeds the tier's length limit
  ST_INTERNAL = 16,       // reference would return util::error::INTERNAL
};

// XDA unit: [7:0] label  [8] has_leaf  [9] has_child  [31:10] base.
// child(node, c) = base(node) ^ c, valid iff label == c and (leaf|child) != 0.
// Bases are unique per parent, which makes the label test sufficient.
struct TrieView {
  const uint32_t *units;
  const uint2 *leaf;   // parallel to units: .x = id | type << 28, .y = score bits
  uint32_t n;
};
struct ModelView {
  TrieView pieces;           // NORMAL / USER_DEFINED / UNUSED and reserved pieces
  TrieView user;             // user-defined symbols (n == 0: none)
  const uint32_t *cm_units;  // charsmap darts units (model data, uploaded verbatim)
  const uint8_t *cm_pool;    // NUL-delimited replacement strings
  const uint32_t *cm_first2; // 65536-bit filter: can a rule start with (b0,b1)?
  const uint32_t *cm_cont2;  // 65536-bit filter: can a character starting (b0,b1) sit INSIDE a rule key / user symbol?
  const uint32_t *cm_first3; // 65536-bit filter by code point U+0800..U+FFFF: does a rule key start with this 3-byte character? (may be null)
  uint32_t cm_n;             // 0: identity normalization
  uint32_t user_first[8];    // 256-bit filter on the first byte of user symbols
  int32_t model_type;
  int32_t unk_id;
  float unk_score;           // min_score_ - kUnkPenalty (unigram_model.cc:550-551)
  float max_score;           // max(FLT_MIN, max NORMAL score) (unigram_model.cc:683-689)
  float smax;                // bound on |score| of any lattice node (for pruning)
  int32_t ring;              // power of two >= max piece bytes + 5
  uint8_t add_dummy_prefix, remove_extra_ws, escape_ws, ws_suffix;
  uint8_t has_unused;        // any UNUSED piece (BPE resegmentation needed)
  uint8_t ascii_plain;       // bytes 0x21..0x7E followed by an ASCII byte never start a rule / user symbol
  uint8_t ws_ctl;            // TAB, LF, FF and CR each normalize to exactly one space whatever follows (word scan: separators)
  uint8_t pad_[1];
};

#if defined(__CUDA_ARCH__)
SPM_HD float __uint_as_float_hd(uint32_t x) { return __uint_as_float(x); }
SPM_HD uint32_t __float_as_uint_hd(float x) { return __float_as_uint(x); }
#else
SPM_HD float __uint_as_float_hd(uint32_t x) { union { uint32_t u; float f; } v; v.u = x; return v.f; }
SPM_HD uint32_t __float_as_uint_hd(float x) { union { uint32_t u; float f; } v; v.f = x; return v.u; }
#endif

// ------------------------------------------------------------------ UTF-8
// src/util.h:389-391
SPM_HD int one_char_len(uint8_t b) { return (int)((0x4322111111111111ull >> ((b >> 4) * 4)) & 0xF); }
SPM_HD bool is_trail(uint8_t x) { return (int8_t)x < -0x40; }  // src/util.h:401
SPM_HD bool is_valid_cp(uint32_t c) { return c < 0xD800u || (c >= 0xE000u && c <= 0x10FFFFu); }

// src/util.cc:187-221 + src/util.h:459-462: returns true and *mblen when the
// prefix is one valid char (a literal U+FFFD counts as valid); else *mblen = 1.
SPM_HD bool valid_decode_utf8(const uint8_t *b, int64_t len, int *mblen) {
  const uint32_t b0 = b[0];
  if (b0 < 0x80) { *mblen = 1; return true; }
  if (len >= 2 && (b0 & 0xE0) == 0xC0) {
    const uint32_t cp = ((b0 & 0x1F) << 6) | (b[1] & 0x3F);
    if (is_trail(b[1]) && cp >= 0x80 && is_valid_cp(cp)) { *mblen = 2; return true; }
  } else if (len >= 3 && (b0 & 0xF0) == 0xE0) {
    const uint32_t cp = ((b0 & 0x0F) << 12) | ((uint32_t)(b[1] & 0x3F) << 6) | (b[2] & 0x3F);
    if (is_trail(b[1]) && is_trail(b[2]) && cp >= 0x800 && is_valid_cp(cp)) { *mblen = 3; return true; }
  } else if (len >= 4 && (b0 & 0xF8) == 0xF0) {
    const uint32_t cp = ((b0 & 0x07) << 18) | ((uint32_t)(b[1] & 0x3F) << 12) |
                        ((uint32_t)(b[2] & 0x3F) << 6) | (b[3] & 0x3F);
    if (is_trail(b[1]) && is_trail(b[2]) && is_trail(b[3]) && cp >= 0x10000 && is_valid_cp(cp)) {
      *mblen = 4; return true;
    }
  }
  *mblen = 1;
  return false;
}

// ------------------------------------------------------------------ tries
#if defined(__CUDA_ARCH__)
#define SPM_LDG(p) __ldg(p)
#else
#define SPM_LDG(p) (*(p))
#endif

SPM_HD bool xda_valid(uint32_t u, uint32_t c) { return (u & 0xFF) == c && (u & 0x300) != 0; }

// Longest key of `t` that is a prefix of s[0..n): 0 when none.
SPM_HD int xda_longest_prefix(const TrieView &t, const uint8_t *s, int64_t n) {
  uint32_t u = SPM_LDG(t.units);
  int best = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (!(u & 0x200)) break;
    const uint32_t c = s[i];
    const uint32_t child = (u >> 10) ^ c;
    u = SPM_LDG(t.units + child);
    if (!xda_valid(u, c)) break;
    if (u & 0x100) best = (int)i + 1;
  }
  return best;
}

// Exact match of s[0..n): leaf word (.x id|type<<28, .y score) or .x = ~0u.
SPM_HD uint2 xda_exact(const TrieView &t, const uint8_t *s, int64_t n) {
  uint32_t u = SPM_LDG(t.units);
  uint32_t pos = 0;
  uint2 miss; miss.x = 0xFFFFFFFFu; miss.y = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (!(u & 0x200)) return miss;
    const uint32_t c = s[i];
    pos = (u >> 10) ^ c;
    u = SPM_LDG(t.units + pos);
    if (!xda_valid(u, c)) return miss;
  }
  if (n == 0 || !(u & 0x100)) return miss;
  return SPM_LDG(t.leaf + pos);
}

// darts-clone unit (data format of the model's precompiled charsmap;
// third_party/darts_clone/darts.h:50-80)
SPM_HD uint32_t du_offset(uint32_t u) { return (u >> 10) << ((u & (1u << 9)) >> 6); }
SPM_HD uint32_t du_label(uint32_t u) { return u & ((1u << 31) | 0xFF); }

// Longest charsmap key that prefixes s[0..n): returns its length (0: none) and
// the pool offset of its replacement (src/normalizer.cc:258-268: the longest
// result wins; results are unique per length).
SPM_HD int charsmap_longest(const ModelView &m, const uint8_t *s, int64_t n, uint32_t *value) {
  const uint32_t *a = m.cm_units;
  uint32_t pos = 0;
  uint32_t u = SPM_LDG(a);
  pos ^= du_offset(u);
  int best = 0;
  for (int64_t i = 0; i < n; ++i) {
    const uint32_t c = s[i];
    pos ^= c;
    if (pos >= m.cm_n) break;  // malformed blob guard
    u = SPM_LDG(a + pos);
    if (du_label(u) != c) break;
    pos ^= du_offset(u);
    if ((u >> 8) & 1) {
      if (pos >= m.cm_n) break;
      best = (int)i + 1;
      *value = SPM_LDG(a + pos) & 0x7FFFFFFFu;
    }
  }
  return best;
}

// normalizer.cc:362-384 restricted to what callers use: longest user-defined
// symbol prefixing s (0: none).
SPM_HD int user_prefix_match(const ModelView &m, const uint8_t *s, int64_t n) {
  if (m.user.n == 0) return 0;
  const uint32_t b0 = s[0];
  if (!((m.user_first[b0 >> 5] >> (b0 & 31)) & 1)) return 0;
  return xda_longest_prefix(m.user, s, n);
}

// ------------------------------------------------------------------ normalizer
// One leftmost-longest step (src/normalizer.cc:231-300).  Returns the number of
// input bytes consumed; the replacement is [*rep, *rep + *rep_len).
SPM_HD int normalize_prefix(const ModelView &m, const uint8_t *in, int64_t n, const uint8_t **rep,
                            int *rep_len) {
  const int ul = user_prefix_match(m, in, n);
  if (ul > 0) { *rep = in; *rep_len = ul; return ul; }  // raw bytes, un-normalised (:239-243)
  if (m.cm_n != 0) {
    const uint32_t b0 = in[0];
    // filter: bit (b0,b1) set iff some rule is b0 alone or starts with b0 b1
    const uint32_t key = (b0 << 8) | (n >= 2 ? (uint32_t)in[1] : 0u);
    bool maybe = true;
    if (m.cm_first2 != nullptr && n >= 2) maybe = (SPM_LDG(m.cm_first2 + (key >> 5)) >> (key & 31)) & 1;
    if (maybe) {
      uint32_t value = 0;
      const int l = charsmap_longest(m, in, n, &value);
      if (l > 0) {
        const uint8_t *r = m.cm_pool + value;
        int rl = 0;
        while (SPM_LDG(r + rl) != 0) ++rl;  // NUL-terminated (:293-296)
        *rep = r; *rep_len = rl;
        return l;
      }
    }
  }
  int mblen = 1;
  if (!valid_decode_utf8(in, n, &mblen)) { *rep = nullptr; *rep_len = 3; return 1; }  // U+FFFD, 1 byte
  *rep = in; *rep_len = mblen;
  return mblen;
}

// Whole-sentence normalisation (src/normalizer.cc:88-211) into out[0..cap).
// Returns the normalized length, or -1 when it does not fit (tier overflow).
// ALIGN: also writes norm_to_orig (n2o[0..len], one entry per normalized byte
// plus the final one; :132-153,:166-176,:191-206) -- ids do not need it, the
// pieces / SentencePieceText output does.
template <bool ALIGN>
SPM_HDN int32_t normalize_sentence_t(const ModelView &m, const uint8_t *in, int64_t n, uint8_t *out,
                                     int64_t cap, int32_t *n2o) {
  if (n == 0) return 0;
  const uint8_t *rep; int rl;
  int32_t consumed = 0;
  if (m.remove_extra_ws) {  // strip leading spaces (:103-112)
    while (n > 0) {
      const int c = normalize_prefix(m, in, n, &rep, &rl);
      if (!(rl == 1 && rep != nullptr && rep[0] == ' ')) break;
      in += c; n -= c; consumed += c;
    }
  }
  if (n == 0) return 0;
  int64_t o = 0;
  const bool esc = m.escape_ws;
  auto put_ws = [&]() -> bool {   // add_ws (:132-153) and the escaped space (:166-172)
    if (esc) {
      if (o + 3 > cap) return false;
      out[o] = 0xE2; out[o + 1] = 0x96; out[o + 2] = 0x81;
      if (ALIGN) { n2o[o] = consumed; n2o[o + 1] = consumed; n2o[o + 2] = consumed; }
      o += 3;
    } else {
      if (o + 1 > cap) return false;
      if (ALIGN) n2o[o] = consumed;
      out[o++] = ' ';
    }
    return true;
  };
  if (!m.ws_suffix && m.add_dummy_prefix && !put_ws()) return -1;
  bool is_prev_space = m.remove_extra_ws;
  while (n > 0) {
    const int c = normalize_prefix(m, in, n, &rep, &rl);
    if (rep == nullptr) {  // malformed byte -> U+FFFD
      if (o + 3 > cap) return -1;
      out[o] = 0xEF; out[o + 1] = 0xBF; out[o + 2] = 0xBD;
      if (ALIGN) { n2o[o] = consumed; n2o[o + 1] = consumed; n2o[o + 2] = consumed; }
      o += 3;
      is_prev_space = false;
    } else {
      int k = 0;
      while (is_prev_space && k < rl && rep[k] == ' ') ++k;  // ConsumePrefix(" ") loop
      if (k < rl) {
        for (; k < rl; ++k) {
          const uint8_t b = rep[k];
          if (esc && b == ' ') {
            if (!put_ws()) return -1;
          } else {
            if (o + 1 > cap) return -1;
            if (ALIGN) n2o[o] = consumed;
            out[o++] = b;
          }
        }
        is_prev_space = rep[rl - 1] == ' ';
      }
    }
    in += c; n -= c; consumed += c;
    if (!m.remove_extra_ws) is_prev_space = false;
  }
  if (m.remove_extra_ws) {  // strip trailing whitespace symbols (:191-201); consumed = n2o[new length]
    if (esc) {
      while (o >= 3 && out[o - 3] == 0xE2 && out[o - 2] == 0x96 && out[o - 1] == 0x81) {
        o -= 3;
        if (ALIGN) consumed = n2o[o];
      }
    } else {
      while (o >= 1 && out[o - 1] == ' ') {
        o -= 1;
        if (ALIGN) consumed = n2o[o];
      }
    }
  }
  if (m.ws_suffix && m.add_dummy_prefix && !put_ws()) return -1;
  if (ALIGN) n2o[o] = consumed;   // :204
  return (int32_t)o;
}
SPM_HDN int32_t normalize_sentence(const ModelView &m, const uint8_t *in, int64_t n, uint8_t *out,
                                   int64_t cap) {
  return normalize_sentence_t<false>(m, in, n, out, cap, nullptr);
}

// ------------------------------------------------------------------ unigram
// Scratch of one sentence for unigram_encode<CAPC>.
struct UniScratch {
  float *ringM;        // [ring]          max backtrace score pending at an end position
  uint32_t *ringCnt;   // [ring]          live candidates at that position
  float *candBt;       // [ring * CAPC]   candidate backtrace scores, in begin order
  uint64_t *candRec;   // [ring * CAPC]   id | len << 32 | prev slot << 48
  uint64_t *back;      // [CAPC][stride]  records of visited positions (slot-major)
  int64_t stride;      // >= n + 1
};
SPM_HD int64_t uni_scratch_words(int ring, int capc, int64_t n) {  // in 4-byte words
  return (int64_t)ring * 2 + (int64_t)ring * capc * 3 + (int64_t)capc * (n + 1) * 2 + 2;
}

SPM_HD uint64_t pack_rec(uint32_t id, uint32_t len, uint32_t prevslot) {
  return (uint64_t)id | ((uint64_t)len << 32) | ((uint64_t)prevslot << 48);
}

// PopulateNodes + Viterbi + backtrace of one normalized sentence
// (src/unigram_model.cc:535-604, :222-261, :705-720) followed by the unk-run
// merge of src/sentencepiece_processor.cc:525-529.
//
// Positions are byte offsets of character starts (characters delimited by
// OneCharLen jumps clamped to the sentence, unigram_model.cc:155-161).  For the
// position p being expanded, every node (p -> e) the reference would insert is
// scored immediately: bt = fl(bt_l* + score) with l* chosen exactly as the
// reference's scan over end_nodes_[p] (ascending begin, strict '>').  Instead
// of the full end list we keep, per pending end position, only candidates that
// can still be chosen by SOME right node:
//   - a candidate whose bt is below the running max M by more than
//     eps = 2^-22 (max(|M|,|bt|) + smax) loses to M for every right score s:
//     rounding is monotonic, so fl(x1+s) <= fl(x2+s) for x1 < x2, and the two sums
//     can only round to the SAME float y when x2 - x1 <= ulp(y) <= 2^-23 |y| with
//     |y| <= (max(|x1|,|x2|) + |s|)(1 + 2^-24) -- eps is twice that bound (round 1
//     used 2^-20, which flagged four times as many sentences in the one-candidate
//     kernel for nothing: 91 % of the 4 KB Japanese lines);
//   - a candidate whose bt equals an earlier kept candidate's bt yields the
//     same sum and loses the first-wins tie.
// Kept candidates stay in begin order, so the scan over them returns the
// reference's choice.  More than CAPC near-ties set ST_CAND_OVERFLOW (the
// caller reruns the sentence with a larger CAPC).
//
// Writes ids (forward order) to out[0..cap) and returns their count; when they
// do not fit, sets ST_IDS_OVERFLOW and still returns the count.  out_end (may be
// null): the byte offset in s where each emitted piece ends (a merged unk run
// ends where its last character ends) -- piece k is s[out_end[k-1], out_end[k]).
// (CAPC is a run-time value here; the wrappers below fix it at compile time for the common
// tier and pass the model's ring size -- which no candidate list can exceed -- for the retry.)
SPM_HD int32_t unigram_encode_impl(const ModelView &m, const uint8_t *s, int32_t n, const UniScratch &sc,
                                   const uint32_t CAPC, int32_t *out, int32_t cap, uint32_t *status,
                                   int32_t *out_end) {
  if (n <= 0) return 0;
  const uint32_t mask = (uint32_t)m.ring - 1;
  for (int i = 0; i < m.ring; ++i) sc.ringCnt[i] = 0;
  sc.ringCnt[0] = 1; sc.ringM[0] = 0.0f; sc.candBt[0] = 0.0f; sc.candRec[0] = pack_rec(0, 0, 0);  // BOS
  uint32_t st = 0;

  auto push = [&](int32_t e, float bt, uint32_t id, uint32_t len, uint32_t prevslot) {
    const uint32_t r = (uint32_t)e & mask;
    const uint32_t cnt = sc.ringCnt[r];
    const uint64_t rec = pack_rec(id, len, prevslot);
    if (cnt == 0) {
      sc.ringM[r] = bt; sc.candBt[r * CAPC] = bt; sc.candRec[r * CAPC] = rec; sc.ringCnt[r] = 1;
      return;
    }
    const float M = sc.ringM[r];
    const float am = fabsf(M), ab = fabsf(bt);
    const float eps = ((am > ab ? am : ab) + m.smax) * 2.384185791015625e-07f;
    if (bt - M > eps) {
      sc.ringM[r] = bt; sc.candBt[r * CAPC] = bt; sc.candRec[r * CAPC] = rec; sc.ringCnt[r] = 1;
    } else if (M - bt > eps) {
      // can never be selected
    } else {
      for (uint32_t k = 0; k < cnt; ++k)
        if (sc.candBt[r * CAPC + k] == bt) return;  // loses the first-wins tie for every s
      if (cnt == CAPC) { st |= ST_CAND_OVERFLOW; return; }
      sc.candBt[r * CAPC + cnt] = bt; sc.candRec[r * CAPC + cnt] = rec; sc.ringCnt[r] = cnt + 1;
      if (bt > M) sc.ringM[r] = bt;
    }
  };

  int32_t p = 0;
  while (p < n) {
    const uint32_t r = (uint32_t)p & mask;
    const uint32_t cnt = sc.ringCnt[r];
    for (uint32_t k = 0; k < cnt; ++k) sc.back[(int64_t)k * sc.stride + p] = sc.candRec[r * CAPC + k];
    // best predecessor for a right node of score sr (unigram_model.cc:229-246)
    auto pick = [&](float sr, uint32_t *slot) {
      float best = sc.candBt[r * CAPC] + sr;
      uint32_t bs = 0;
      for (uint32_t k = 1; k < cnt; ++k) {
        const float v = sc.candBt[r * CAPC + k] + sr;
        if (v > best) { best = v; bs = k; }
      }
      *slot = bs;
      return best;
    };
    int cl = one_char_len(s[p]);
    if (cl > n - p) cl = n - p;
    int32_t nb = p + cl;   // end of the character that contains the byte being walked
    int nchars = 1;
    bool has_single = false;
    uint32_t u = SPM_LDG(m.pieces.units);
    for (int32_t i = p; i < n; ++i) {
      if (!(u & 0x200)) break;
      const uint32_t c = s[i];
      const uint32_t child = (u >> 10) ^ c;
      u = SPM_LDG(m.pieces.units + child);
      if (!xda_valid(u, c)) break;
      if (i == nb) {  // byte i starts the next character
        int l2 = one_char_len((uint8_t)c);
        if (l2 > n - i) l2 = n - i;
        nb = i + l2;
        ++nchars;
      }
      if (u & 0x100) {
        const uint2 lf = SPM_LDG(m.pieces.leaf + child);
        const uint32_t type = lf.x >> 28;
        if (type == T_NORMAL || type == T_USER_DEFINED) {  // UNUSED skipped (:586); reserved not in trie_
          // user-defined bonus: float product, double add, float store (:590-591)
          const float sr = type == T_USER_DEFINED ? (float)((double)((float)nchars * m.max_score) + 1.0)
                                                  : __uint_as_float_hd(lf.y);
          uint32_t slot;
          const float bt = pick(sr, &slot);
          push(nb, bt, lf.x & 0x0FFFFFFFu, (uint32_t)(nb - p), slot);
          if (nchars == 1) has_single = true;
        }
      }
    }
    if (!has_single) {  // UNK node of one character (:598-602)
      uint32_t slot;
      const float bt = pick(m.unk_score, &slot);
      push(p + cl, bt, (uint32_t)m.unk_id, (uint32_t)cl, slot);
    }
    sc.ringCnt[r] = 0;
    p += cl;
  }
  // EOS (score 0) picks among the candidates ending at n
  const uint32_t r = (uint32_t)n & mask;
  const uint32_t cnt = sc.ringCnt[r];
  for (uint32_t k = 0; k < cnt; ++k) sc.back[(int64_t)k * sc.stride + n] = sc.candRec[r * CAPC + k];
  uint32_t eos_slot = 0;
  {
    float best = sc.candBt[r * CAPC] + 0.0f;
    for (uint32_t k = 1; k < cnt; ++k) {
      const float v = sc.candBt[r * CAPC + k] + 0.0f;
      if (v > best) { best = v; eos_slot = k; }
    }
  }
  // backtrace twice: count (with unk-run merge), then write from the back
  int32_t total = 0;
  for (int pass = 0; pass < 2; ++pass) {
    int32_t pos = n, idx = total;
    uint32_t slot = eos_slot;
    int64_t pending = -1;
    int32_t pend_end = n;
    while (pos > 0) {
      const uint64_t rec = sc.back[(int64_t)slot * sc.stride + pos];
      const int32_t id = (int32_t)(uint32_t)rec;
      const bool merge = pending >= 0 && id == m.unk_id && (int32_t)pending == m.unk_id;
      if (pending >= 0 && !merge) {
        if (pass == 0) ++total;
        else { out[--idx] = (int32_t)pending; if (out_end) out_end[idx] = pend_end; }
      }
      if (!merge) pend_end = pos;
      pending = id;
      pos -= (int32_t)((rec >> 32) & 0xFFFF);
      slot = (uint32_t)(rec >> 48) & 0xFF;
    }
    if (pending >= 0) {
      if (pass == 0) ++total;
      else { out[--idx] = (int32_t)pending; if (out_end) out_end[idx] = pend_end; }
    }
    if (pass == 0 && total > cap) { st |= ST_IDS_OVERFLOW; break; }
  }
  *status |= st;
  return total;
}
template <int CAPC>
SPM_HDN int32_t unigram_encode(const ModelView &m, const uint8_t *s, int32_t n, const UniScratch &sc,
                               int32_t *out, int32_t cap, uint32_t *status, int32_t *out_end = nullptr) {
  return unigram_encode_impl(m, s, n, sc, (uint32_t)CAPC, out, cap, status, out_end);
}
// Any capacity up to 255 (the back records keep the predecessor's slot in 8 bits).  With
// capc >= ring no sentence can overflow: the candidates pending at one end position begin at
// distinct earlier positions of the ring.
SPM_HDN int32_t unigram_encode_capc(const ModelView &m, const uint8_t *s, int32_t n, const UniScratch &sc,
                                    int capc, int32_t *out, int32_t cap, uint32_t *status,
                                    int32_t *out_end = nullptr) {
  return unigram_encode_impl(m, s, n, sc, (uint32_t)capc, out, cap, status, out_end);
}

// ------------------------------------------------------------------ BPE
// Scratch of one sentence for bpe_encode (4-byte words): symbols (5 ns), agenda
// heap (4 cap), UNUSED-merge log (3 cap), with ns <= n symbols and cap = 3 ns + 4.
SPM_HD int64_t bpe_scratch_words(int64_t n) { return 5 * n + 7 * (3 * n + 4) + 2; }

// src/bpe_model.cc:37-199 followed by the post-processing of
// src/sentencepiece_processor.cc:488-551 (unk-run merge; a CONTROL id leaves
// normalized bytes unconsumed => INTERNAL error, no ids).
// Returns the id count (ids forward in out), or -1 for the INTERNAL case.
// out_end (may be null): as in unigram_encode.
SPM_HDN int32_t bpe_encode(const ModelView &m, const uint8_t *s, int32_t n, int32_t *w, int32_t *out,
                           int32_t cap, uint32_t *status, int32_t *out_end = nullptr) {
  if (n <= 0) return 0;
  int32_t *begin = w, *size = w + n, *prev = w + 2 * n, *next = w + 3 * n, *freeze = w + 4 * n;
  const int32_t hcap = 3 * n + 4;
  float *hscore = (float *)(w + 5 * n);
  int32_t *hleft = w + 5 * n + hcap, *hright = hleft + hcap, *hsize = hright + hcap;
  int32_t *lbegin = hsize + hcap, *lsize = lbegin + hcap, *lleft = lsize + hcap;
  int32_t hn = 0, ln = 0, ns = 0;

  // split into symbols: one char each, or one frozen user-defined symbol (:121-131)
  for (int32_t p = 0; p < n;) {
    int mb = user_prefix_match(m, s + p, n - p);
    const int fr = mb > 0;
    if (!fr) { mb = one_char_len(s[p]); if (mb > n - p) mb = n - p; }
    begin[ns] = p; size[ns] = mb; prev[ns] = ns - 1; next[ns] = ns + 1; freeze[ns] = fr;
    ++ns; p += mb;
  }
  next[ns - 1] = -1;

  // agenda order (:55-61): larger score first, then smaller left index
  auto before = [&](int a, int b) {
    return hscore[a] > hscore[b] || (hscore[a] == hscore[b] && hleft[a] < hleft[b]);
  };
  auto hswap = [&](int a, int b) {
    float fs = hscore[a]; hscore[a] = hscore[b]; hscore[b] = fs;
    int32_t t = hleft[a]; hleft[a] = hleft[b]; hleft[b] = t;
    t = hright[a]; hright[a] = hright[b]; hright[b] = t;
    t = hsize[a]; hsize[a] = hsize[b]; hsize[b] = t;
  };
  // MaybeAddNewSymbolPair (:94-118)
  auto maybe_add = [&](int32_t l, int32_t r) {
    if (l < 0 || r < 0 || freeze[l] || freeze[r]) return;
    const int32_t sz = size[l] + size[r];
    const uint2 lf = xda_exact(m.pieces, s + begin[l], sz);
    if (lf.x == 0xFFFFFFFFu) return;
    const uint32_t type = lf.x >> 28;
    if (!(type == T_NORMAL || type == T_USER_DEFINED || type == T_UNUSED)) return;  // pieces_ only
    if (hn >= hcap) { *status |= ST_INTERNAL; return; }
    int i = hn++;
    hscore[i] = __uint_as_float_hd(lf.y); hleft[i] = l; hright[i] = r; hsize[i] = sz;
    while (i > 0) {
      const int par = (i - 1) >> 1;
      if (!before(i, par)) break;
      hswap(i, par); i = par;
    }
    if (type == T_UNUSED && ln < hcap) { lbegin[ln] = begin[l]; lsize[ln] = sz; lleft[ln] = size[l]; ++ln; }
  };
  for (int32_t i = 1; i < ns; ++i) maybe_add(i - 1, i);
  while (hn > 0) {  // main loop (:143-169)
    const int32_t l = hleft[0], r = hright[0], sz = hsize[0];
    --hn;
    if (hn > 0) {
      hswap(0, hn);
      int i = 0;
      for (;;) {
        int a = 2 * i + 1, b = a + 1, best = i;
        if (a < hn && before(a, best)) best = a;
        if (b < hn && before(b, best)) best = b;
        if (best == i) break;
        hswap(i, best); i = best;
      }
    }
    if (size[l] == 0 || size[r] == 0 || size[l] + size[r] != sz) continue;  // stale (:148-152)
    size[l] = sz;
    next[l] = next[r];
    if (next[r] >= 0) prev[next[r]] = l;
    size[r] = 0;
    maybe_add(prev[l], l);
    maybe_add(l, next[l]);
  }

  // emit: PieceToId (model_interface.cc:87-97), UNUSED pieces split back through
  // the last-recorded merge of the same string (rev_merge, :171-189)
  int32_t total = 0;
  bool is_prev_unk = false, control_seen = false;
  auto emit = [&](int32_t id, uint32_t type, int32_t end) {
    const bool is_unk = type == T_UNKNOWN;
    if (type == T_CONTROL) {
      control_seen = true;
    } else if (!(is_prev_unk && is_unk)) {
      if (total < cap) { out[total] = id; if (out_end) out_end[total] = end; }
      ++total;
    } else if (out_end && total - 1 < cap) {
      out_end[total - 1] = end;   // the unk run grows (sentencepiece_processor.cc:525-529)
    }
    is_prev_unk = is_unk;
  };
  int32_t *stb = (int32_t *)hscore, *sts = hleft;  // explicit stack reuses the (now empty) agenda
  for (int32_t k = 0; k != -1; k = next[k]) {
    int sp = 0;
    stb[0] = begin[k]; sts[0] = size[k]; sp = 1;
    while (sp > 0) {
      --sp;
      const int32_t b = stb[sp], z = sts[sp];
      const uint2 lf = xda_exact(m.pieces, s + b, z);
      int32_t id = m.unk_id;
      uint32_t type = T_UNKNOWN;
      if (lf.x != 0xFFFFFFFFu) { id = (int32_t)(lf.x & 0x0FFFFFFFu); type = lf.x >> 28; }
      int32_t split = -1;
      if (type == T_UNUSED) {
        for (int32_t j = ln - 1; j >= 0 && split < 0; --j) {
          if (lsize[j] != z) continue;
          bool eq = true;
          for (int32_t q = 0; q < z && eq; ++q) eq = s[lbegin[j] + q] == s[b + q];
          if (eq) split = lleft[j];
        }
      }
      if (split < 0) { emit(id, type, b + z); continue; }
      if (sp + 2 > hcap) { *status |= ST_INTERNAL; break; }
      stb[sp] = b + split; sts[sp] = z - split; ++sp;   // right first, left on top
      stb[sp] = b; sts[sp] = split; ++sp;
    }
  }
  if (control_seen) { *status |= ST_INTERNAL; return -1; }
  if (total > cap) *status |= ST_IDS_OVERFLOW;
  return total;
}


// ------------------------------------------------------------------ extra options
// ApplyExtraOptions (src/sentencepiece_processor.cc:945-979) folded into one
// layout: final = pre[0..npre) ++ (flip ? reversed(core) : core) ++ suf[0..nsuf).
struct ExtraPlan {
  int32_t npre, nsuf, flip;
  int32_t pre[16], suf[16];
};
SPM_HD ExtraPlan make_extra_plan(uint32_t opts, int32_t bos_id, int32_t eos_id) {
  ExtraPlan p;
  p.npre = p.nsuf = p.flip = 0;
  for (uint32_t o = opts; o & 3u; o >>= 2) {
    const uint32_t c = o & 3u;
    if (c == 1u) {  // bos: insert at the front
      for (int i = p.npre; i > 0; --i) p.pre[i] = p.pre[i - 1];
      p.pre[0] = bos_id; ++p.npre;
    } else if (c == 2u) {  // eos: append
      p.suf[p.nsuf++] = eos_id;
    } else {  // reverse the whole sequence
      ExtraPlan q;
      q.npre = p.nsuf; q.nsuf = p.npre; q.flip = p.flip ^ 1;
      for (int i = 0; i < p.nsuf; ++i) q.pre[i] = p.suf[p.nsuf - 1 - i];
      for (int i = 0; i < p.npre; ++i) q.suf[i] = p.pre[p.npre - 1 - i];
      p = q;
    }
  }
  return p;
}
// number of final ids of a sentence whose model output has `core` ids
// (core < 0: the reference returned an error and no ids at all)
SPM_HD int32_t final_count(const ExtraPlan &p, int32_t core) { return core < 0 ? 0 : core + p.npre + p.nsuf; }
SPM_HD int32_t final_id_at(const ExtraPlan &p, const int32_t *core_ids, int32_t core, int32_t j) {
  if (j < p.npre) return p.pre[j];
  j -= p.npre;
  if (j < core) return core_ids[p.flip ? core - 1 - j : j];
  return p.suf[j - core];
}


// SentencePieceText view of final piece j (PopulateSentencePieceText + ApplyExtraOptions,
// src/sentencepiece_processor.cc:488-551,945-979): the id, its span [nbegin, nend) in the
// normalized text and the span [begin, end) of its surface in the input (norm_to_orig
// at both ends).  The bos / eos pieces of the extra options carry no span: nbegin = nend
// = -1, begin = end = 0 (unset proto fields).
struct PieceSpan { int32_t id, begin, end, nbegin, nend; };
SPM_HD PieceSpan final_piece_at(const ExtraPlan &p, const int32_t *core_ids, const int32_t *core_end,
                                const int32_t *n2o, int32_t core, int32_t j) {
  PieceSpan r;
  r.begin = r.end = 0; r.nbegin = r.nend = -1;
  if (j < p.npre) { r.id = p.pre[j]; return r; }
  j -= p.npre;
  if (j >= core) { r.id = p.suf[j - core]; return r; }
  const int32_t k = p.flip ? core - 1 - j : j;
  r.id = core_ids[k];
  r.nbegin = k > 0 ? core_end[k - 1] : 0;
  r.nend = core_end[k];
  r.begin = n2o[r.nbegin];
  r.end = n2o[r.nend];
  return r;
}


// ---- spans from ids: the fast half of the pieces path --------------------------------
// The id path (word memo / Viterbi / merge kernels) yields the final ids of a sentence but
// not where each piece lies.  Every piece other than the unknown one has a fixed string, so
// walking the ids over the normalized text recovers the boundaries; only a run of unknown
// characters has no fixed length.  Its end e* is found by elimination: every character of
// the run has no usable one-character piece (unigram: PopulateNodes adds the UNK node only
// then, src/unigram_model.cc:596-602; BPE: PieceToId of the symbol misses,
// src/model_interface.cc:87-97) and the next piece's string starts at e*.  e* itself always
// satisfies both, so when exactly ONE position does, it is e*; with two or more (or any
// inconsistency: a piece that does not fit, text left over) the function returns false and
// the sentence is redone by the exact tier, which records the piece ends.
struct PieceTable {
  const int32_t *plen;    // bytes of piece id
  const int32_t *olen;    // bytes of its surface in a plain sentence (every U+2581 = one space): plen - 2 * count(U+2581)
  const uint32_t *poff;   // offset of its string in pool
  const uint8_t *pool;
};
SPM_HD bool unk_eligible(const ModelView &m, const uint8_t *c, int cl) {
  const uint2 lf = xda_exact(m.pieces, c, cl);
  if (lf.x == 0xFFFFFFFFu) return true;
  const uint32_t type = lf.x >> 28;
  return m.model_type == M_UNIGRAM ? !(type == T_NORMAL || type == T_USER_DEFINED) : type == T_UNKNOWN;
}
// s[0..nl): normalized sentence, n2o its norm_to_orig; fids[0..fc): the FINAL ids of the
// sentence (extra options applied).  Writes the four span arrays at the same final positions
// (entries >= room are skipped).
SPM_HDN bool spans_from_ids(const ModelView &m, const PieceTable &pt, const ExtraPlan &plan, const uint8_t *s,
                            int32_t nl, const int32_t *n2o, const int32_t *fids, int32_t fc, int32_t *begin,
                            int32_t *end, int32_t *nbegin, int32_t *nend, int64_t room) {
  if (fc == 0) return true;   // nothing to place (empty input without extra options, or the reference's error)
  const int32_t core = fc - plan.npre - plan.nsuf;
  if (core < 0) return false;
  for (int32_t j = 0; j < fc; ++j) {
    if (j >= plan.npre && j < plan.npre + core) continue;
    if (j < room) { begin[j] = 0; end[j] = 0; nbegin[j] = -1; nend[j] = -1; }
  }
  int32_t pos = 0;
  int32_t ob = core > 0 ? n2o[0] : 0;   // norm_to_orig[pos]
  for (int32_t k = 0; k < core;) {
    // Known pieces go eight at a time: their ids, lengths and norm_to_orig entries are
    // independent loads (one thread serves a whole sentence, so this is where the
    // memory-level parallelism comes from); an unknown piece is settled on its own.
    constexpr int G = 8;
    const int32_t g = core - k < G ? core - k : G;
    int32_t idv[G], ev[G];
    bool has_unk = false;
    for (int u = 0; u < G; ++u)
      if (u < g) { idv[u] = fids[plan.npre + (plan.flip ? core - 1 - (k + u) : k + u)]; has_unk |= idv[u] == m.unk_id; }
    if (!has_unk) {
      const int32_t start = pos;
      for (int u = 0; u < G; ++u) if (u < g) ev[u] = SPM_LDG(pt.plen + idv[u]);
      for (int u = 0; u < G; ++u) if (u < g) { pos += ev[u]; ev[u] = pos; }   // piece ends
      if (pos > nl) return false;
      int32_t ov[G];
      for (int u = 0; u < G; ++u) if (u < g) ov[u] = n2o[ev[u]];
      for (int u = 0; u < G; ++u) {
        if (u >= g) break;
        const int32_t j = plan.npre + (plan.flip ? core - 1 - (k + u) : k + u);
        if (j < room) {
          nbegin[j] = u ? ev[u - 1] : start; nend[j] = ev[u];
          begin[j] = u ? ov[u - 1] : ob; end[j] = ov[u];
        }
      }
      ob = ov[g - 1];
      k += g;
      continue;
    }
    const int32_t j = plan.npre + (plan.flip ? core - 1 - k : k);
    const int32_t id = idv[0];
    int32_t e;
    if (id != m.unk_id) {
      e = pos + SPM_LDG(pt.plen + id);
      if (e > nl) return false;
    } else {
      const uint8_t *np = nullptr;
      int32_t npl = 0;
      if (k + 1 < core) {
        const int32_t nid = fids[plan.npre + (plan.flip ? core - 2 - k : k + 1)];
        if (nid == m.unk_id) return false;   // unknown runs are merged: cannot follow each other
        np = pt.pool + SPM_LDG(pt.poff + nid);
        npl = SPM_LDG(pt.plen + nid);
      }
      int32_t found = -1, q = pos;
      while (q < nl) {
        int cl = one_char_len(s[q]);
        if (cl > nl - q) cl = nl - q;
        if (!unk_eligible(m, s + q, cl)) break;
        q += cl;
        bool match;
        if (np == nullptr) {
          match = q == nl;
        } else {
          match = q + npl <= nl;
          for (int32_t b = 0; match && b < npl; ++b) match = s[q + b] == SPM_LDG(np + b);
        }
        if (match) {
          if (found >= 0) return false;   // two candidate ends: only the Viterbi tie-break knows
          found = q;
        }
      }
      if (found < 0) return false;
      e = found;
    }
    const int32_t oe = n2o[e];
    if (j < room) { nbegin[j] = pos; nend[j] = e; begin[j] = ob; end[j] = oe; }
    pos = e;
    ob = oe;
    ++k;
  }
  return pos == nl;
}

// ---- the plain sentence: the pieces path without the normalizer ------------------------
// Most sentences of most corpora are printable ASCII words separated by single spaces.  When the
// model leaves such bytes alone (ascii_plain) and escapes whitespace, the normalized text is the
// input with U+2581 for every space (and in front, with add_dummy_prefix), and norm_to_orig is
// affine between spaces: orig(dummy U+2581) = 0 for all its bytes, a space's U+2581 maps to the
// space, every other byte to itself.  So neither the normalized text nor norm_to_orig has to be
// made to place the pieces: a piece of plen normalized bytes with w U+2581 covers plen - 2 w input
// bytes (olen), one less when its first U+2581 is the dummy prefix.
SPM_HD bool pc_plain_model(const ModelView &m) {
  return m.ascii_plain && m.escape_ws && !m.ws_suffix && m.user.n == 0;
}
// Normalized length of the plain sentence s[0..n), or -1 when it is not one (a byte outside
// 0x20..0x7E, a leading / trailing / repeated space, or empty).  s is read in aligned 8-byte
// words within [lo, hi) (fast_load8_guard's contract is not needed: bytes outside [0, n) are masked).
SPM_HDN int32_t pc_plain_length(const ModelView &m, const uint8_t *s, int64_t n) {
  if (n <= 0 || n > (1 << 24)) return -1;
  if (s[0] == 0x20u || s[n - 1] == 0x20u) return -1;
  int32_t spaces = 0;
  uint32_t prev = 0;
  for (int64_t i = 0; i < n; ++i) {
    const uint32_t c = s[i];
    if (c < 0x20u || c > 0x7Eu) return -1;
    if (c == 0x20u) { if (prev == 0x20u) return -1; ++spaces; }
    prev = c;
  }
  return (int32_t)n + 2 * spaces + (m.add_dummy_prefix ? 3 : 0);
}
// The normalized text of a plain sentence: out[0 .. pc_plain_length).
SPM_HD void pc_plain_expand(const ModelView &m, const uint8_t *s, int64_t n, uint8_t *out) {
  int64_t o = 0;
  if (m.add_dummy_prefix) { out[0] = 0xE2; out[1] = 0x96; out[2] = 0x81; o = 3; }
  for (int64_t i = 0; i < n; ++i) {
    if (s[i] == 0x20u) { out[o] = 0xE2; out[o + 1] = 0x96; out[o + 2] = 0x81; o += 3; }
    else out[o++] = s[i];
  }
}
// spans_from_ids for a plain sentence of n input bytes and nl normalized bytes.  False (the exact
// tier redoes the sentence) when an unknown piece is among the ids or the lengths do not add up.
SPM_HDN bool spans_plain_from_ids(const ModelView &m, const PieceTable &pt, const ExtraPlan &plan, int32_t n, int32_t nl,
                                  const int32_t *fids, int32_t fc, int32_t *begin, int32_t *end, int32_t *nbegin,
                                  int32_t *nend, int64_t room) {
  if (fc == 0) return nl == 0;
  const int32_t core = fc - plan.npre - plan.nsuf;
  if (core < 0) return false;
  for (int32_t j = 0; j < fc; ++j) {
    if (j >= plan.npre && j < plan.npre + core) continue;
    if (j < room) { begin[j] = 0; end[j] = 0; nbegin[j] = -1; nend[j] = -1; }
  }
  int32_t pos = 0, opos = 0;
  for (int32_t k = 0; k < core;) {
    constexpr int G = 8;   // independent loads, as in spans_from_ids
    const int32_t g = core - k < G ? core - k : G;
    int32_t idv[G], pl[G], ol[G];
    bool has_unk = false;
    for (int u = 0; u < G; ++u)
      if (u < g) { idv[u] = fids[plan.npre + (plan.flip ? core - 1 - (k + u) : k + u)]; has_unk |= idv[u] == m.unk_id; }
    if (has_unk) return false;
    for (int u = 0; u < G; ++u) if (u < g) { pl[u] = SPM_LDG(pt.plen + idv[u]); ol[u] = SPM_LDG(pt.olen + idv[u]); }
    if (k == 0 && m.add_dummy_prefix) ol[0] -= 1;      // the dummy prefix stands for no input byte
    for (int u = 0; u < G; ++u) {
      if (u >= g) break;
      const int32_t j = plan.npre + (plan.flip ? core - 1 - (k + u) : k + u);
      if (j < room) { nbegin[j] = pos; nend[j] = pos + pl[u]; begin[j] = opos; end[j] = opos + ol[u]; }
      pos += pl[u];
      opos += ol[u];
    }
    k += g;
  }
  return pos == nl && opos == n;
}

}  // namespace spmb
