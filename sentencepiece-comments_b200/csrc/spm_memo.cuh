// spm_memo.cuh -- the word-memo tier: every distinct whitespace-delimited word of
// a batch is segmented once, every occurrence copies the ids.
//
// Why this is exact.  The tier is only used for models in which no piece carries
// U+2581 at a non-initial position.  Every piece that covers a U+2581 character of
// the normalized text then starts there, so every lattice path (unigram,
// src/unigram_model.cc:535-604) and every merge (BPE, src/bpe_model.cc:37-199) is
// confined to one "word" = [U+2581, next U+2581).
//   * BPE: the merge order inside a word depends on pair scores only, so the pieces
//     of a word are a pure function of its bytes.
//   * Unigram: in exact arithmetic the best path of the sentence is the concatenation
//     of the best paths of its words.  In the reference's float arithmetic the
//     backtrace score B accumulated before the word shifts every candidate of the
//     word by the same amount but rounds each one differently, so a decision inside
//     the word is context-free only while its margin exceeds the rounding noise:
//     with k = characters of the word + 1 additions per candidate,
//         margin > 2 k 2^-24 (|B| + mag_best + mag_other)
//     (each float addition errs by at most 2^-24 of its result, both in the
//     sentence and in the stand-alone evaluation).  memo_unigram_word() therefore
//     returns, next to the ids, bmax = the largest |B| for which EVERY decision of
//     the word (on or off the best path) keeps that margin with a 2x safety factor;
//     memo_emit() tracks an upper bound of |B| along the sentence and sends the
//     sentence to the Viterbi tier (fast_unigram) the moment a word's bmax is
//     exceeded.  Exactly tied candidates have margin 0 and are never memoised.
//
// The table is built per call (no state survives spm_cuda_encode_batch), keyed by
// the EXACT bytes of the word (<= kMemoMaxBytes after the U+2581 prefix; longer
// words send their sentence to the Viterbi tier), claimed with one 128-bit CAS.
// Everything is `__host__ __device__` so tests/host_core.cc runs the same code.
#pragma once
#include "spm_fast.cuh"

namespace spmb {

constexpr int kMemoKeyBytes = 14;      // word bytes held by one key
constexpr int kMemoTailBytes = 10;     // bytes held by the key of a tail entry (next to the head's slot)
constexpr int kMemoMaxBytes = kMemoKeyBytes + kMemoTailBytes;   // word bytes after the U+2581 prefix
constexpr int kMemoInlineIds = 5;      // ids kept inside the result record
constexpr int kMemoOverflowIds = 20;   // ids 4.. of a longer word (overflow row of the word's list index)
constexpr int kMemoMaxIds = kMemoInlineIds - 1 + kMemoOverflowIds;
constexpr int kMemoProbes = 24;        // linear probes before giving up
constexpr uint32_t kMemoReady = 0x100u;   // meta: word encoded
constexpr uint32_t kMemoBad = 0x200u;     // meta: the word cannot be memoised
constexpr uint32_t kMemoFirstUnk = 0x400u;  // meta: first id is the unknown id
constexpr uint32_t kMemoLastUnk = 0x800u;   // meta: last id is the unknown id

// Key of a word of <= 14 bytes: meta | bytes 0..6  /  bytes 7..13 | meta << 56, with
// meta = length << 1 | has-prefix (1..29).  A longer word is keyed in two steps: a HEAD
// entry (meta 0x40 | prefix) for its first 14 bytes, then the word's own TAIL entry
// (meta 0x80 | rest length) whose key holds the head's slot (3 bytes) and the <= 10 bytes
// that are left.  Both halves of every key carry the non-zero meta byte; (0, 0) = empty.
struct alignas(16) MemoKey {
  uint64_t k0, k1;
};
// What an occurrence of the word copies: one 32-byte sector.
struct alignas(32) MemoRes {
  uint32_t meta;       // id count | kMemo* flags
  float bmax;          // unigram: largest |backtrace score at the word start| that keeps the word context-free
  float sabs;          // unigram: bound of |best-path score of the word|
  int32_t ids[kMemoInlineIds];   // count <= 5: all ids; else ids 0..3 and the overflow row in ids[4]
};

struct MemoTable {
  MemoKey *keys;
  MemoRes *res;
  uint32_t mask;        // slots - 1 (power of two)
  uint32_t *new_list;   // slots claimed in this pass
  uint32_t *n_new;      // their count (may exceed max_new: the excess is never encoded)
  uint32_t max_new;
  int32_t *overflow;    // max_new rows of kMemoOverflowIds ids
  uint32_t *n_ovf;      // rows handed out so far -- per CALL, not per pass: results outlive the pass (chunk) that made them
};

// slot lists are written once and read once: keep them out of the way of the table in L2
SPM_HD void memo_store_slot(uint32_t *p, uint32_t v) {
#if defined(__CUDA_ARCH__)
  __stcs(p, v);
#else
  *p = v;
#endif
}

#ifndef SPM_MEMO_HASH32        // 32-bit multiply-add hash (default) or three 64-bit multiplies (A/B builds: 2.60 -> 2.55 ms)
#define SPM_MEMO_HASH32 1
#endif
SPM_HD uint32_t memo_hash(uint64_t k0, uint64_t k1) {
#if SPM_MEMO_HASH32
  uint32_t h = (uint32_t)k0 * 0x9E3779B1u + (uint32_t)(k0 >> 32) * 0x85EBCA77u + (uint32_t)k1 * 0xC2B2AE3Du +
               (uint32_t)(k1 >> 32) * 0x27D4EB2Fu;
  h ^= h >> 15; h *= 0x2C1B3C6Du;
  h ^= h >> 12; h *= 0x297A2D39u;
  return h ^ (h >> 15);
#else
  uint64_t h = k0 * 0x9E3779B97F4A7C15ull ^ (k1 + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full;
  h ^= h >> 29;
  h *= 0xBF58476D1CE4E5B9ull;
  return (uint32_t)(h >> 32);
#endif
}

// compare-and-swap of the 16-byte key against "empty"; returns the previous key
SPM_HD void memo_claim(MemoKey *e, uint64_t k0, uint64_t k1, uint64_t *o0, uint64_t *o1) {
#if defined(__CUDA_ARCH__)
  asm volatile(
      "{\n\t.reg .b128 c, s, d;\n\t"
      "mov.b128 c, {%3, %4};\n\tmov.b128 s, {%5, %6};\n\t"
      "atom.relaxed.gpu.global.cas.b128 d, [%2], c, s;\n\t"
      "mov.b128 {%0, %1}, d;\n\t}"
      : "=l"(*o0), "=l"(*o1)
      : "l"(e), "l"(0ull), "l"(0ull), "l"(k0), "l"(k1)
      : "memory");
#else
  *o0 = e->k0; *o1 = e->k1;
  if (*o0 == 0 && *o1 == 0) { e->k0 = k0; e->k1 = k1; }
#endif
}

// Slot of the word with key (k0, k1), inserting it when absent; 0xFFFFFFFF when the
// probe sequence is exhausted.  A (cached, possibly stale or torn) plain load decides the
// common case: both key halves carry the non-zero meta byte, so a half-written key can
// never compare equal, and keys are immutable once written; "looks empty" is settled
// by the CAS itself.
SPM_HD void memo_load_key(const MemoKey *e, uint64_t *e0, uint64_t *e1) {
#if defined(__CUDA_ARCH__)
  asm volatile("ld.global.v2.u64 {%0, %1}, [%2];" : "=l"(*e0), "=l"(*e1) : "l"(e) : "memory");
#else
  *e0 = e->k0; *e1 = e->k1;
#endif
}
// The probe sequence from `slot`, whose entry (e0, e1) has been loaded already.
SPM_HD uint32_t memo_find_or_insert_from(const MemoTable &tb, uint64_t k0, uint64_t k1, uint32_t slot, uint64_t e0,
                                         uint64_t e1) {
  for (int probe = 0;;) {
    if (e0 == k0 && e1 == k1) return slot;
    if (e0 == 0 && e1 == 0) {
      memo_claim(tb.keys + slot, k0, k1, &e0, &e1);
      if (e0 == 0 && e1 == 0) {
#if defined(__CUDA_ARCH__)
        const uint32_t idx = atomicAdd(tb.n_new, 1u);
#else
        const uint32_t idx = (*tb.n_new)++;
#endif
        if (idx < tb.max_new) tb.new_list[idx] = slot;
        else tb.res[slot].meta = 0;      // beyond the list: never encoded, never served (res is NOT zeroed per call:
                                         // every claimed slot gets its meta here or from the words kernel)
        return slot;
      }
      if (e0 == k0 && e1 == k1) return slot;
    }
    if (++probe == kMemoProbes) return 0xFFFFFFFFu;
    slot = (slot + 1) & tb.mask;
    memo_load_key(tb.keys + slot, &e0, &e1);
  }
}
SPM_HD uint32_t memo_find_or_insert(const MemoTable &tb, uint64_t k0, uint64_t k1) {
  const uint32_t slot = memo_hash(k0, k1) & tb.mask;
  uint64_t e0, e1;
  memo_load_key(tb.keys + slot, &e0, &e1);
  return memo_find_or_insert_from(tb, k0, k1, slot, e0, e1);
}
// The same in two halves, so that a caller can key and issue the first load of its NEXT word before it
// waits for the load of the current one (k_memo_scan_words: the probe is an L2 / DRAM round trip).
struct MemoProbe {
  uint64_t k0, k1, e0, e1;
  uint32_t slot;
};
SPM_HD void memo_probe_issue(const MemoTable &tb, uint64_t k0, uint64_t k1, MemoProbe *p) {
  p->k0 = k0; p->k1 = k1;
  p->slot = memo_hash(k0, k1) & tb.mask;
  memo_load_key(tb.keys + p->slot, &p->e0, &p->e1);
}
SPM_HD uint32_t memo_probe_finish(const MemoTable &tb, const MemoProbe &p) {
  return memo_find_or_insert_from(tb, p.k0, p.k1, p.slot, p.e0, p.e1);
}

// One out-of-line copy for the general scan: MemoSink looks words up at five places, and with the probe
// loop inlined at each of them k_memo_scan stalled 11.5 of 26 cycles per instruction on instruction fetch.
#if defined(__CUDACC__)
__host__ __device__ __noinline__
#else
inline
#endif
uint32_t memo_find_or_insert_shared(const MemoTable &tb, uint64_t k0, uint64_t k1) {
  return memo_find_or_insert(tb, k0, k1);
}

// Output sink of fast_normalize_scan(): the normalized characters of a sentence go
// straight into word keys (the normalized text itself is only materialised for the
// sentences the memo cannot serve).  slots[] receives the table slot of every word.
struct MemoSink {
  MemoTable tb;
  uint32_t *slots;
  int32_t cap;              // capacity of slots[]
  int32_t nw;
  uint64_t c0, c1;          // head: content bytes 0..7 / 8..13; tail: the tail key without its meta byte
  int len, prefix;
  bool tail;                // the current key is the tail of a long word
  bool started;             // the current word has a prefix or a byte
  bool fail;                // a word cannot be keyed (too long, table full, too many words)
  int32_t pending;          // bare U+2581 words not yet written to slots[]
  static constexpr bool kRuns = true;
  static constexpr bool kFirst3 = false;   // (the word memo's scan sees mostly ASCII: the extra test only costs it, 9.2 -> 9.7 ms on noisy text)
  SPM_HD void init(const MemoTable &t, uint32_t *s, int32_t c) {
    tb = t; slots = s; cap = c; nw = 0; c0 = 0; c1 = 0; len = 0; prefix = 0;
    tail = false; started = false; fail = false; pending = 0;
  }
  SPM_HD bool room(int64_t) const { return true; }
  SPM_HD void at(int32_t) {}
  SPM_HD bool done() const { return fail; }
  SPM_HD uint32_t lookup(uint64_t mb) {
    if (tail) return memo_find_or_insert_shared(tb, c0 | mb, c1 | (mb << 56));
    return memo_find_or_insert_shared(tb, mb | (c0 << 8), (c0 >> 56) | (c1 << 8) | (mb << 56));
  }
  SPM_HD void close_word() {
    const uint32_t slot = lookup(tail ? (uint64_t)(0x80 | len) : (uint64_t)((len << 1) | prefix));
    if (slot == 0xFFFFFFFFu || nw >= cap) { fail = true; return; }
    slots[nw++] = slot;
  }
  // words that are a bare U+2581 wait here: at the end of the sentence they are trailing
  // whitespace, which remove_extra_whitespaces strips (src/normalizer.cc:191-201)
  SPM_HD void flush_bare() {
    for (; pending > 0 && !fail; --pending) {
      const uint32_t slot = memo_find_or_insert_shared(tb, 1ull, 1ull << 56);   // length 0, prefix 1
      if (slot == 0xFFFFFFFFu || nw >= cap) { fail = true; return; }
      slots[nw++] = slot;
    }
  }
  SPM_HD bool bare() const { return prefix == 1 && len == 0 && !tail; }
  SPM_HD void put_byte(uint64_t b) {
    if (pending) flush_bare();
    if (!tail && len == kMemoKeyBytes) {       // the word goes on: key its head, start the tail
      const uint32_t hs = lookup((uint64_t)(0x40 | prefix));
      if (hs == 0xFFFFFFFFu) { fail = true; return; }
      c0 = (uint64_t)hs << 8; c1 = 0; len = 0; tail = true;
    }
    if (tail) {
      if (len >= kMemoTailBytes) { fail = true; return; }
      if (len < 4) c0 |= b << (8 * (len + 4));
      else c1 |= b << (8 * (len - 4));
    } else {
      if (len < 8) c0 |= b << (8 * len);
      else c1 |= b << (8 * (len - 8));
    }
    ++len; started = true;
  }
  // up to `run` (1..8) bytes of one word at once (the low bytes of w); returns how many were taken
  SPM_HD int take(uint64_t w, int run) {
    if (fail || tail) return 0;
    if (pending) { flush_bare(); if (fail) return 0; }
    if (run > kMemoKeyBytes - len) run = kMemoKeyBytes - len;
    if (run <= 0) return 0;
    const uint64_t wm = run == 8 ? w : (w & ((1ull << (8 * run)) - 1));
    if (len < 8) {
      c0 |= wm << (8 * len);
      if (len + run > 8) c1 |= wm >> (8 * (8 - len));
    } else {
      c1 |= wm << (8 * (len - 8));
    }
    len += run; started = true;
    return run;
  }
  SPM_HD void put(uint32_t v, int k) {  // one character, k in 1..4 bytes, little-endian in v
    if (fail) return;
    if (k == 3 && v == 0x8196E2u) {     // U+2581: the next word starts here
      if (started) {
        if (bare()) ++pending; else close_word();
      }
      c0 = 0; c1 = 0; len = 0; prefix = 1; started = true; tail = false;
      return;
    }
    for (int j = 0; j < k && !fail; ++j) put_byte((v >> (8 * j)) & 0xFFu);
  }
  // end of the sentence.  Returns the word count or -1.
  SPM_HD int32_t finish(bool strip_trailing) {
    if (fail) return -1;
    if (started) {
      if (bare()) ++pending; else close_word();
    }
    if (strip_trailing) pending = 0;
    flush_bare();
    return fail ? -1 : nw;
  }
};

// The common sentence, without the general normalizer: printable ASCII words separated by
// single spaces, no leading / trailing space (HostModel::ascii_plain: such bytes pass the
// normalizer untouched; a space becomes U+2581 = the start of the next word).  One loop
// iteration takes a run of up to 8 word bytes with SWAR and, at a space, keys the word.
// s[0..n) must be readable in aligned 8-byte words up to 16 bytes past the end.
// Returns the word count, -1 when the memo cannot serve the sentence, or -2 when the
// sentence is not of this shape (the general scan, fast_normalize_scan + MemoSink, decides).
// relax (memo_relax_of): 1 = the model removes extra whitespace (leading, trailing and repeated spaces vanish
// before the dummy prefix is added, src/normalizer.cc:88-211), so such spaces only separate words here; 2 = and a
// TAB, LF, FF and CR each normalize to exactly one space (ModelView::ws_ctl), so they are such separators too; 0 = they make the
// sentence unfit (-2).
SPM_HD int memo_relax_of(const ModelView &m) { return m.remove_extra_ws ? (m.ws_ctl ? 2 : 1) : 0; }
SPM_HD bool memo_is_ctl_space(uint32_t b) { return b >= 0x09u && b <= 0x0Du && b != 0x0Bu; }
SPM_HDN int32_t memo_scan_plain(const MemoTable &tb, const uint8_t *s, int32_t n, int prefix0, uint32_t *slots,
                                int32_t cap, int relax = 0) {
  const uint8_t *q = (const uint8_t *)((uintptr_t)s & ~(uintptr_t)7);
  int r = (int)(s - q);
  uint64_t a0 = fast_load8(q), a1 = fast_load8(q + 8);
  int32_t p = 0, nw = 0;
  // key bytes of the current word in (c0, c1); pos = where the next byte goes: a head key
  // fills bytes 0..13, the tail key of a long word bytes 4..13 (after the head's slot)
  uint64_t c0 = 0, c1 = 0;
  int pos = 0;
  bool tail = false;
  uint64_t prefix = (uint64_t)prefix0;
  for (;;) {
    const uint64_t w = fast_window(a0, a1, r);
    const int32_t rem = n - p;
    if (relax && pos == 0 && !tail && ((w & 0xFFu) == 0x20u || (relax == 2 && memo_is_ctl_space((uint32_t)w & 0xFFu)))) {   // a separator where a word would begin: skip it
      p += 1; r += 1;
      if (p >= n) break;
      if (r >= 8) { r -= 8; a0 = a1; q += 8; a1 = fast_load8(q + 8); }
      continue;
    }
    // 0x80 marks bytes < 0x21 or > 0x7E (borrows / carries only disturb bytes above the first mark)
    const uint64_t np = (((w - 0x2121212121212121ull) & ~w) | ((w + 0x0101010101010101ull) | w)) & 0x8080808080808080ull;
    int run = np ? (ffs64(np) - 1) >> 3 : 8;
    bool endw = false;
    int skip = 0;
    if (run >= rem) {
      run = rem; endw = true;                    // the sentence ends inside the window
    } else if (run == 8) {
      run = 7;                                   // keep the successor in sight
    } else {
      const uint32_t b = (uint32_t)(w >> (8 * run)) & 0xFFu;
      if (b != 0x20u && !(relax == 2 && memo_is_ctl_space(b))) return -2;   // control / non-ASCII: rules may apply
      endw = true; skip = 1;
      if (run + 1 >= rem && !relax) return -2;   // trailing space
    }
    if (run == 0) return -2;                     // leading / repeated space
    uint64_t wm = w & (0xFFFFFFFFFFFFFFFFull >> (8 * (8 - run)));
    int k = run;
    if (pos + k > 14) {                          // a long word: key its head, go on in a tail key
      if (tail) return -1;                       // longer than kMemoMaxBytes
      const int h = 14 - pos;
      if (h > 0) {
        const uint64_t hm = wm & (0xFFFFFFFFFFFFFFFFull >> (8 * (8 - h)));
        if (pos < 8) { c0 |= hm << (8 * pos); if (pos + h > 8) c1 |= hm >> (8 * (8 - pos)); }
        else c1 |= hm << (8 * (pos - 8));
        wm >>= 8 * h; k -= h;
      }
      const uint64_t mb = 0x40ull | prefix;
      const uint32_t hs = memo_find_or_insert(tb, mb | (c0 << 8), (c0 >> 56) | (c1 << 8) | (mb << 56));
      if (hs == 0xFFFFFFFFu) return -1;
      c0 = (uint64_t)hs << 8; c1 = 0; pos = 4; tail = true;
    }
    if (pos < 8) { c0 |= wm << (8 * pos); if (pos + k > 8) c1 |= wm >> (8 * (8 - pos)); }
    else c1 |= wm << (8 * (pos - 8));
    pos += k;
    p += run + skip; r += run + skip;
    if (endw) {
      uint32_t slot;
      if (tail) {
        const uint64_t mb = 0x80ull | (uint64_t)(pos - 4);
        slot = memo_find_or_insert(tb, c0 | mb, c1 | (mb << 56));
      } else {
        const uint64_t mb = ((uint64_t)pos << 1) | prefix;
        slot = memo_find_or_insert(tb, mb | (c0 << 8), (c0 >> 56) | (c1 << 8) | (mb << 56));
      }
      if (slot == 0xFFFFFFFFu || nw >= cap) return -1;
      memo_store_slot(slots + nw++, slot);
      if (p >= n) break;
      c0 = 0; c1 = 0; pos = 0; tail = false; prefix = 1;
    }
    if (r >= 8) { r -= 8; a0 = a1; q += 8; a1 = fast_load8(q + 8); }
  }
  return nw;
}

// ---- the same scan, byte-parallel (k_memo_scan_words): the text of 32 consecutive sentences is
// swept 8 bytes per lane; every lane reports, as 8-bit masks over its bytes, where words start,
// where they end and which bytes make a sentence unfit for this scan.  A word is a maximal
// run of bytes 0x21..0x7E inside one sentence; `b9` bit j says that a sentence starts at byte
// j of the chunk (bit 8: at the first byte after it; the end of the warp's text counts as a
// start), `lo`/`hi` clip the chunk to the warp's text ([lo, hi) of 0..8), `pb`/`nb` are the
// bytes before and after the chunk.  A sentence is unfit (memo_scan_plain's -2) when it holds
// a byte outside 0x20..0x7E, or -- with relax == 0 -- a space that is leading, trailing or follows a space
// (relax: see memo_scan_plain; TAB, LF, FF and CR count as spaces with relax == 2).
struct SwMasks {
  uint32_t ws, we, bad, b;   // b: the sentence starts among the chunk's 8 bytes
};
SPM_HD uint32_t sw_gather(uint64_t f) {   // bit 7 of byte j -> bit j
  return (uint32_t)(((f >> 7) * 0x0102040810204080ull) >> 56);
}
SPM_HD SwMasks sw_chunk_masks(uint64_t w, uint32_t pb, uint32_t nb, uint32_t b9, int lo, int hi, int relax = 0) {
  const uint64_t H = 0x8080808080808080ull, L7 = 0x7F7F7F7F7F7F7F7Full;
  const uint64_t lw = w & L7;
  const uint64_t ge21 = (lw | H) - 0x2121212121212121ull;   // bit 7: low seven bits >= 0x21 (no borrow leaves a byte)
  const uint64_t le7e = ~(lw + 0x0101010101010101ull);      // bit 7: low seven bits <= 0x7E (no carry leaves a byte)
  const uint64_t x = w ^ 0x2020202020202020ull;
  const uint64_t spf = ~(((x & L7) + L7) | x) & H;          // byte == 0x20
  lo = lo < 0 ? 0 : (lo > 8 ? 8 : lo);
  hi = hi < 0 ? 0 : (hi > 8 ? 8 : hi);
  const uint32_t valid = hi > lo ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
  const uint32_t good = sw_gather(ge21 & le7e & ~w & H) & valid;
  uint32_t sp = sw_gather(spf);
  uint32_t other = valid & ~good & ~sp;                     // neither a word byte nor a space
  if (relax == 2 && other != 0) {                           // ... unless TAB, LF, FF, CR: 0x09..0x0D without 0x0B.
    // (Only looked at when there is such a byte at all: clean text never pays for it.)
    const uint64_t ge09 = (lw | H) - 0x0909090909090909ull;   // bit 7: low seven bits >= 0x09
    const uint64_t le0d = ~(lw + 0x7272727272727272ull);      // bit 7: low seven bits <= 0x0D
    const uint64_t xb = w ^ 0x0B0B0B0B0B0B0B0Bull;
    const uint64_t eq0b = ~(((xb & L7) + L7) | xb);           // bit 7: byte == 0x0B
    sp |= sw_gather(ge09 & le0d & ~w & ~eq0b & H);
    other = valid & ~good & ~sp;
  }
  const uint32_t pgood = (pb >= 0x21u && pb <= 0x7Eu) ? 1u : 0u, psp = pb == 0x20u ? 1u : 0u;   // (psp: relax == 0 only)
  const uint32_t ngood = (nb >= 0x21u && nb <= 0x7Eu) ? 0x80u : 0u;
  const uint32_t bcur = b9 & 0xFFu, bnext = (b9 >> 1) & 0xFFu;
  const uint32_t gprev = (good << 1) | pgood, gnext = (good >> 1) | ngood;
  SwMasks m;
  m.b = bcur;
  m.ws = good & (bcur | ~gprev) & 0xFFu;
  m.we = good & (bnext | ~gnext) & 0xFFu;
  // (relax, see memo_scan_plain: leading, trailing and repeated spaces are no reason to leave the sentence out)
  m.bad = (other | (relax ? 0u : (sp & valid & (bcur | bnext | (sp << 1) | psp)))) & 0xFFu;
  return m;
}

// Slot of the word s[0..n) (n >= 1 bytes of 0x21..0x7E, s readable in aligned 8-byte words up to
// 32 + 7 bytes past its start): exactly the keys memo_scan_plain() builds -- one head key up to
// 14 bytes, head + tail key up to 24.  0xFFFFFFFF: the word cannot be keyed (too long, table full).
// Key of a word of 1..14 bytes (the one-key case of sw_word_slot below).
SPM_HD void sw_short_key(const uint8_t *s, int n, uint64_t prefix, uint64_t *k0, uint64_t *k1) {
  const uint8_t *q = (const uint8_t *)((uintptr_t)s & ~(uintptr_t)7);
  const int r = (int)(s - q);
  const uint64_t a0 = fast_load8(q), a1 = fast_load8(q + 8), a2 = fast_load8(q + 16);
  uint64_t w0 = fast_window(a0, a1, r), w1 = fast_window(a1, a2, r);
  if (n < 8) { w0 &= ~0ull >> (8 * (8 - n)); w1 = 0; }
  else w1 &= n == 8 ? 0ull : ~0ull >> (8 * (16 - n));
  const uint64_t mb = ((uint64_t)n << 1) | prefix;
  *k0 = mb | (w0 << 8);
  *k1 = (w0 >> 56) | (w1 << 8) | (mb << 56);
}
SPM_HD uint32_t sw_word_slot(const MemoTable &tb, const uint8_t *s, int n, uint64_t prefix) {
  if (n > kMemoMaxBytes) return 0xFFFFFFFFu;
  const uint8_t *q = (const uint8_t *)((uintptr_t)s & ~(uintptr_t)7);
  const int r = (int)(s - q);
  const uint64_t a0 = fast_load8(q), a1 = fast_load8(q + 8), a2 = fast_load8(q + 16), a3 = fast_load8(q + 24);
  uint64_t w0 = fast_window(a0, a1, r), w1 = fast_window(a1, a2, r);
  if (n <= kMemoKeyBytes) {
    if (n < 8) { w0 &= ~0ull >> (8 * (8 - n)); w1 = 0; }
    else if (n < 16) w1 &= n == 8 ? 0ull : ~0ull >> (8 * (16 - n));
    const uint64_t mb = ((uint64_t)n << 1) | prefix;
    return memo_find_or_insert(tb, mb | (w0 << 8), (w0 >> 56) | (w1 << 8) | (mb << 56));
  }
  const uint64_t w2 = fast_window(a2, a3, r);
  const uint64_t hb = 0x40ull | prefix;
  const uint64_t h1 = w1 & 0x0000FFFFFFFFFFFFull;
  const uint32_t hs = memo_find_or_insert(tb, hb | (w0 << 8), (w0 >> 56) | (h1 << 8) | (hb << 56));
  if (hs == 0xFFFFFFFFu) return hs;
  const int nt = n - kMemoKeyBytes;                        // 1..10 bytes go into the tail key
  uint64_t t0 = (w1 >> 48) | (w2 << 16), t1 = w2 >> 48;    // word bytes 14..21, 22..23
  if (nt < 8) { t0 &= ~0ull >> (8 * (8 - nt)); t1 = 0; }
  else t1 &= nt == 8 ? 0ull : (nt == 9 ? 0xFFull : 0xFFFFull);
  const uint64_t mb = 0x80ull | (uint64_t)nt;
  return memo_find_or_insert(tb, ((uint64_t)hs << 8) | (t0 << 32) | mb, (t0 >> 32) | (t1 << 32) | (mb << 56));
}

// Bytes of the word in slot `slot`: U+2581 prefix (when keyed with one) + content, zero
// padded to `cap` (>= 40) bytes.  Returns the byte count; 0 for a head entry (not a word).
SPM_HD int memo_word_bytes(const MemoKey *keys, uint32_t slot, uint8_t *out, int cap) {
  uint64_t k0 = keys[slot].k0, k1 = keys[slot].k1;
  const int mb = (int)(k0 & 0xFFu);
  int nb = 0;
  if (mb & 0x40) {
    nb = 0;
  } else {
    uint64_t t0 = 0, t1 = 0;
    int tl = 0;
    if (mb & 0x80) {   // tail: fetch the head (immutable once claimed)
      t0 = k0; t1 = k1; tl = mb & 0x3F;
      const MemoKey *h = keys + ((k0 >> 8) & 0xFFFFFFu);
      k0 = h->k0; k1 = h->k1;
    }
    const int hb = (int)(k0 & 0xFFu);
    const int len = (hb & 0x40) ? kMemoKeyBytes : (hb >> 1), prefix = hb & 1;
    if (prefix) { out[0] = 0xE2; out[1] = 0x96; out[2] = 0x81; nb = 3; }
    for (int j = 0; j < len; ++j)
      out[nb++] = (uint8_t)(j < 7 ? (k0 >> (8 * (j + 1))) : (k1 >> (8 * (j - 7))));
    for (int j = 0; j < tl; ++j)
      out[nb++] = (uint8_t)(j < 4 ? (t0 >> (8 * (j + 4))) : (t1 >> (8 * (j - 4))));
  }
  for (int j = nb; j < cap; ++j) out[j] = 0;
  return nb;
}

// Stores ids (forward order), flags and bounds of an encoded word; `row` = its index in new_list.
SPM_HD void memo_store(const MemoTable &tb, uint32_t row, const int32_t *ids, int32_t cnt, bool bad,
                       int32_t unk_id, float bmax, float sabs) {
  MemoRes *e = tb.res + tb.new_list[row];
  uint32_t meta = kMemoReady;
  if (bad || cnt <= 0 || cnt > kMemoMaxIds) { meta |= kMemoBad; cnt = 0; }
  if (cnt > kMemoInlineIds) {       // an overflow row of its own, kept until the end of the call
#if defined(__CUDA_ARCH__)
    const uint32_t orow = atomicAdd(tb.n_ovf, 1u);
#else
    const uint32_t orow = (*tb.n_ovf)++;
#endif
    if (orow >= tb.max_new) { meta |= kMemoBad; cnt = 0; }
    else {
      for (int j = 0; j < kMemoInlineIds - 1; ++j) e->ids[j] = ids[j];
      e->ids[kMemoInlineIds - 1] = (int32_t)orow;
      for (int j = kMemoInlineIds - 1; j < cnt; ++j)
        tb.overflow[(size_t)orow * kMemoOverflowIds + (j - (kMemoInlineIds - 1))] = ids[j];
    }
  } else {
    for (int j = 0; j < cnt; ++j) e->ids[j] = ids[j];
  }
  if (cnt > 0 && ids[0] == unk_id) meta |= kMemoFirstUnk;
  if (cnt > 0 && ids[cnt - 1] == unk_id) meta |= kMemoLastUnk;
  e->bmax = bmax;
  e->sabs = sabs;
  e->meta = meta | (uint32_t)cnt;
}
// a word whose segmentation does not depend on its context (BPE)
SPM_HD void memo_store_ids(const MemoTable &tb, uint32_t row, const int32_t *ids, int32_t cnt, bool bad,
                           int32_t unk_id) {
  memo_store(tb, row, ids, cnt, bad, unk_id, 3.0e38f, 0.0f);
}

// Unigram: Viterbi of one stand-alone word with the margin bookkeeping described above.
// Same lattice as fast_unigram() (nodes of the fast view + the one-character UNK node,
// candidates of an end position in begin order, strict '>').
SPM_HDN void memo_unigram_word(const FastView &fv, bool neg_only, const MemoTable &tb, uint32_t row) {
  constexpr int kMaxChars = kMemoMaxBytes + 1;
  const float NEG = -3.0e38f;
  uint8_t txt[40];
  const int nb = memo_word_bytes(tb.keys, tb.new_list[row], txt, 40);
  if (nb == 0) { tb.res[tb.new_list[row]].meta = kMemoReady | kMemoBad; return; }   // a head entry
  float b1[kMaxChars + 1], b2[kMaxChars + 1];
  uint32_t rec[kMaxChars + 1];
  int cstart[kMaxChars + 1];
  int m = 0;
  for (int p = 0; p < nb; p += one_char_len(txt[p])) cstart[m++] = p;
  cstart[m] = nb;
  for (int k = 0; k <= m; ++k) { b1[k] = NEG; b2[k] = NEG; rec[k] = 0; }
  b1[0] = 0.0f;
  const uint32_t root = fv.units[0];
  for (int ci = 0; ci < m; ++ci) {
    const float btp = b1[ci];
    uint32_t u = root;
    int k = 0;
    bool single = false;
    for (int j = cstart[ci]; j < nb; ++j) {
      const uint32_t c = txt[j];
      if (!(u & 0x200u) || c == 0) break;
      const uint32_t child = (u >> 10) ^ c;
      const uint32_t v = fv.units[child];
      if ((v & 0xFFu) != c || (v & 0x300u) == 0) break;
      u = v;
      if ((c & 0xC0u) != 0x80u) ++k;
      if (v & 0x100u) {
        const float val = btp + fv.score[child];
        const int t = ci + k;
        if (val > b1[t]) { b2[t] = b1[t]; b1[t] = val; rec[t] = child | ((uint32_t)k << 24); }
        else if (val > b2[t]) b2[t] = val;
        if (k == 1) single = true;
      }
    }
    if (!single) {
      const float val = btp + fv.unk_score;
      const int t = ci + 1;
      if (val > b1[t]) { b2[t] = b1[t]; b1[t] = val; rec[t] = kFastUnk | (1u << 24); }
      else if (val > b2[t]) b2[t] = val;
    }
  }
  // margins -> bmax
  const float K = (float)(m + 1) * 2.384185791015625e-07f;   // k 2^-22 = 2 x (2 k 2^-24)
  const float magw = (float)m * fv.smax;
  float bmax = 3.0e38f;
  for (int t = 1; t <= m; ++t) {
    if (b2[t] <= NEG) continue;                 // a single candidate: nothing to decide
    const float gap = b1[t] - b2[t];
    // all scores <= 0: partial sums are monotone, |other| = |best| + gap
    const float allow = neg_only ? (gap * (1.0f - K) / K - 2.0f * fabsf(b1[t])) : (gap / K - 2.0f * magw);
    if (allow < bmax) bmax = allow;
  }
  // backtrace (reverse), unknown runs collapse (sentencepiece_processor.cc:525-529)
  int32_t rev[kMaxChars];
  int cnt = 0;
  bool prev_unk = false;
  for (int ci = m; ci > 0;) {
    const uint32_t node = rec[ci] & 0x00FFFFFFu;
    const bool is_unk = node == kFastUnk;
    if (!(is_unk && prev_unk)) rev[cnt++] = is_unk ? fv.unk_id : (int32_t)fv.leaf_id[node];
    prev_unk = is_unk;
    ci -= (int)(rec[ci] >> 24);
  }
  int32_t fwd[kMaxChars];
  for (int j = 0; j < cnt; ++j) fwd[j] = rev[cnt - 1 - j];
  memo_store(tb, row, fwd, cnt, !(bmax > 0.0f), fv.unk_id, bmax,
             (neg_only ? fabsf(b1[m]) : magw) * 1.001f + 1.0e-3f);
}

// Ids of one sentence from the slots of its words (forward order, at most `cap`).  Returns
// the id count, or -1 when a word is not memoised / not context-free at this point of the
// sentence.  (Sequential statement of what k_memo_emit does with one 16-lane tile.)
SPM_HDN int32_t memo_emit(const MemoTable &tb, const uint32_t *slots, int32_t nw, bool unigram, int32_t *out,
                          int32_t cap) {
  int32_t cnt = 0;
  float bub = 0.0f;          // upper bound of |backtrace score| at the start of the next word
  bool prev_unk = false;
  for (int32_t w = 0; w < nw; ++w) {
    const MemoRes *e = tb.res + slots[w];
    const uint32_t meta = e->meta;
    if ((meta & (kMemoReady | kMemoBad)) != kMemoReady) return -1;
    if (unigram) {
      if (!(bub < e->bmax)) return -1;
      bub += e->sabs;
    }
    const int k = (int)(meta & 0xFFu);
    int j = 0;
    if (prev_unk && (meta & kMemoFirstUnk)) j = 1;   // an unknown run continues across the boundary
    if (cnt + (k - j) > cap) return -1;
    for (; j < k; ++j) {
      if (k <= kMemoInlineIds || j < kMemoInlineIds - 1) out[cnt++] = e->ids[j];
      else out[cnt++] = tb.overflow[(size_t)e->ids[kMemoInlineIds - 1] * kMemoOverflowIds + (j - (kMemoInlineIds - 1))];
    }
    prev_unk = (meta & kMemoLastUnk) != 0;
  }
  return cnt;
}

}  // namespace spmb
