// TEST HARNESS ONLY (built by tests/conftest.py into tests/_build/, never
// shipped, never loaded by the product package): compiles the kernels'
// __host__ __device__ core (spm_core.cuh) and the host model loader for the
// CPU, so the exact arithmetic that runs on the GPU can be compared with the
// oracle in a container without a GPU.  The product library has no host
// encode path.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "spm_core.cuh"
#include "spm_fast.cuh"
#include "spm_memo.cuh"
#include "spm_estep.cuh"
#include "spm_model.h"

using namespace spmb;

namespace {
struct Handle {
  HostModel hm;
  ModelView mv;
  PieceTable pt;
};
void make_view(Handle *h) {
  HostModel &m = h->hm;
  ModelView &v = h->mv;
  memset(&v, 0, sizeof(v));
  v.pieces.units = m.pieces.units.data(); v.pieces.leaf = m.pieces.leaf.data(); v.pieces.n = (uint32_t)m.pieces.units.size();
  v.user.units = m.user.units.data(); v.user.leaf = m.user.leaf.data(); v.user.n = (uint32_t)m.user.units.size();
  v.cm_units = m.cm_units; v.cm_pool = m.cm_pool; v.cm_n = m.cm_n;
  v.cm_first2 = m.cm_first2.empty() ? nullptr : m.cm_first2.data();
  v.cm_cont2 = m.cm_cont2.empty() ? nullptr : m.cm_cont2.data();
  v.cm_first3 = m.cm_first3.empty() ? nullptr : m.cm_first3.data();
  memcpy(v.user_first, m.user_first, sizeof(v.user_first));
  v.model_type = m.model_type; v.unk_id = m.unk_id;
  v.unk_score = m.min_score - 10.0f; v.max_score = m.max_score; v.smax = m.smax; v.ring = m.ring;
  v.add_dummy_prefix = m.add_dummy_prefix; v.remove_extra_ws = m.remove_extra_ws;
  v.escape_ws = m.escape_ws; v.ws_suffix = m.ws_suffix; v.has_unused = m.has_unused;
  v.ascii_plain = m.ascii_plain; v.ws_ctl = m.ws_ctl;
  h->pt.plen = m.plen.data(); h->pt.olen = m.plen.data() + m.plen.size() / 2; h->pt.poff = m.poff.data(); h->pt.pool = (const uint8_t *)m.ppool.data();
}
}  // namespace

extern "C" {

void *hc_load(const void *data, size_t len, int *code) {
  Handle *h = new Handle;
  *code = build_host_model(data, len, &h->hm);
  if (*code) { delete h; return nullptr; }
  make_view(h);
  return h;
}
void hc_free(void *p) { delete (Handle *)p; }
int hc_expansion(void *p) { return ((Handle *)p)->hm.expansion; }
// bit 0: unigram fast tier usable, bit 1: BPE fast tier usable, bit 2: word memo usable,
// bit 3: all node scores <= 0
int hc_fast_flags(void *p) {
  const HostModel &m = ((Handle *)p)->hm;
  return (m.fast_ok ? 1 : 0) | (m.bpe_fast_ok ? 2 : 0) | (m.memo_ok ? 4 : 0) | (m.neg_only ? 8 : 0);
}
int hc_trie_units(void *p) { return (int)((Handle *)p)->hm.pieces.units.size(); }
void hc_set_piece_types(void *p, const uint8_t *t, int n) {
  Handle *h = (Handle *)p;
  for (int i = 0; i < n && i < (int)h->hm.type.size(); ++i) h->hm.type[i] = t[i];
  refresh_leaf_types(&h->hm);
  make_view(h);
}

int64_t hc_normalize(void *p, const uint8_t *in, int64_t n, uint8_t *out, int64_t cap) {
  return normalize_sentence(((Handle *)p)->mv, in, n, out, cap);
}

// fast-tier normaliser (spm_fast.cuh) on a copy of the text placed at an odd address
// inside a guard-banded buffer; returns the length or -1 (does not fit: general tier).
int64_t hc_normalize_fast(void *p, const uint8_t *in, int64_t n, uint8_t *out, int64_t cap, int32_t *nchars) {
  Handle *h = (Handle *)p;
  if (n == 0) return 0;  // the kernels never see empty sentences
  std::vector<uint64_t> ibuf((n + 64) / 8 + 2, 0x5A5A5A5A5A5A5A5Aull);  // junk around the text must not leak in
  uint8_t *src = (uint8_t *)ibuf.data() + 11;
  memcpy(src, in, n);
  const int32_t fcap = (int32_t)(3 * n + 72);
  std::vector<uint64_t> obuf((fcap + kNormPad + 16) / 8 + 2, 0x7777777777777777ull);
  int32_t nc = 0;
  const int32_t nl = fast_normalize(h->mv, h->mv.cm_first2, src, n, src, src + n, (uint8_t *)obuf.data(), fcap, &nc);
  if (nchars) *nchars = nc;
  if (nl < 0) return -1;
  const uint8_t *o = (const uint8_t *)obuf.data();
  for (int k = 0; k < kNormPad; ++k) if (o[nl + k] != 0) return -2;  // padding contract
  if (nl > cap) return -3;
  memcpy(out, o, nl);
  return nl;
}

// k_normalize_long's scheme, sequentially: pieces of ~chunk raw bytes cut at norm_find_cut
// positions, each normalized on its own, concatenated and finished.  Returns the length, -1
// (general tier), or < -1 on a broken contract.
int64_t hc_normalize_chunked(void *p, const uint8_t *in, int64_t n, int chunk, uint8_t *out, int64_t cap) {
  Handle *h = (Handle *)p;
  const ModelView &mv = h->mv;
  if (n == 0) return 0;
  std::vector<uint64_t> ibuf((n + 64) / 8 + 2, 0x5A5A5A5A5A5A5A5Aull);
  uint8_t *src = (uint8_t *)ibuf.data() + 5;
  memcpy(src, in, n);
  const int nchunks = (int)((n + chunk - 1) / chunk);
  std::vector<int64_t> cuts(nchunks + 1);
  for (int i = 0; i <= nchunks; ++i)
    cuts[i] = i == 0 ? 0 : (i == nchunks ? n : norm_find_cut(mv, mv.cm_first2, src, (int64_t)i * chunk, n));
  std::vector<uint64_t> tmp((3 * n + 88 * (int64_t)nchunks + 96) / 8 + 2, 0x7777777777777777ull);
  std::vector<int32_t> olen(nchunks, 0), onch(nchunks, 0);
  bool bad = false, started = false;
  for (int i = 0; i < nchunks; ++i) {
    if (cuts[i + 1] < cuts[i]) return -4;
    if (cuts[i + 1] == cuts[i]) continue;
    bool content = false;
    const bool ok = norm_piece(mv, mv.cm_first2, src, cuts[i], cuts[i + 1], src, src + n,
                               (uint8_t *)tmp.data() + norm_piece_slot(cuts[i], i), i, &olen[i], &onch[i], &content);
    if (i == 0 && content) started = true;
    if (!ok || (i == 0 && !content && nchunks > 1)) bad = true;
  }
  if (bad) return -1;
  int64_t total = 0, nch = 0;
  for (int i = 0; i < nchunks; ++i) { total += olen[i]; nch += onch[i]; }
  if (total + 3 > 3 * n + 72) return -1;
  std::vector<uint8_t> o(3 * n + 72 + kNormPad + 16, 0x33);
  int64_t at = 0;
  for (int i = 0; i < nchunks; ++i) {
    memcpy(o.data() + at, (const uint8_t *)tmp.data() + norm_piece_slot(cuts[i], i), olen[i]);
    at += olen[i];
  }
  const int32_t nl = norm_finish(mv, o.data(), (int32_t)total, (int32_t)nch, started, n);
  if (nl < 0) return -1;
  if (nl > cap) return -3;
  memcpy(out, o.data(), nl);
  return nl;
}

// capc: candidate capacity of the unigram tier to emulate (1, 4 or 16; other values: with the
// library's retry at min(ring, 255) on overflow); 0: the fast tiers
// with the library's routing; -1: the word memo in front of them (sentences it served carry
// status bit 20).  status[n] receives the per-sentence status bits.
int hc_encode_batch(void *p, const uint8_t *text, const int64_t *off, int64_t n, uint32_t opts, int capc_arg,
                    int32_t *ids, int64_t cap, int64_t *id_off, uint32_t *status) {
  Handle *h = (Handle *)p;
  const ModelView &mv = h->mv;
  const bool use_memo = capc_arg == -1 && h->hm.memo_ok;
  const int capc = capc_arg == -1 ? 0 : capc_arg;
  const bool fast_side = getenv("HC_FAST_NO_SIDE") == nullptr;   // near ties of the fast tier in line (default) or flagged
  // word memo state of this call (small table: probe-exhaustion paths get exercised too)
  std::vector<MemoKey> mkeys;
  std::vector<MemoRes> mres;
  std::vector<uint32_t> mnew;
  std::vector<int32_t> movf;
  uint32_t m_n_new = 0, m_done = 0, m_n_ovf = 0;
  MemoTable tb;
  memset(&tb, 0, sizeof(tb));
  FastView mfv;
  memset(&mfv, 0, sizeof(mfv));
  if (use_memo) {
    const size_t slots = getenv("HC_MEMO_SMALL") ? (1 << 10) : (1 << 19);
    mkeys.resize(slots);
    mres.resize(slots);
    memset((void *)mkeys.data(), 0, slots * sizeof(MemoKey));
    memset((void *)mres.data(), 0, slots * sizeof(MemoRes));
    mnew.resize(slots / 2);
    movf.resize(mnew.size() * kMemoOverflowIds);
    tb.keys = mkeys.data(); tb.res = mres.data(); tb.mask = (uint32_t)slots - 1; tb.new_list = mnew.data();
    tb.n_new = &m_n_new; tb.max_new = (uint32_t)mnew.size(); tb.overflow = movf.data(); tb.n_ovf = &m_n_ovf;
    mfv.units = mfv.units_hot = h->hm.funits.data();
    mfv.score = mfv.score_hot = h->hm.fscore.data();
    mfv.leaf_id = h->hm.fleaf.data();
    mfv.n_units = (uint32_t)h->hm.funits.size();
    mfv.k_units = 0; mfv.k_scores = 0;
    mfv.ring = h->hm.max_piece_chars + 1;
    mfv.unk_score = mv.unk_score; mfv.smax = mv.smax; mfv.unk_id = mv.unk_id;
  }
  const ExtraPlan plan = make_extra_plan(opts, piece_to_id(h->hm, h->hm.bos_piece), piece_to_id(h->hm, h->hm.eos_piece));
  std::vector<uint8_t> norm;
  std::vector<int32_t> words, core;
  id_off[0] = 0;
  for (int64_t i = 0; i < n; ++i) {
    const int64_t len = off[i + 1] - off[i];
    const int64_t ncap = (int64_t)h->hm.expansion * len + 8;
    norm.resize(ncap);
    uint32_t st = 0;
    int32_t nl = normalize_sentence(mv, text + off[i], len, norm.data(), ncap);
    int32_t cnt = 0;
    if (nl < 0) { st |= ST_NORM_OVERFLOW; nl = 0; }
    core.resize(nl + 1);
    if (mv.model_type == M_UNIGRAM && capc == 0 && len > 0) {
      // same routing as the library: fast normaliser first, general tier when it does not fit
      std::vector<uint8_t> fn(3 * len + 72 + 8);
      int32_t fnc = 0;
      const int64_t fl = hc_normalize_fast(h, text + off[i], len, fn.data(), 3 * len + 72, &fnc);
      if (fl < -1) return 100 - (int)fl;
      if (fl >= 0 && (fl != nl || memcmp(fn.data(), norm.data(), nl) != 0)) return 99;
    }
    bool memo_served = false;
    if (use_memo && len > 0) {
      // k_memo_scan: the raw sentence straight into word keys (no normalized text)
      std::vector<uint32_t> slots(len / 2 + 4);
      int32_t nw = -2;
      if (mv.ascii_plain && mv.escape_ws && len < kLongMaxRaw) {   // k_memo_scan_plain
        std::vector<uint64_t> pbuf(len / 8 + 5, 0x2020202020202020ull);   // what follows must not matter
        uint8_t *ps = (uint8_t *)pbuf.data() + (i % 8);                    // any alignment
        memcpy(ps, text + off[i], len);
        nw = memo_scan_plain(tb, ps, (int32_t)len, mv.add_dummy_prefix ? 1 : 0, slots.data(), (int32_t)(len / 2 + 4),
                             memo_relax_of(mv));
        if (nw != -2) st |= 1u << 23;
      }
      if (nw == -2) {                                                  // k_memo_scan
        MemoSink sink;
        sink.init(tb, slots.data(), (int32_t)(len / 2 + 4));
        int32_t nc = 0;
        bool content = false;
        fast_normalize_scan(mv, mv.cm_first2, text + off[i], len, text + off[0], text + off[n], sink, &nc, &content);
        nw = sink.finish(mv.remove_extra_ws != 0);
      }
      for (; m_done < m_n_new && m_done < tb.max_new; ++m_done) {   // the "words" kernels
        if (mv.model_type == M_UNIGRAM) {
          memo_unigram_word(mfv, h->hm.neg_only, tb, m_done);
        } else {
          uint64_t wt[8];
          const int nb = memo_word_bytes(tb.keys, tb.new_list[m_done], (uint8_t *)wt, 64);
          std::vector<uint32_t> scr(kBpeWordsPerThread, 0);
          BpeScratch bs;
          bs.stride = 1;
          bs.node = scr.data(); bs.ps = (float *)(scr.data() + kBpeMaxSyms);
          bs.text = scr.data() + 2 * kBpeMaxSyms; bs.cb = bs.text + kBpeMaxBytes / 4;
          uint32_t txt2[2] = {0, 0};
          int32_t wid[kBpeMaxSyms];
          uint32_t wst = 0;
          FastView bfv = mfv;
          bfv.ring = 0; bfv.unk_score = 0; bfv.smax = 0;
          const int32_t c = nb > 0 ? fast_bpe(bfv, (const uint8_t *)wt, nb, bs, txt2, wid, &wst) : 0;
          memo_store_ids(tb, m_done, wid, c, wst != 0 || nb == 0, mv.unk_id);
        }
      }
      if (nw >= 0) {
        std::vector<int32_t> mid(len + kFastExtraChars);
        const int32_t c = memo_emit(tb, slots.data(), nw, mv.model_type == M_UNIGRAM, mid.data(),
                                    (int32_t)len + kFastExtraChars);
        if (c >= 0) {
          cnt = c; memo_served = true; st |= 1u << 20;
          core.resize(c + 1);
          memcpy(core.data(), mid.data(), c * sizeof(int32_t));
        } else {
          st |= 1u << 22;    // a word was not context-free here / not memoisable
        }
      } else {
        st |= 1u << 21;         // a word could not be keyed
      }
    }
    if (memo_served) {
      // ids are in core[] already
    } else if (mv.model_type == M_UNIGRAM && capc == 0) {
      // fast tier (spm_fast.cuh): padded, 8-byte aligned text; flagged sentences fall
      // back to the exact tier exactly as the library does
      if (!h->hm.fast_ok) return 12;
      FastView fv;
      fv.units_hot_saddr = fv.score_hot_saddr = 0;
      fv.units = fv.units_hot = h->hm.funits.data();
      fv.score = fv.score_hot = h->hm.fscore.data();
      fv.leaf_id = h->hm.fleaf.data();
      fv.n_units = (uint32_t)h->hm.funits.size();
      fv.k_units = fv.n_units / 3; fv.k_scores = fv.n_units / 5;  // exercise both the hot and the cold branch
      fv.ring = h->hm.max_piece_chars + 1;
      fv.unk_score = mv.unk_score; fv.smax = mv.smax; fv.unk_id = mv.unk_id;
      std::vector<uint64_t> tbuf((nl + kNormPad + 15) / 8 + 1, 0);
      memcpy(tbuf.data(), norm.data(), nl);
      std::vector<float> rbt(fv.ring);
      std::vector<uint32_t> rbk(fv.ring), back(nl + 12), txt(2);
      std::vector<int32_t> rev(nl + 1);
      std::vector<uint32_t> side(kFastSideWords, 0xDEADBEEFu);   // (the function initialises what it reads)
      cnt = fast_unigram(fv, (const uint8_t *)tbuf.data(), nl, rbt.data(), rbk.data(), txt.data(), 1, (uint32_t *)(((uintptr_t)back.data() + 15) & ~(uintptr_t)15), rev.data(), &st,
                         fast_side ? side.data() : nullptr);
      if (!(st & ST_CAND_OVERFLOW) && fast_side) st |= 1u << 24;   // (served with the side area in use)
      if (st & ST_CAND_OVERFLOW) {
        // reported; the result comes from the exact tier as in the library: 8 candidates per
        // position, then the retry with room for every candidate
        st = ST_CAND_OVERFLOW;
        for (int pass = 0; pass < 2; ++pass) {
          const int cc = pass == 0 ? 8 : (mv.ring < 255 ? mv.ring : 255);
          words.assign(uni_scratch_words(mv.ring, cc, nl) + 2, 0);
          UniScratch sc;
          int32_t *w = words.data();
          sc.ringM = (float *)w; w += mv.ring;
          sc.ringCnt = (uint32_t *)w; w += mv.ring;
          sc.candBt = (float *)w; w += mv.ring * cc;
          w += ((w - words.data()) & 1);
          sc.candRec = (uint64_t *)w; w += 2 * mv.ring * cc;
          sc.back = (uint64_t *)w;
          sc.stride = nl + 1;
          uint32_t st2 = 0;
          cnt = unigram_encode_capc(mv, norm.data(), nl, sc, cc, core.data(), nl, &st2);
          if (!(st2 & ST_CAND_OVERFLOW)) break;
          if (pass == 0) st |= 1u << 23;
        }
      } else {
        for (int32_t j = 0; j < cnt; ++j) core[j] = rev[cnt - 1 - j];
      }
    } else if (mv.model_type == M_BPE && capc == 0 && h->hm.bpe_fast_ok) {
      // fast BPE tier (spm_fast.cuh) with the library's routing: too-long words -> general tier
      FastView fv;
      fv.units_hot_saddr = fv.score_hot_saddr = 0;
      fv.units = fv.units_hot = h->hm.funits.data();
      fv.score = fv.score_hot = h->hm.fscore.data();
      fv.leaf_id = h->hm.fleaf.data();
      fv.n_units = (uint32_t)h->hm.funits.size();
      fv.k_units = fv.n_units / 3; fv.k_scores = fv.n_units / 5;
      fv.ring = 0; fv.unk_score = 0; fv.smax = 0; fv.unk_id = mv.unk_id;
      std::vector<uint64_t> tbuf((nl + kNormPad + 15) / 8 + 1, 0);
      memcpy(tbuf.data(), norm.data(), nl);
      std::vector<uint32_t> scr(kBpeWordsPerThread, 0);
      BpeScratch bs;
      bs.stride = 1;
      bs.node = scr.data(); bs.ps = (float *)(scr.data() + kBpeMaxSyms);
      bs.text = scr.data() + 2 * kBpeMaxSyms; bs.cb = bs.text + kBpeMaxBytes / 4;
      uint32_t txt2[2] = {0, 0};
      cnt = fast_bpe(fv, (const uint8_t *)tbuf.data(), nl, bs, txt2, core.data(), &st);
      if (st & ST_TOO_LONG) {
        words.assign(bpe_scratch_words(nl), 0);
        uint32_t st2 = 0;
        cnt = bpe_encode(mv, norm.data(), nl, words.data(), core.data(), nl, &st2);
        st |= st2;
      }
    } else if (mv.model_type == M_UNIGRAM) {
      words.assign(uni_scratch_words(mv.ring, capc, nl), 0);
      UniScratch sc;
      int32_t *w = words.data();
      sc.ringM = (float *)w; w += mv.ring;
      sc.ringCnt = (uint32_t *)w; w += mv.ring;
      sc.candBt = (float *)w; w += mv.ring * capc;
      w += ((w - words.data()) & 1);
      sc.candRec = (uint64_t *)w; w += 2 * mv.ring * capc;
      sc.back = (uint64_t *)w;
      sc.stride = nl + 1;
      if (capc == 1) cnt = unigram_encode<1>(mv, norm.data(), nl, sc, core.data(), nl, &st);
      else if (capc == 4) cnt = unigram_encode<4>(mv, norm.data(), nl, sc, core.data(), nl, &st);
      else if (capc == 16) cnt = unigram_encode<16>(mv, norm.data(), nl, sc, core.data(), nl, &st);
      else {
        // any other capacity: the library's flow (gen_encode_one + run_capc_retry) -- a sentence that
        // meets more near-tied candidates than `capc` runs again with room for every candidate
        cnt = unigram_encode_capc(mv, norm.data(), nl, sc, capc, core.data(), nl, &st);
        if (st & ST_CAND_OVERFLOW) {
          const int big = mv.ring < 255 ? mv.ring : 255;
          words.assign(uni_scratch_words(mv.ring, big, nl) + 2, 0);
          w = words.data();
          sc.ringM = (float *)w; w += mv.ring;
          sc.ringCnt = (uint32_t *)w; w += mv.ring;
          sc.candBt = (float *)w; w += mv.ring * big;
          w += ((w - words.data()) & 1);
          sc.candRec = (uint64_t *)w; w += 2 * mv.ring * big;
          sc.back = (uint64_t *)w;
          st = 0;
          cnt = unigram_encode_capc(mv, norm.data(), nl, sc, big, core.data(), nl, &st);
          st |= 1u << 23;   // reported: the retry ran
        }
      }
    } else {
      words.assign(bpe_scratch_words(nl), 0);
      cnt = bpe_encode(mv, norm.data(), nl, words.data(), core.data(), nl, &st);
    }
    status[i] = st;
    const int32_t fc = final_count(plan, cnt);
    if (id_off[i] + fc > cap) return 8;
    for (int32_t j = 0; j < fc; ++j) ids[id_off[i] + j] = final_id_at(plan, core.data(), cnt, j);
    id_off[i + 1] = id_off[i] + fc;
  }
  return 0;
}

// The pieces path (spm_pieces.inc: k_pc_normalize -> k_gen_encode with piece ends -> k_pc_emit)
// on the CPU: aligned normalizer, exact tier with out_end, final_piece_at.  Same CSR layout
// as spm_cuda_encode_batch_pieces.
int hc_encode_pieces_batch(void *p, const uint8_t *text, const int64_t *off, int64_t n, uint32_t opts, int32_t *ids,
                           int32_t *begin, int32_t *end, int32_t *nbegin, int32_t *nend, int64_t cap, int64_t *id_off,
                           uint8_t *norm_out, int64_t norm_cap, int64_t *norm_off, int64_t *n_flagged, int64_t *n_plain) {
  Handle *h = (Handle *)p;
  const ModelView &mv = h->mv;
  const ExtraPlan plan = make_extra_plan(opts, piece_to_id(h->hm, h->hm.bos_piece), piece_to_id(h->hm, h->hm.eos_piece));
  std::vector<uint8_t> norm;
  std::vector<int32_t> words, core, ends, n2o;
  id_off[0] = 0;
  norm_off[0] = 0;
  for (int64_t i = 0; i < n; ++i) {
    const int64_t len = off[i + 1] - off[i];
    const int64_t need = ((int64_t)h->hm.expansion * len + 8 + 7) & ~(int64_t)7;
    norm.assign(need, 0);
    n2o.assign(need, -1);
    uint32_t st = 0;
    int32_t nl = normalize_sentence_t<true>(mv, text + off[i], len, norm.data(), need - 1, n2o.data());
    if (nl < 0) return 90;
    if (len > 0) {
      // the fast tier's aligned scan (k_pc_normalize tries it first) must give the same text and map
      std::vector<uint64_t> ibuf((len + 64) / 8 + 2, 0x5A5A5A5A5A5A5A5Aull);
      uint8_t *src = (uint8_t *)ibuf.data() + 3 + (i % 8);
      memcpy(src, text + off[i], len);
      const int32_t fcap = (int32_t)(3 * len + 72);
      std::vector<uint64_t> obuf((fcap + kNormPad + 16) / 8 + 2, 0x7777777777777777ull);
      std::vector<int32_t> fmap(fcap + kNormPad + 16, -7);
      const int32_t fl = fast_normalize_aligned(mv, mv.cm_first2, src, len, src, src + len, (uint8_t *)obuf.data(), fcap,
                                                fmap.data());
      if (fl >= 0) {
        if (fl != nl || memcmp(obuf.data(), norm.data(), nl) != 0) return 96;
        for (int32_t k = 0; nl > 0 && k <= nl; ++k) if (fmap[k] != n2o[k]) return 97;
      }
    }
    core.assign(nl + 1, 0);
    ends.assign(nl + 1, 0);
    int32_t cnt = 0;
    if (mv.model_type == M_UNIGRAM) {
      words.assign(uni_scratch_words(mv.ring, 8, nl), 0);
      UniScratch sc;
      int32_t *w = words.data();
      sc.ringM = (float *)w; w += mv.ring;
      sc.ringCnt = (uint32_t *)w; w += mv.ring;
      sc.candBt = (float *)w; w += mv.ring * 8;
      w += ((w - words.data()) & 1);
      sc.candRec = (uint64_t *)w; w += 2 * mv.ring * 8;
      sc.back = (uint64_t *)w;
      sc.stride = nl + 1;
      cnt = unigram_encode<8>(mv, norm.data(), nl, sc, core.data(), nl, &st, ends.data());
    } else {
      words.assign(bpe_scratch_words(nl), 0);
      cnt = bpe_encode(mv, norm.data(), nl, words.data(), core.data(), nl, &st, ends.data());
    }
    if (st & (ST_CAND_OVERFLOW | ST_IDS_OVERFLOW)) return 91;
    const int32_t fc = final_count(plan, cnt);
    if (id_off[i] + fc > cap || norm_off[i] + nl > norm_cap) return 8;
    for (int32_t j = 0; j < fc; ++j) {
      const PieceSpan ps = final_piece_at(plan, core.data(), ends.data(), n2o.data(), cnt, j);
      const int64_t q = id_off[i] + j;
      ids[q] = ps.id; begin[q] = ps.begin; end[q] = ps.end; nbegin[q] = ps.nbegin; nend[q] = ps.nend;
    }
    if (n_flagged != nullptr) {
      // the fast half (k_pc_spans): spans recovered from the final ids alone; where it claims
      // success it must agree with the exact tier's piece ends
      std::vector<int32_t> fb(fc + 1), fe(fc + 1), fnb(fc + 1), fne(fc + 1);
      const bool ok = spans_from_ids(mv, h->pt, plan, norm.data(), nl, n2o.data(), ids + id_off[i], fc, fb.data(),
                                     fe.data(), fnb.data(), fne.data(), fc);
      if (!ok) {
        ++*n_flagged;
      } else {
        for (int32_t j = 0; j < fc; ++j) {
          const int64_t q = id_off[i] + j;
          if (fb[j] != begin[q] || fe[j] != end[q] || fnb[j] != nbegin[q] || fne[j] != nend[q]) return 95;
        }
      }
    }
    // the plain-sentence path (k_pc_prepare / k_pc_spans / k_pc_copy_norm): same length, same text, and --
    // where it claims success -- the exact tier's spans
    if (pc_plain_model(mv)) {
      const int32_t pl = pc_plain_length(mv, text + off[i], len);
      {   // the kernel's eight-bytes-at-a-time version, at every alignment and with foreign bytes around the sentence
        std::vector<uint64_t> pbuf((len + 40) / 8 + 2, 0x2020202020802020ull);
        uint8_t *src = (uint8_t *)pbuf.data() + 8 + (i % 8);
        memcpy(src, text + off[i], len);
        if (pc_plain_length_swar(mv, src, len, (const uint8_t *)pbuf.data(), (const uint8_t *)(pbuf.data() + pbuf.size())) != pl) return 93;
      }
      if (pl >= 0) {
        if (pl != nl) return 98;
        std::vector<uint8_t> ex(pl + 8, 0);
        pc_plain_expand(mv, text + off[i], len, ex.data());
        if (memcmp(ex.data(), norm.data(), nl) != 0) return 99;
        std::vector<int32_t> fb(fc + 1), fe(fc + 1), fnb(fc + 1), fne(fc + 1);
        if (spans_plain_from_ids(mv, h->pt, plan, (int32_t)len, nl, ids + id_off[i], fc, fb.data(), fe.data(), fnb.data(),
                                 fne.data(), fc)) {
          for (int32_t j = 0; j < fc; ++j) {
            const int64_t q = id_off[i] + j;
            if (fb[j] != begin[q] || fe[j] != end[q] || fnb[j] != nbegin[q] || fne[j] != nend[q]) return 94;
          }
          if (n_plain != nullptr) ++*n_plain;
        }
      }
    }
    memcpy(norm_out + norm_off[i], norm.data(), nl);
    id_off[i + 1] = id_off[i] + fc;
    norm_off[i + 1] = norm_off[i] + nl;
  }
  return 0;
}

// Loader verdicts of a model: bit 0 ascii_plain, bit 1 ws_ctl, bit 2 memo_ok, bit 3 remove_extra_ws,
// bit 4 a rule filter for 3-byte characters exists.
int hc_model_flags(void *p) {
  Handle *h = (Handle *)p;
  return (h->hm.ascii_plain ? 1 : 0) | (h->hm.ws_ctl ? 2 : 0) | (h->hm.memo_ok ? 4 : 0) |
         (h->hm.remove_extra_ws ? 8 : 0) | (h->hm.cm_first3.empty() ? 0 : 16);
}

// cm_first3 must not hide a rule: for every 3-byte character U+0800..U+FFFF (surrogates left out) whose bit
// is clear, NormalizePrefix -- whatever follows the character -- returns the character itself.  Also what the
// loader claims with ws_ctl: TAB, LF, FF, CR -> one space.  Returns the number of violations.
int64_t hc_check_filters(void *p) {
  Handle *h = (Handle *)p;
  const ModelView &mv = h->mv;
  int64_t bad = 0;
  if (mv.cm_first3 != nullptr) {
    const uint8_t follow[5] = {0x00, 0x20, 0x61, 0xCC, 0xE3};   // end / space / letter / a combining mark's lead / another CJK lead
    for (uint32_t cp = 0x800; cp < 0x10000; ++cp) {
      if (cp >= 0xD800 && cp < 0xE000) continue;
      if ((mv.cm_first3[cp >> 5] >> (cp & 31)) & 1) continue;
      for (int f = 0; f < 5; ++f) {
        uint8_t in[8] = {(uint8_t)(0xE0 | (cp >> 12)), (uint8_t)(0x80 | ((cp >> 6) & 63)), (uint8_t)(0x80 | (cp & 63)),
                         follow[f], 0x81, 0x82, 0, 0};
        const int n = f == 0 ? 3 : 6;
        const uint8_t *rep = nullptr;
        int rl = 0;
        const int c = normalize_prefix(mv, in, n, &rep, &rl);
        if (!(c == 3 && rl == 3 && rep != nullptr && rep[0] == in[0] && rep[1] == in[1] && rep[2] == in[2])) ++bad;
      }
    }
  }
  if (h->hm.ws_ctl) {
    const uint8_t ctl[4] = {0x09, 0x0A, 0x0C, 0x0D};
    for (int k = 0; k < 4; ++k)
      for (int x = 0; x < 256; ++x) {
        uint8_t in[4] = {ctl[k], (uint8_t)x, 0, 0};
        const uint8_t *rep = nullptr;
        int rl = 0;
        const int c = normalize_prefix(mv, in, 2, &rep, &rl);
        if (!(c == 1 && rl == 1 && rep != nullptr && rep[0] == ' ')) ++bad;
      }
  }
  return bad;
}

// estep_softplus() against libm on m points of [-50, 0]: the largest absolute difference.
double hc_lse_table_error(int64_t m) {
  std::vector<double> lse;
  build_lse_table(&lse);
  double worst = 0;
  for (int64_t i = 0; i <= m; ++i) {
    const double d = -50.0 * (double)i / (double)m;
    const double e = fabs(estep_softplus(lse.data(), d) - (double)log1pl(expl((long double)d)));
    if (e > worst) worst = e;
  }
  return worst;
}

// E-step of the unigram trainer: the kernel's per-word function (spm_estep.cuh) on the CPU.
int hc_estep(const char *pieces, const float *scores, int V, const uint8_t *words, const int64_t *off,
             const int64_t *freq, int64_t n, double all_freq, double *expected, double *obj_ntok, float *logz) {
  std::vector<std::string> ps;
  std::vector<float> sc(scores, scores + V);
  const char *q = pieces;
  for (int i = 0; i < V; ++i) { ps.emplace_back(q); q += ps.back().size() + 1; }
  HostTrie trie;
  float min_score = 0;
  std::string err;
  const int rc = build_estep_trie(ps, sc, &trie, &min_score, &err);
  if (rc) return rc;
  EstepView ev;
  std::vector<double> lse;
  build_lse_table(&lse);
  ev.lse = lse.data();
  ev.units_hot = nullptr; ev.k_units = 0;
  ev.units = trie.units.data(); ev.leaf = trie.leaf.data();
  ev.unk_score = min_score - 10.0f; ev.unk_id = 0; ev.id_mask = 0x0FFFFFFFu;
  for (int64_t i = 0; i < n; ++i) {
    const int len = (int)(off[i + 1] - off[i]);
    if (len == 0) { if (logz) logz[i] = 0; continue; }
    std::vector<float> A(len + 1), B(len + 1);
    std::vector<int32_t> C(len + 1);
    int32_t nt = 0;
    const float f = (float)freq[i];
    const uint8_t *w = words + off[i];
    const float z = estep_word(ev, [w](int j) -> uint32_t { return w[j]; }, len, f, A.data(), B.data(), C.data(), 1,
                               [&](uint32_t x, double v) { expected[x & ev.id_mask] += v; }, &nt);
    obj_ntok[0] -= (double)(f * z) / all_freq;
    obj_ntok[1] += nt;
    if (logz) logz[i] = z;
  }
  return 0;
}


// The byte-parallel word scan (k_memo_scan_words: sw_chunk_masks + sw_word_slot, lanes replayed
// one after another) against memo_scan_plain() on the same table: same word count, same
// slots, no key that the per-sentence scan did not make.  text/off: n sentences; groups of 32
// are laid out like a warp's staging buffer at every alignment.  Returns 0, or 100000000 * kind +
// sentence index of the first difference.
int64_t hc_scan_words_check(const uint8_t *text, const int64_t *off, int64_t n, int prefix0, int relax, int64_t *n_plain) {
  const size_t slots_n = 1 << 18;
  std::vector<MemoKey> mkeys(slots_n);
  std::vector<MemoRes> mres(slots_n);      // (a slot claimed beyond max_new gets its meta word written)
  std::vector<uint32_t> mnew(slots_n / 2);
  memset((void *)mkeys.data(), 0, slots_n * sizeof(MemoKey));
  uint32_t n_new = 0;
  MemoTable tb;
  memset(&tb, 0, sizeof(tb));
  tb.keys = mkeys.data(); tb.res = mres.data(); tb.mask = (uint32_t)slots_n - 1; tb.new_list = mnew.data();
  tb.n_new = &n_new; tb.max_new = (uint32_t)mnew.size();
  *n_plain = 0;
  for (int64_t tw = 0; tw < n; tw += 32) {
    const int64_t te = std::min<int64_t>(tw + 32, n);
    const int64_t a = off[tw], b = off[te];
    const int a_rel = (int)(tw / 32 % 16);
    const int b_rel = a_rel + (int)(b - a);
    if (b_rel > 5120 - 48) continue;
    std::vector<uint64_t> sbuf(5120 / 8 + 8, 0xA5A5A5A5A5A5A5A5ull);      // what lies around the text must not matter
    uint8_t *stage = (uint8_t *)sbuf.data();
    memcpy(stage + a_rel, text + a, b - a);
    // reference: one sentence at a time
    std::vector<std::vector<uint32_t>> want(32);
    int32_t want_nw[32];
    bool all_ok = true;
    for (int64_t t = tw; t < te; ++t) {
      const int64_t len = off[t + 1] - off[t];
      want[t - tw].assign(len / 2 + 4, 0);
      want_nw[t - tw] = len > 0 ? memo_scan_plain(tb, stage + a_rel + (off[t] - a), (int32_t)len, prefix0,
                                                  want[t - tw].data(), (int32_t)(len / 2 + 4), relax) : 0;
      if (want_nw[t - tw] < 0) all_ok = false; else ++*n_plain;
    }
    const uint32_t new_before = n_new;
    // the warp
    int sstart[32];
    std::vector<uint8_t> bmap((5120 / 32 + 4) * 4, 0);
    for (int l = 0; l < 32; ++l) {
      sstart[l] = tw + l < te ? a_rel + (int)(off[tw + l] - a) : b_rel;
      bmap[sstart[l] >> 3] |= (uint8_t)(1u << (sstart[l] & 7));
    }
    bmap[b_rel >> 3] |= (uint8_t)(1u << (b_rel & 7));
    auto owner = [&](int pos) { int t = 0; for (int st = 16; st; st >>= 1) if (sstart[t + st] <= pos) t += st; return t; };
    std::vector<int> wpos, wend;
    uint32_t flags[32] = {0};
    for (int g0 = a_rel & ~7; g0 < b_rel; g0 += 256)
      for (int l = 0; l < 32; ++l) {
        const int g = g0 + l * 8;
        if (g >= b_rel) continue;
        uint64_t w;
        memcpy(&w, stage + g, 8);
        const uint32_t pb = g > 0 ? stage[g - 1] : 0u, nb = stage[g + 8];
        const uint32_t b9 = (uint32_t)bmap[g >> 3] | ((uint32_t)bmap[(g >> 3) + 1] << 8);
        const SwMasks m = sw_chunk_masks(w, pb, nb, b9, a_rel - g, b_rel - g, relax);
        for (uint32_t x = m.ws; x; x &= x - 1) wpos.push_back(g + __builtin_ctz(x));
        for (uint32_t x = m.we; x; x &= x - 1) wend.push_back(g + __builtin_ctz(x));
        for (uint32_t x = m.bad; x; x &= x - 1) flags[owner(g + __builtin_ctz(x))] |= 2u;
      }
    if (wpos.size() != wend.size()) return 100000000ll + tw;
    std::vector<std::vector<uint32_t>> got(32);
    int first[33];
    for (int l = 0; l < 32; ++l) first[l] = (int)(std::lower_bound(wpos.begin(), wpos.end(), sstart[l]) - wpos.begin());
    first[32] = (int)wpos.size();
    for (size_t k = 0; k < wpos.size(); ++k) {
      const int s = wpos[k], len = wend[k] - s + 1;
      if (len < 1) return 200000000ll + tw;
      const int u = owner(s);
      const uint32_t slot = sw_word_slot(tb, stage + s, len, (int)k == first[u] ? (uint64_t)prefix0 : 1ull);   // the first word of its sentence
      if (slot == 0xFFFFFFFFu) flags[u] |= 1u;
      if ((int)got[u].size() != (int)k - first[u]) return 300000000ll + tw + u;
      got[u].push_back(slot);
    }
    for (int64_t t = tw; t < te; ++t) {
      const int l = (int)(t - tw);
      int32_t nw = first[l + 1] - first[l];
      if (flags[l] & 2u) nw = -2; else if (flags[l] & 1u) nw = -1;
      const int32_t w = want_nw[l];
      if (w == -2 && nw != -2) return 400000000ll + t;
      if (w == -1 && nw >= 0) return 500000000ll + t;
      if (w >= 0) {
        if (nw != w) return 600000000ll + t;
        for (int k = 0; k < w; ++k)
          if (got[l][k] != want[l][k]) return 700000000ll + t;
      }
    }
    if (all_ok && n_new != new_before) return 800000000ll + tw;
  }
  return 0;
}

}  // extern "C"
