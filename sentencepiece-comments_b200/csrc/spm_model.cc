#include "spm_model.h"
#include "spm_estep.cuh"

#include <float.h>
#include <math.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <queue>

namespace spmb {
namespace {

// ------------------------------------------------------------ proto2 wire reader
struct Reader {
  const uint8_t *p, *e;
  bool ok;
  Reader(const void *d, size_t n) : p((const uint8_t *)d), e((const uint8_t *)d + n), ok(true) {}
  bool more() const { return ok && p < e; }
  uint64_t varint() {
    uint64_t v = 0;
    for (int sh = 0; sh < 70 && p < e; sh += 7) {
      const uint8_t b = *p++;
      if (sh < 64) v |= (uint64_t)(b & 0x7F) << sh;
      if (!(b & 0x80)) return v;
    }
    ok = false;
    return 0;
  }
  struct Field { int num, wt; uint64_t v; const uint8_t *s; size_t n; };
  Field next() {
    Field f = {0, 0, 0, nullptr, 0};
    const uint64_t key = varint();
    if (!ok) return f;
    f.num = (int)(key >> 3);
    f.wt = (int)(key & 7);
    if (f.wt == 0) {
      f.v = varint();
    } else if (f.wt == 1) {
      if (e - p < 8) { ok = false; return f; }
      memcpy(&f.v, p, 8); p += 8;
    } else if (f.wt == 5) {
      if (e - p < 4) { ok = false; return f; }
      uint32_t x; memcpy(&x, p, 4); f.v = x; p += 4;
    } else if (f.wt == 2) {
      const uint64_t l = varint();
      if (!ok || (uint64_t)(e - p) < l) { ok = false; return f; }
      f.s = p; f.n = (size_t)l; p += l;
    } else {
      ok = false;  // groups are not used by sentencepiece_model.proto
    }
    return f;
  }
};

bool parse_proto(const void *data, size_t len, HostModel *m) {
  Reader r(data, len);
  while (r.more()) {
    const Reader::Field f = r.next();
    if (!r.ok) return false;
    if (f.wt != 2) continue;
    Reader q(f.s, f.n);
    if (f.num == 1) {  // repeated SentencePiece pieces
      std::string pc;
      float sc = 0.0f;
      uint32_t ty = T_NORMAL;
      while (q.more()) {
        const Reader::Field g = q.next();
        if (!q.ok) return false;
        if (g.num == 1 && g.wt == 2) pc.assign((const char *)g.s, g.n);
        else if (g.num == 2 && g.wt == 5) { const uint32_t x = (uint32_t)g.v; memcpy(&sc, &x, 4); }
        else if (g.num == 3 && g.wt == 0) ty = (uint32_t)g.v;
      }
      m->piece.push_back(pc);
      m->score.push_back(sc);
      m->type.push_back((uint8_t)ty);
    } else if (f.num == 2) {  // TrainerSpec
      while (q.more()) {
        const Reader::Field g = q.next();
        if (!q.ok) return false;
        if (g.num == 3 && g.wt == 0) m->model_type = (int)g.v;
        else if (g.num == 24 && g.wt == 0) m->ws_suffix = g.v != 0;
        else if (g.num == 44 && g.wt == 2) m->unk_surface = std::string((const char *)g.s, g.n).c_str();  // .c_str(): sentencepiece_processor.cc:676
        else if (g.num == 45 && g.wt == 2) m->unk_piece.assign((const char *)g.s, g.n);
        else if (g.num == 46 && g.wt == 2) m->bos_piece.assign((const char *)g.s, g.n);
        else if (g.num == 47 && g.wt == 2) m->eos_piece.assign((const char *)g.s, g.n);
        else if (g.num == 48 && g.wt == 2) m->pad_piece.assign((const char *)g.s, g.n);
      }
    } else if (f.num == 4) {  // SelfTestData { repeated Sample samples = 1 { input = 1, expected = 2 } }
      while (q.more()) {
        const Reader::Field g = q.next();
        if (!q.ok) return false;
        if (g.num != 1 || g.wt != 2) continue;
        Reader u(g.s, g.n);
        std::string in, ex;
        while (u.more()) {
          const Reader::Field h = u.next();
          if (!u.ok) return false;
          if (h.num == 1 && h.wt == 2) in.assign((const char *)h.s, h.n);
          else if (h.num == 2 && h.wt == 2) ex.assign((const char *)h.s, h.n);
        }
        m->self_test_input.push_back(in);
        m->self_test_expected.push_back(ex);
      }
    } else if (f.num == 3) {  // NormalizerSpec
      while (q.more()) {
        const Reader::Field g = q.next();
        if (!q.ok) return false;
        if (g.num == 2 && g.wt == 2) m->charsmap.assign((const char *)g.s, g.n);
        else if (g.num == 3 && g.wt == 0) m->add_dummy_prefix = g.v != 0;
        else if (g.num == 4 && g.wt == 0) m->remove_extra_ws = g.v != 0;
        else if (g.num == 5 && g.wt == 0) m->escape_ws = g.v != 0;
      }
    }
  }
  return r.ok;
}

// ------------------------------------------------------------ XDA builder
// Flat double array with unique bases (label test suffices) and hot-first
// placement: nodes are laid out in order of descending subtree weight, so the
// units the walks touch most sit at the low indices (one compact, cache- and
// shared-memory-friendly prefix of the array).
struct TNode {
  std::map<uint8_t, int> kids;
  int value = -1;  // index into the key list
  double weight = 0;
};

struct Key { std::string s; uint32_t leaf_x; uint32_t leaf_y; double weight; };

bool build_xda(const std::vector<Key> &keys, HostTrie *out, std::string *err) {
  out->units.clear();
  out->leaf.clear();
  if (keys.empty()) return true;
  std::vector<TNode> nodes(1);
  for (size_t k = 0; k < keys.size(); ++k) {
    int cur = 0;
    nodes[0].weight += keys[k].weight;
    for (unsigned char ch : keys[k].s) {
      auto it = nodes[cur].kids.find(ch);
      int nx;
      if (it == nodes[cur].kids.end()) {
        nx = (int)nodes.size();
        nodes[cur].kids[ch] = nx;
        nodes.emplace_back();
      } else {
        nx = it->second;
      }
      nodes[nx].weight += keys[k].weight;
      cur = nx;
    }
    nodes[cur].value = (int)k;
  }
  // Text positions inside a word also start walks: every character-boundary
  // suffix of a key is a fragment the walks see about as often as the key, so
  // its path through the trie earns the key's weight too.
  for (size_t k = 0; k < keys.size(); ++k) {
    const std::string &s = keys[k].s;
    for (size_t b = 1; b < s.size(); ++b) {
      if (((unsigned char)s[b] & 0xC0) == 0x80) continue;
      int cur = 0;
      for (size_t i = b; i < s.size(); ++i) {
        auto it = nodes[cur].kids.find((unsigned char)s[i]);
        if (it == nodes[cur].kids.end()) break;
        cur = it->second;
        nodes[cur].weight += keys[k].weight;
      }
    }
  }
  std::vector<uint32_t> &units = out->units;
  std::vector<uint2> &leaf = out->leaf;
  std::vector<uint8_t> used, base_used;
  std::vector<int> block_free;
  auto grow = [&]() {
    units.resize(units.size() + 256, 0);
    used.resize(units.size(), 0);
    base_used.resize(units.size(), 0);
    block_free.push_back(256);
  };
  grow();
  used[0] = 1; block_free[0]--;
  base_used[0] = 1;  // base 0 would make the root (label 0 at index 0) its own child on a NUL byte
  std::vector<uint32_t> pos_of(nodes.size(), 0);
  units[0] = 0;
  typedef std::pair<double, int> QE;
  std::priority_queue<QE> pq;
  pq.push(QE(nodes[0].weight, 0));
  size_t lo_block = 0;
  while (!pq.empty()) {
    const int nd = pq.top().second;
    pq.pop();
    TNode &N = nodes[nd];
    if (N.kids.empty()) continue;
    std::vector<uint8_t> labels;
    for (auto &kv : N.kids) labels.push_back(kv.first);
    uint32_t base = 0;
    bool found = false;
    for (size_t B = lo_block; !found; ++B) {
      if (B == block_free.size()) grow();
      if (block_free[B] < (int)labels.size()) continue;
      const uint32_t bb = (uint32_t)B << 8;
      for (uint32_t x = 0; x < 256 && !found; ++x) {
        const uint32_t b = bb | x;
        if (base_used[b]) continue;
        bool fit = true;
        for (uint8_t c : labels)
          if (used[b ^ c]) { fit = false; break; }
        if (fit) { base = b; found = true; }
      }
    }
    while (lo_block + 1 < block_free.size() &&
           (block_free[lo_block] == 0 || block_free.size() - lo_block > 16)) ++lo_block;
    if (base >= (1u << 22)) { *err = "piece trie exceeds 4M units"; return false; }
    base_used[base] = 1;
    units[pos_of[nd]] |= (base << 10) | 0x200;
    for (auto &kv : N.kids) {
      const uint32_t cp = base ^ kv.first;
      used[cp] = 1;
      block_free[cp >> 8]--;
      pos_of[kv.second] = cp;
      units[cp] = kv.first | (nodes[kv.second].value >= 0 ? 0x100u : 0u);
      pq.push(QE(nodes[kv.second].weight, kv.second));
    }
  }
  leaf.assign(units.size(), uint2{0xFFFFFFFFu, 0});
  for (size_t i = 1; i < nodes.size(); ++i)
    if (nodes[i].value >= 0) {
      leaf[pos_of[i]].x = keys[nodes[i].value].leaf_x;
      leaf[pos_of[i]].y = keys[nodes[i].value].leaf_y;
    }
  return true;
}

uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

bool is_trie_type(uint32_t t) { return t == T_NORMAL || t == T_USER_DEFINED || t == T_UNUSED; }

// charsmap statistics: first-two-bytes filter and worst-case expansion.
void scan_charsmap(HostModel *m) {
  m->cm_first2.assign(65536 / 32, 0);
  m->cm_first3.assign(65536 / 32, 0);
  m->cm_cont2.assign(65536 / 32, 0);
  // user-defined symbols are matched before the charsmap (src/normalizer.cc:240-249)
  for (size_t i = 0; i < m->piece.size(); ++i) {
    if (m->type[i] != T_USER_DEFINED) continue;
    const std::string &pc = m->piece[i];
    for (size_t k = one_char_len((uint8_t)pc[0]); k < pc.size(); k += one_char_len((uint8_t)pc[k])) {
      const uint32_t b0 = (uint8_t)pc[k];
      if (b0 < 0x80 || k + 1 >= pc.size()) { for (uint32_t x = 0; x < 256; ++x) m->cm_cont2[((b0 << 8) | x) >> 5] |= 1u << (((b0 << 8) | x) & 31); }
      else { const uint32_t b1 = (uint8_t)pc[k + 1]; m->cm_cont2[((b0 << 8) | b1) >> 5] |= 1u << (((b0 << 8) | b1) & 31); }
    }
  }
  m->expansion = 3;  // ' ' -> U+2581 and a malformed byte -> U+FFFD are both 1 -> 3
  if (m->cm_n == 0) { m->cm_first2.clear(); m->cm_first3.clear(); return; }   // (cm_cont2 keeps the user-symbol marks)
  const uint32_t *a = m->cm_units;
  const uint32_t n = m->cm_n;
  // cm_cont2: first two bytes of every character that sits at a NON-initial position of some
  // key (prefixes of keys included: harmless over-marking).  A character outside this set can
  // never be swallowed by a match that started earlier: a safe place to cut a long line.
  auto mark_cont = [&](uint32_t b0, uint32_t b1) { m->cm_cont2[((b0 << 8) | b1) >> 5] |= 1u << (((b0 << 8) | b1) & 31); };
  struct Item { uint32_t pos; int depth; uint32_t b0, b1; int lead; };
  std::vector<Item> stack;
  stack.push_back(Item{du_offset(a[0]), 0, 0, 0, -1});
  while (!stack.empty()) {
    const Item it = stack.back();
    stack.pop_back();
    for (uint32_t c = 0; c < 256; ++c) {
      const uint32_t p = it.pos ^ c;
      if (p >= n) continue;
      const uint32_t u = a[p];
      if (du_label(u) != c) continue;
      const uint32_t np = p ^ du_offset(u);
      const int depth = it.depth + 1;
      const uint32_t b0 = depth == 1 ? c : it.b0, b1 = depth == 2 ? c : it.b1;
      int lead = -1;
      if (depth > 1) {
        if (c < 0x80) { for (uint32_t x = 0; x < 256; ++x) mark_cont(c, x); }
        else if (c >= 0xC0) lead = (int)c;
        else if (it.lead >= 0) mark_cont((uint32_t)it.lead, c);
      }
      if (depth == 2) m->cm_first2[((b0 << 8) | b1) >> 5] |= 1u << (((b0 << 8) | b1) & 31);
      // cm_first3: the first2 filter passes a whole block of 64 code points when one of them starts a rule (kana,
      // CJK: nearly every block); keyed by the code point itself it passes the one character.  (b0, b1, b2) ->
      // (b0 & 15) << 12 | (b1 & 63) << 6 | (b2 & 63); a key that ends inside its first character marks the block.
      if ((b0 & 0xF0u) == 0xE0u) {
        auto mark3 = [&](uint32_t x1, uint32_t x2) {
          const uint32_t cp = ((b0 & 15u) << 12) | ((x1 & 63u) << 6) | (x2 & 63u);
          m->cm_first3[cp >> 5] |= 1u << (cp & 31);
        };
        if (depth == 3) mark3(it.b1, c);
        if (((u >> 8) & 1) && depth == 2) for (uint32_t x = 0; x < 64; ++x) mark3(b1, x);
        if (((u >> 8) & 1) && depth == 1) for (uint32_t x = 0; x < 4096; ++x) mark3(x >> 6, x & 63);
      }
      if ((u >> 8) & 1) {
        if (depth == 1)
          for (uint32_t x = 0; x < 256; ++x) m->cm_first2[((b0 << 8) | x) >> 5] |= 1u << (((b0 << 8) | x) & 31);
        if (np < n) {
          const uint32_t v = a[np] & 0x7FFFFFFFu;
          if (v < m->cm_pool_size) {
            int outb = 0;
            for (const uint8_t *q = m->cm_pool + v; q < m->cm_pool + m->cm_pool_size && *q; ++q)
              outb += (*q == ' ') ? 3 : 1;
            const int ratio = (outb + depth - 1) / depth;
            if (ratio > m->expansion) m->expansion = ratio;
          }
        }
      }
      if (depth < 64 && np < n) stack.push_back(Item{np, depth, b0, b1, lead});
    }
  }
}

}  // namespace

// log(exp(d) + 1) on [-kLseLo, 0]: per interval the polynomial of degree kLseCoef - 1 through the function's
// values (long double) at the Chebyshev nodes, in powers of t = d - centre.
void build_lse_table(std::vector<double> *tab) {
  constexpr int N = kLseN, C = kLseCoef;
  tab->assign((size_t)N * C, 0.0);
  const long double h = (long double)kLseLo / N, pi = 3.14159265358979323846264338327950288L;
  for (int k = 0; k < N; ++k) {
    const long double centre = -((long double)k + 0.5L) * h;
    long double a[C][C + 1];
    for (int j = 0; j < C; ++j) {
      const long double u = cosl((2 * j + 1) * pi / (2 * C));   // node in [-1, 1]; t = u h / 2
      long double pw = 1.0L;
      for (int q = 0; q < C; ++q) { a[j][q] = pw; pw *= u; }
      a[j][C] = log1pl(expl(centre + u * h / 2));
    }
    for (int col = 0; col < C; ++col) {   // Gauss-Jordan with partial pivoting
      int piv = col;
      for (int r = col + 1; r < C; ++r) if (fabsl(a[r][col]) > fabsl(a[piv][col])) piv = r;
      for (int q = 0; q <= C; ++q) std::swap(a[col][q], a[piv][q]);
      for (int r = 0; r < C; ++r) {
        if (r == col) continue;
        const long double f = a[r][col] / a[col][col];
        for (int q = col; q <= C; ++q) a[r][q] -= f * a[col][q];
      }
    }
    long double scale = 1.0L;   // coefficient of u^q -> of t^q: / (h/2)^q
    for (int q = 0; q < C; ++q) {
      (*tab)[(size_t)k * C + q] = (double)(a[q][C] / a[q][q] / scale);
      scale *= h / 2;
    }
  }
}

int build_estep_trie(const std::vector<std::string> &pieces, const std::vector<float> &scores, HostTrie *out,
                     float *min_score, std::string *err, std::vector<int32_t> *hot_ids) {
  if (pieces.empty() || pieces.size() != scores.size()) { *err = "no pieces are loaded."; return kInternal; }
  if (pieces.size() >= (1u << 28)) { *err = "vocabulary too large"; return kUnimplemented; }
  std::vector<Key> keys;
  std::map<std::string, int> seen;
  float mn = FLT_MAX;
  // hot slots (spm_estep.cuh): the kEstepHot - 1 pieces with the highest scores, when ids fit 20 bits
  std::vector<uint32_t> hot(pieces.size(), 0);
  if (hot_ids) {
    hot_ids->assign(1, -1);                                  // slot 0 = "not hot"
    if (pieces.size() <= (1u << 20)) {
      std::vector<int32_t> order(pieces.size());
      for (size_t i = 0; i < order.size(); ++i) order[i] = (int32_t)i;
      const size_t k = std::min<size_t>(order.size(), kEstepHot - 1);
      std::partial_sort(order.begin(), order.begin() + k, order.end(), [&](int32_t a, int32_t b) {
        return scores[a] > scores[b] || (scores[a] == scores[b] && a < b);
      });
      for (size_t j = 0; j < k; ++j) { hot[order[j]] = (uint32_t)(j + 1); hot_ids->push_back(order[j]); }
    }
  }
  for (size_t i = 0; i < pieces.size(); ++i) {
    if (pieces[i].empty()) { *err = "piece must not be empty."; return kInternal; }
    if (!seen.emplace(pieces[i], (int)i).second) { *err = pieces[i] + " is already defined."; return kInternal; }
    mn = std::min(mn, scores[i]);
    const uint32_t x = hot_ids ? ((uint32_t)i | (hot[i] << 20)) : ((uint32_t)i | ((uint32_t)T_NORMAL << 28));
    keys.push_back(Key{pieces[i], x, f2u(scores[i]), exp((double)scores[i])});
  }
  *min_score = mn;
  if (!build_xda(keys, out, err)) return kUnimplemented;
  return kOk;
}

int piece_to_id(const HostModel &m, const std::string &piece) {
  // reserved_id_map_ and pieces_ are disjoint key sets over all pieces, so a
  // linear probe by string is the same lookup; only used at option-parse time.
  for (size_t i = 0; i < m.piece.size(); ++i)
    if (m.piece[i] == piece) return (int)i;
  return m.unk_id;
}

static int char_count(const std::string &s) {
  int n = 0;
  for (size_t i = 0; i < s.size();) { i += one_char_len((uint8_t)s[i]); ++n; }
  return n;
}

void refresh_leaf_types(HostModel *m) {
  m->has_unused = false;
  for (auto t : m->type) if (t == T_UNUSED) m->has_unused = true;
  for (auto &lf : m->pieces.leaf)
    if (lf.x != 0xFFFFFFFFu) {
      const uint32_t id = lf.x & 0x0FFFFFFFu;
      lf.x = id | ((uint32_t)m->type[id] << 28);
    }
  // fast view (unigram): a node is a hit iff PopulateNodes would insert it
  // (src/unigram_model.cc:582-595: UNUSED skipped; reserved pieces are not in trie_)
  const size_t n = m->pieces.units.size();
  m->funits = m->pieces.units;
  m->fscore.assign(n, 0.0f);
  m->fleaf.assign(n, 0);
  for (size_t i = 0; i < n; ++i) {
    const uint2 lf = m->pieces.leaf[i];
    if (lf.x == 0xFFFFFFFFu) continue;
    const uint32_t id = lf.x & 0x0FFFFFFFu, type = lf.x >> 28;
    m->fleaf[i] = id;
    if (type == T_NORMAL) {
      m->fscore[i] = m->score[id];
    } else if (type == T_USER_DEFINED) {
      // length * max_score_ + 1.0: float product, double add, float store (:590-591)
      const float len = (float)char_count(m->piece[id]);
      volatile float prod = len * m->max_score;
      m->fscore[i] = (float)((double)prod + 1.0);
    } else {
      m->funits[i] &= ~0x100u;
      // a node that is neither a hit nor a parent must stop the walk
      if (!(m->funits[i] & 0x200u)) m->funits[i] = 0;
    }
  }
  // fast_bpe() preconditions: escaped whitespace as a word prefix, no piece carries U+2581
  // at a non-initial position (no merge crosses a word boundary), no frozen (USER_DEFINED)
  // symbols, no UNUSED pieces to re-split, no single-character reserved piece (PieceToId of
  // a lone character consults the reserved map first, src/model_interface.cc:87-97).
  bool ok = m->model_type == M_BPE && m->text_valid && m->escape_ws && !m->ws_suffix &&
            m->pieces.units.size() < 0x00FFFFFFu;
  for (size_t i = 0; i < m->piece.size() && ok; ++i) {
    const std::string &pc = m->piece[i];
    const uint8_t t = m->type[i];
    if (t == T_USER_DEFINED || t == T_UNUSED) ok = false;
    if (t != T_NORMAL && char_count(pc) == 1) ok = false;
    if (pc.find("\xE2\x96\x81", 1) != std::string::npos) ok = false;
  }
  m->bpe_fast_ok = ok;
  // word memo (spm_memo.cuh): every node that covers a U+2581 must start there
  bool inner = false;
  for (size_t i = 0; i < m->piece.size() && !inner; ++i) {
    const uint8_t t = m->type[i];
    if ((t == T_NORMAL || t == T_USER_DEFINED) && m->piece[i].find("\xE2\x96\x81", 1) != std::string::npos)
      inner = true;
  }
  m->memo_ok = (m->model_type == M_UNIGRAM ? m->fast_ok : m->bpe_fast_ok) && !inner && !m->ws_suffix;
  bool neg = m->min_score - 10.0f <= 0.0f;
  for (size_t i = 0; i < n && neg; ++i)
    if ((m->funits[i] & 0x100u) && !(m->fscore[i] <= 0.0f)) neg = false;
  m->neg_only = neg;
}

int build_host_model(const void *data, size_t len, HostModel *m) {
  if (data == nullptr || len == 0) { m->error = "model is empty."; return kInternal; }
  if (!parse_proto(data, len, m)) { m->error = "could not parse ModelProto"; return kInternal; }
  if (m->model_type != M_UNIGRAM && m->model_type != M_BPE) {
    m->error = "model_type WORD/CHAR is outside the GPU batch-encode path";
    return kUnimplemented;
  }
  // empty names fall back to the defaults (src/model_interface.cc:30-48)
  if (m->unk_piece.empty()) m->unk_piece = "<unk>";
  if (m->bos_piece.empty()) m->bos_piece = "<s>";
  if (m->eos_piece.empty()) m->eos_piece = "</s>";
  if (m->pad_piece.empty()) m->pad_piece = "<pad>";
  const size_t V = m->piece.size();
  if (V == 0) { m->error = "no pieces are loaded."; return kInternal; }
  if (V >= (1u << 28)) { m->error = "vocabulary too large"; return kUnimplemented; }
  // InitializePieces (src/model_interface.cc:101-144)
  std::map<std::string, int> seen_normal, seen_reserved;
  m->unk_id = -1;
  for (size_t i = 0; i < V; ++i) {
    if (m->piece[i].empty()) { m->error = "piece must not be empty."; return kInternal; }
    auto &dst = is_trie_type(m->type[i]) ? seen_normal : seen_reserved;
    if (!dst.emplace(m->piece[i], (int)i).second) { m->error = m->piece[i] + " is already defined."; return kInternal; }
    if (m->type[i] == T_UNKNOWN) {
      if (m->unk_id >= 0) { m->error = "unk is already defined."; return kInternal; }
      m->unk_id = (int)i;
    }
    m->max_piece_bytes = std::max<int>(m->max_piece_bytes, (int)m->piece[i].size());
    m->max_piece_chars = std::max<int>(m->max_piece_chars, char_count(m->piece[i]));
  }
  if (m->unk_id < 0) { m->error = "unk is not defined."; return kInternal; }
  if (m->max_piece_bytes > 4000) { m->error = "piece longer than 4000 bytes"; return kUnimplemented; }
  // min/max over NORMAL (src/unigram_model.cc:682-689; max starts at FLT_MIN)
  m->min_score = FLT_MAX;
  m->max_score = FLT_MIN;
  for (size_t i = 0; i < V; ++i)
    if (m->type[i] == T_NORMAL) {
      m->min_score = std::min(m->min_score, m->score[i]);
      m->max_score = std::max(m->max_score, m->score[i]);
    }
  // piece trie.  A string present both as a reserved and as a normal piece keeps
  // the reserved id for PieceToId (reserved map is consulted first) -- the
  // reference allows that combination, we do not.
  std::vector<Key> keys, ukeys;
  for (size_t i = 0; i < V; ++i) {
    if (!is_trie_type(m->type[i]) && seen_normal.count(m->piece[i])) {
      m->error = m->piece[i] + " is defined both as a reserved and a normal piece";
      return kUnimplemented;
    }
    // weight: unigram log-prob, or BPE merge rank (score = -rank)
    const double w = m->model_type == M_UNIGRAM ? exp((double)m->score[i]) : 1.0 / (2.0 + fabs((double)m->score[i]));
    keys.push_back(Key{m->piece[i], (uint32_t)i | ((uint32_t)m->type[i] << 28), f2u(m->score[i]), w});
    if (m->type[i] == T_USER_DEFINED) {
      ukeys.push_back(Key{m->piece[i], (uint32_t)i | ((uint32_t)T_USER_DEFINED << 28), 0, 1.0});
      m->user_first[(uint8_t)m->piece[i][0] >> 5] |= 1u << ((uint8_t)m->piece[i][0] & 31);
    }
  }
  if (!build_xda(keys, &m->pieces, &m->error)) return kUnimplemented;
  if (!build_xda(ukeys, &m->user, &m->error)) return kUnimplemented;
  refresh_leaf_types(m);
  // charsmap blob (src/normalizer.cc:318-337)
  m->cm_n = 0;
  if (!m->charsmap.empty()) {
    uint32_t tb = 0;
    if (m->charsmap.size() <= 4) { m->error = "Blob for normalization rule is broken."; return kInternal; }
    memcpy(&tb, m->charsmap.data(), 4);
    if (tb >= m->charsmap.size() || tb < 4 || (tb & 3)) { m->error = "Blob for normalization rule is broken."; return kInternal; }
    m->cm_units = (const uint32_t *)(m->charsmap.data() + 4);
    m->cm_n = tb / 4;
    m->cm_pool = (const uint8_t *)m->charsmap.data() + 4 + tb;
    m->cm_pool_size = (uint32_t)(m->charsmap.size() - 4 - tb);
  }
  scan_charsmap(m);
  // printable ASCII passes through the normalizer untouched when no rule / user symbol can
  // start at such a byte as long as the NEXT byte is ASCII too (fast_normalize_scan's runs)
  m->ascii_plain = true;
  for (uint32_t b = 0x21; b <= 0x7E && m->ascii_plain; ++b) {
    if ((m->user_first[b >> 5] >> (b & 31)) & 1) m->ascii_plain = false;
    if (!m->cm_first2.empty())
      for (uint32_t x = 0; x < 0x80; ++x)
        if ((m->cm_first2[((b << 8) | x) >> 5] >> (((b << 8) | x) & 31)) & 1) m->ascii_plain = false;
  }
  // TAB, LF, FF and CR are word separators for the byte-parallel word scan when the normalizer turns each of them
  // into exactly one space whatever follows (nmt_nfkc does; text with CRLF line ends otherwise takes the general
  // scan sentence by sentence): normalize_prefix() itself is asked, on host copies of the tables.
  {
    ModelView v;
    memset(&v, 0, sizeof(v));
    v.user.units = m->user.units.data(); v.user.leaf = m->user.leaf.data(); v.user.n = (uint32_t)m->user.units.size();
    v.cm_units = m->cm_units; v.cm_pool = m->cm_pool; v.cm_n = m->cm_n;
    v.cm_first2 = m->cm_first2.empty() ? nullptr : m->cm_first2.data();
    memcpy(v.user_first, m->user_first, sizeof(v.user_first));
    m->ws_ctl = m->cm_n != 0;
    for (int ctl : {0x09, 0x0A, 0x0C, 0x0D})
    for (int x = -1; x < 256 && m->ws_ctl; ++x) {     // -1: the byte is the last one
      uint8_t in[8] = {(uint8_t)ctl, (uint8_t)(x < 0 ? 0 : x), 0, 0, 0, 0, 0, 0};
      const uint8_t *rep = nullptr;
      int rl = 0;
      const int c = normalize_prefix(v, in, x < 0 ? 1 : 2, &rep, &rl);
      if (!(c == 1 && rl == 1 && rep != nullptr && rep[0] == ' ')) m->ws_ctl = false;
    }
  }
  m->ring = 8;
  while (m->ring < m->max_piece_bytes + 5) m->ring <<= 1;
  float smax = fabsf(m->min_score - 10.0f);
  for (size_t i = 0; i < V; ++i) smax = std::max(smax, fabsf(m->score[i]));
  smax = std::max(smax, (float)m->max_piece_bytes * fabsf(m->max_score) + 1.0f);
  m->smax = smax;
  // The fast tier counts characters by lead bytes, which equals the reference's
  // OneCharLen jumps only on valid UTF-8.  Normalized text is valid UTF-8 unless a
  // user-defined symbol (copied through raw) or a charsmap replacement is not.
  auto valid_utf8 = [](const uint8_t *b, size_t n) {
    for (size_t i = 0; i < n;) {
      int mb = 1;
      if (!valid_decode_utf8(b + i, (int64_t)(n - i), &mb)) return false;
      i += mb;
    }
    return true;
  };
  bool text_valid = true;
  for (size_t i = 0; i < V; ++i)
    if (m->type[i] == T_USER_DEFINED && !valid_utf8((const uint8_t *)m->piece[i].data(), m->piece[i].size())) text_valid = false;
  for (uint32_t a = 0; a < m->cm_pool_size && text_valid;) {
    uint32_t e = a;
    while (e < m->cm_pool_size && m->cm_pool[e] != 0) ++e;
    if (!valid_utf8(m->cm_pool + a, e - a)) text_valid = false;
    a = e + 1;
  }
  m->text_valid = text_valid;
  m->fast_ok = m->model_type == M_UNIGRAM && text_valid && m->pieces.units.size() < 0x00FFFFFFu &&
               m->max_piece_chars < 200;
  refresh_leaf_types(m);  // (re)derives bpe_fast_ok as well
  // piece strings by id (spans_from_ids of the pieces path, spm_core.cuh)
  // (plen[V + i]: bytes of the piece's surface in a plain-ASCII sentence, where every U+2581 stands for one
  // space -- PieceTable::olen, spans_fit_from_ids)
  m->plen.resize(2 * V);
  m->poff.resize(V);
  m->ppool.clear();
  for (size_t i = 0; i < V; ++i) {
    m->plen[i] = (int32_t)m->piece[i].size();
    int32_t nws = 0;
    for (size_t b = 0; b + 2 < m->piece[i].size(); ++b)
      if ((uint8_t)m->piece[i][b] == 0xE2 && (uint8_t)m->piece[i][b + 1] == 0x96 && (uint8_t)m->piece[i][b + 2] == 0x81) ++nws;
    m->plen[V + i] = m->plen[i] - 2 * nws;
    m->poff[i] = (uint32_t)m->ppool.size();
    m->ppool += m->piece[i];
  }
  m->ppool += '\0';
  return kOk;
}

}  // namespace spmb
