// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// C-ABI shim around the UNMODIFIED reference SentencePieceProcessor, compiled
// from the sources where they lie under /root/reference (see build_ref.py).
// The result, oracle/_ref/libspm_ref.so, is the parity checker and the
// "reference" CPU baseline.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load it.
//
// Reference interfaces wrapped here:
//   SentencePieceProcessor::LoadFromSerializedProto  src/sentencepiece_processor.h:203
//   SentencePieceProcessor::Encode(ids)              src/sentencepiece_processor.h:244
//   SentencePieceProcessor::SetEncodeExtraOptions    src/sentencepiece_processor.h:210
//   SentencePieceProcessor::SetVocabulary            src/sentencepiece_processor.h:224
//   SentencePieceProcessor::EncodeAsSerializedProto  src/sentencepiece_processor.h:381 (.cc:752-758)
//   SentencePieceProcessor::Decode(ids, string*)     src/sentencepiece_processor.h:252 (.cc:670-733)
//   normalizer::Normalizer::Normalize                src/normalizer.h
//   SentencePieceTrainer::Train(args)                src/sentencepiece_trainer.h:59
//   normalizer::Builder::CompileCharsMap             src/builder.h (fixture charsmaps for the golden tests)
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "normalizer.h"
#include "sentencepiece.pb.h"
#include "sentencepiece_model.pb.h"
#include "sentencepiece_processor.h"
#ifdef REF_WITH_TRAINER
#include "builder.h"
#include "sentencepiece_trainer.h"
#include "unigram_model_trainer.h"
#endif

namespace {
struct RefHandle {
  sentencepiece::SentencePieceProcessor sp;
  std::unique_ptr<sentencepiece::normalizer::Normalizer> norm;
};
thread_local std::string g_err;
}  // namespace

extern "C" {

const char *ref_last_error() { return g_err.c_str(); }

void *ref_load(const void *model, size_t len) {
  auto *h = new RefHandle;
  auto st = h->sp.LoadFromSerializedProto(
      std::string(static_cast<const char *>(model), len));
  if (!st.ok()) {
    g_err = st.ToString();
    delete h;
    return nullptr;
  }
  h->norm.reset(new sentencepiece::normalizer::Normalizer(
      h->sp.model_proto().normalizer_spec(), h->sp.model_proto().trainer_spec()));
  return h;
}

void ref_free(void *p) { delete static_cast<RefHandle *>(p); }

int ref_set_encode_extra_options(void *p, const char *opts) {
  auto st = static_cast<RefHandle *>(p)->sp.SetEncodeExtraOptions(opts);
  if (!st.ok()) g_err = st.ToString();
  return static_cast<int>(st.code());
}

// pieces: NUL-separated list of n valid pieces; n < 0 => ResetVocabulary.
int ref_set_vocabulary(void *p, const char *pieces, int n) {
  auto *h = static_cast<RefHandle *>(p);
  sentencepiece::util::Status st;
  if (n < 0) {
    st = h->sp.ResetVocabulary();
  } else {
    std::vector<std::string> v;
    const char *q = pieces;
    for (int i = 0; i < n; ++i) {
      v.emplace_back(q);
      q += v.back().size() + 1;
    }
    st = h->sp.SetVocabulary(v);
  }
  if (!st.ok()) g_err = st.ToString();
  return static_cast<int>(st.code());
}

int ref_piece_size(void *p) { return static_cast<RefHandle *>(p)->sp.GetPieceSize(); }
int ref_piece_type(void *p, int id) {
  return static_cast<RefHandle *>(p)->sp.model_proto().pieces(id).type();
}

// Batch encode with `threads` std::threads (Encode is const / re-entrant,
// src/unigram_model.cc:710).  Two-call protocol: ids == nullptr => only
// id_offsets is filled (sizes); else ids must hold id_offsets[n] entries.
// Returns 0, or the first non-OK status code.
int ref_encode_batch(void *p, const char *text, const int64_t *off, int64_t n,
                     int threads, int32_t *ids, int64_t cap, int64_t *id_off) {
  auto *h = static_cast<RefHandle *>(p);
  if (threads < 1) threads = 1;
  std::vector<std::vector<int>> out(n);
  std::vector<int> codes(threads, 0);
  auto work = [&](int t) {
    const int64_t lo = n * t / threads, hi = n * (t + 1) / threads;
    for (int64_t i = lo; i < hi; ++i) {
      auto st = h->sp.Encode(std::string(text + off[i], off[i + 1] - off[i]), &out[i]);
      if (!st.ok() && codes[t] == 0) codes[t] = static_cast<int>(st.code());
    }
  };
  if (threads == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < threads; ++t) th.emplace_back(work, t);
    for (auto &x : th) x.join();
  }
  id_off[0] = 0;
  for (int64_t i = 0; i < n; ++i) id_off[i + 1] = id_off[i] + (int64_t)out[i].size();
  if (ids != nullptr) {
    if (id_off[n] > cap) return 8;  // RESOURCE_EXHAUSTED
    for (int64_t i = 0; i < n; ++i)
      if (!out[i].empty())
        std::memcpy(ids + id_off[i], out[i].data(), out[i].size() * sizeof(int));
  }
  for (int c : codes)
    if (c) return c;
  return 0;
}

// Timed variant for the CPU baseline: encodes and discards, returns total ids.
int64_t ref_encode_count(void *p, const char *text, const int64_t *off, int64_t n,
                         int threads) {
  auto *h = static_cast<RefHandle *>(p);
  if (threads < 1) threads = 1;
  std::vector<int64_t> tot(threads, 0);
  auto work = [&](int t) {
    const int64_t lo = n * t / threads, hi = n * (t + 1) / threads;
    std::vector<int> ids;
    for (int64_t i = lo; i < hi; ++i) {
      h->sp.Encode(std::string(text + off[i], off[i + 1] - off[i]), &ids).IgnoreError();
      tot[t] += (int64_t)ids.size();
    }
  };
  std::vector<std::thread> th;
  for (int t = 0; t < threads; ++t) th.emplace_back(work, t);
  for (auto &x : th) x.join();
  int64_t s = 0;
  for (auto v : tot) s += v;
  return s;
}

// Encode(input, SentencePieceText*) per sentence, `threads` std::threads; the
// serialized SentencePieceText protos go back to back into out (CSR out_off[n+1]).
// Two-call protocol as above: out == nullptr => sizes only.  count_only: nothing is
// kept (timed CPU baseline of the pieces path), returns the total proto bytes.
int64_t ref_encode_proto_batch(void *p, const char *text, const int64_t *off, int64_t n, int threads,
                               char *out, int64_t cap, int64_t *out_off, int count_only) {
  auto *h = static_cast<RefHandle *>(p);
  if (threads < 1) threads = 1;
  std::vector<std::string> res(count_only ? 0 : n);
  std::vector<int64_t> tot(threads, 0);
  auto work = [&](int t) {
    const int64_t lo = n * t / threads, hi = n * (t + 1) / threads;
    for (int64_t i = lo; i < hi; ++i) {
      sentencepiece::SentencePieceText spt;
      auto st = h->sp.Encode(std::string(text + off[i], off[i + 1] - off[i]), &spt);
      if (!st.ok()) spt.Clear();
      if (count_only) tot[t] += (int64_t)spt.pieces_size();
      else res[i] = spt.SerializeAsString();
    }
  };
  std::vector<std::thread> th;
  for (int t = 0; t < threads; ++t) th.emplace_back(work, t);
  for (auto &x : th) x.join();
  if (count_only) {
    int64_t s = 0;
    for (auto v : tot) s += v;
    return s;
  }
  out_off[0] = 0;
  for (int64_t i = 0; i < n; ++i) out_off[i + 1] = out_off[i] + (int64_t)res[i].size();
  if (out != nullptr) {
    if (out_off[n] > cap) return -8;
    for (int64_t i = 0; i < n; ++i) std::memcpy(out + out_off[i], res[i].data(), res[i].size());
  }
  return out_off[n];
}

// Decode(ids, std::string*) per sentence; CSR in, CSR out (same two-call protocol).
int64_t ref_decode_batch(void *p, const int32_t *ids, const int64_t *id_off, int64_t n, int threads,
                         char *out, int64_t cap, int64_t *out_off, int count_only) {
  auto *h = static_cast<RefHandle *>(p);
  if (threads < 1) threads = 1;
  std::vector<std::string> res(count_only ? 0 : n);
  std::vector<int64_t> tot(threads, 0);
  auto work = [&](int t) {
    const int64_t lo = n * t / threads, hi = n * (t + 1) / threads;
    std::string text;
    for (int64_t i = lo; i < hi; ++i) {
      std::vector<int> v(ids + id_off[i], ids + id_off[i + 1]);
      h->sp.Decode(v, &text).IgnoreError();
      if (count_only) tot[t] += (int64_t)text.size();
      else res[i] = text;
    }
  };
  std::vector<std::thread> th;
  for (int t = 0; t < threads; ++t) th.emplace_back(work, t);
  for (auto &x : th) x.join();
  if (count_only) {
    int64_t s = 0;
    for (auto v : tot) s += v;
    return s;
  }
  out_off[0] = 0;
  for (int64_t i = 0; i < n; ++i) out_off[i + 1] = out_off[i] + (int64_t)res[i].size();
  if (out != nullptr) {
    if (out_off[n] > cap) return -8;
    for (int64_t i = 0; i < n; ++i) std::memcpy(out + out_off[i], res[i].data(), res[i].size());
  }
  return out_off[n];
}

// Normalizer only.  out must hold cap bytes; n2o (may be null) cap+1 entries.
// Returns the normalized length, or -1 when cap is too small.
int64_t ref_normalize(void *p, const char *text, int64_t len, char *out, int64_t cap,
                      int64_t *n2o) {
  auto *h = static_cast<RefHandle *>(p);
  std::string norm;
  std::vector<size_t> map;
  h->norm->SetPrefixMatcher(nullptr);
  auto st = h->norm->Normalize(std::string(text, len), &norm, &map);
  if (!st.ok()) {
    g_err = st.ToString();
    return -2;
  }
  if ((int64_t)norm.size() > cap) return -1;
  std::memcpy(out, norm.data(), norm.size());
  if (n2o)
    for (size_t i = 0; i < map.size(); ++i) n2o[i] = (int64_t)map[i];
  return (int64_t)norm.size();
}

#ifdef REF_WITH_TRAINER
// rules: n pairs of NUL-terminated UTF-8 strings (src, trg).  Writes the
// precompiled charsmap blob (src/builder.cc:161-225) and returns its size.
int64_t ref_compile_charsmap(const char *rules, int n, char *out, int64_t cap) {
  using namespace sentencepiece;
  normalizer::Builder::CharsMap cm;
  const char *q = rules;
  for (int i = 0; i < n; ++i) {
    std::string a(q); q += a.size() + 1;
    std::string b(q); q += b.size() + 1;
    normalizer::Builder::Chars ca, cb;
    for (const char32 c : string_util::UTF8ToUnicodeText(a)) ca.push_back(c);
    for (const char32 c : string_util::UTF8ToUnicodeText(b)) cb.push_back(c);
    cm[ca] = cb;
  }
  std::string blob;
  auto st = normalizer::Builder::CompileCharsMap(cm, &blob);
  if (!st.ok()) { g_err = st.ToString(); return -1; }
  if ((int64_t)blob.size() > cap) return -2;
  std::memcpy(out, blob.data(), blob.size());
  return (int64_t)blob.size();
}

int ref_train(const char *args) {
  auto st = sentencepiece::SentencePieceTrainer::Train(args);
  if (!st.ok()) g_err = st.ToString();
  return static_cast<int>(st.code());
}

// Unigram E-step over (word, freq) with the reference TrainerModel
// (src/unigram_model_trainer.cc:237-287, single thread, stride order).
// pieces: NUL-separated; scores[V].  words: NUL-separated normalized words.
// Outputs expected[V] (float, reference accumulation order), logz[n] (per
// word Z, float), obj, ntokens.
int ref_estep(const char *pieces, const float *scores, int V, const char *words,
              const int64_t *freq, int64_t n, float *expected, float *logz,
              double *obj, int64_t *ntokens) {
  using namespace sentencepiece;
  TrainerSpec ts;
  NormalizerSpec ns;
  unigram::TrainerModel model(ts, ns);
  unigram::TrainerModel::SentencePieces sps;
  const char *q = pieces;
  for (int i = 0; i < V; ++i) {
    sps.emplace_back(std::string(q), scores[i]);
    q += sps.back().first.size() + 1;
  }
  model.SetSentencePieces(std::move(sps));
  std::vector<float> exp(V, 0.0f);
  int64_t all_freq = 0;
  for (int64_t i = 0; i < n; ++i) all_freq += freq[i];
  float o = 0.0f;
  int64_t nt = 0;
  unigram::Lattice lattice;
  const char *w = words;
  for (int64_t i = 0; i < n; ++i) {
    const size_t wl = std::strlen(w);
    lattice.SetSentence(absl::string_view(w, wl));
    model.PopulateNodes(&lattice);
    const float Z = lattice.PopulateMarginal(freq[i], &exp);
    nt += lattice.Viterbi().size();
    if (logz) logz[i] = Z;
    o -= Z / all_freq;
    w += wl + 1;
  }
  std::memcpy(expected, exp.data(), V * sizeof(float));
  *obj = o;
  *ntokens = nt;
  return 0;
}

// ref_estep plus expected64[V]: the float64 re-accumulation of the reference's OWN posteriors
// (SURVEY.md H9).  The reference sums freq * posterior of every node into a float vector over
// the whole corpus (:321), whose rounding error on large counts exceeds 1e-5; here each word's
// PopulateMarginal runs a second time into a zeroed float scratch, and the word's contribution
// per piece id (a handful of node posteriors) is added to a double.  expected[] stays exactly
// what the reference accumulates.
int ref_estep64(const char *pieces, const float *scores, int V, const char *words,
                const int64_t *freq, int64_t n, float *expected, float *logz,
                double *obj, int64_t *ntokens, double *expected64) {
  using namespace sentencepiece;
  TrainerSpec ts;
  NormalizerSpec ns;
  unigram::TrainerModel model(ts, ns);
  unigram::TrainerModel::SentencePieces sps;
  const char *q = pieces;
  for (int i = 0; i < V; ++i) {
    sps.emplace_back(std::string(q), scores[i]);
    q += sps.back().first.size() + 1;
  }
  model.SetSentencePieces(std::move(sps));
  std::vector<float> exp(V, 0.0f), tmp(V, 0.0f);
  for (int i = 0; i < V; ++i) expected64[i] = 0.0;
  int64_t all_freq = 0;
  for (int64_t i = 0; i < n; ++i) all_freq += freq[i];
  float o = 0.0f;
  int64_t nt = 0;
  unigram::Lattice lattice;
  const char *w = words;
  for (int64_t i = 0; i < n; ++i) {
    const size_t wl = std::strlen(w);
    lattice.SetSentence(absl::string_view(w, wl));
    model.PopulateNodes(&lattice);
    const float Z = lattice.PopulateMarginal(freq[i], &exp);
    lattice.PopulateMarginal(freq[i], &tmp);
    for (int pos = 0; pos < lattice.size(); ++pos) {
      for (const auto *node : lattice.begin_nodes(pos)) {
        if (node->id >= 0 && tmp[node->id] != 0.0f) {
          expected64[node->id] += (double)tmp[node->id];
          tmp[node->id] = 0.0f;
        }
      }
    }
    nt += lattice.Viterbi().size();
    if (logz) logz[i] = Z;
    o -= Z / all_freq;
    w += wl + 1;
  }
  std::memcpy(expected, exp.data(), V * sizeof(float));
  *obj = o;
  *ntokens = nt;
  return 0;
}
#endif

}  // extern "C"
