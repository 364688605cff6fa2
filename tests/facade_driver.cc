// Test driver (test infrastructure): include/spm_cuda_processor.h compiled against the UNMODIFIED
// reference header (/root/reference/src/sentencepiece_processor.h) and linked with the reference
// build oracle/_ref/libspm_ref.so and the product library libspm_cuda.so.  Every call below goes
// through a `SentencePieceProcessor *`, the way existing callers of the reference hold it; the
// base class (the reference itself) is the checker.  Returns the number of mismatches, or a
// negative code with a message in err.
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#define SPM_CUDA_PROCESSOR_WITH_PROTO 1
#include "spm_cuda_processor.h"

using sentencepiece::SentencePieceProcessor;

static void set_err(char *err, int cap, const std::string &m) {
  if (err && cap > 0) { std::snprintf(err, (size_t)cap, "%s", m.c_str()); }
}

extern "C" int facade_run(const char *model, size_t model_len, const char *text, const int64_t *off, int64_t n,
                          const char *extra_options, const char *vocab_nul_separated, int n_vocab, char *err,
                          int errcap) {
  std::unique_ptr<SentencePieceProcessor> cuda(new sentencepiece::CudaSentencePieceProcessor(0));
  SentencePieceProcessor base;
  // not loaded yet: same status as the reference
  {
    std::vector<int> ids;
    const auto a = cuda->Encode("x", &ids), b = base.Encode("x", &ids);
    if (a.code() != b.code()) { set_err(err, errcap, "unloaded Encode: status differs"); return -2; }
  }
  const sentencepiece::util::min_string_view mv(model, model_len);
  auto st = cuda->LoadFromSerializedProto(mv);
  if (!st.ok()) { set_err(err, errcap, std::string("cuda Load: ") + st.ToString()); return -1; }
  st = base.LoadFromSerializedProto(mv);
  if (!st.ok()) { set_err(err, errcap, std::string("base Load: ") + st.ToString()); return -1; }
  if (extra_options && *extra_options) {
    const auto a = cuda->SetEncodeExtraOptions(extra_options), b = base.SetEncodeExtraOptions(extra_options);
    if (a.code() != b.code()) { set_err(err, errcap, "SetEncodeExtraOptions: status differs"); return -3; }
    if (!a.ok()) return 0;
  }
  if (n_vocab >= 0) {
    std::vector<std::string> vocab;
    const char *q = vocab_nul_separated;
    for (int i = 0; i < n_vocab; ++i) { vocab.emplace_back(q); q += vocab.back().size() + 1; }
    const auto a = cuda->SetVocabulary(vocab), b = base.SetVocabulary(vocab);
    if (a.code() != b.code() || !a.ok()) { set_err(err, errcap, "SetVocabulary: " + a.ToString()); return -4; }
  }
  int bad = 0;
  std::string first;
  auto note = [&](const std::string &what, int64_t i) {
    if (bad++ == 0) first = what + " differs at sentence " + std::to_string(i);
  };
  std::vector<int> all;
  for (int64_t i = 0; i < n; ++i) {
    const std::string s(text + off[i], (size_t)(off[i + 1] - off[i]));
    const std::vector<int> want = base.EncodeAsIds(s);
    if (cuda->EncodeAsIds(s) != want) note("EncodeAsIds", i);
    std::vector<int> got;
    if (!cuda->Encode(s, &got).ok() || got != want) note("Encode(ids)", i);
    all.insert(all.end(), want.begin(), want.end());
    if (i % 16 == 0) {     // the slower proto-based calls on a sample
      if (cuda->EncodeAsPieces(s) != base.EncodeAsPieces(s)) note("EncodeAsPieces", i);
      if (cuda->EncodeAsSerializedProto(s) != base.EncodeAsSerializedProto(s)) note("EncodeAsSerializedProto", i);
      if (cuda->DecodeIds(want) != base.DecodeIds(want)) note("DecodeIds", i);
    }
  }
  // the batch entry point of the subclass
  {
    auto *c = static_cast<sentencepiece::CudaSentencePieceProcessor *>(cuda.get());
    std::vector<int> ids;
    std::vector<int64_t> io;
    st = c->EncodeBatch(text, off, n, &ids, &io);
    if (!st.ok()) { set_err(err, errcap, std::string("EncodeBatch: ") + st.ToString()); return -5; }
    if (ids != all || (int64_t)io.size() != n + 1 || io[n] != (int64_t)all.size()) note("EncodeBatch", -1);
  }
  if (n_vocab >= 0) {
    cuda->ResetVocabulary();
    base.ResetVocabulary();
    for (int64_t i = 0; i < n && i < 200; ++i) {
      const std::string s(text + off[i], (size_t)(off[i + 1] - off[i]));
      if (cuda->EncodeAsIds(s) != base.EncodeAsIds(s)) note("EncodeAsIds after ResetVocabulary", i);
    }
  }
  if (cuda->GetPieceSize() != base.GetPieceSize() || cuda->unk_id() != base.unk_id()) note("model introspection", -1);
  if (bad) set_err(err, errcap, first);
  return bad;
}
