// spm_cuda_processor.h -- the reference-side binding of the B200 encoder (C++11, header only).
//
// sentencepiece::CudaSentencePieceProcessor derives from the reference's
// sentencepiece::SentencePieceProcessor (src/sentencepiece_processor.h:176-447, every method
// virtual) and routes the encode path through the C ABI of spm_cuda.h.  Code that holds a
// `SentencePieceProcessor *` keeps working unchanged: Load / LoadFromSerializedProto / Encode /
// EncodeAsIds / SetEncodeExtraOptions / SetVocabulary keep their signatures, error codes and
// messages (util::Status carries the C ABI's util::error::Code and spm_cuda_last_error()).
//
// Only the reference's PUBLIC header is needed (no common.h / util.h internals), so the class
// builds inside or outside the reference tree:
//     g++ -std=c++11 -I<reference>/src -I<this repo>/include app.cc -lsentencepiece -lspm_cuda
// With -DSPM_CUDA_PROCESSOR_WITH_PROTO (and the generated sentencepiece.pb.h on the include path)
// Encode(input, SentencePieceText*) -- hence EncodeAsPieces / EncodeAsSerializedProto -- goes
// through spm_cuda_encode_batch_pieces too; without it those stay on the base class.
// tests/facade_driver.cc compiles this file against the unmodified reference and compares both
// classes through base-class pointers (tests/test_facade.py).
#ifndef SPM_CUDA_PROCESSOR_H_
#define SPM_CUDA_PROCESSOR_H_

#include <algorithm>
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

#include "sentencepiece_processor.h"
#ifdef SPM_CUDA_PROCESSOR_WITH_PROTO
#include "sentencepiece.pb.h"
#endif
#include "spm_cuda.h"

namespace sentencepiece {

class CudaSentencePieceProcessor : public SentencePieceProcessor {
 public:
  explicit CudaSentencePieceProcessor(int device = 0) : device_(device) {}
  ~CudaSentencePieceProcessor() override { spm_cuda_model_free(h_); }

  // Load(filename) / Load(istream) of the base class end in LoadFromSerializedProto or
  // Load(unique_ptr<ModelProto>); the serialized form is what the GPU loader reads.
  util::Status LoadFromSerializedProto(util::min_string_view serialized) override {
    util::Status st = SentencePieceProcessor::LoadFromSerializedProto(serialized);   // keeps model_proto(), id <-> piece maps
    if (!st.ok()) return st;
    spm_cuda_model *h = nullptr;
    const int rc = spm_cuda_model_load(serialized.data(), serialized.size(), device_, &h);
    if (rc) return Error(rc);
    spm_cuda_model_free(h_);
    h_ = h;
    extra_ = 0;
    return util::Status();
  }

  // "bos:eos:reverse" in application order (src/sentencepiece_processor.cc:945-1023)
  util::Status SetEncodeExtraOptions(util::min_string_view opts) override {
    util::Status st = SentencePieceProcessor::SetEncodeExtraOptions(opts);   // validates names and ids
    if (!st.ok()) return st;
    uint32_t packed = 0;
    int k = 0;
    const std::string s(opts.data(), opts.size());
    for (size_t a = 0; a <= s.size() && !s.empty();) {
      size_t b = s.find(':', a);
      if (b == std::string::npos) b = s.size();
      const std::string o = s.substr(a, b - a);
      const uint32_t code = o == "bos" ? SPM_CUDA_OPT_BOS : o == "eos" ? SPM_CUDA_OPT_EOS : SPM_CUDA_OPT_REVERSE;
      if (k >= 16) return util::Status(util::error::INVALID_ARGUMENT, "too many extra options");
      packed |= code << (2 * k++);
      a = b + 1;
    }
    extra_ = packed;
    return util::Status();
  }

  util::Status SetVocabulary(const std::vector<std::string> &valid_vocab) override {
    util::Status st = SentencePieceProcessor::SetVocabulary(valid_vocab);
    return st.ok() ? PushPieceTypes() : st;
  }
  util::Status ResetVocabulary() override {
    util::Status st = SentencePieceProcessor::ResetVocabulary();
    return st.ok() ? PushPieceTypes() : st;
  }
  util::Status LoadVocabulary(util::min_string_view filename, int threshold) override {
    util::Status st = SentencePieceProcessor::LoadVocabulary(filename, threshold);   // ends in SetVocabulary (virtual)
    return st;
  }

  // One sentence = a one-row batch.
  util::Status Encode(util::min_string_view input, std::vector<int> *ids) const override {
    if (ids == nullptr) return util::Status(util::error::INTERNAL, "output container is null");
    ids->clear();
    if (h_ == nullptr) return SentencePieceProcessor::status();
    const int64_t off[2] = {0, static_cast<int64_t>(input.size())};
    return EncodeBatch(input.data(), off, 1, ids, nullptr);
  }

  // The batch entry point: what the line loop of spm_encode (src/spm_encode_main.cc:183-193) and
  // the TF op loop (tensorflow/sentencepiece_processor_ops.cc:288-304) become.  Sentence i is
  // text[offsets[i], offsets[i+1]); ids of sentence i are (*ids)[(*id_offsets)[i] .. (*id_offsets)[i+1]).
  util::Status EncodeBatch(const char *text, const int64_t *offsets, int64_t n, std::vector<int> *ids,
                           std::vector<int64_t> *id_offsets) const {
    if (ids == nullptr) return util::Status(util::error::INTERNAL, "output container is null");
    if (h_ == nullptr) return SentencePieceProcessor::status();
    static_assert(sizeof(int) == sizeof(int32_t), "ids are 32-bit");
    std::vector<int64_t> io(n + 1, 0);
    ids->resize(static_cast<size_t>(std::max<int64_t>(64, (offsets[n] - offsets[0]) / 2 + 4 * n)));
    int rc = spm_cuda_encode_batch(h_, text, offsets, n, extra_, ids->data(), static_cast<int64_t>(ids->size()),
                                   io.data(), SPM_CUDA_HOST, nullptr);
    if (rc == util::error::RESOURCE_EXHAUSTED) {   // io[n] holds the capacity to retry with
      ids->resize(static_cast<size_t>(io[n]));
      rc = spm_cuda_encode_batch(h_, text, offsets, n, extra_, ids->data(), static_cast<int64_t>(ids->size()),
                                 io.data(), SPM_CUDA_HOST, nullptr);
    }
    if (rc) { ids->clear(); return Error(rc); }
    ids->resize(static_cast<size_t>(io[n]));
    if (id_offsets) *id_offsets = std::move(io);
    return util::Status();
  }

  util::Status Decode(const std::vector<int> &ids, std::string *detokenized) const override {
    if (detokenized == nullptr) return util::Status(util::error::INTERNAL, "output container is null");
    detokenized->clear();
    if (h_ == nullptr) return SentencePieceProcessor::status();
    const int64_t io[2] = {0, static_cast<int64_t>(ids.size())};
    int64_t to[2] = {0, 0};
    std::string out(ids.size() * 8 + 64, '\0');
    int rc = spm_cuda_decode_batch(h_, ids.data(), io, 1, &out[0], static_cast<int64_t>(out.size()), to, SPM_CUDA_HOST,
                                   nullptr);
    if (rc == util::error::RESOURCE_EXHAUSTED) {
      out.assign(static_cast<size_t>(to[1]), '\0');
      rc = spm_cuda_decode_batch(h_, ids.data(), io, 1, &out[0], static_cast<int64_t>(out.size()), to, SPM_CUDA_HOST,
                                 nullptr);
    }
    if (rc) return Error(rc);
    detokenized->assign(out.data(), static_cast<size_t>(to[1]));
    return util::Status();
  }

#ifdef SPM_CUDA_PROCESSOR_WITH_PROTO
  // Encode(input, SentencePieceText*): pieces, surfaces and byte spans (.cc:488-573)
  util::Status Encode(util::min_string_view input, SentencePieceText *spt) const override {
    if (spt == nullptr) return util::Status(util::error::INTERNAL, "output proto is null");
    spt->Clear();
    if (h_ == nullptr) return SentencePieceProcessor::status();
    const int64_t off[2] = {0, static_cast<int64_t>(input.size())};
    int64_t cap = static_cast<int64_t>(input.size()) + 40, ncap = 3 * static_cast<int64_t>(input.size()) + 64;
    std::vector<int> id, b, e, nb, ne;
    std::string norm;
    int64_t io[2] = {0, 0}, no[2] = {0, 0};
    int rc = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
      id.resize(cap); b.resize(cap); e.resize(cap); nb.resize(cap); ne.resize(cap);
      norm.assign(static_cast<size_t>(ncap), '\0');
      rc = spm_cuda_encode_batch_pieces(h_, input.data(), off, 1, extra_, id.data(), b.data(), e.data(), nb.data(),
                                        ne.data(), cap, io, &norm[0], ncap, no, SPM_CUDA_HOST, nullptr);
      if (rc != util::error::RESOURCE_EXHAUSTED) break;
      cap = std::max(cap, io[1]);
      ncap = std::max(ncap, no[1]);
    }
    if (rc) return Error(rc);
    spt->set_text(input.data(), input.size());
    for (int64_t k = 0; k < io[1]; ++k) {
      auto *sp = spt->add_pieces();
      sp->set_id(id[k]);
      if (nb[k] < 0) { sp->set_piece(IdToPiece(id[k])); continue; }        // bos / eos of the extra options
      sp->set_piece(norm.data() + nb[k], static_cast<size_t>(ne[k] - nb[k]));
      sp->set_surface(input.data() + b[k], static_cast<size_t>(e[k] - b[k]));
      sp->set_begin(b[k]);
      sp->set_end(e[k]);
    }
    return util::Status();
  }
  // (the base class's Encode(input, std::vector<std::string>*) calls Encode(input, &spt): virtual)
#endif

  const spm_cuda_model *handle() const { return h_; }

 private:
  static util::Status Error(int rc) {
    return util::Status(static_cast<util::error::Code>(rc), spm_cuda_last_error());
  }
  // SetVocabulary flips NORMAL <-> UNUSED in the base class's ModelProto; mirror it on the device.
  util::Status PushPieceTypes() {
    if (h_ == nullptr) return util::Status();
    const int n = spm_cuda_model_piece_size(h_);
    std::vector<uint8_t> types(static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) {
      int t = spm_cuda_model_piece_type(h_, i);
      if (t == 1 || t == 5) t = SentencePieceProcessor::IsUnused(i) ? 5 : 1;   // ModelProto::SentencePiece::Type
      types[static_cast<size_t>(i)] = static_cast<uint8_t>(t);
    }
    const int rc = spm_cuda_model_set_piece_types(h_, types.data(), n);
    return rc ? Error(rc) : util::Status();
  }

  spm_cuda_model *h_ = nullptr;
  uint32_t extra_ = 0;
  int device_ = 0;
};

}  // namespace sentencepiece
#endif  // SPM_CUDA_PROCESSOR_H_
