// Host-side model loader: reads an unmodified SentencePiece .model protobuf and
// flattens it into the arrays the kernels use (replaces the load-time work of
// src/sentencepiece_processor.cc:119-157, src/model_interface.cc:101-144,
// src/unigram_model.cc:624-695 and src/normalizer.cc:61-80).
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "spm_core.cuh"

namespace spmb {

// util::error::Code values (src/sentencepiece_processor.h:101-119)
enum StatusCode {
  kOk = 0, kInvalidArgument = 3, kNotFound = 5, kResourceExhausted = 8, kOutOfRange = 11,
  kUnimplemented = 12, kInternal = 13,
};

struct HostTrie {
  std::vector<uint32_t> units;
  std::vector<uint2> leaf;
};

struct HostModel {
  // ---- parsed (src/sentencepiece_model.proto)
  std::vector<std::string> piece;
  std::vector<float> score;
  std::vector<uint8_t> type;
  int model_type = M_UNIGRAM;
  bool ws_suffix = false, add_dummy_prefix = true, remove_extra_ws = true, escape_ws = true;
  std::string charsmap;  // precompiled_charsmap blob
  std::string unk_piece = "<unk>", bos_piece = "<s>", eos_piece = "</s>", pad_piece = "<pad>";
  // ModelProto.self_test_data: Load refuses a model whose samples do not reproduce
  // (src/sentencepiece_processor.cc:134-154)
  std::vector<std::string> self_test_input, self_test_expected;
  std::string unk_surface = " \xE2\x81\x87 ";  // TrainerSpec.unk_surface (sentencepiece_model.proto:184), used by Decode
  // ---- derived
  int unk_id = -1;
  float min_score = 0, max_score = 0;
  HostTrie pieces, user;
  std::vector<uint32_t> cm_first2;
  std::vector<uint32_t> cm_cont2;   // 64 Kbit: characters that occur inside (not at the start of) a rule key / user symbol
  uint32_t user_first[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const uint32_t *cm_units = nullptr;  // into charsmap
  uint32_t cm_n = 0;
  const uint8_t *cm_pool = nullptr;
  uint32_t cm_pool_size = 0;
  int max_piece_bytes = 0;
  int ring = 8;
  float smax = 0;
  int expansion = 3;  // upper bound of normalized bytes per input byte
  bool has_unused = false;
  // ---- fast view of the piece trie for the unigram fast tier (spm_fast.cuh)
  std::vector<uint32_t> funits;   // pieces.units with has_leaf only where PopulateNodes inserts a node
  std::vector<float> fscore;      // node score (USER_DEFINED bonus pre-computed)
  std::vector<uint32_t> fleaf;    // node -> piece id
  int max_piece_chars = 0;
  bool fast_ok = false;           // the model fits the fast tier's packed records
  bool text_valid = true;         // normalized text is always valid UTF-8
  bool bpe_fast_ok = false;       // BPE model whose merges never cross a word boundary (fast_bpe)
  bool ascii_plain = false;       // see build_host_model
  bool ws_ctl = false;            // see build_host_model
  std::vector<uint32_t> cm_first3;   // 65536-bit filter: does a rule key start with this 3-byte character (U+0800..U+FFFF)?
  bool memo_ok = false;           // no lattice node / merge crosses a U+2581: words can be memoised (spm_memo.cuh)
  bool neg_only = false;          // every node score of the fast view (and the UNK score) is <= 0
  // ---- piece strings by id (pieces path)
  std::vector<int32_t> plen;
  std::vector<uint32_t> poff;
  std::string ppool;
  std::string error;
};

// Parses + builds everything; returns a StatusCode and fills m->error.
int build_host_model(const void *data, size_t len, HostModel *m);

// Re-derives the leaf tables after piece types changed (SetVocabulary,
// src/sentencepiece_processor.cc:203-231; min/max scores stay as loaded).
void refresh_leaf_types(HostModel *m);

// Piece trie of a TrainerModel for the E-step (src/unigram_model_trainer.cc:97-119: every
// piece NORMAL, ids = positions, trie over all of them).  leaf = (id, score bits).
// *min_score = min over the scores.  Returns a StatusCode and fills *err.
// hot_ids (may be null): slot -> piece id of the hot pieces packed into the leaf words (slot 0: -1).
// estep_softplus()'s table (spm_estep.cuh): kLseN x kLseCoef doubles.
void build_lse_table(std::vector<double> *tab);
int build_estep_trie(const std::vector<std::string> &pieces, const std::vector<float> &scores, HostTrie *out,
                     float *min_score, std::string *err, std::vector<int32_t> *hot_ids = nullptr);

// id of a piece by string as ModelInterface::PieceToId (model_interface.cc:87-97)
int piece_to_id(const HostModel &m, const std::string &piece);

}  // namespace spmb
