/* spm_cuda.h -- C ABI of the B200 batch encoder for SentencePiece models.
 *
 * This is the drop-in boundary for the reference's encode path.  The reference
 * (SentencePiece 0.1.82) has no C ABI or batch entry point of its own; each
 * function below names the C++ interface it stands in for, so that a
 * `sentencepiece::SentencePieceProcessor`-shaped facade (see
 * include/spm_cuda_processor.h and INTEGRATION.md) can be built on top with
 * unchanged signatures.
 *
 * Error convention: every int-returning function returns a
 * sentencepiece::util::error::Code value (src/sentencepiece_processor.h:101-119;
 * 0 = OK).  spm_cuda_last_error() returns the thread-local message of the
 * last failure.  Nothing aborts the process.
 *
 * There is no CPU fallback: all calls fail with UNAVAILABLE-style INTERNAL
 * errors when no CUDA device is usable.
 *
 * Threading: a handle owns ONE device workspace, so calls on the same handle are
 * serialised by a mutex inside the library (they are safe from any thread, they
 * do not overlap).  Concurrency = one handle per stream / thread: handles are
 * independent (tools/config5.py streams two blocks at a time through two handles
 * on one GPU).  Every call runs on the handle's device and restores the caller's
 * current CUDA device before it returns.
 */
#ifndef SPM_CUDA_H_
#define SPM_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct spm_cuda_model spm_cuda_model;

/* memory_space values */
#define SPM_CUDA_HOST 0   /* text, offsets, ids, id_offsets are host pointers   */
#define SPM_CUDA_DEVICE 1 /* ... are device pointers on the model's device      */

/* Extra options (SetEncodeExtraOptions "bos:eos:reverse",
 * src/sentencepiece_processor.cc:945-1023) are a list applied in order; they
 * are packed two bits per entry, first entry in the low bits, 0 terminates. */
#define SPM_CUDA_OPT_BOS 1u
#define SPM_CUDA_OPT_EOS 2u
#define SPM_CUDA_OPT_REVERSE 3u

/* Replaces SentencePieceProcessor::LoadFromSerializedProto
 * (src/sentencepiece_processor.h:203, .cc:103-157): parses an unmodified
 * serialized ModelProto, builds the piece trie / charsmap tables and uploads
 * them to `device`.  UNIGRAM and BPE models only (WORD/CHAR: UNIMPLEMENTED). */
int spm_cuda_model_load(const void *serialized_model, size_t len, int device, spm_cuda_model **out);

void spm_cuda_model_free(spm_cuda_model *model);

/* Replaces the effect of SetVocabulary / ResetVocabulary
 * (src/sentencepiece_processor.h:224-228, .cc:203-231): overwrite the piece
 * types (ModelProto::SentencePiece::Type values, one byte per piece).  Must
 * not run concurrently with encode calls on the same model. */
int spm_cuda_model_set_piece_types(spm_cuda_model *model, const uint8_t *types, int n);

/* Per-handle knobs.  The reference has no counterpart (its processor has no tiers); they exist
 * so that tests and measurements can reach, on small batches, the code paths a real batch only
 * takes at scale, and switch single tiers off.  Results never depend on them.
 *   "memo"              1   word-memo tier on / off
 *   "memo_bits"         23  log2 of the most memo-table slots one call may use (12..26)
 *   "host_chunks"       0   SPM_CUDA_HOST calls: number of pipelined chunks (0: 12 unigram / 5 BPE)
 *   "host_chunk_min"    2^20  ... for calls of at least this many sentences
 *   "max_chunk"         0   most sentences per chunk (0: 2^24 with a fast tier, 2^20 without)
 *   "warp"              1   warp-per-line kernel for long / near-tied lines on / off
 *   "warp_route"        262144  leftover lists up to this size go to the warp kernel as a whole (0: never)
 *   "scan_per_sentence" 0   thread-per-sentence word scan instead of the byte-parallel one
 *   "emit_blocks"       0   CTAs of the memo emit kernel per SM (0: occupancy query)
 *   "pieces_exact"      0   encode_batch_pieces: every sentence through the exact tier
 *   "pieces_plain"      1   encode_batch_pieces: plain sentences (printable ASCII, single spaces) skip the normalizer
 *   "bpe_route"         8192  BPE: leftover lists up to this size go to the general tier as a whole (0: never)
 *   "fast_side"         1   thread-per-sentence Viterbi kernel keeps near-tied candidates itself (0: hands the sentence on)
 *   "mid_lines"         0   lines of 4..8 KiB take the thread-per-sentence kernel instead of the warp kernel
 * NOT_FOUND for an unknown name.  Must not run concurrently with calls on the same model. */
int spm_cuda_model_set_option(spm_cuda_model *model, const char *name, int64_t value);
int spm_cuda_model_get_option(const spm_cuda_model *model, const char *name, int64_t *value);

/* Frees the device workspace kept from earlier calls (it grows to the largest batch seen and is
 * otherwise reused); model tables stay loaded.  No counterpart in the reference (host heap). */
int spm_cuda_model_release_workspace(spm_cuda_model *model);

/* Replaces a loop of SentencePieceProcessor::Encode(input, std::vector<int>*)
 * (src/sentencepiece_processor.h:244, .cc:319-330) over n sentences.
 *   text, text_offsets[n+1] : sentence i is text[text_offsets[i], text_offsets[i+1])
 *   extra_options           : packed option list (see above)
 *   ids[ids_capacity]       : output ids, CSR values
 *   id_offsets[n+1]         : output CSR offsets; id_offsets[n] = total ids
 * Returns RESOURCE_EXHAUSTED (8) when ids_capacity is too small (id_offsets is
 * still complete, so id_offsets[n] is the capacity to retry with), INTERNAL
 * (13) if some sentence hit the reference's "all normalized characters are not
 * consumed" error (that sentence gets zero ids, as the reference returns).
 * cuda_stream is a cudaStream_t (NULL = default stream); the call returns
 * after the stream work for it has completed. */
int spm_cuda_encode_batch(const spm_cuda_model *model, const char *text, const int64_t *text_offsets,
                          int64_t n, uint32_t extra_options, int32_t *ids, int64_t ids_capacity,
                          int64_t *id_offsets, int memory_space, void *cuda_stream);

/* Model introspection (SentencePieceProcessor::GetPieceSize / IdToPiece /
 * GetScore / PieceToId / unk_id ..., src/sentencepiece_processor.h:397-432). */
int spm_cuda_model_piece_size(const spm_cuda_model *model);
const char *spm_cuda_model_piece(const spm_cuda_model *model, int id, int *len);
float spm_cuda_model_score(const spm_cuda_model *model, int id);
int spm_cuda_model_piece_type(const spm_cuda_model *model, int id);
int spm_cuda_model_piece_to_id(const spm_cuda_model *model, const char *piece, size_t len);
int spm_cuda_model_type(const spm_cuda_model *model);  /* TrainerSpec::ModelType */
int spm_cuda_model_unk_id(const spm_cuda_model *model);
int spm_cuda_model_bos_id(const spm_cuda_model *model);
int spm_cuda_model_eos_id(const spm_cuda_model *model);
int spm_cuda_model_pad_id(const spm_cuda_model *model);

/* Validation half of SetEncodeExtraOptions / ParseExtraOptions
 * (src/sentencepiece_processor.cc:991-1023): INTERNAL when a requested bos/eos
 * piece resolves to the unknown id. */
int spm_cuda_model_check_extra_options(const spm_cuda_model *model, uint32_t extra_options);

/* Statistics of the last encode call on this model (for bench.py / profiling):
 * stats[0] kernel launches, [1] sentences sent to the exact (general) tier,
 * [2] device-side milliseconds between first and last launch (CUDA events on
 * the caller's stream), [3] bytes of workspace currently allocated, [4] device
 * milliseconds spent in the dominant stage, summed over its launches (CUDA
 * events around each stage, same stream), [5] number of those event pairs,
 * [6] sentences, [7] ids produced, [8] which stage is the dominant one and
 * [9..14] milliseconds of every stage: 0 Viterbi / merge kernels (or the general
 * tier for a model without a fast tier), 1 normalizer, 2 word-memo collect,
 * 3 word-memo words, 4 word-memo emit, 5 compaction; [15] sentences the word memo
 * left to the Viterbi / merge kernels.  n <= 16. */
int spm_cuda_last_stats(const spm_cuda_model *model, double *stats, int n);

const char *spm_cuda_last_error(void);

/* ---- Line splitting (the step before the path) ----------------
 * A raw file buffer -> the CSR input of spm_cuda_encode_batch, with the lines
 * std::getline hands to spm_encode (src/filesystem.cc:42-44,
 * src/spm_encode_main.cc:183-193): a line ends at '\n' (dropped; a preceding '\r'
 * stays), a last line without '\n' counts when it is not empty.
 * text (>= len bytes) receives the bytes without the newlines, offsets[0..n] the CSR
 * offsets into it, *n_lines = n.  RESOURCE_EXHAUSTED when n > capacity (*n_lines is set).
 * The model handle only supplies the device and its workspace. */
int spm_cuda_split_lines(const spm_cuda_model *model, const char *buf, int64_t len, char *text, int64_t *offsets,
                         int64_t capacity, int64_t *n_lines, int memory_space, void *cuda_stream);

/* The same line splitting for callers that cannot allocate device memory (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE
 * REFERENCE.  Device pointers only (the
 * SPM_CUDA_DEVICE form of spm_cuda_split_lines, same outputs, same errors); the scratch of the
 * two kernels -- per-tile newline counts and their scan -- lives in a file-scope __device__ array
 * of 786,448 bytes that every GPU loading the library holds until exit, so the call makes no
 * cudaMalloc / cudaFree / allocating library call.  Limit: len <= 268,435,456 bytes (256 MiB);
 * a longer buffer is refused with OUT_OF_RANGE and a message naming the limit.  All calls on one
 * GPU share the array: calls that overlap in time on the same GPU (different streams or threads)
 * race; serialise them.  model may be NULL (the current device is used). */
int spm_cuda_split_lines_noalloc(const spm_cuda_model *model, const char *buf, int64_t len, char *text,
                                 int64_t *offsets, int64_t capacity, int64_t *n_lines, void *cuda_stream);

/* ---- Pieces / SentencePieceText (the step after the path) ------
 * Replaces a loop of SentencePieceProcessor::Encode(input, SentencePieceText*)
 * (src/sentencepiece_processor.h:258, .cc:560-573 -> PopulateSentencePieceText :488-551 ->
 * ApplyExtraOptions :945-979), i.e. what EncodeAsPieces / EncodeAsSerializedProto return.
 * Per final piece k of sentence i (k in [id_offsets[i], id_offsets[i+1])):
 *   ids[k]                       SentencePiece.id
 *   begin[k], end[k]             SentencePiece.begin / .end: byte span of the surface in the INPUT
 *                                sentence (norm_to_orig at both ends, src/normalizer.cc:132-206)
 *   nbegin[k], nend[k]           byte span of the piece in the NORMALIZED sentence; the piece string
 *                                is norm_text[norm_offsets[i] + nbegin[k] .. + nend[k]) -- for an
 *                                unknown piece that is the merged run of unknown characters
 *                                (.cc:525-529), for every other piece it equals IdToPiece(id).
 *   A bos / eos piece added by the extra options has nbegin = nend = -1 and begin = end = 0
 *   (the reference leaves surface / begin / end unset); its string is IdToPiece(id).
 * norm_text / norm_offsets[n+1]: the normalized sentences as CSR (Normalizer::Normalize).
 * The five per-piece arrays share ids_capacity.  RESOURCE_EXHAUSTED when ids_capacity or
 * norm_capacity is too small (id_offsets[n] / norm_offsets[n] hold the sizes to retry with);
 * other errors as spm_cuda_encode_batch. */
int spm_cuda_encode_batch_pieces(const spm_cuda_model *model, const char *text, const int64_t *text_offsets,
                                 int64_t n, uint32_t extra_options, int32_t *ids, int32_t *begin, int32_t *end,
                                 int32_t *nbegin, int32_t *nend, int64_t ids_capacity, int64_t *id_offsets,
                                 char *norm_text, int64_t norm_capacity, int64_t *norm_offsets, int memory_space,
                                 void *cuda_stream);

/* ---- Batch Decode ----------------------------------------------
 * Replaces a loop of SentencePieceProcessor::Decode(const std::vector<int>&, std::string*)
 * (src/sentencepiece_processor.h:252, .cc:670-733): CONTROL pieces decode to "", the unknown
 * id to TrainerSpec.unk_surface, every other piece to its string with U+2581 -> ' ', after
 * dropping one leading U+2581 while the text is still empty.  CSR ids in, CSR text out
 * (text_offsets[n+1]).  OUT_OF_RANGE for an id outside [0, GetPieceSize());
 * RESOURCE_EXHAUSTED when text_capacity is too small (text_offsets[n] = bytes needed). */
int spm_cuda_decode_batch(const spm_cuda_model *model, const int32_t *ids, const int64_t *id_offsets, int64_t n,
                          char *text, int64_t text_capacity, int64_t *text_offsets, int memory_space,
                          void *cuda_stream);

/* ---- Dense / sparse layouts of a CSR result ---------------------
 * The output shapes of the reference's only batch API, the TF ops
 * (tensorflow/sentencepiece_processor_ops.cc:330-428), from the DEVICE-resident CSR that
 * spm_cuda_encode_batch(SPM_CUDA_DEVICE) leaves behind (all pointers are device pointers,
 * except max_len):
 *   csr_max_len   *max_len (host) = longest row              (dense_shape[1], :339-343)
 *   csr_to_dense  values[n, max_len] padded with `pad`, seq_len[n]      (:391-428)
 *   csr_to_sparse indices[nnz, 2] = (row, column) of every id; values = ids (:345-378) */
int spm_cuda_csr_max_len(const spm_cuda_model *model, const int64_t *id_offsets, int64_t n, int64_t *max_len,
                         void *cuda_stream);
int spm_cuda_csr_to_dense(const spm_cuda_model *model, const int32_t *ids, const int64_t *id_offsets, int64_t n,
                          int32_t pad, int32_t *values, int64_t max_len, int32_t *seq_len, void *cuda_stream);
int spm_cuda_csr_to_sparse(const spm_cuda_model *model, const int64_t *id_offsets, int64_t n, int64_t *indices,
                           void *cuda_stream);

/* ---- Unigram-trainer E-step (optional kernel) ----------------
 * Replaces the per-thread loop of unigram::Trainer::RunEStep
 * (src/unigram_model_trainer.cc:237-287): for every (word, freq) build the lattice of a
 * TrainerModel over the given pieces (all NORMAL, unk id 0, UNK score = min score - 10;
 * src/unigram_model_trainer.cc:97-119), Lattice::PopulateMarginal
 * (src/unigram_model.cc:272-328) and the Viterbi token count.
 *
 * pieces: V NUL-terminated strings back to back; scores[V].
 * words:  UTF-8 bytes of the n words, word i = [word_offsets[i], word_offsets[i+1]);
 *         freq[i] > 0.  all_freq = the sum of freq over ALL shards (divisor of obj).
 * expected[V] (double) += freq * marginal of every node, by piece id;
 * obj_ntokens[0] -= sum freq * logZ / all_freq;  obj_ntokens[1] += Viterbi tokens.
 * Both are ACCUMULATED into (the caller zeroes them; a sharded run sums them over the
 * ranks with one all-reduce).  logz (optional, may be NULL): per-word log-likelihood.
 * Results agree with the reference to 1e-5 relative (double accumulation here, float
 * in the reference). */
typedef struct spm_cuda_estep spm_cuda_estep;
int spm_cuda_estep_create(const char *pieces, const float *scores, int V, int device, spm_cuda_estep **out);
void spm_cuda_estep_free(spm_cuda_estep *estep);
int spm_cuda_estep_vocab_size(const spm_cuda_estep *estep);
int spm_cuda_estep_run(const spm_cuda_estep *estep, const char *words, const int64_t *word_offsets,
                       const int64_t *freq, int64_t n, double all_freq, double *expected, double *obj_ntokens,
                       float *logz, int memory_space, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* SPM_CUDA_H_ */
