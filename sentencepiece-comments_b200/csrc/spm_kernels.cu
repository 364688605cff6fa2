// spm_kernels.cu -- CUDA kernels (sm_100a) and the C ABI of include/spm_cuda.h.
//
// Pipeline of one spm_cuda_encode_batch call (sentences are independent units;
// nothing here is GEMM-shaped, the work is byte/trie/scan traffic):
//   k_normalize   one thread per sentence: charsmap leftmost-longest rewrite,
//                 whitespace policy                      (src/normalizer.cc:88-300)
//   k_scan_*      exclusive scans (scratch offsets, CSR offsets)
//   k_encode      one thread per sentence: trie walk + on-the-fly Viterbi with
//                 pruned candidate lists, or BPE agenda (src/unigram_model.cc:
//                 535-604,222-261; src/bpe_model.cc:37-199), unk-run merge
//   k_compact     one warp per sentence: extra options + CSR write
//                                          (src/sentencepiece_processor.cc:319-330,945-979)
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/spm_cuda.h"
#include "spm_core.cuh"
#include "spm_model.h"

using namespace spmb;

namespace {

thread_local std::string g_err;

#define CUDA_TRY(expr)                                                                    \
  do {                                                                                    \
    cudaError_t e_ = (expr);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      g_err = std::string(#expr) + ": " + cudaGetErrorString(e_);                         \
      return kInternal;                                                                   \
    }                                                                                     \
  } while (0)

constexpr int kCapc = 8;          // near-tie candidates kept per position by the exact tier
constexpr int kThreads = 128;

// ---------------------------------------------------------------- kernels
__device__ __forceinline__ int64_t norm_base(const int64_t *off, int64_t c0, int64_t i, int exp_) {
  return (int64_t)exp_ * (off[i] - off[c0]) + 8 * (i - c0);
}

__global__ void __launch_bounds__(kThreads)
k_normalize(ModelView mv, const uint8_t *text, const int64_t *off, int64_t c0, int64_t cn, int exp_,
            uint8_t *norm, int32_t *nlen, int32_t *words, uint32_t *status) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= cn) return;
  const int64_t i = c0 + t;
  const int64_t len = off[i + 1] - off[i];
  const int64_t cap = (int64_t)exp_ * len + 8;
  int32_t nl = normalize_sentence(mv, text + off[i], len, norm + norm_base(off, c0, i, exp_), cap);
  uint32_t st = 0;
  if (nl < 0) { nl = 0; st = ST_NORM_OVERFLOW; }
  nlen[t] = nl;
  status[t] = st;
  int64_t w = mv.model_type == M_UNIGRAM ? uni_scratch_words(mv.ring, kCapc, nl) : bpe_scratch_words(nl);
  words[t] = (int32_t)((w + 1) & ~(int64_t)1);  // keep 8-byte alignment of each region
}

__global__ void __launch_bounds__(kThreads)
k_encode(ModelView mv, const int64_t *off, int64_t c0, int64_t cn, int exp_, const uint8_t *norm,
         const int32_t *nlen, const int64_t *woff, int32_t *scratch, int32_t *stage, int32_t *cnt,
         uint32_t *status, uint32_t *status_or) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= cn) return;
  const int64_t i = c0 + t;
  const int64_t nb = norm_base(off, c0, i, exp_);
  const int32_t nl = nlen[t];
  uint32_t st = status[t];
  int32_t c = 0;
  int32_t *w = scratch + woff[t];
  if (st == 0) {
    if (mv.model_type == M_UNIGRAM) {
      UniScratch sc;
      sc.ringM = (float *)w; w += mv.ring;
      sc.ringCnt = (uint32_t *)w; w += mv.ring;
      sc.candBt = (float *)w; w += mv.ring * kCapc;
      sc.candRec = (uint64_t *)w; w += 2 * mv.ring * kCapc;
      sc.back = (uint64_t *)w;
      sc.stride = nl + 1;
      c = unigram_encode<kCapc>(mv, norm + nb, nl, sc, stage + nb, nl, &st);
    } else {
      c = bpe_encode(mv, norm + nb, nl, w, stage + nb, nl, &st);
    }
  }
  cnt[t] = c;
  status[t] = st;
  if (st) atomicOr(status_or, st);
}

// ---- exclusive scan of f(in[i]) over one chunk: int32 -> int64, with carry
struct IdentityOp { __device__ int64_t operator()(int32_t v) const { return v; } };
struct FinalCountOp {
  int32_t extra;
  __device__ int64_t operator()(int32_t v) const { return v < 0 ? 0 : v + extra; }
};

constexpr int kScanThreads = 256, kScanItems = 4, kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ int64_t block_exclusive_scan(int64_t v, int64_t *total) {
  __shared__ int64_t warp_sums[kScanThreads / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int64_t inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int64_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
    if (lane >= d) inc += y;
  }
  __syncthreads();
  if (lane == 31) warp_sums[wid] = inc;
  __syncthreads();
  int64_t base = 0, tot = 0;
#pragma unroll
  for (int k = 0; k < kScanThreads / 32; ++k) {
    const int64_t s = warp_sums[k];
    if (k < wid) base += s;
    tot += s;
  }
  *total = tot;
  return base + inc - v;
}

template <class Op>
__global__ void __launch_bounds__(kScanThreads) k_scan_partial(const int32_t *in, int64_t n, Op op, int64_t *partial) {
  const int64_t b0 = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int64_t s = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k)
    if (b0 + k < n) s += op(in[b0 + k]);
  int64_t tot;
  block_exclusive_scan(s, &tot);
  if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

// single block: partial[] -> exclusive offsets (+ carry in); carry out
__global__ void __launch_bounds__(kScanThreads) k_scan_top(int64_t *partial, int64_t nb, int64_t *carry, int use_carry) {
  __shared__ int64_t running;
  if (threadIdx.x == 0) running = use_carry ? *carry : 0;
  __syncthreads();
  for (int64_t b = 0; b < nb; b += kScanThreads) {
    const int64_t i = b + threadIdx.x;
    const int64_t v = i < nb ? partial[i] : 0;
    int64_t tot;
    const int64_t ex = block_exclusive_scan(v, &tot);
    const int64_t base = running;
    if (i < nb) partial[i] = base + ex;
    __syncthreads();
    if (threadIdx.x == 0) running = base + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) *carry = running;
}

template <class Op>
__global__ void __launch_bounds__(kScanThreads)
k_scan_final(const int32_t *in, int64_t n, Op op, const int64_t *partial, const int64_t *carry, int64_t *out) {
  const int64_t b0 = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int64_t v[kScanItems], s = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    v[k] = b0 + k < n ? op(in[b0 + k]) : 0;
    s += v[k];
  }
  int64_t tot;
  int64_t ex = block_exclusive_scan(s, &tot) + partial[blockIdx.x];
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    if (b0 + k < n) out[b0 + k] = ex;
    ex += v[k];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = *carry;
}

// one warp per sentence: final layout (extra options) into the CSR values
__global__ void __launch_bounds__(256)
k_compact(ExtraPlan plan, const int64_t *off, int64_t c0, int64_t cn, int exp_, const int32_t *stage,
          const int32_t *cnt, const int64_t *id_off, int32_t *ids, int64_t capacity) {
  const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (t >= cn) return;
  const int64_t i = c0 + t;
  const int32_t core = cnt[t];
  const int32_t fc = final_count(plan, core);
  const int32_t *src = stage + norm_base(off, c0, i, exp_);
  const int64_t dst = id_off[i];
  for (int32_t j = lane; j < fc; j += 32)
    if (dst + j < capacity) ids[dst + j] = final_id_at(plan, src, core, j);
}

// ---------------------------------------------------------------- host side
struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return kOk;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    CUDA_TRY(cudaMalloc(&p, want));
    cap = want;
    return kOk;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

}  // namespace

struct spm_cuda_model {
  HostModel hm;
  int device = 0;
  ModelView mv;
  DevBuf d_units, d_leaf, d_uunits, d_uleaf, d_cm_units, d_cm_pool, d_first2;
  // workspace of encode calls (guarded by mu)
  mutable std::mutex mu;
  mutable DevBuf w_text, w_off, w_ids, w_idoff, w_norm, w_stage, w_nlen, w_words, w_cnt, w_status,
      w_woff, w_scratch, w_partial, w_scalars;
  mutable std::vector<int64_t> h_off;
  mutable double stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  mutable cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  mutable std::vector<cudaEvent_t> kev;  // event pairs around the dominant (encode) kernel launches
  int bos_id = 0, eos_id = 0, pad_id = 0;
};

namespace {

int upload(DevBuf *b, const void *src, size_t bytes) {
  if (bytes == 0) return kOk;
  int rc = b->ensure(bytes);
  if (rc) return rc;
  CUDA_TRY(cudaMemcpy(b->p, src, bytes, cudaMemcpyHostToDevice));
  return kOk;
}

int upload_model(spm_cuda_model *M) {
  HostModel &hm = M->hm;
  int rc;
  if ((rc = upload(&M->d_units, hm.pieces.units.data(), hm.pieces.units.size() * 4))) return rc;
  if ((rc = upload(&M->d_leaf, hm.pieces.leaf.data(), hm.pieces.leaf.size() * 8))) return rc;
  if ((rc = upload(&M->d_uunits, hm.user.units.data(), hm.user.units.size() * 4))) return rc;
  if ((rc = upload(&M->d_uleaf, hm.user.leaf.data(), hm.user.leaf.size() * 8))) return rc;
  if ((rc = upload(&M->d_cm_units, hm.cm_units, (size_t)hm.cm_n * 4))) return rc;
  if ((rc = upload(&M->d_cm_pool, hm.cm_pool, hm.cm_pool_size))) return rc;
  if ((rc = upload(&M->d_first2, hm.cm_first2.data(), hm.cm_first2.size() * 4))) return rc;
  ModelView &v = M->mv;
  memset(&v, 0, sizeof(v));
  v.pieces.units = (const uint32_t *)M->d_units.p; v.pieces.leaf = (const uint2 *)M->d_leaf.p;
  v.pieces.n = (uint32_t)hm.pieces.units.size();
  v.user.units = (const uint32_t *)M->d_uunits.p; v.user.leaf = (const uint2 *)M->d_uleaf.p;
  v.user.n = (uint32_t)hm.user.units.size();
  v.cm_units = (const uint32_t *)M->d_cm_units.p; v.cm_pool = (const uint8_t *)M->d_cm_pool.p;
  v.cm_first2 = hm.cm_first2.empty() ? nullptr : (const uint32_t *)M->d_first2.p;
  v.cm_n = hm.cm_n;
  memcpy(v.user_first, hm.user_first, sizeof(v.user_first));
  v.model_type = hm.model_type; v.unk_id = hm.unk_id;
  v.unk_score = hm.min_score - 10.0f;  // kUnkPenalty, src/unigram_model.cc:550-551
  v.max_score = hm.max_score; v.smax = hm.smax; v.ring = hm.ring;
  v.add_dummy_prefix = hm.add_dummy_prefix; v.remove_extra_ws = hm.remove_extra_ws;
  v.escape_ws = hm.escape_ws; v.ws_suffix = hm.ws_suffix; v.has_unused = hm.has_unused;
  return kOk;
}

template <class Op>
int run_scan(const spm_cuda_model *M, cudaStream_t st, const int32_t *in, int64_t n, Op op, int64_t *out,
             int64_t *carry, int use_carry, double *launches) {
  const int64_t nb = (n + kScanTile - 1) / kScanTile;
  int rc = M->w_partial.ensure((size_t)(nb + 1) * 8);
  if (rc) return rc;
  int64_t *partial = (int64_t *)M->w_partial.p;
  k_scan_partial<Op><<<(unsigned)nb, kScanThreads, 0, st>>>(in, n, op, partial);
  k_scan_top<<<1, kScanThreads, 0, st>>>(partial, nb, carry, use_carry);
  k_scan_final<Op><<<(unsigned)nb, kScanThreads, 0, st>>>(in, n, op, partial, carry, out);
  *launches += 3;
  CUDA_TRY(cudaGetLastError());
  return kOk;
}

}  // namespace

// ---------------------------------------------------------------- C ABI
extern "C" {

const char *spm_cuda_last_error(void) { return g_err.c_str(); }

int spm_cuda_model_load(const void *serialized_model, size_t len, int device, spm_cuda_model **out) {
  if (out == nullptr) { g_err = "output container is null"; return kInternal; }
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    g_err = "no CUDA device available (this library has no CPU path)";
    return kInternal;
  }
  if (device < 0 || device >= ndev) { g_err = "invalid CUDA device ordinal"; return kInvalidArgument; }
  spm_cuda_model *M = new spm_cuda_model;
  int rc = build_host_model(serialized_model, len, &M->hm);
  if (rc) { g_err = M->hm.error; delete M; return rc; }
  M->device = device;
  M->bos_id = piece_to_id(M->hm, M->hm.bos_piece);
  M->eos_id = piece_to_id(M->hm, M->hm.eos_piece);
  M->pad_id = piece_to_id(M->hm, M->hm.pad_piece);
  if (cudaSetDevice(device) != cudaSuccess) { g_err = "cudaSetDevice failed"; delete M; return kInternal; }
  rc = upload_model(M);
  if (rc) { spm_cuda_model_free(M); return rc; }
  cudaEventCreate(&M->ev0);
  cudaEventCreate(&M->ev1);
  *out = M;
  return kOk;
}

void spm_cuda_model_free(spm_cuda_model *M) {
  if (!M) return;
  cudaSetDevice(M->device);
  DevBuf *bufs[] = {&M->d_units, &M->d_leaf, &M->d_uunits, &M->d_uleaf, &M->d_cm_units, &M->d_cm_pool,
                    &M->d_first2, &M->w_text, &M->w_off, &M->w_ids, &M->w_idoff, &M->w_norm, &M->w_stage,
                    &M->w_nlen, &M->w_words, &M->w_cnt, &M->w_status, &M->w_woff, &M->w_scratch,
                    &M->w_partial, &M->w_scalars};
  for (DevBuf *b : bufs) b->release();
  if (M->ev0) cudaEventDestroy(M->ev0);
  if (M->ev1) cudaEventDestroy(M->ev1);
  for (cudaEvent_t e : M->kev) cudaEventDestroy(e);
  delete M;
}

int spm_cuda_model_set_piece_types(spm_cuda_model *M, const uint8_t *types, int n) {
  if (!M || !types) { g_err = "null argument"; return kInvalidArgument; }
  if (n != (int)M->hm.type.size()) { g_err = "types length must equal the piece count"; return kInvalidArgument; }
  std::lock_guard<std::mutex> lk(M->mu);
  for (int i = 0; i < n; ++i) {
    // SetVocabulary only flips NORMAL <-> UNUSED (src/sentencepiece_processor.cc:211-227)
    const uint8_t a = M->hm.type[i], b = types[i];
    const bool ok = a == b || ((a == T_NORMAL || a == T_UNUSED) && (b == T_NORMAL || b == T_UNUSED));
    if (!ok) { g_err = "only NORMAL <-> UNUSED transitions are supported"; return kInvalidArgument; }
  }
  for (int i = 0; i < n; ++i) M->hm.type[i] = types[i];
  refresh_leaf_types(&M->hm);
  CUDA_TRY(cudaSetDevice(M->device));
  CUDA_TRY(cudaMemcpy(M->d_leaf.p, M->hm.pieces.leaf.data(), M->hm.pieces.leaf.size() * 8, cudaMemcpyHostToDevice));
  M->mv.has_unused = M->hm.has_unused;
  return kOk;
}

int spm_cuda_model_piece_size(const spm_cuda_model *M) { return (int)M->hm.piece.size(); }
const char *spm_cuda_model_piece(const spm_cuda_model *M, int id, int *len) {
  if (id < 0 || id >= (int)M->hm.piece.size()) { if (len) *len = 0; return ""; }
  if (len) *len = (int)M->hm.piece[id].size();
  return M->hm.piece[id].c_str();
}
float spm_cuda_model_score(const spm_cuda_model *M, int id) {
  return (id < 0 || id >= (int)M->hm.score.size()) ? 0.0f : M->hm.score[id];
}
int spm_cuda_model_piece_type(const spm_cuda_model *M, int id) {
  return (id < 0 || id >= (int)M->hm.type.size()) ? 0 : M->hm.type[id];
}
int spm_cuda_model_piece_to_id(const spm_cuda_model *M, const char *piece, size_t len) {
  return piece_to_id(M->hm, std::string(piece, len));
}
int spm_cuda_model_type(const spm_cuda_model *M) { return M->hm.model_type; }
int spm_cuda_model_unk_id(const spm_cuda_model *M) { return M->hm.unk_id; }
// bos_id()/eos_id()/pad_id(): the id only when it is a CONTROL piece, else -1
// (src/sentencepiece_processor.cc:917-939)
static int control_or_minus1(const spm_cuda_model *M, int id) {
  return (id >= 0 && id < (int)M->hm.type.size() && M->hm.type[id] == T_CONTROL) ? id : -1;
}
int spm_cuda_model_bos_id(const spm_cuda_model *M) { return control_or_minus1(M, M->bos_id); }
int spm_cuda_model_eos_id(const spm_cuda_model *M) { return control_or_minus1(M, M->eos_id); }
int spm_cuda_model_pad_id(const spm_cuda_model *M) { return control_or_minus1(M, M->pad_id); }

// ParseExtraOptions validation (src/sentencepiece_processor.cc:991-1023)
int spm_cuda_model_check_extra_options(const spm_cuda_model *M, uint32_t opts) {
  if (!M) { g_err = "model is null"; return kInvalidArgument; }
  for (uint32_t o = opts; o & 3u; o >>= 2) {
    const uint32_t c = o & 3u;
    if (c == SPM_CUDA_OPT_BOS && M->hm.type[M->bos_id] == T_UNKNOWN) {
      g_err = "id for `" + M->hm.bos_piece + "` is not defined.";
      return kInternal;
    }
    if (c == SPM_CUDA_OPT_EOS && M->hm.type[M->eos_id] == T_UNKNOWN) {
      g_err = "id for `" + M->hm.eos_piece + "` is not defined.";
      return kInternal;
    }
  }
  return kOk;
}

int spm_cuda_last_stats(const spm_cuda_model *M, double *stats, int n) {
  if (!M || !stats) return kInvalidArgument;
  std::lock_guard<std::mutex> lk(M->mu);
  for (int i = 0; i < n && i < 8; ++i) stats[i] = M->stats[i];
  return kOk;
}

int spm_cuda_encode_batch(const spm_cuda_model *M, const char *text, const int64_t *text_offsets, int64_t n,
                          uint32_t extra_options, int32_t *ids, int64_t ids_capacity, int64_t *id_offsets,
                          int memory_space, void *cuda_stream) {
  if (!M) { g_err = "model is null"; return kInvalidArgument; }
  if (n < 0 || !text_offsets || !id_offsets || (ids_capacity > 0 && !ids)) {
    g_err = "output container is null";  // CHECK_OR_RETURN_STATUS_STL, src/sentencepiece_processor.cc:276-279
    return kInternal;
  }
  if (memory_space != SPM_CUDA_HOST && memory_space != SPM_CUDA_DEVICE) { g_err = "bad memory_space"; return kInvalidArgument; }
  std::lock_guard<std::mutex> lk(M->mu);
  CUDA_TRY(cudaSetDevice(M->device));
  cudaStream_t st = (cudaStream_t)cuda_stream;
  const bool host = memory_space == SPM_CUDA_HOST;
  double launches = 0;
  int rc;

  // offsets on the host (chunk planning) and on the device
  M->h_off.resize(n + 1);
  const int64_t *d_off;
  const uint8_t *d_text;
  if (host) {
    memcpy(M->h_off.data(), text_offsets, (n + 1) * 8);
  } else {
    CUDA_TRY(cudaMemcpyAsync(M->h_off.data(), text_offsets, (n + 1) * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
  }
  const int64_t *ho = M->h_off.data();
  for (int64_t i = 0; i < n; ++i)
    if (ho[i + 1] < ho[i]) { g_err = "text_offsets must be non-decreasing"; return kInvalidArgument; }
  const int64_t text_lo = ho[0], text_hi = ho[n];
  if (host) {
    if ((rc = M->w_off.ensure((n + 1) * 8))) return rc;
    if ((rc = M->w_text.ensure((size_t)(text_hi - text_lo) + 16))) return rc;
    if ((rc = M->w_idoff.ensure((n + 1) * 8))) return rc;
    CUDA_TRY(cudaMemcpyAsync(M->w_off.p, text_offsets, (n + 1) * 8, cudaMemcpyHostToDevice, st));
    if (text_hi > text_lo)
      CUDA_TRY(cudaMemcpyAsync(M->w_text.p, text + text_lo, text_hi - text_lo, cudaMemcpyHostToDevice, st));
    d_off = (const int64_t *)M->w_off.p;
    d_text = (const uint8_t *)M->w_text.p - text_lo;  // indexed by absolute offsets
  } else {
    d_off = text_offsets;
    d_text = (const uint8_t *)text;
  }
  int64_t *d_idoff = host ? (int64_t *)M->w_idoff.p : id_offsets;
  int32_t *d_ids = ids;
  int64_t d_cap = ids_capacity;
  if (host) {
    // bound: one id per normalized character plus the extra options
    d_cap = (text_hi - text_lo) * (int64_t)M->hm.expansion + 40 * n + 64;
    if (d_cap > ids_capacity + 1) d_cap = ids_capacity + 1;
    if ((rc = M->w_ids.ensure((size_t)(d_cap > 0 ? d_cap : 1) * 4))) return rc;
    d_ids = (int32_t *)M->w_ids.p;
    if (d_cap > ids_capacity) d_cap = ids_capacity;
  }

  if ((rc = M->w_scalars.ensure(64))) return rc;
  int64_t *d_carry = (int64_t *)M->w_scalars.p;        // running CSR total
  int64_t *d_wtotal = d_carry + 1;                      // scratch words of the chunk
  uint32_t *d_stor = (uint32_t *)(d_carry + 2);         // OR of all status words
  CUDA_TRY(cudaMemsetAsync(M->w_scalars.p, 0, 64, st));
  CUDA_TRY(cudaEventRecord(M->ev0, st));

  const ExtraPlan plan = make_extra_plan(extra_options, M->bos_id, M->eos_id);
  const int exp_ = M->hm.expansion;
  const int64_t kNormBudget = 256ll << 20, kMaxChunk = 1 << 20;
  int64_t c0 = 0;
  size_t nk = 0;
  while (c0 < n) {
    int64_t c1 = c0, bytes = 0;
    while (c1 < n && c1 - c0 < kMaxChunk) {
      const int64_t need = (int64_t)exp_ * (ho[c1 + 1] - ho[c1]) + 8;
      if (c1 > c0 && bytes + need > kNormBudget) break;
      bytes += need;
      ++c1;
    }
    const int64_t cn = c1 - c0;
    if ((rc = M->w_norm.ensure(bytes + 16))) return rc;
    if ((rc = M->w_stage.ensure(bytes * 4 + 16))) return rc;
    if ((rc = M->w_nlen.ensure(cn * 4))) return rc;
    if ((rc = M->w_words.ensure(cn * 4))) return rc;
    if ((rc = M->w_cnt.ensure(cn * 4))) return rc;
    if ((rc = M->w_status.ensure(cn * 4))) return rc;
    if ((rc = M->w_woff.ensure((cn + 1) * 8))) return rc;
    const unsigned grid = (unsigned)((cn + kThreads - 1) / kThreads);
    k_normalize<<<grid, kThreads, 0, st>>>(M->mv, d_text, d_off, c0, cn, exp_, (uint8_t *)M->w_norm.p,
                                           (int32_t *)M->w_nlen.p, (int32_t *)M->w_words.p,
                                           (uint32_t *)M->w_status.p);
    launches += 1;
    if ((rc = run_scan(M, st, (const int32_t *)M->w_words.p, cn, IdentityOp(), (int64_t *)M->w_woff.p, d_wtotal, 0,
                       &launches)))
      return rc;
    int64_t wtotal = 0;
    CUDA_TRY(cudaMemcpyAsync(&wtotal, d_wtotal, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if ((rc = M->w_scratch.ensure((size_t)wtotal * 4 + 16))) return rc;
    if (M->kev.size() < 2 * (nk + 1)) {
      cudaEvent_t a, b;
      CUDA_TRY(cudaEventCreate(&a));
      CUDA_TRY(cudaEventCreate(&b));
      M->kev.push_back(a); M->kev.push_back(b);
    }
    CUDA_TRY(cudaEventRecord(M->kev[2 * nk], st));
    k_encode<<<grid, kThreads, 0, st>>>(M->mv, d_off, c0, cn, exp_, (const uint8_t *)M->w_norm.p,
                                        (const int32_t *)M->w_nlen.p, (const int64_t *)M->w_woff.p,
                                        (int32_t *)M->w_scratch.p, (int32_t *)M->w_stage.p,
                                        (int32_t *)M->w_cnt.p, (uint32_t *)M->w_status.p, d_stor);
    CUDA_TRY(cudaEventRecord(M->kev[2 * nk + 1], st));
    ++nk;
    launches += 1;
    FinalCountOp fop;
    fop.extra = plan.npre + plan.nsuf;
    if ((rc = run_scan(M, st, (const int32_t *)M->w_cnt.p, cn, fop, d_idoff + c0, d_carry, 1, &launches))) return rc;
    const unsigned cgrid = (unsigned)((cn * 32 + 255) / 256);
    k_compact<<<cgrid, 256, 0, st>>>(plan, d_off, c0, cn, exp_, (const int32_t *)M->w_stage.p,
                                     (const int32_t *)M->w_cnt.p, d_idoff, d_ids, d_cap);
    launches += 1;
    CUDA_TRY(cudaGetLastError());
    c0 = c1;
  }
  CUDA_TRY(cudaEventRecord(M->ev1, st));
  int64_t hs[3] = {0, 0, 0};
  CUDA_TRY(cudaMemcpyAsync(hs, M->w_scalars.p, 24, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  const int64_t total = hs[0];
  const uint32_t stor = (uint32_t)hs[2];
  if (n == 0) {
    const int64_t z = 0;
    if (host) id_offsets[0] = 0;
    else CUDA_TRY(cudaMemcpyAsync(id_offsets, &z, 8, cudaMemcpyHostToDevice, st));
  }
  if (host && n > 0) {
    CUDA_TRY(cudaMemcpyAsync(id_offsets, d_idoff, (n + 1) * 8, cudaMemcpyDeviceToHost, st));
    const int64_t ncopy = total < ids_capacity ? total : ids_capacity;
    if (ncopy > 0) CUDA_TRY(cudaMemcpyAsync(ids, d_ids, ncopy * 4, cudaMemcpyDeviceToHost, st));
  }
  CUDA_TRY(cudaStreamSynchronize(st));
  float ms = 0;
  cudaEventElapsedTime(&ms, M->ev0, M->ev1);
  M->stats[0] = launches;
  M->stats[1] = (double)n;  // v1: every sentence goes through the exact tier
  M->stats[2] = ms;
  double kms = 0;
  for (size_t k = 0; k < nk; ++k) {
    float t = 0;
    cudaEventElapsedTime(&t, M->kev[2 * k], M->kev[2 * k + 1]);
    kms += t;
  }
  M->stats[4] = kms;
  M->stats[5] = (double)nk;
  M->stats[6] = (double)n;
  M->stats[7] = (double)total;
  size_t ws = 0;
  const DevBuf *bufs[] = {&M->w_text, &M->w_off, &M->w_ids, &M->w_idoff, &M->w_norm, &M->w_stage, &M->w_nlen,
                          &M->w_words, &M->w_cnt, &M->w_status, &M->w_woff, &M->w_scratch, &M->w_partial, &M->w_scalars};
  for (const DevBuf *b : bufs) ws += b->cap;
  M->stats[3] = (double)ws;
  if (stor & (ST_NORM_OVERFLOW | ST_CAND_OVERFLOW | ST_IDS_OVERFLOW)) {
    g_err = "internal capacity exceeded in the exact tier (status " + std::to_string(stor) + ")";
    return kInternal;
  }
  if (total > ids_capacity) {
    g_err = "ids_capacity too small: " + std::to_string(total) + " ids";
    return kResourceExhausted;
  }
  if (stor & ST_INTERNAL) {
    g_err = "all normalized characters are not consumed.";  // src/sentencepiece_processor.cc:543
    return kInternal;
  }
  return kOk;
}

}  // extern "C"
