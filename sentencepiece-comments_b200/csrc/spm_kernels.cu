// spm_kernels.cu -- CUDA kernels (sm_100a) and the C ABI of include/spm_cuda.h.
//
// One spm_cuda_encode_batch call is cut into chunks of sentences; per chunk
// (sentences are independent units; nothing here is GEMM-shaped, the work is
// byte / trie / hash / scan traffic).  A sentence is finished by the first tier
// that can prove its result exact:
//
//   TIER 1, WORD MEMO (spm_memo.cuh; models without U+2581 inside pieces)
//   k_memo_scan_words  raw text of 32 sentences staged through shared memory; lanes
//                      sweep it 8 bytes each (word starts / ends as bit masks), then
//                      one lane per word keys and probes; other sentences are listed
//                      (k_memo_scan_plain: the older thread-per-sentence form)
//   k_memo_scan        general normalizer writing word keys for the listed ones
//                                                     (src/normalizer.cc:88-300)
//   k_memo_words_*     every NEW word once: stand-alone Viterbi with margin
//                      bookkeeping / BPE merges of the word
//   k_memo_emit        tile per sentence, lane per word: copy the ids, carry the
//                      bound of the backtrace score; leftovers are listed
//   TIER 2, VITERBI / MERGE KERNELS over the leftovers (all sentences without memo)
//   k_hist/k_binscan/k_scatter  counting sort by raw length (descending)
//   k_normalize_fast   normalized text into a padded, aligned per-sentence slot
//   k_encode_fast      persistent CTAs, one thread per sentence: trie walk + Viterbi
//                      with the hot trie prefix, the DP ring and the text window in
//                      shared memory / registers (spm_fast.cuh;
//                      src/unigram_model.cc:535-604,222-261)
//   k_encode_bpe_fast  word-wise exact merges (src/bpe_model.cc:37-199)
//   TIER 3, GENERAL TIER (what tier 2 flags: near-tied Viterbi candidates, heavy
//                 expansion, lines >= 64 KiB; every sentence of other models)
//   k_gen_normalize / k_gen_encode(_block)  exact candidate lists / BPE agenda in
//                      global or shared scratch (spm_core.cuh)
//   ALL
//   k_validate         offsets checked on the device
//   k_scan_*           exclusive scans (scratch offsets, CSR offsets)
//   k_compact          extra options + CSR write
//                      (src/sentencepiece_processor.cc:319-330,945-979)
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/spm_cuda.h"
#include "spm_core.cuh"
#include "spm_fast.cuh"
#include "spm_memo.cuh"
#include "spm_estep.cuh"
#include "spm_warp.cuh"
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

// Every ABI entry runs on the model's device and hands the caller's current device back on
// every exit path (a host framework with several GPUs in one process keeps its own).
struct DeviceGuard {
  int prev = -1, dev = -1;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int device) : dev(device) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    if (prev >= 0 && prev != dev) cudaSetDevice(prev);
  }
};
#define DEVICE_GUARD(device)       \
  DeviceGuard device_guard_(device); \
  CUDA_TRY(device_guard_.err)

constexpr int kCapc = 8;          // near-tie candidates kept per position by the exact tier
constexpr int kThreads = 128;     // general-tier block size
#ifndef SPM_FAST_THREADS
#define SPM_FAST_THREADS 1024
#endif
constexpr int kFastThreads = SPM_FAST_THREADS; // fast-tier block size (one persistent CTA per SM)
constexpr int kBpeThreads = 512;  // BPE fast-tier block size
constexpr uint32_t ST_GEN = 1u << 16;   // sentence was encoded by the general tier (ids in gstage, forward order)
constexpr uint32_t ST_FWD = 1u << 17;   // fast-tier ids were staged in forward order (BPE)
constexpr uint32_t ST_PACKED = 1u << 18;  // word memo: ids forward, packed with the group's (k_memo_emit)

// device-side counters of one chunk
enum { C_WORK = 0, C_NFAST = 1, C_NGEN = 2, C_STATUS_OR = 3, C_WORK_LONG = 4, C_NLONG = 5,
       C_NVIT = 6, C_WORK_VIT = 7, C_NVITL = 8, C_WORK_VITL = 9, C_NNEW = 10, C_NSLOW = 11, C_NRETRY = 12,
       C_NMID = 13, C_WORK_MID = 14, C_COUNT = 16 };
// k_publish snapshot (int64 slots): counters | scalars | running id total of chunk k
constexpr int kPubScal = C_COUNT, kPubTot = C_COUNT + 8;

// ---------------------------------------------------------------- general tier
// Sentence t of the list is sentence list[t] of the chunk (list == nullptr: t).
__global__ void __launch_bounds__(kThreads)
k_gen_prepare(const int64_t *off, int64_t c0, const int32_t *list, const uint32_t *count, int64_t t0, int64_t tn,
              int exp_, int32_t *need) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= tn) return;
  int32_t v = 0;
  if (count == nullptr || t0 + t < (int64_t)*count) {
    const int64_t i = c0 + (list ? list[t0 + t] : t0 + t);
    v = (int32_t)(((int64_t)exp_ * (off[i + 1] - off[i]) + 8 + 7) & ~(int64_t)7);
  }
  need[t] = v;
}

__global__ void __launch_bounds__(kThreads)
k_gen_normalize(ModelView mv, const uint8_t *text, const int64_t *off, int64_t c0, const int32_t *list,
                const uint32_t *count, int64_t t0, int64_t tn, const int32_t *need, const int64_t *noff,
                uint8_t *norm, int32_t *nlen, int32_t *words, uint32_t *gstatus) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= tn) return;
  int32_t nl = 0, w = 0;
  uint32_t st = 0;
  if (count == nullptr || t0 + t < (int64_t)*count) {
    const int64_t i = c0 + (list ? list[t0 + t] : t0 + t);
    nl = normalize_sentence(mv, text + off[i], off[i + 1] - off[i], norm + noff[t], need[t]);
    if (nl < 0) { nl = 0; st = ST_NORM_OVERFLOW; }
    const int64_t ww = mv.model_type == M_UNIGRAM ? uni_scratch_words(mv.ring, kCapc, nl) : bpe_scratch_words(nl);
    w = (int32_t)((ww + 1) & ~(int64_t)1);  // keep 8-byte alignment of each region
  }
  nlen[t] = nl;
  words[t] = w;
  gstatus[t] = st;
}

// status_or[0]: OR of the status words; status_or[1]: some sentence met more than kCapc near-tied
// candidates and waits for the retry with room for every candidate (run_capc_retry).
__device__ __forceinline__ void gen_encode_one(const ModelView &mv, int64_t i, int64_t nb, int32_t nl, uint32_t st,
                                               int32_t *w, const uint8_t *norm, int32_t *gstage, int32_t *cnt,
                                               uint32_t *status, int64_t *srcoff, uint32_t *status_or,
                                               int32_t *gends = nullptr, int capc = kCapc) {
  int32_t c = 0;
  if (st == 0) {
    if (mv.model_type == M_UNIGRAM) {
      UniScratch sc;
      sc.ringM = (float *)w; w += mv.ring;
      sc.ringCnt = (uint32_t *)w; w += mv.ring;
      sc.candBt = (float *)w; w += mv.ring * capc;
      sc.candRec = (uint64_t *)w; w += 2 * mv.ring * capc;
      sc.back = (uint64_t *)w;
      sc.stride = nl + 1;
      if (capc == kCapc)
        c = unigram_encode<kCapc>(mv, norm + nb, nl, sc, gstage + nb, nl, &st, gends ? gends + nb : nullptr);
      else
        c = unigram_encode_capc(mv, norm + nb, nl, sc, capc, gstage + nb, nl, &st, gends ? gends + nb : nullptr);
    } else {
      c = bpe_encode(mv, norm + nb, nl, w, gstage + nb, nl, &st, gends ? gends + nb : nullptr);
    }
  }
  cnt[i] = c;
  status[i] = st | ST_GEN;
  srcoff[i] = nb;
  if (capc == kCapc && (st & ST_CAND_OVERFLOW)) {
    atomicOr(status_or + 1, 1u);
    st &= ~ST_CAND_OVERFLOW;
  }
  if (st) atomicOr(status_or, st);
}

// The retry: listed sentence t takes a scratch slice for `capc` candidates per position from
// a bump counter (plan), then runs the exact tier again (retry).  rbase[t] < 0: not flagged.
__device__ __forceinline__ int64_t retry_words(const ModelView &mv, int capc, int32_t nl) {
  return (uni_scratch_words(mv.ring, capc, nl) + 1) & ~(int64_t)1;   // 8-byte aligned slices
}
__global__ void __launch_bounds__(kThreads)
k_gen_retry_plan(ModelView mv, const int32_t *list, int64_t tn, const int32_t *nlen, const uint32_t *status,
                 int capc, int64_t *rbase, unsigned long long *total) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= tn) return;
  const int64_t i = list ? list[t] : t;
  int64_t b = -1;
  if (status[i] & ST_CAND_OVERFLOW) b = (int64_t)atomicAdd(total, (unsigned long long)retry_words(mv, capc, nlen[t]));
  rbase[t] = b;
}
__global__ void __launch_bounds__(kThreads)
k_gen_retry(ModelView mv, const int32_t *list, int64_t tn, const int64_t *noff, const uint8_t *norm,
            const int32_t *nlen, const int64_t *rbase, int capc, int32_t *scratch, int32_t *gstage, int32_t *cnt,
            uint32_t *status, int64_t *srcoff, uint32_t *status_or, int32_t *gends) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= tn || rbase[t] < 0) return;
  const int64_t i = list ? list[t] : t;
  gen_encode_one(mv, i, noff[t], nlen[t], 0u, scratch + rbase[t], norm, gstage, cnt, status, srcoff, status_or, gends,
                 capc);
}

// throughput variant: one thread per listed sentence, scratch in global memory
__global__ void __launch_bounds__(kThreads)
k_gen_encode(ModelView mv, const int32_t *list, const uint32_t *count, int64_t t0, int64_t tn,
             const int64_t *noff, const uint8_t *norm, const int32_t *nlen, const int64_t *woff, int64_t wbase,
             int32_t *scratch, int32_t *gstage, const uint32_t *gstatus, int32_t *cnt, uint32_t *status, int64_t *srcoff,
             uint32_t *status_or, int32_t *gends) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= tn) return;
  if (count != nullptr && t0 + t >= (int64_t)*count) return;
  const int64_t i = list ? list[t0 + t] : t0 + t;  // chunk-relative sentence index
  gen_encode_one(mv, i, noff[t], nlen[t], gstatus[t], scratch + (woff[t] - wbase), norm, gstage, cnt, status, srcoff,
                 status_or, gends);
}

// latency variant for short lists (the sentences the fast tier flags): one block per
// sentence, one working thread, scratch in shared memory when it fits -- the dependent
// accesses of the exact tier then cost ~30 cycles instead of an L2 round trip each.
constexpr int kGenSmemWords = 12288;   // 48 KB
__global__ void __launch_bounds__(32)
k_gen_encode_block(ModelView mv, const int32_t *list, int64_t tn, const int64_t *noff, const uint8_t *norm,
                   const int32_t *nlen, const int32_t *words, const int64_t *woff, int32_t *scratch, int32_t *gstage,
                   const uint32_t *gstatus, int32_t *cnt, uint32_t *status, int64_t *srcoff, uint32_t *status_or) {
  __shared__ __align__(16) int32_t sm[kGenSmemWords];
  const int64_t t = blockIdx.x;
  if (t >= tn || threadIdx.x != 0) return;
  const int64_t i = list ? list[t] : t;
  int32_t *w = words[t] <= kGenSmemWords ? sm : scratch + woff[t];
  gen_encode_one(mv, i, noff[t], nlen[t], gstatus[t], w, norm, gstage, cnt, status, srcoff, status_or);
}

// ---------------------------------------------------------------- fast tier
// Slot of sentence t (chunk-relative) in the normalized buffer: 16-byte aligned, room
// for 3 * len + 72 normalized bytes, kNormPad zero bytes and word-granular stores.
__device__ __forceinline__ int64_t fast_norm_base(const int64_t *off, int64_t c0, int64_t t) {
  return ((3 * (off[c0 + t] - off[c0]) + 15) & ~(int64_t)15) + 160 * t;
}
// Slot in the id staging buffer: len + kFastExtraChars ids (ids <= characters).
__device__ __forceinline__ int64_t fast_stage_base(const int64_t *off, int64_t c0, int64_t t) {
  return (off[c0 + t] - off[c0]) + kFastExtraChars * t;
}

// Sort key = raw length.  Empty sentences produce nothing; sentences of kFastMaxRaw bytes
// or more go straight to the general tier.  Block-private histogram in shared memory,
// one global atomic per non-empty bin and block.
constexpr int kSortThreads = 1024, kSortItems = 8, kSortTile = kSortThreads * kSortItems;
__global__ void __launch_bounds__(kSortThreads)
k_hist(const int64_t *off, int64_t c0, int64_t cn, const int32_t *list, int list_cnt, int32_t *cnt,
       uint32_t *status, uint32_t *hist, int32_t *gen_list, int32_t *long_list, int32_t *mid_list, uint32_t *counters) {
  // mid_list (may be null): lines of kFastMaxRaw <= bytes < kMidMaxRaw get a list of their own -- normalized in
  // pieces like the long ones, encoded by k_encode_fast like the short ones
  // list == nullptr: all cn sentences of the chunk; else the counters[list_cnt] listed ones
  // (what the word memo left over; their cnt / status are initialised already)
  __shared__ uint32_t sh[kFastMaxRaw];
  for (int k = threadIdx.x; k < kFastMaxRaw; k += kSortThreads) sh[k] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * kSortTile;
  const int64_t m = list ? (int64_t)counters[list_cnt] : cn;
#pragma unroll
  for (int j = 0; j < kSortItems; ++j) {
    const int64_t i = base + j * kSortThreads + threadIdx.x;
    if (i < m) {
      const int64_t t = list ? list[i] : i;
      const int64_t len = off[c0 + t + 1] - off[c0 + t];
      if (!list) {
        cnt[t] = 0;
        status[t] = 0;
      }
      if (len >= kLongMaxRaw) gen_list[atomicAdd(&counters[C_NGEN], 1u)] = (int32_t)t;
      else if (len >= kFastMaxRaw && len < kMidMaxRaw && mid_list) mid_list[atomicAdd(&counters[C_NMID], 1u)] = (int32_t)t;
      else if (len >= kFastMaxRaw) long_list[atomicAdd(&counters[C_NLONG], 1u)] = (int32_t)t;
      else if (len > 0) atomicAdd(&sh[len], 1u);
    }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < kFastMaxRaw; k += kSortThreads)
    if (sh[k]) atomicAdd(&hist[k], sh[k]);
}

// hist[1..kFastMaxRaw-1] -> first slot of each bin in DESCENDING key order; total -> counters[C_NFAST]
__global__ void __launch_bounds__(1024) k_binscan(uint32_t *hist, uint32_t *bincur, uint32_t *counters) {
  __shared__ uint32_t wsum[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  constexpr int kPer = kFastMaxRaw / 1024;
  uint32_t v[kPer], sum = 0;
#pragma unroll
  for (int j = 0; j < kPer; ++j) {  // thread 0 owns the longest keys
    const int key = kFastMaxRaw - 1 - (tid * kPer + j);
    v[j] = key >= 1 ? hist[key] : 0;
    sum += v[j];
  }
  uint32_t inc = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
    if (lane >= d) inc += y;
  }
  if (lane == 31) wsum[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    uint32_t s = wsum[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, s, d);
      if (lane >= d) s += y;
    }
    wsum[lane] = s;
  }
  __syncthreads();
  uint32_t run = (wid ? wsum[wid - 1] : 0) + inc - sum;
#pragma unroll
  for (int j = 0; j < kPer; ++j) {
    const int key = kFastMaxRaw - 1 - (tid * kPer + j);
    if (key >= 1) { bincur[key] = run; hist[key] = 0; }
    run += v[j];
  }
  if (tid == 1023) counters[C_NFAST] = run;
}

// Same tiles as k_hist: the block reserves a range of every bin it touches with one global
// atomic, then ranks its sentences inside the range with shared-memory atomics.
__global__ void __launch_bounds__(kSortThreads)
k_scatter(const int64_t *off, int64_t c0, int64_t cn, const int32_t *list, int list_cnt, uint32_t *bincur,
          int32_t *perm, const uint32_t *counters) {
  __shared__ uint32_t sh[kFastMaxRaw];
  for (int k = threadIdx.x; k < kFastMaxRaw; k += kSortThreads) sh[k] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * kSortTile;
  const int64_t m = list ? (int64_t)counters[list_cnt] : cn;
  int32_t len[kSortItems], who[kSortItems];
  uint32_t rank[kSortItems];
#pragma unroll
  for (int j = 0; j < kSortItems; ++j) {
    const int64_t i = base + j * kSortThreads + threadIdx.x;
    len[j] = 0;
    if (i < m) {
      const int64_t t = list ? list[i] : i;
      const int64_t l = off[c0 + t + 1] - off[c0 + t];
      who[j] = (int32_t)t;
      if (l > 0 && l < kFastMaxRaw) { len[j] = (int32_t)l; rank[j] = atomicAdd(&sh[l], 1u); }
    }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < kFastMaxRaw; k += kSortThreads)
    if (sh[k]) sh[k] = atomicAdd(&bincur[k], sh[k]);
  __syncthreads();
#pragma unroll
  for (int j = 0; j < kSortItems; ++j)
    if (len[j]) perm[sh[len[j]] + rank[j]] = who[j];
}

// One thread per sentence, in sorted order (lanes of a warp: equal raw lengths).
// Block shape of k_normalize_fast, measured on 2.2 M Japanese lines (the kernel's time inside the normalize
// stage): 256 threads 9.4 ms, 512 or 1024 threads 5.7 ms, 128 threads 16 ms; persistent blocks (SPM_NORM_BLOCKS
// per SM, tiles dealt round-robin) 5 x 256: 12.9, 2 x 512: 7.6, 3 x 512: 6.6 ms -- the grid over the whole
// chunk with 512 threads is what ships.
#ifndef SPM_NORM_THREADS
#define SPM_NORM_THREADS 512
#endif
constexpr int kNormThreads = SPM_NORM_THREADS;
#ifndef SPM_NORM_BLOCKS
#define SPM_NORM_BLOCKS (1 << 20)      // blocks of k_normalize_fast per SM at most (no cap)
#endif
__global__ void __launch_bounds__(kNormThreads)
k_normalize_fast(ModelView mv, const uint8_t *text, const int64_t *off, int64_t c0, const uint8_t *lo,
                 const uint8_t *hi, const int32_t *perm, int count_idx, uint8_t *norm, int32_t *nlen,
                 int32_t *gen_list, uint32_t *counters) {
  __shared__ uint32_t first2[2048];
  // (the grid covers the chunk at most; the list is usually far shorter -- what the word memo left over -- and
  // a block without work must not pay for the 8 KB table)
  const int64_t count = (int64_t)counters[count_idx];
  if ((int64_t)blockIdx.x * blockDim.x >= count) return;
  if (mv.cm_first2 != nullptr)
    for (int k = threadIdx.x; k < 2048; k += kNormThreads) first2[k] = mv.cm_first2[k];
  __syncthreads();
  // persistent blocks: tile k of the sorted list goes to block k mod gridDim, the table is staged once
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < count; idx += (int64_t)gridDim.x * blockDim.x) {
    const int32_t t = perm[idx];
    const int64_t i = c0 + t;
    const int64_t len = off[i + 1] - off[i];
    int32_t nc = 0;
    int32_t nl = fast_normalize(mv, first2, text + off[i], len, lo, hi, norm + fast_norm_base(off, c0, t),
                                (int32_t)(3 * len + 72), &nc);
    if (nl < 0 || nc > len + kFastExtraChars) {
      gen_list[atomicAdd(&counters[C_NGEN], 1u)] = t;
      nl = -1;
    }
    nlen[t] = nl;
  }
}

// Long lines: one BLOCK per line, its threads share the pieces of ~kNormChunk raw bytes.  Pieces start
// at positions norm_find_cut() proves safe, are normalized independently into a temporary
// slot each (fast_normalize_scan with mid = true), then the block concatenates them into the
// line's slot of `norm` and finishes the line as fast_normalize() does (trailing whitespace,
// suffix, zero padding).  A 33 KB line takes the time of three 128-byte pieces instead of 20 ms.
// (Pieces of 512 bytes, one per thread: a 4..8 KB line -- most long lines -- kept 8..16 of the block's
// 128 threads busy; 14.6 ms for the 350 MB of long lines in 2.2 M lines of Japanese text.)
#ifndef SPM_NORM_LONG_THREADS
#define SPM_NORM_LONG_THREADS 128
#endif
constexpr int kNormLongThreads = SPM_NORM_LONG_THREADS, kNormMaxChunks = kLongMaxRaw / kNormChunk;   // 512 pieces at most
__device__ __forceinline__ int64_t norm_tmp_base(const int64_t *off, int64_t c0, int64_t t) {
  const int64_t o = off[c0 + t] - off[c0];
  return ((3 * o + 7) & ~(int64_t)7) + 88 * (o / kNormChunk) + 96 * t;
}
__global__ void __launch_bounds__(kNormLongThreads)
k_normalize_long(ModelView mv, const uint8_t *text, const int64_t *off, int64_t c0, const uint8_t *lo,
                 const uint8_t *hi, const int32_t *list, int count_idx, uint8_t *tmp, uint8_t *norm, int32_t *nlen,
                 int32_t *gen_list, uint32_t *counters) {
  __shared__ uint32_t first2[2048];
  __shared__ int64_t cuts[kNormMaxChunks + 1];
  __shared__ int32_t olen[kNormMaxChunks], obase[kNormMaxChunks + 1], onch[kNormMaxChunks];
  __shared__ int bad, started, total_len, total_nch;
  if (blockIdx.x >= counters[count_idx]) return;
  if (mv.cm_first2 != nullptr)
    for (int k = threadIdx.x; k < 2048; k += kNormLongThreads) first2[k] = mv.cm_first2[k];
  const int32_t t = list[blockIdx.x];
  const int64_t o0 = off[c0 + t], len = off[c0 + t + 1] - o0;
  const uint8_t *in = text + o0;
  uint8_t *line_tmp = tmp + norm_tmp_base(off, c0, t);
  const int i = threadIdx.x;
  const int nchunks = (int)((len + kNormChunk - 1) / kNormChunk);   // <= kNormMaxChunks: len < kLongMaxRaw
  if (i == 0) { bad = 0; started = 0; cuts[0] = 0; cuts[nchunks] = len; }
  __syncthreads();
  for (int k = i; k < nchunks; k += kNormLongThreads)
    if (k > 0) cuts[k] = norm_find_cut(mv, first2, in, (int64_t)k * kNormChunk, len);
  __syncthreads();
  for (int k = i; k < nchunks; k += kNormLongThreads) {
    olen[k] = 0; onch[k] = 0;
    const int64_t a = cuts[k], b = cuts[k + 1];
    if (b > a) {
      bool content = false;
      const bool ok = norm_piece(mv, first2, in, a, b, lo, hi, line_tmp + norm_piece_slot(a, k), k, &olen[k], &onch[k],
                                 &content);
      if (k == 0 && content) started = 1;
      // pieces after the first assume that the line's content has started
      if (!ok || (k == 0 && !content && nchunks > 1)) bad = 1;
    }
  }
  __syncthreads();
  if (i < 32) {   // exclusive scan of the piece lengths by one warp
    int32_t run = 0, nc = 0;
    for (int k0 = 0; k0 < nchunks; k0 += 32) {
      const int k = k0 + i;
      const int32_t v = k < nchunks ? olen[k] : 0;
      int32_t inc = v, c = k < nchunks ? onch[k] : 0;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (i >= d) inc += y;
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, d);
      if (k < nchunks) obase[k] = run + inc - v;
      run += __shfl_sync(0xFFFFFFFFu, inc, 31);
      nc += c;
    }
    if (i == 0) {
      total_len = run; total_nch = nc;
      if (run + 3 > 3 * len + 72) bad = 1;
    }
  }
  __syncthreads();
  uint8_t *out = norm + fast_norm_base(off, c0, t);
  if (!bad) {
    // a warp per piece: 32 consecutive bytes per pass
    const int lane = i & 31, wid = i >> 5;
    for (int k = wid; k < nchunks; k += kNormLongThreads / 32) {
      const int32_t n_k = olen[k];
      if (n_k == 0) continue;
      const uint8_t *src = line_tmp + norm_piece_slot(cuts[k], k);
      uint8_t *dst = out + obase[k];
      for (int32_t j = lane; j < n_k; j += 32) dst[j] = src[j];
    }
  }
  __syncthreads();
  if (i == 0) {
    int32_t nl = bad ? -1 : norm_finish(mv, out, total_len, total_nch, started != 0, len);
    if (nl < 0) gen_list[atomicAdd(&counters[C_NGEN], 1u)] = t;
    nlen[t] = nl;
    total_len = nl;
  }
  __syncthreads();
  // zero padding: up to the next 8-byte boundary + kNormPad bytes
  const int32_t nl = total_len;
  if (nl >= 0) {
    const int32_t z1 = ((nl + 7) & ~7) + kNormPad;
    for (int32_t j = nl + i; j < z1; j += kNormLongThreads) out[j] = 0;
  }
}

// Checks the offsets of the whole batch on the device (no O(n) work on the host):
// scal[3] |= 1 when offsets decrease, scal[6] = longest sentence, scal[7] = longest
// sentence shorter than kFastMaxRaw, scal[1] = longest sentence shorter than kLongMaxRaw
// (both size the back-pointer arrays).
__global__ void __launch_bounds__(256) k_validate(const int64_t *off, int64_t n, unsigned long long *scal) {
  long long mx = 0, mxf = 0, mxl = 0;
  int bad = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const long long len = off[i + 1] - off[i];
    if (len < 0) bad = 1;
    if (len > mx) mx = len;
    if (len < kFastMaxRaw && len > mxf) mxf = len;
    if (len < kLongMaxRaw && len > mxl) mxl = len;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, d));
    mxf = max(mxf, __shfl_xor_sync(0xFFFFFFFFu, mxf, d));
    mxl = max(mxl, __shfl_xor_sync(0xFFFFFFFFu, mxl, d));
    bad |= __shfl_xor_sync(0xFFFFFFFFu, bad, d);
  }
  if ((threadIdx.x & 31) == 0) {
    if (bad) atomicOr(&scal[3], 1ull);
    if (mx > 0) atomicMax(&scal[6], (unsigned long long)mx);
    if (mxf > 0) atomicMax(&scal[7], (unsigned long long)mxf);
    if (mxl > 0) atomicMax(&scal[1], (unsigned long long)mxl);
  }
}

// Host-visible snapshot of the small device-side scalars.  The host reads them from
// MAPPED pinned memory after a stream sync, so no tiny copy ever queues behind the bulk
// PCIe transfers of the copy-in / copy-out streams.
// out[0..C_COUNT) = counters, out[kPubScal..+8) = scalars, out[kPubTot + slot] = scalars[0]
// (running id total).
__global__ void k_publish(const uint32_t *counters, const int64_t *scalars, volatile int64_t *out, int slot) {
  const int k = threadIdx.x;
  if (k < C_COUNT) out[k] = counters ? (int64_t)counters[k] : 0;
  else if (k < C_COUNT + 8) out[k] = scalars[k - C_COUNT];
  if (k == 0 && slot >= 0) out[kPubTot + slot] = scalars[0];
}

struct FastParams {
  const uint32_t *units;
  const float *score;
  const uint32_t *leaf_id;
  uint32_t k_units, k_scores, n_units;
  int32_t ring;
  float unk_score, smax;
  int32_t unk_id;
};

// Work units of the persistent encode kernels: `lpu` consecutive list entries per warp fetch.
// The length-sorted main pass always takes 32; the long-line pass has few, unsorted entries
// and spreads them over every resident warp (a warp with 4 busy lanes is as fast as one with
// 32 -- the kernels are latency bound).
__device__ __forceinline__ uint32_t fast_lanes_per_unit(uint32_t count, int warps_per_cta) {
  if (warps_per_cta * 32 == (int)blockDim.x) return 32u;
  const uint32_t warps = gridDim.x * (uint32_t)warps_per_cta;
  const uint32_t per = (count + warps - 1) / warps;
  return per < 1u ? 1u : (per > 32u ? 32u : per);
}

// Persistent CTAs (one per SM).  Shared memory: hot units | hot scores | ring_bt[R][T] | ring_bk[R][T].
__global__ void __launch_bounds__(kFastThreads, 1)
k_encode_fast(FastParams fp, const int64_t *off, int64_t c0, const uint8_t *norm, const int32_t *nlen,
              const int32_t *perm, int count_idx, int work_idx, int max_warps, uint32_t *back, int64_t back_rows,
              int32_t *stage, int32_t *cnt, uint32_t *status, int32_t *flag_list, int flag_idx, uint32_t *counters,
              uint32_t warp_route, int route_idx, int side_words) {
  // side_words: kFastSideWords at the end of every thread's back_rows hold the near-tied candidates
  // (fast_unigram), or 0: a sentence with a near tie is flagged.  route_idx: the list whose size decides the
  // routing below (the mid-length pass follows the main list's decision)
  // flag_list / flag_idx: where sentences with near-tied candidates go -- the list of k_encode_warp
  // (keeps kWarpCapc candidates per position) when that kernel follows, else the general tier's
  extern __shared__ __align__(16) uint32_t smem[];
  // A short list (what the word memo leaves over: a few thousand sentences with rare words) is
  // latency-bound here -- one thread walks a whole sentence through cold trie paths -- so it is
  // handed to k_encode_warp as a whole: 32 lanes per sentence, 6 x shorter on 6 k sentences.
  if (warp_route != 0 && counters[route_idx] <= warp_route) {
    const uint32_t nf = counters[count_idx];
    for (uint32_t i = blockIdx.x * kFastThreads + threadIdx.x; i < nf; i += gridDim.x * kFastThreads)
      flag_list[atomicAdd(&counters[flag_idx], 1u)] = perm[i];
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&counters[C_NRETRY], nf);
    return;
  }
  uint32_t *units_hot = smem;
  float *score_hot = (float *)(smem + fp.k_units);
  float *ring_bt = (float *)(smem + fp.k_units + fp.k_scores);
  uint32_t *ring_bk = (uint32_t *)(ring_bt + (size_t)fp.ring * kFastThreads);
  uint32_t *txt = ring_bk + (size_t)fp.ring * kFastThreads;   // [2][kFastThreads] text prefetch staging
  for (uint32_t k = threadIdx.x; k < fp.k_units; k += kFastThreads) units_hot[k] = fp.units[k];
  for (uint32_t k = threadIdx.x; k < fp.k_scores; k += kFastThreads) score_hot[k] = fp.score[k];
  __syncthreads();
  FastView fv;
  fv.units = fp.units; fv.score = fp.score; fv.leaf_id = fp.leaf_id;
  fv.units_hot = units_hot; fv.score_hot = score_hot;
  fv.units_hot_saddr = (uint32_t)__cvta_generic_to_shared(units_hot);
  fv.score_hot_saddr = (uint32_t)__cvta_generic_to_shared(score_hot);
  fv.k_units = fp.k_units; fv.k_scores = fp.k_scores; fv.n_units = fp.n_units;
  fv.ring = fp.ring; fv.unk_score = fp.unk_score; fv.smax = fp.smax; fv.unk_id = fp.unk_id;
  const uint32_t nfast = counters[count_idx];
  const int lane = threadIdx.x & 31;
  if ((int)(threadIdx.x >> 5) >= max_warps) return;   // long lines run on a few warps per SM (back arrays are big)
  const uint32_t lpu = fast_lanes_per_unit(nfast, max_warps);
  uint32_t *myback = back + ((int64_t)blockIdx.x * max_warps * 32 + threadIdx.x) * back_rows;
  for (;;) {
    uint32_t w = 0;
    if (lane == 0) w = atomicAdd(&counters[work_idx], 1u);
    w = __shfl_sync(0xFFFFFFFFu, w, 0);
    const uint32_t idx = w * lpu + lane;
    if (w * lpu >= nfast) break;
    if ((uint32_t)lane < lpu && idx < nfast) {
      const int32_t t = perm[idx];
      uint32_t st = 0;
      const int32_t nl = nlen[t];
      const int32_t c = nl <= 0 ? 0 :
                        fast_unigram(fv, norm + fast_norm_base(off, c0, t), nl, ring_bt + threadIdx.x,
                                     ring_bk + threadIdx.x, txt + threadIdx.x, kFastThreads, myback,
                                     stage + fast_stage_base(off, c0, t), &st,
                                     side_words ? myback + (back_rows - side_words) : nullptr);
      if (st) {
        flag_list[atomicAdd(&counters[flag_idx], 1u)] = t;
        if (flag_idx != C_NGEN) atomicAdd(&counters[C_NRETRY], 1u);
      } else {
        cnt[t] = c;
      }
    }
    __syncwarp();
  }
}

// The long lines in order of descending length (buckets of 256 bytes; one block): the persistent
// warps of k_encode_warp then start with the longest lines, which bound its run time.
__global__ void __launch_bounds__(1024)
k_sort_long(const int64_t *off, int64_t c0, const int32_t *list, int count_idx, int32_t *sorted,
            const uint32_t *counters) {
  constexpr int kBuckets = kLongMaxRaw / 256;
  __shared__ uint32_t hist[kBuckets];
  const uint32_t n = counters[count_idx];
  for (int k = threadIdx.x; k < kBuckets; k += 1024) hist[k] = 0;
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < n; i += 1024) {
    const int64_t t = list[i];
    atomicAdd(&hist[kBuckets - 1 - (int)((off[c0 + t + 1] - off[c0 + t]) >> 8)], 1u);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int k = 0; k < kBuckets; ++k) { const uint32_t c = hist[k]; hist[k] = run; run += c; }
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < n; i += 1024) {
    const int32_t t = list[i];
    sorted[atomicAdd(&hist[kBuckets - 1 - (int)((off[c0 + t + 1] - off[c0 + t]) >> 8)], 1u)] = t;
  }
}

// Long lines (kFastMaxRaw <= bytes < kLongMaxRaw) of unigram models: one WARP per line, exact
// candidate lists in shared memory (spm_warp.cuh).  Persistent warps pull lines from the list.
constexpr int kWarpWarps = 4;
template <int HITS>
__global__ void __launch_bounds__(kWarpWarps * 32)
k_encode_warp(FastParams fp, const int64_t *off, int64_t c0, const uint8_t *norm, const int32_t *nlen,
              const int32_t *list, int count_idx, int work_idx, uint32_t *back, int64_t stride, int32_t *stage,
              int32_t *cnt, int32_t *gen_list, uint32_t *counters) {
  __shared__ WarpScratch<HITS> ws[kWarpWarps];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  FastView fv;
  fv.units = fp.units; fv.score = fp.score; fv.leaf_id = fp.leaf_id;
  fv.units_hot = fp.units; fv.score_hot = fp.score; fv.units_hot_saddr = 0; fv.score_hot_saddr = 0;
  fv.k_units = 0; fv.k_scores = 0; fv.n_units = fp.n_units;
  fv.ring = fp.ring; fv.unk_score = fp.unk_score; fv.smax = fp.smax; fv.unk_id = fp.unk_id;
  const uint32_t count = counters[count_idx];
  uint32_t *myback = back + ((int64_t)blockIdx.x * kWarpWarps + wid) * kWarpCapc * stride;
  for (;;) {
    uint32_t w = 0;
    if (lane == 0) w = atomicAdd(&counters[work_idx], 1u);
    w = __shfl_sync(0xFFFFFFFFu, w, 0);
    if (w >= count) break;
    const int32_t t = list[w];
    const int32_t nl = nlen[t];
    if (nl <= 0) {                         // empty, or routed to the general tier by the normalizer (which sets cnt later)
      if (lane == 0) cnt[t] = 0;
      continue;
    }
    int32_t c = -1;
    if (nl + 2 <= stride)
      c = warp_unigram(fv, ws[wid], norm + fast_norm_base(off, c0, t), nl, myback, stride,
                       stage + fast_stage_base(off, c0, t));
    if (lane == 0) {
      if (c < 0) gen_list[atomicAdd(&counters[C_NGEN], 1u)] = t;
      else cnt[t] = c;
    }
    __syncwarp();
  }
}

// BPE fast tier: same scheduling as k_encode_fast.  Shared memory: hot units | hot scores |
// node[S][T] | ps[S][T] | text[B/4][T] | cb[..][T] | txt[2][T]  (T = kBpeThreads).
__global__ void __launch_bounds__(kBpeThreads, 1)
k_encode_bpe_fast(FastParams fp, const int64_t *off, int64_t c0, const uint8_t *norm, const int32_t *nlen,
                  const int32_t *perm, int count_idx, int work_idx, int max_warps, int32_t *stage, int32_t *cnt,
                  uint32_t *status, int32_t *gen_list, uint32_t *counters, uint32_t gen_route) {
  extern __shared__ __align__(16) uint32_t smem[];
  // A short list (what the word memo leaves over: a few hundred sentences with words it cannot key) is
  // pure latency here -- one thread merges a whole sentence word by word, 1.7 ms whatever the count --
  // so it goes to the general tier as a whole: one block per sentence, agenda in shared memory.
  if (gen_route != 0 && counters[count_idx] <= gen_route) {
    const uint32_t nf = counters[count_idx];
    for (uint32_t i = blockIdx.x * kBpeThreads + threadIdx.x; i < nf; i += gridDim.x * kBpeThreads)
      gen_list[atomicAdd(&counters[C_NGEN], 1u)] = perm[i];
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&counters[C_NRETRY], nf);
    return;
  }
  uint32_t *units_hot = smem;
  float *score_hot = (float *)(smem + fp.k_units);
  uint32_t *scr = smem + fp.k_units + fp.k_scores;
  for (uint32_t k = threadIdx.x; k < fp.k_units; k += kBpeThreads) units_hot[k] = fp.units[k];
  for (uint32_t k = threadIdx.x; k < fp.k_scores; k += kBpeThreads) score_hot[k] = fp.score[k];
  __syncthreads();
  FastView fv;
  fv.units = fp.units; fv.score = fp.score; fv.leaf_id = fp.leaf_id;
  fv.units_hot = units_hot; fv.score_hot = score_hot;
  fv.units_hot_saddr = (uint32_t)__cvta_generic_to_shared(units_hot);
  fv.score_hot_saddr = (uint32_t)__cvta_generic_to_shared(score_hot);
  fv.k_units = fp.k_units; fv.k_scores = fp.k_scores; fv.n_units = fp.n_units;
  fv.ring = 0; fv.unk_score = 0; fv.smax = 0; fv.unk_id = fp.unk_id;
  BpeScratch bs;
  bs.stride = kBpeThreads;
  bs.node = scr + threadIdx.x;
  bs.ps = (float *)(bs.node + kBpeMaxSyms * kBpeThreads);
  bs.text = bs.node + 2 * kBpeMaxSyms * kBpeThreads;
  bs.cb = bs.text + (kBpeMaxBytes / 4) * kBpeThreads;
  uint32_t *txt = bs.cb + ((kBpeMaxSyms + 4) / 4) * kBpeThreads;
  const uint32_t nfast = counters[count_idx];
  const int lane = threadIdx.x & 31;
  if ((int)(threadIdx.x >> 5) >= max_warps) return;
  const uint32_t lpu = fast_lanes_per_unit(nfast, max_warps);
  for (;;) {
    uint32_t w = 0;
    if (lane == 0) w = atomicAdd(&counters[work_idx], 1u);
    w = __shfl_sync(0xFFFFFFFFu, w, 0);
    const uint32_t idx = w * lpu + lane;
    if (w * lpu >= nfast) break;
    if ((uint32_t)lane < lpu && idx < nfast) {
      const int32_t t = perm[idx];
      uint32_t st = 0;
      const int32_t nl = nlen[t];
      const int32_t c = nl <= 0 ? 0 : fast_bpe(fv, norm + fast_norm_base(off, c0, t), nl, bs, txt,
                                               stage + fast_stage_base(off, c0, t), &st);
      if (st) {
        gen_list[atomicAdd(&counters[C_NGEN], 1u)] = t;
      } else {
        cnt[t] = c;
        status[t] = ST_FWD;
      }
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------- word-memo tier (spm_memo.cuh)
// Slots of the words of sentence t: at most len / 2 + 2 words come from the raw text
// itself; a charsmap expansion that adds more makes the sink give up on the sentence.
__device__ __forceinline__ int64_t memo_wslot_base(const int64_t *off, int64_t c0, int64_t t) {
  return ((off[c0 + t] - off[c0]) >> 1) + 4 * t;
}

// k_memo_scan_plain: one thread per sentence in ORIGINAL order -- the 32 sentences of a warp
// are one contiguous piece of the text buffer, which the warp copies to shared memory with
// coalesced 16-byte loads before every lane runs memo_scan_plain() on its own sentence:
// raw bytes straight into word keys, no normalized text is written.  Sentences of any other
// shape (or warps whose text does not fit the staging buffer) are listed for k_memo_scan.
// nwords[t] = word count, -1 = the memo cannot serve this sentence.
constexpr int kScanWarps = 8, kScanWarpBytes = 5120;
constexpr int kMemoThreads = 256;
__global__ void __launch_bounds__(kScanWarps * 32)
k_memo_scan_plain(ModelView mv, MemoTable tb, const uint8_t *text, const int64_t *off, int64_t c0, int64_t cn,
                  const uint8_t *lo, const uint8_t *hi, uint32_t *wslots, uint32_t *wbase, uint32_t half,
                  int32_t *nwords, int32_t *cnt, uint32_t *status, int32_t *slow_list, uint32_t *counters) {
  __shared__ __align__(16) uint8_t stage[kScanWarps][kScanWarpBytes];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t tw = ((int64_t)blockIdx.x * kScanWarps + wid) * 32;
  if (tw >= cn) return;
  const int64_t t = tw + lane;
  const int64_t a = off[c0 + tw], b = off[c0 + min(tw + 32, cn)];
  const uint8_t *src = (const uint8_t *)((uintptr_t)(text + a) & ~(uintptr_t)15);
  const int64_t span = (text + b) - src;
  const bool staged = span <= kScanWarpBytes - 32 && mv.ascii_plain && mv.escape_ws;
  if (staged) {
    for (int64_t i = lane * 16; i < span + 16; i += 512) {   // + 16: the loop reads past the last sentence
      const uint8_t *p = src + i;
      uint4 v;
      if (p >= lo && p + 16 <= hi) {
        v = __ldcs(reinterpret_cast<const uint4 *>(p));      // streamed once: do not displace the memo table in L2
      } else {
        uint8_t tmp[16];
        for (int j = 0; j < 16; ++j) tmp[j] = (p + j >= lo && p + j < hi) ? p[j] : 0;
        memcpy(&v, tmp, 16);
      }
      *reinterpret_cast<uint4 *>(&stage[wid][i]) = v;
    }
    __syncwarp();
  }
  if (t >= cn) return;
  const int64_t o0 = off[c0 + t], len = off[c0 + t + 1] - o0;
  int32_t nw = 0;
  if (len >= kLongMaxRaw) nw = -1;
  else if (len > 0 && !staged) nw = -2;
  else if (len > 0)
    nw = memo_scan_plain(tb, &stage[wid][0] + ((text + o0) - src), (int32_t)len, mv.add_dummy_prefix ? 1 : 0,
                         wslots + memo_wslot_base(off, c0, t), (int32_t)(len >> 1) + 4, memo_relax_of(mv));
  if (nw == -2) slow_list[atomicAdd(&counters[C_NSLOW], 1u)] = (int32_t)t;
  wbase[t] = (uint32_t)memo_wslot_base(off, c0, t) + (nw == -2 ? half : 0u);
  nwords[t] = nw;
  cnt[t] = 0;
  status[t] = 0;
}

// k_memo_scan_words: the same scan with the lanes of a warp working on BYTES and on WORDS instead
// of on sentences (one thread per sentence runs at the pace of the longest of 32 sentences and
// diverges at every word end: 14 active threads per warp instruction).  The warp stages the
// contiguous text of its 32 sentences as before, marks the sentence starts in a bitmap, then
//   sweep: 8 bytes per lane and step (sw_chunk_masks): word starts / word ends / unfit bytes as
//          bit masks, placed in two shared lists by a warp prefix sum (k-th start pairs with
//          the k-th end);
//   words: one lane per word: sentence of the word (binary search over the 32 starts), key from
//          the staged bytes (sw_word_slot), 16-byte probe / 128-bit CAS, slot stored at
//          (warp's slot list)[k]: coalesced, dense; wbase[t] tells k_memo_emit where the words
//          of sentence t begin.
// A warp whose text holds more than kSwCap words runs memo_scan_plain per lane instead.
#ifndef SPM_SW_BYTES          // build-time knobs of the measurements in profiles/r1_v21_summary.md
#define SPM_SW_BYTES 4096
#endif
// 32 sentences of the bench corpus: 3.2 KB of text and 512 words on average; 7.3 KB of shared
// memory per warp -> 7 CTAs = 28 warps per SM
constexpr int kSwWarps = 4, kSwBytes = SPM_SW_BYTES, kSwCap = 640, kSwLong = 64, kSwMapWords = kSwBytes / 32 + 4;
struct SwWarp {
  __align__(16) uint8_t stage[kSwBytes];
  uint32_t bmap[kSwMapWords];        // bit i: a sentence starts at stage byte i (or the warp's text ends there)
  uint16_t wpos[kSwCap], wend[kSwCap];   // first (| 0x8000: it begins its sentence) / last byte of word k
  uint16_t wlong[kSwLong];           // words of 15..24 bytes (two dependent probes): keyed after the others
  uint16_t sstart[32];               // first byte of the sentence of lane l
  uint16_t fw[32];                   // words before it
  uint32_t flags[32];                // 1: a word cannot be keyed (-1), 2: not a plain sentence (-2)
};
__device__ __forceinline__ int sw_owner(const uint16_t *sstart, int pos) {   // last sentence starting at or before pos
  int t = 0;
#pragma unroll
  for (int step = 16; step; step >>= 1)
    if ((int)sstart[t + step] <= pos) t += step;
  return t;
}
__global__ void __launch_bounds__(kSwWarps * 32)
k_memo_scan_words(ModelView mv, MemoTable tb, const uint8_t *text, const int64_t *off, int64_t c0, int64_t cn,
                  const uint8_t *lo, const uint8_t *hi, uint32_t *wslots, uint32_t *wbase, uint32_t half,
                  int32_t *nwords, int32_t *cnt, uint32_t *status, int32_t *slow_list, uint32_t *counters) {
  __shared__ SwWarp shw[kSwWarps];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  SwWarp &sh = shw[wid];
  const int64_t tw = ((int64_t)blockIdx.x * kSwWarps + wid) * 32;
  if (tw >= cn) return;
  const int64_t t = tw + lane;
  const bool live = t < cn;
  const int64_t a = off[c0 + tw], b = off[c0 + min(tw + 32, cn)];
  const int64_t o0 = live ? off[c0 + t] : b, o1 = live ? off[c0 + t + 1] : b, len = o1 - o0;
  const uint8_t *src = (const uint8_t *)((uintptr_t)(text + a) & ~(uintptr_t)15);
  const int64_t span = (text + b) - src;
  const bool staged = span <= kSwBytes - 48 && mv.ascii_plain && mv.escape_ws;   // sw_word_slot reads 32 + 7 bytes
  int32_t nw = 0;
  // where k_memo_emit finds the slots of my sentence: the general scan and the per-lane scan keep one
  // list per sentence in the upper half of wslots, the words of a swept warp lie side by side
  uint32_t wb = live ? half + (uint32_t)memo_wslot_base(off, c0, t) : 0u;
  if (!staged) {
    if (len >= kLongMaxRaw) nw = -1;
    else if (len > 0) nw = -2;
  } else {
    for (int64_t i = lane * 16; i < span + 16; i += 512) {   // + 16: the sweep looks one byte past the text
      const uint8_t *p = src + i;
      uint4 v;
      if (p >= lo && p + 16 <= hi) {
        v = __ldcs(reinterpret_cast<const uint4 *>(p));      // streamed once: do not displace the memo table in L2
      } else {
        uint8_t tmp[16];
        for (int j = 0; j < 16; ++j) tmp[j] = (p + j >= lo && p + j < hi) ? p[j] : 0;
        memcpy(&v, tmp, 16);
      }
      *reinterpret_cast<uint4 *>(&sh.stage[i]) = v;
    }
    const int a_rel = (int)((text + a) - src), b_rel = (int)span, s_rel = (int)((text + o0) - src);
    for (int j = lane; j < (b_rel >> 5) + 2; j += 32) sh.bmap[j] = 0;
    sh.sstart[lane] = (uint16_t)s_rel;
    sh.flags[lane] = 0;
    __syncwarp();
    atomicOr(&sh.bmap[s_rel >> 5], 1u << (s_rel & 31));
    if (lane == 31) atomicOr(&sh.bmap[b_rel >> 5], 1u << (b_rel & 31));
    __syncwarp();
    // ---- sweep
    const int relax = memo_relax_of(mv);   // leading / trailing / repeated spaces (and tabs) only separate words
    const uint8_t *bm8 = reinterpret_cast<const uint8_t *>(sh.bmap);
    int total = 0, total_e = 0;
    bool over = false;
    for (int g0 = a_rel & ~7; g0 < b_rel; g0 += 256) {
      const int g = g0 + lane * 8;
      SwMasks m;
      m.ws = 0; m.we = 0; m.bad = 0; m.b = 0;
      if (g < b_rel) {
        const uint64_t w = *reinterpret_cast<const uint64_t *>(&sh.stage[g]);
        const uint32_t pb = g > 0 ? sh.stage[g - 1] : 0u, nb = sh.stage[g + 8];
        const uint32_t b9 = (uint32_t)bm8[g >> 3] | ((uint32_t)bm8[(g >> 3) + 1] << 8);
        m = sw_chunk_masks(w, pb, nb, b9, a_rel - g, b_rel - g, relax);
      }
      const uint32_t mine = (uint32_t)__popc(m.ws) | ((uint32_t)__popc(m.we) << 16);
      uint32_t inc = mine;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += y;
      }
      const uint32_t all = __shfl_sync(0xFFFFFFFFu, inc, 31);
      if (total + (int)(all & 0xFFFFu) > kSwCap) { over = true; break; }
      int ks = total + (int)((inc - mine) & 0xFFFFu), ke = total_e + (int)((inc - mine) >> 16);
      // (bit 15 -- the word begins its sentence: the dummy prefix decides its first key bit -- is set below)
      for (uint32_t x = m.ws; x; x &= x - 1) sh.wpos[ks++] = (uint16_t)(g + __ffs(x) - 1);
      for (uint32_t x = m.we; x; x &= x - 1) sh.wend[ke++] = (uint16_t)(g + __ffs(x) - 1);
      for (uint32_t x = m.bad; x; x &= x - 1) atomicOr(&sh.flags[sw_owner(sh.sstart, g + __ffs(x) - 1)], 2u);
      total += (int)(all & 0xFFFFu);
      total_e += (int)(all >> 16);
    }
    __syncwarp();
    if (over) {
      if (len > 0)
        nw = memo_scan_plain(tb, &sh.stage[s_rel], (int32_t)len, mv.add_dummy_prefix ? 1 : 0, wslots + wb,
                             (int32_t)(len >> 1) + 4, relax);
    } else {
      // words before my sentence = starts listed below my first byte
      int first = 0;
      for (int n = total; n > 0;) {
        const int half = n >> 1;
        if ((int)(sh.wpos[first + half] & 0x7FFFu) < s_rel) { first += half + 1; n -= half + 1; }
        else n = half;
      }
      sh.fw[lane] = (uint16_t)first;
      const int next = __shfl_down_sync(0xFFFFFFFFu, first, 1);
      nw = (lane == 31 ? total : next) - first;
      // the first word of a sentence (not necessarily at its first byte: leading spaces) carries bit 15
      if (nw > 0) sh.wpos[first] |= 0x8000u;
      __syncwarp();
      // ---- words
      const uint32_t wbw = (uint32_t)memo_wslot_base(off, c0, tw);   // the warp's words: wl[0 .. total)
      uint32_t *wl = wslots + wbw;
      const uint64_t prefix0 = mv.add_dummy_prefix ? 1ull : 0ull;
      // A word of more than 14 bytes takes two dependent probes (head key, tail key): one such word
      // among 32 would hold the whole warp for two more round trips, so these wait in a list.
      int nlong = 0;
      // Software pipeline over the words: the key of word k + 32 is built and the first load of its probe
      // issued BEFORE the probe of word k is waited for, so the key arithmetic (~80 instructions, shared
      // loads) runs under the L2 / DRAM round trip of the probe instead of after it.
      MemoProbe pc;
      int s_c = 0;
      bool have_c = false, later_c = false;
      auto prep = [&](int k, MemoProbe *p, int *s_out, bool *have, bool *later) {
        *have = false; *later = false;
        if (k < total) {
          const uint32_t e = sh.wpos[k];
          const int s = (int)(e & 0x7FFFu), n = (int)sh.wend[k] - s + 1;
          *s_out = s;
          if (n <= kMemoKeyBytes) {
            uint64_t k0w, k1w;
            sw_short_key(&sh.stage[s], n, (e >> 15) ? prefix0 : 1ull, &k0w, &k1w);
            memo_probe_issue(tb, k0w, k1w, p);
            *have = true;
          } else if (n <= kMemoMaxBytes) {
            *later = true;
          } else {
            atomicOr(&sh.flags[sw_owner(sh.sstart, s)], 1u);      // longer than a key pair holds
          }
        }
      };
      prep(lane, &pc, &s_c, &have_c, &later_c);
      for (int k0 = 0; k0 < total; k0 += 32) {
        const int k = k0 + lane;
        MemoProbe pn;
        int s_n = 0;
        bool have_n, later_n;
        prep(k + 32, &pn, &s_n, &have_n, &later_n);
        if (have_c) {
          const uint32_t slot = memo_probe_finish(tb, pc);
          if (slot == 0xFFFFFFFFu) atomicOr(&sh.flags[sw_owner(sh.sstart, s_c)], 1u);   // rare: the owner is only looked up here
          else memo_store_slot(wl + k, slot);
        }
        const bool later = later_c;
        const uint32_t bl = __ballot_sync(0xFFFFFFFFu, later);
        if (bl) {
          const int at = nlong + __popc(bl & ((1u << lane) - 1u));
          if (later && at < kSwLong) sh.wlong[at] = (uint16_t)k;
          nlong += __popc(bl);
          if (nlong > kSwLong) {   // no room: key it now
            if (later && at >= kSwLong) {
              const uint32_t e = sh.wpos[k];
              const int s = (int)(e & 0x7FFFu), n = (int)sh.wend[k] - s + 1;
              const uint32_t slot = sw_word_slot(tb, &sh.stage[s], n, (e >> 15) ? prefix0 : 1ull);
              if (slot == 0xFFFFFFFFu) atomicOr(&sh.flags[sw_owner(sh.sstart, s)], 1u);
              else memo_store_slot(wl + k, slot);
            }
            nlong = kSwLong;
          }
        }
        pc = pn; s_c = s_n; have_c = have_n; later_c = later_n;
      }
      __syncwarp();
      for (int j = lane; j < nlong; j += 32) {
        const int k = sh.wlong[j];
        const uint32_t e = sh.wpos[k];
        const int s = (int)(e & 0x7FFFu), n = (int)sh.wend[k] - s + 1;
        const uint32_t slot = sw_word_slot(tb, &sh.stage[s], n, (e >> 15) ? prefix0 : 1ull);
        if (slot == 0xFFFFFFFFu) atomicOr(&sh.flags[sw_owner(sh.sstart, s)], 1u);
        else memo_store_slot(wl + k, slot);
      }
      __syncwarp();
      const uint32_t f = sh.flags[lane];
      if (f & 2u) nw = -2;
      else if (f & 1u) nw = -1;
      else wb = wbw + (uint32_t)first;
    }
  }
  if (!live) return;
  if (nw == -2) slow_list[atomicAdd(&counters[C_NSLOW], 1u)] = (int32_t)t;
  wbase[t] = wb;
  nwords[t] = nw;
  cnt[t] = 0;
  status[t] = 0;
}

// The sentences listed for the general scan, in order of descending raw length (bins of 8 bytes): the 32
// threads of a warp of k_memo_scan then walk sentences of one length instead of running at the pace
// of the longest of 32 random ones.  Counting sort: block-private histograms, one block scans the bins.
constexpr int kLsBins = 512, kLsThreads = 256;
__device__ __forceinline__ int ls_bin(const int64_t *off, int64_t c0, int32_t t) {
  const int64_t len = off[c0 + t + 1] - off[c0 + t];
  const int b = (int)(len >> 3);
  return kLsBins - 1 - (b < kLsBins ? b : kLsBins - 1);      // bin 0 = longest
}
__global__ void __launch_bounds__(kLsThreads)
k_ls_hist(const int64_t *off, int64_t c0, const int32_t *list, const uint32_t *counters, int count_idx, uint32_t *hist) {
  __shared__ uint32_t sh[kLsBins];
  for (int k = threadIdx.x; k < kLsBins; k += kLsThreads) sh[k] = 0;
  __syncthreads();
  const uint32_t n = counters[count_idx];
  for (uint32_t i = blockIdx.x * kLsThreads + threadIdx.x; i < n; i += gridDim.x * kLsThreads)
    atomicAdd(&sh[ls_bin(off, c0, list[i])], 1u);
  __syncthreads();
  for (int k = threadIdx.x; k < kLsBins; k += kLsThreads)
    if (sh[k]) atomicAdd(&hist[k], sh[k]);
}
__global__ void __launch_bounds__(kLsBins) k_ls_scan(uint32_t *hist, uint32_t *cursor) {
  __shared__ uint32_t wsum[kLsBins / 32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t v = hist[tid];
  uint32_t inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
    if (lane >= d) inc += y;
  }
  if (lane == 31) wsum[wid] = inc;
  __syncthreads();
  uint32_t base = 0;
  for (int k = 0; k < wid; ++k) base += wsum[k];
  cursor[tid] = base + inc - v;
  hist[tid] = 0;                      // ready for the next chunk
}
__global__ void __launch_bounds__(kLsThreads)
k_ls_scatter(const int64_t *off, int64_t c0, const int32_t *list, const uint32_t *counters, int count_idx,
             uint32_t *cursor, int32_t *sorted) {
  const uint32_t n = counters[count_idx];
  for (uint32_t i = blockIdx.x * kLsThreads + threadIdx.x; i < n; i += gridDim.x * kLsThreads) {
    const int32_t t = list[i];
    sorted[atomicAdd(&cursor[ls_bin(off, c0, t)], 1u)] = t;
  }
}

// k_memo_scan: the general scan (fast_normalize_scan + MemoSink: every normalizer rule,
// user symbols, whitespace policy, long words) over the sentences k_memo_scan_plain listed.
#ifndef SPM_MEMO_SCAN_THREADS
#define SPM_MEMO_SCAN_THREADS 256
#endif
constexpr int kMemoScanThreads = SPM_MEMO_SCAN_THREADS;
__global__ void __launch_bounds__(kMemoScanThreads)
k_memo_scan(ModelView mv, MemoTable tb, const uint8_t *text, const int64_t *off, int64_t c0, const uint8_t *lo,
            const uint8_t *hi, const int32_t *slow_list, uint32_t *wslots, int32_t *nwords,
            const uint32_t *counters) {
  __shared__ uint32_t first2[2048];
  if ((int64_t)blockIdx.x * blockDim.x >= (int64_t)counters[C_NSLOW]) return;   // (see k_normalize_fast)
  if (mv.cm_first2 != nullptr)
    for (int k = threadIdx.x; k < 2048; k += kMemoScanThreads) first2[k] = mv.cm_first2[k];
  __syncthreads();
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)counters[C_NSLOW]) return;
  const int64_t t = slow_list[idx];
  const int64_t o0 = off[c0 + t], len = off[c0 + t + 1] - o0;
  MemoSink sink;
  sink.init(tb, wslots + memo_wslot_base(off, c0, t), (int32_t)(len >> 1) + 4);
  int32_t nc = 0;
  bool content = false;
  fast_normalize_scan(mv, first2, text + o0, len, lo, hi, sink, &nc, &content);
  nwords[t] = sink.finish(mv.remove_extra_ws != 0);
}

__device__ __forceinline__ FastView memo_cold_view(const FastParams &fp) {
  FastView fv;
  fv.units = fp.units; fv.score = fp.score; fv.leaf_id = fp.leaf_id;
  fv.units_hot = fp.units; fv.score_hot = fp.score;     // nothing staged in shared memory here
  fv.units_hot_saddr = 0; fv.score_hot_saddr = 0;
  fv.k_units = 0; fv.k_scores = 0; fv.n_units = fp.n_units;
  fv.ring = fp.ring; fv.unk_score = fp.unk_score; fv.smax = fp.smax; fv.unk_id = fp.unk_id;
  return fv;
}

// One thread per word claimed in this pass.
__global__ void __launch_bounds__(kMemoThreads)
k_memo_words_unigram(FastParams fp, int neg_only, MemoTable tb) {
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t n_new = min(*tb.n_new, tb.max_new);
  if (idx >= n_new) return;
  const FastView fv = memo_cold_view(fp);
  memo_unigram_word(fv, neg_only != 0, tb, idx);
}

// BPE: the word text goes to a 64-byte global slot first (fast_bpe prefetches its text with
// cp.async from global memory); the merges run in the next kernel.
__global__ void __launch_bounds__(kMemoThreads) k_memo_words_text(MemoTable tb, uint8_t *wtext, int32_t *wlen) {
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t n_new = min(*tb.n_new, tb.max_new);
  if (idx >= n_new) return;
  __align__(16) uint8_t buf[64];
  const int nb = memo_word_bytes(tb.keys, tb.new_list[idx], buf, 64);
  uint4 *dst = reinterpret_cast<uint4 *>(wtext + (size_t)idx * 64);
  const uint4 *src = reinterpret_cast<const uint4 *>(buf);
#pragma unroll
  for (int j = 0; j < 4; ++j) dst[j] = src[j];
  wlen[idx] = nb;
}

constexpr int kMemoBpeThreads = 128;
__global__ void __launch_bounds__(kMemoBpeThreads)
k_memo_words_bpe(FastParams fp, MemoTable tb, const uint8_t *wtext, const int32_t *wlen) {
  __shared__ uint32_t scr[(kBpeWordsPerThread + 2) * kMemoBpeThreads];
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t n_new = min(*tb.n_new, tb.max_new);
  if (idx >= n_new) return;
  FastView fv = memo_cold_view(fp);
  fv.ring = 0; fv.unk_score = 0; fv.smax = 0;
  BpeScratch bs;
  bs.stride = kMemoBpeThreads;
  bs.node = scr + threadIdx.x;
  bs.ps = (float *)(bs.node + kBpeMaxSyms * kMemoBpeThreads);
  bs.text = bs.node + 2 * kBpeMaxSyms * kMemoBpeThreads;
  bs.cb = bs.text + (kBpeMaxBytes / 4) * kMemoBpeThreads;
  uint32_t *txt = bs.cb + ((kBpeMaxSyms + 4) / 4) * kMemoBpeThreads;
  int32_t ids[kBpeMaxSyms];
  uint32_t st = 0;
  const int32_t nb = wlen[idx];
  const int32_t c = nb > 0 ? fast_bpe(fv, wtext + (size_t)idx * 64, nb, bs, txt, ids, &st) : 0;
  memo_store_ids(tb, idx, ids, c, st != 0 || nb == 0, fv.unk_id);
}

// k_memo_emit: one WARP per group of 32 consecutive sentences, one lane per WORD of the group's flat word
// list (memo_emit() of spm_memo.cuh restated with warp scans).  k_memo_scan_words lays the slots of a
// group's words side by side, so 32 lanes fetch 32 slots with one coalesced load and gather the 32-byte
// result records of their words; what a sentence needs in order -- the bound of the backtrace score
// (exclusive sum of sabs), the unknown run that continues across a word boundary -- is a SEGMENTED warp
// scan over the flat list, and the id positions are one plain prefix sum over the whole group: the ids
// of the group's sentences land back to back ("packed") from the staging slot of its first sentence, so
// that k_compact streams whole groups.  A sentence with a word that is not memoised / not context-free
// here joins the list of the Viterbi tier (its ids in the packed area are never read).
// spos[t]: where the ids of sentence t start, relative to the group's first staging slot.
constexpr int kEmitWarps = 8;
#ifndef SPM_EMIT_MINB          // resident CTAs per SM the register allocation aims at (A/B builds)
#define SPM_EMIT_MINB 5
#endif
__global__ void __launch_bounds__(kEmitWarps * 32, SPM_EMIT_MINB)
k_memo_emit(MemoTable tb, int unigram, const int64_t *off, int64_t c0, int64_t cn, const uint32_t *wslots,
            const uint32_t *wbase, const int32_t *nwords, int32_t *stage, int32_t *cnt, uint32_t *status,
            int32_t *spos, int32_t *vit_list, uint32_t *counters, int list_idx) {
  __shared__ int32_t s_pos[kEmitWarps][32];
  __shared__ uint32_t s_bad[kEmitWarps];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const unsigned FULL = 0xFFFFFFFFu;
  const int64_t groups = (cn + 31) >> 5;
  for (int64_t g = (int64_t)blockIdx.x * kEmitWarps + wid; g < groups; g += (int64_t)gridDim.x * kEmitWarps) {
    const int64_t tw = g << 5, t = tw + lane;
    const bool live = t < cn;
    const int32_t nw_raw = live ? nwords[t] : 0;
    const int32_t nw = nw_raw > 0 ? nw_raw : 0;
    const uint32_t wb = nw > 0 ? wbase[t] : 0u;
    int32_t cap = 0;
    if (live) cap = (int32_t)(off[c0 + t + 1] - off[c0 + t]) + kFastExtraChars;
    // exclusive prefix of the word counts: flat word f of the group belongs to the last sentence s with pre[s] <= f
    int32_t pinc = nw;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int32_t y = __shfl_up_sync(FULL, pinc, d);
      if (lane >= d) pinc += y;
    }
    const int32_t pre = pinc - nw, total = __shfl_sync(FULL, pinc, 31);
    const int64_t gbase = fast_stage_base(off, c0, tw);
    const int32_t lim = (int32_t)min((int64_t)0x7FFFFFFF, fast_stage_base(off, c0, min(tw + 32, cn)) - gbase);
    int32_t *out = stage + gbase;
    if (lane == 0) s_bad[wid] = 0;
    s_pos[wid][lane] = 0;
    __syncwarp();
    int32_t run_ids = 0;           // ids placed by the words before this step (uniform)
    float carry_bub = 0.0f;        // of the sentence that continues from the step before
    bool carry_unk = false;
    for (int32_t f0 = 0; f0 < total; f0 += 32) {
      const int32_t f = f0 + lane;
      const bool active = f < total;
      int s = 0;
#pragma unroll
      for (int step = 16; step; step >>= 1) {
        const int32_t v = __shfl_sync(FULL, pre, (s + step) & 31);
        if (s + step < 32 && v <= f) s += step;
      }
      const int32_t my_pre = __shfl_sync(FULL, pre, s);
      const uint32_t my_wb = __shfl_sync(FULL, wb, s);
      const bool head = active && f == my_pre;
      uint32_t meta = kMemoReady;
      float bmax = 3.0e38f, sabs = 0.0f;
      int4 lo4 = make_int4(0, 0, 0, 0), hi4 = make_int4(0, 0, 0, 0);
      if (active) {
        const uint32_t slot = __ldcs(wslots + my_wb + (uint32_t)(f - my_pre));   // slots: streamed
        const int4 *r = reinterpret_cast<const int4 *>(tb.res + slot);
        lo4 = __ldg(r); hi4 = __ldg(r + 1);
        meta = (uint32_t)lo4.x; bmax = __int_as_float(lo4.y); sabs = __int_as_float(lo4.z);
      }
      bool ok = (meta & (kMemoReady | kMemoBad)) == kMemoReady;
      // inclusive scan of sabs inside my sentence: lanes from the head of my sentence (or lane 0 when
      // it began in an earlier step) up to me
      const uint32_t heads = __ballot_sync(FULL, head);
      const uint32_t below = heads & (0xFFFFFFFFu >> (31 - lane));
      const int seg0 = below ? 31 - __clz(below) : 0;
      const bool continued = below == 0;                    // my sentence began before this step
      float inc = sabs;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const float y = __shfl_up_sync(FULL, inc, d);
        if (lane - d >= seg0) inc += y;
      }
      if (continued) inc += carry_bub;
      if (unigram && active && !(inc - sabs < bmax)) ok = false;
      // an unknown run continues across a word boundary: drop the repeated id
      const bool last_unk = active && (meta & kMemoLastUnk) != 0;
      bool prev_unk = __shfl_up_sync(FULL, last_unk, 1);
      if (lane == 0) prev_unk = carry_unk;
      if (head) prev_unk = false;
      const int k = active ? (int)(meta & 0xFFu) : 0;
      const int drop = (active && prev_unk && (meta & kMemoFirstUnk)) ? 1 : 0;
      int pos = k - drop;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int y = __shfl_up_sync(FULL, pos, d);
        if (lane >= d) pos += y;
      }
      const int step_ids = __shfl_sync(FULL, pos, 31);
      pos += run_ids - (k - drop);                           // exclusive, group-wide
      if (head) s_pos[wid][s] = pos;
      if (!ok) atomicOr(&s_bad[wid], 1u << s);
      if (k > 0 && pos + k - drop <= lim) {
        const int32_t v[kMemoInlineIds] = {lo4.w, hi4.x, hi4.y, hi4.z, hi4.w};
        if (k <= kMemoInlineIds) {
#pragma unroll
          for (int j = 0; j < kMemoInlineIds; ++j)
            if (j >= drop && j < k) __stcs(out + pos + j - drop, v[j]);
        } else {
          const int32_t *ov = tb.overflow + (size_t)v[kMemoInlineIds - 1] * kMemoOverflowIds;
#pragma unroll
          for (int j = 0; j < kMemoInlineIds - 1; ++j)
            if (j >= drop) __stcs(out + pos + j - drop, v[j]);
          for (int j = kMemoInlineIds - 1; j < k; ++j) __stcs(out + pos + j - drop, ov[j - (kMemoInlineIds - 1)]);
        }
      }
      run_ids += step_ids;
      carry_bub = __shfl_sync(FULL, inc, 31);
      carry_unk = __shfl_sync(FULL, last_unk, 31);
    }
    __syncwarp();
    // per sentence: its ids are [my_pos, start of the next sentence with words) of the packed area
    const uint32_t badmask = s_bad[wid];
    const int32_t my_pos = nw > 0 ? s_pos[wid][lane] : 0;
    int32_t nxt = nw > 0 ? my_pos : 0x7FFFFFFF;            // inclusive suffix minimum
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int32_t y = __shfl_down_sync(FULL, nxt, d);
      if (lane + d < 32) nxt = min(nxt, y);
    }
    int32_t end = __shfl_down_sync(FULL, nxt, 1);
    if (lane == 31) end = 0x7FFFFFFF;
    end = min(end, run_ids);
    const int32_t count = nw > 0 ? end - my_pos : 0;
    const bool bad = live && (nw_raw < 0 || ((badmask >> lane) & 1u) || count > cap || my_pos + count > lim);
    if (live && !bad) {
      cnt[t] = count;
      status[t] = ST_FWD | ST_PACKED;
      spos[t] = my_pos;
    }
    const uint32_t bm = __ballot_sync(FULL, bad);
    if (bm) {
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(&counters[list_idx], (uint32_t)__popc(bm));
      base = __shfl_sync(FULL, base, 0);
      if (bad) vit_list[base + __popc(bm & ((1u << lane) - 1u))] = (int32_t)t;
    }
    __syncwarp();
  }
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

// Final layout (extra options) into the CSR values, one warp per group of 32 sentences.  Each lane
// reads the header of one sentence (coalesced).  A group the word memo served as a whole lies
// packed in `stage` from the group's first slot, in CSR order: one streaming copy.  Otherwise
// the warp copies the 32 sentences one after another: memo-served ones from the packed area
// (forward), Viterbi / merge tier ones from their own slot in `vstage` (unigram: reversed),
// general-tier ones from `gstage` (forward).
__global__ void __launch_bounds__(256)
k_compact(const __grid_constant__ ExtraPlan plan, const int64_t *off, int64_t c0, int64_t cn, const int32_t *stage,
          const int32_t *vstage, const int32_t *gstage, const int64_t *srcoff, const uint32_t *status,
          const int32_t *cnt, const int32_t *spos, const int64_t *id_off, int32_t *ids, int64_t capacity) {
  const int lane = threadIdx.x & 31;
  const unsigned FULL = 0xFFFFFFFFu;
  const int64_t t0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32;
  if (t0 >= cn) return;
  const int64_t t = t0 + lane;
  int32_t my_core = -1;
  int64_t my_src = 0, my_dst = 0;
  int my_gen = 0;      // 0: tier 2, reversed; 1: general tier; 2: tier 2, forward; 3: word memo, packed
  bool packed = true;
  const int64_t gbase = fast_stage_base(off, c0, t0);
  if (t < cn) {
    const uint32_t st = status[t];
    my_core = cnt[t];
    packed = (st & ST_PACKED) != 0;
    my_gen = (st & ST_GEN) != 0 ? 1 : (packed ? 3 : ((st & ST_FWD) ? 2 : 0));
    my_src = my_gen == 1 ? srcoff[t] : (packed ? gbase + spos[t] : fast_stage_base(off, c0, t));
    my_dst = id_off[c0 + t];
  }
  const bool simple = plan.npre == 0 && plan.nsuf == 0;
  const int m = (int)min((int64_t)32, cn - t0);
  if (simple && plan.flip == 0 && __all_sync(FULL, packed)) {
    // the whole group: ids [gbase, gbase + total) -> [id_off[t0], id_off[t0 + m])
    const int64_t dst = __shfl_sync(FULL, my_dst, 0);
    const int64_t total = id_off[c0 + t0 + m] - dst;
    const int32_t *src = stage + gbase;
    if (dst + total <= capacity) {
      // 16-byte stores: peel to the alignment of the destination, then every lane moves 4 ids per
      // step (four independent loads in flight; the source is aligned to 4 bytes only)
      const int64_t head = min(total, (int64_t)((4 - (((uintptr_t)(ids + dst) >> 2) & 3)) & 3));
      if (lane < head) __stcs(ids + dst + lane, __ldcs(src + lane));
      const int64_t body = (total - head) >> 2;
      const int32_t *s4 = src + head;
      int4 *d4 = reinterpret_cast<int4 *>(ids + dst + head);
      for (int64_t q = lane; q < body; q += 32) {
        int4 v;
        v.x = __ldcs(s4 + 4 * q); v.y = __ldcs(s4 + 4 * q + 1); v.z = __ldcs(s4 + 4 * q + 2); v.w = __ldcs(s4 + 4 * q + 3);
        __stcs(d4 + q, v);
      }
      const int64_t done = head + 4 * body;
      if (done + lane < total) __stcs(ids + dst + done + lane, __ldcs(src + done + lane));
    } else {
      for (int64_t q = lane; q < total; q += 32)
        if (dst + q < capacity) ids[dst + q] = src[q];
    }
    return;
  }
  for (int j = 0; j < m; ++j) {
    const int32_t core = __shfl_sync(FULL, my_core, j);
    const int gen = __shfl_sync(FULL, my_gen, j);
    const int64_t so = __shfl_sync(FULL, my_src, j);
    const int64_t dst = __shfl_sync(FULL, my_dst, j);
    const int32_t *src = (gen == 1 ? gstage : (gen == 3 ? stage : vstage)) + so;
    if (simple) {
      const bool rev = (plan.flip != 0) == (gen != 0);   // unigram fast-tier staging is already reversed
      for (int32_t q = lane; q < core; q += 32)
        if (dst + q < capacity) ids[dst + q] = src[rev ? core - 1 - q : q];
    } else {
      ExtraPlan pl = plan;
      if (!gen) pl.flip ^= 1;
      const int32_t fc = final_count(pl, core);
      for (int32_t q = lane; q < fc; q += 32)
        if (dst + q < capacity) ids[dst + q] = final_id_at(pl, src, core, q);
    }
  }
}

// ---------------------------------------------------------------- line splitting (the step before the path)
// A raw file buffer -> CSR (text without the '\n' bytes, offsets), lines as std::getline
// yields them to spm_encode (src/filesystem.cc:42-44, src/spm_encode_main.cc:183-193): a
// line ends at '\n' (dropped; a '\r' before it stays in the line), a last line without
// '\n' counts when it is not empty.  Tiles of kNlTile bytes; pass 1 counts the newlines of
// a tile, pass 2 (after the scan of the counts) ranks them and moves the bytes.
constexpr int kNlThreads = 256, kNlPer = 16, kNlTile = kNlThreads * kNlPer;
__global__ void __launch_bounds__(kNlThreads) k_nl_count(const uint8_t *buf, int64_t len, int32_t *tile_count) {
  const int64_t p0 = (int64_t)blockIdx.x * kNlTile + (int64_t)threadIdx.x * kNlPer;
  int c = 0;
#pragma unroll
  for (int k = 0; k < kNlPer; ++k)
    if (p0 + k < len && buf[p0 + k] == '\n') ++c;
  int64_t tot;
  block_exclusive_scan((int64_t)c, &tot);
  if (threadIdx.x == 0) tile_count[blockIdx.x] = (int32_t)tot;
}
__global__ void __launch_bounds__(kNlThreads)
k_nl_scatter(const uint8_t *buf, int64_t len, const int64_t *tile_base, uint8_t *text, int64_t *offsets,
             int64_t capacity) {
  const int64_t p0 = (int64_t)blockIdx.x * kNlTile + (int64_t)threadIdx.x * kNlPer;
  uint8_t b[kNlPer];
  int c = 0;
#pragma unroll
  for (int k = 0; k < kNlPer; ++k) {
    b[k] = p0 + k < len ? buf[p0 + k] : 0;
    if (p0 + k < len && b[k] == '\n') ++c;
  }
  int64_t tot;
  int64_t rank = block_exclusive_scan((int64_t)c, &tot) + tile_base[blockIdx.x];   // newlines before p0
#pragma unroll
  for (int k = 0; k < kNlPer; ++k) {
    const int64_t p = p0 + k;
    if (p >= len) break;
    if (b[k] == '\n') {
      ++rank;                                             // line `rank` starts after this byte
      if (rank <= capacity) offsets[rank] = p + 1 - rank;
    } else {
      text[p - rank] = b[k];
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) offsets[0] = 0;
}

// ---------------------------------------------------------------- unigram trainer E-step (spm_estep.cuh)
// One thread per word, words taken in order of descending length (counting sort, so the 32 words
// of a warp run equally long).  Words of up to 23 bytes keep their forward / backward
// / Viterbi tables in shared memory ([position][thread]: no local-memory traffic); longer ones
// (text without spaces) use a slice of a global scratch buffer.  expected[] is accumulated in
// double: pieces with a hot slot (the kEstepHot - 1 highest scores, which collect most of the mass)
// in a per-CTA shared table flushed once at the end, the rest with one global atomicAdd per node;
// the log-likelihood and the token count are reduced per warp first.
constexpr int kEstepThreads = 640, kEstepBins = 64;   // ONE persistent CTA per SM (20 warps)
constexpr int kEstepPos = 24;      // words of up to 23 bytes keep their tables (and their bytes) in shared memory
constexpr int kEstepHotUnits = 14336;   // units of the top of the trie kept in shared memory (56 KB)
// hot table | LogSumExp polynomials | hot units | A, BV (float), C, word (u8): 223,232 bytes
constexpr int estep_smem_bytes(int pos) {
  return kEstepHot * 8 + kLseN * kLseCoef * 8 + kEstepHotUnits * 4 + pos * kEstepThreads * 10;
}
__global__ void __launch_bounds__(256)
k_estep_hist(const int64_t *off, int64_t n, uint32_t *hist) {
  __shared__ uint32_t sh[kEstepBins];
  if (threadIdx.x < kEstepBins) sh[threadIdx.x] = 0;
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t len = off[i + 1] - off[i];
    atomicAdd(&sh[kEstepBins - 1 - (int)min(len < 0 ? (int64_t)0 : len, (int64_t)(kEstepBins - 1))], 1u);
  }
  __syncthreads();
  if (threadIdx.x < kEstepBins && sh[threadIdx.x]) atomicAdd(&hist[threadIdx.x], sh[threadIdx.x]);
}
__global__ void k_estep_binscan(uint32_t *hist, uint32_t *cursor) {   // one warp, 2 bins per lane
  const int lane = threadIdx.x;
  const uint32_t a = hist[2 * lane], b = hist[2 * lane + 1];
  uint32_t inc = a + b;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
    if (lane >= d) inc += y;
  }
  cursor[2 * lane] = inc - a - b;
  cursor[2 * lane + 1] = inc - b;
}
__global__ void __launch_bounds__(256)
k_estep_scatter(const int64_t *off, int64_t n, uint32_t *cursor, uint32_t *perm) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t len = off[i + 1] - off[i];
    const int bin = kEstepBins - 1 - (int)min(len < 0 ? (int64_t)0 : len, (int64_t)(kEstepBins - 1));
    perm[atomicAdd(&cursor[bin], 1u)] = (uint32_t)i;
  }
}

// POS = table entries per thread in shared memory (9 bytes each): what bounds the resident warps.
// One launch over the whole sorted list [*range_lo or 0, *range_hi).  (Tried: three launches, one
// per length class with 12 / 20 / 32 entries -- more resident warps for the short words, but the
// tails of the classes no longer overlap: 1.25 ms against 1.16 ms.)
template <int POS>
__global__ void __launch_bounds__(kEstepThreads, 1)
k_estep(EstepView ev, const int32_t *hot_ids, const uint8_t *words, const int64_t *off, const int64_t *freq,
        const uint32_t *range_lo, const uint32_t *range_hi, const uint32_t *perm, double inv_all_freq, float *scratch,
        double *expected, double *obj_ntok, float *logz) {
  extern __shared__ __align__(16) uint8_t estep_smem[];     // hot table | A | beta / Viterbi | C
  constexpr int kTab = POS * kEstepThreads;
  const int64_t lo = range_lo ? (int64_t)*range_lo : 0, n = (int64_t)*range_hi;
  if (lo + (int64_t)blockIdx.x * blockDim.x >= n) return;
  double *s_hot = reinterpret_cast<double *>(estep_smem);
  double *s_lse = s_hot + kEstepHot;
  uint32_t *s_units = reinterpret_cast<uint32_t *>(s_lse + kLseN * kLseCoef);
  float *sA = reinterpret_cast<float *>(s_units + kEstepHotUnits), *sB = sA + kTab;
  uint8_t *sC = reinterpret_cast<uint8_t *>(sB + kTab), *sW = sC + kTab;
  for (int k = threadIdx.x; k < kEstepHot; k += kEstepThreads) s_hot[k] = 0.0;
  // (with most of the SM's memory carved out as shared, L1 is ~30 KB: whatever every word touches sits here)
  for (int k = threadIdx.x; k < kLseN * kLseCoef; k += kEstepThreads) s_lse[k] = ev.lse[k];
  for (uint32_t k = threadIdx.x; k < ev.k_units; k += kEstepThreads) s_units[k] = ev.units[k];
  ev.lse = s_lse;
  ev.units_hot = s_units;
  __syncthreads();
  double obj = 0.0, ntok = 0.0;
  // persistent CTAs: tile k of the sorted list goes to CTA k mod gridDim (every CTA gets long and short
  // words alike), so the hot table is zeroed and flushed once per CTA, not once per 128 words
  for (int64_t idx = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = perm[idx];
    const int64_t o = off[i], len = off[i + 1] - o;
    float z = 0.0f;
    int32_t nt = 0;
    if (len > 0 && len < (1 << 20)) {
      const float f = (float)freq[i];       // PopulateMarginal takes float freq (src/unigram_model.cc:272)
      auto add = [&](uint32_t x, double v) {
        const uint32_t slot = ev.id_mask == 0xFFFFFu ? x >> 20 : 0u;
        if (slot) atomicAdd(&s_hot[slot], v);
        else atomicAdd(&expected[x & ev.id_mask], v);
      };
      const uint8_t *w = words + o;
      if (len < POS) {
        uint8_t *myw = sW + threadIdx.x;
        for (int j = 0; j < (int)len; ++j) myw[j * kEstepThreads] = w[j];
        z = estep_word(ev, [myw](int j) -> uint32_t { return myw[j * kEstepThreads]; }, (int)len, f, sA + threadIdx.x,
                       sB + threadIdx.x, sC + threadIdx.x, kEstepThreads, add, &nt);
      } else {
        float *base = scratch + 4 * ((o - off[0]) + i);
        z = estep_word(ev, [w](int j) -> uint32_t { return w[j]; }, (int)len, f, base, base + (len + 1),
                       (int32_t *)(base + 2 * (len + 1)), 1, add, &nt);
      }
      // objs[n] -= Z / all_sentence_freq with Z = freq * log-likelihood (:262-266)
      obj -= (double)(f * z) * inv_all_freq;
      ntok += (double)nt;
    } else if (len != 0) {
      obj = __longlong_as_double(0x7FF8000000000000ll);   // a word of 1 MiB or more: the objective comes back NaN
    }
    if (logz) logz[i] = z;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    obj += __shfl_xor_sync(0xFFFFFFFFu, obj, d);
    ntok += __shfl_xor_sync(0xFFFFFFFFu, ntok, d);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&obj_ntok[0], obj);
    atomicAdd(&obj_ntok[1], ntok);
  }
  __syncthreads();
  for (int k = threadIdx.x + 1; k < kEstepHot; k += kEstepThreads) {
    const double v = s_hot[k];
    if (v != 0.0) atomicAdd(&expected[hot_ids[k]], v);
  }
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

// Per-handle knobs (spm_cuda_model_set_option).  The defaults are what ships; tests and
// measurements switch tiers off / shrink tables through the setter to reach paths that real
// batches only take at scale.  The SPM_CUDA_* environment variables of round 1 are read ONCE
// PER LOAD as initial values (tools/gpu_variants.sh style A/B runs), never latched process-wide.
struct Options {
  int64_t memo = 1;               // word-memo tier
  int64_t memo_bits = 23;         // log2 of the most table slots a call may use (12..26)
  int64_t host_chunks = 0;        // host-memory calls: chunk count (0: 12 unigram / 5 BPE)
  int64_t host_chunk_min = 1 << 20;   // ... for calls of at least this many sentences
  int64_t max_chunk = 0;          // most sentences per chunk (0: 2^24 fast models, 2^20 others)
  int64_t warp = 1;               // k_encode_warp for long / flagged lines
  int64_t scan_per_sentence = 0;  // k_memo_scan_plain instead of k_memo_scan_words
  int64_t warp_route = 262144;    // leftover lists up to this size go to k_encode_warp as a whole (0: never)
  int64_t emit_blocks = 0;        // CTAs of k_memo_emit per SM (0: occupancy query)
  int64_t pieces_exact = 0;       // pieces path: every sentence through the exact tier
  int64_t pieces_plain = 1;       // pieces path: plain sentences (printable ASCII, single spaces) skip the normalizer
  int64_t fast_side = 1;          // k_encode_fast keeps near-tied candidates itself (0: flags the sentence, round 1)
  int64_t mid_lines = 0;          // lines of 4..8 KiB: second pass of k_encode_fast instead of the warp kernel (measured: a loss, see below)
  int64_t bpe_route = 8192;       // BPE leftover lists up to this size go to the general tier as a whole (0: never)
};

struct spm_cuda_model {
  HostModel hm;
  int device = 0;
  Options opt;
  ModelView mv;
  FastParams fp;
  bool fast = false;       // a fast tier is usable
  int fast_kind = 0;       // 1: unigram (k_encode_fast), 2: BPE (k_encode_bpe_fast)
  int sm_count = 1;
  size_t fast_smem = 0;
  DevBuf d_units, d_leaf, d_uunits, d_uleaf, d_cm_units, d_cm_pool, d_first2, d_cont2, d_funits, d_fscore, d_fleaf, d_first3;
  // workspace of encode calls (guarded by mu)
  mutable std::mutex mu;
  mutable bool lshist_zeroed = false;
  mutable DevBuf w_lshist, w_text, w_off, w_ids, w_idoff, w_norm, w_stage, w_vstage, w_spos, w_nlen, w_nchars, w_cnt, w_status, w_srcoff,
      w_perm, w_genlist, w_longlist, w_midlist, w_wslots, w_wbase, w_nwords, w_vit, w_vitl, w_memo, w_mres, w_movf, w_wback, w_ntmp, w_newlist, w_wtext, w_wlen, w_hist, w_counters, w_back, w_partial, w_scalars, g_need, g_noff, g_norm, g_stage,
      g_nlen, g_words, g_woff, g_scratch, g_status, g_rbase, g_rscratch;
  mutable DevBuf px[20];   // workspace of the pieces / decode / dense calls (spm_pieces.inc)
  mutable int emit_blocks = 0;      // resident CTAs of k_memo_emit per SM (occupancy query, first use)
  mutable uint32_t last_stor = 0;   // OR of the status words of the last encode_batch_locked call
  mutable bool pt_ready = false;    // piece strings uploaded (px[PX_PLEN..PX_PPOOL])
  mutable bool dec_ready = false;   // decode tables uploaded (px[PX_DECREC], px[PX_DECPOOL])
  mutable int dec_maxlen = 1;       // longest decoded piece in bytes
  mutable std::vector<int64_t> h_woff;
  mutable double stats[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  mutable cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  mutable std::vector<cudaEvent_t> kev;  // event pairs around the stages of a call
  mutable std::vector<int> kev_tag;
  mutable size_t nk = 0;
  // host-memory calls: copy-in / copy-out streams so that PCIe traffic overlaps the kernels
  mutable cudaStream_t s_in = nullptr, s_out = nullptr;
  mutable std::vector<cudaEvent_t> ev_in, ev_done;
  mutable cudaEvent_t ev_start = nullptr, ev_out = nullptr, ev_mid = nullptr;
  mutable int64_t *h_pin = nullptr;   // mapped pinned memory: snapshots written by k_publish
  mutable int64_t *d_pin = nullptr;   // its device address
  int bos_id = 0, eos_id = 0, pad_id = 0;
  std::vector<DevBuf *> workspace_bufs() const {   // all_bufs() without the model tables (d_*)
    std::vector<DevBuf *> v = all_bufs();
    v.erase(v.begin(), v.begin() + 12);
    return v;
  }
  std::vector<DevBuf *> all_bufs() const {
    spm_cuda_model *m = const_cast<spm_cuda_model *>(this);
    return {&m->d_units, &m->d_leaf, &m->d_uunits, &m->d_uleaf, &m->d_cm_units, &m->d_cm_pool, &m->d_first2, &m->d_cont2,
            &m->d_funits, &m->d_fscore, &m->d_fleaf, &m->d_first3, &m->w_text, &m->w_off, &m->w_ids, &m->w_idoff, &m->w_norm,
            &m->w_stage, &m->w_vstage, &m->w_spos, &m->w_lshist, &m->w_nlen, &m->w_nchars, &m->w_cnt, &m->w_status, &m->w_srcoff, &m->w_perm,
            &m->w_genlist, &m->w_longlist, &m->w_midlist, &m->w_wslots, &m->w_wbase, &m->w_nwords, &m->w_vit, &m->w_vitl, &m->w_memo, &m->w_mres, &m->w_movf, &m->w_wback, &m->w_ntmp,
            &m->w_newlist, &m->w_wtext, &m->w_wlen, &m->w_hist, &m->w_counters, &m->w_back, &m->w_partial, &m->w_scalars, &m->g_need,
            &m->g_noff, &m->g_norm, &m->g_stage, &m->g_nlen, &m->g_words, &m->g_woff, &m->g_scratch, &m->g_status, &m->g_rbase, &m->g_rscratch,
            &m->px[0], &m->px[1], &m->px[2], &m->px[3], &m->px[4], &m->px[5], &m->px[6], &m->px[7], &m->px[8], &m->px[9],
            &m->px[10], &m->px[11], &m->px[12], &m->px[13], &m->px[14], &m->px[15], &m->px[16], &m->px[17], &m->px[18],
            &m->px[19]};
  }
};

namespace {

int upload(DevBuf *b, const void *src, size_t bytes) {
  if (bytes == 0) return kOk;
  int rc = b->ensure(bytes);
  if (rc) return rc;
  CUDA_TRY(cudaMemcpy(b->p, src, bytes, cudaMemcpyHostToDevice));
  return kOk;
}

int upload_fast_view(spm_cuda_model *M) {
  HostModel &hm = M->hm;
  int rc;
  if ((rc = upload(&M->d_funits, hm.funits.data(), hm.funits.size() * 4))) return rc;
  if ((rc = upload(&M->d_fscore, hm.fscore.data(), hm.fscore.size() * 4))) return rc;
  if ((rc = upload(&M->d_fleaf, hm.fleaf.data(), hm.fleaf.size() * 4))) return rc;
  return kOk;
}

// Chooses the fast tier the model can use (it can change with SetVocabulary: UNUSED pieces
// switch BPE to the general tier) and sizes its shared memory: hot trie prefix + per-thread state.
int configure_fast(spm_cuda_model *M) {
  HostModel &hm = M->hm;
  int rc;
  M->fast = false;
  M->fast_kind = 0;
  int smem_max = 0;
  CUDA_TRY(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, M->device));
  const int R = hm.max_piece_chars + 1;
  size_t state_bytes = 0;
  int kind = 0;
  if (hm.fast_ok) { kind = 1; state_bytes = (size_t)R * kFastThreads * 8 + 2 * kFastThreads * 4; }  // DP rings + text staging
  else if (hm.bpe_fast_ok) { kind = 2; state_bytes = (size_t)(kBpeWordsPerThread + 2) * kBpeThreads * 4; }
  if (kind == 0 || state_bytes + 32768 > (size_t)smem_max) return kOk;
  if ((rc = upload_fast_view(M))) return rc;
  const size_t n_units = hm.funits.size();
  const size_t hot_words = ((size_t)smem_max - state_bytes - 256) / 4;
  size_t ku = std::min(n_units, hot_words * 2 / 3 & ~(size_t)3);
  size_t ks = std::min(n_units, (hot_words - ku) & ~(size_t)3);
  if (ks == n_units) ku = std::min(n_units, (hot_words - ks) & ~(size_t)3);
  FastParams &fp = M->fp;
  fp.units = (const uint32_t *)M->d_funits.p; fp.score = (const float *)M->d_fscore.p;
  fp.leaf_id = (const uint32_t *)M->d_fleaf.p;
  fp.k_units = (uint32_t)ku; fp.k_scores = (uint32_t)ks; fp.n_units = (uint32_t)n_units;
  fp.ring = R; fp.unk_score = M->mv.unk_score; fp.smax = M->mv.smax; fp.unk_id = M->mv.unk_id;
  M->fast_smem = (ku + ks) * 4 + state_bytes;
  if (kind == 1)
    CUDA_TRY(cudaFuncSetAttribute(k_encode_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)M->fast_smem));
  else
    CUDA_TRY(cudaFuncSetAttribute(k_encode_bpe_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)M->fast_smem));
  M->fast = true;
  M->fast_kind = kind;
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
  if ((rc = upload(&M->d_cont2, hm.cm_cont2.data(), hm.cm_cont2.size() * 4))) return rc;
  if ((rc = upload(&M->d_first3, hm.cm_first3.data(), hm.cm_first3.size() * 4))) return rc;
  ModelView &v = M->mv;
  memset(&v, 0, sizeof(v));
  v.pieces.units = (const uint32_t *)M->d_units.p; v.pieces.leaf = (const uint2 *)M->d_leaf.p;
  v.pieces.n = (uint32_t)hm.pieces.units.size();
  v.user.units = (const uint32_t *)M->d_uunits.p; v.user.leaf = (const uint2 *)M->d_uleaf.p;
  v.user.n = (uint32_t)hm.user.units.size();
  v.cm_units = (const uint32_t *)M->d_cm_units.p; v.cm_pool = (const uint8_t *)M->d_cm_pool.p;
  v.cm_first2 = hm.cm_first2.empty() ? nullptr : (const uint32_t *)M->d_first2.p;
  v.cm_cont2 = hm.cm_cont2.empty() ? nullptr : (const uint32_t *)M->d_cont2.p;
  v.cm_first3 = hm.cm_first3.empty() ? nullptr : (const uint32_t *)M->d_first3.p;
  v.cm_n = hm.cm_n;
  memcpy(v.user_first, hm.user_first, sizeof(v.user_first));
  v.model_type = hm.model_type; v.unk_id = hm.unk_id;
  v.unk_score = hm.min_score - 10.0f;  // kUnkPenalty, src/unigram_model.cc:550-551
  v.max_score = hm.max_score; v.smax = hm.smax; v.ring = hm.ring;
  v.add_dummy_prefix = hm.add_dummy_prefix; v.remove_extra_ws = hm.remove_extra_ws;
  v.escape_ws = hm.escape_ws; v.ws_suffix = hm.ws_suffix; v.has_unused = hm.has_unused;
  v.ascii_plain = hm.ascii_plain; v.ws_ctl = hm.ws_ctl;

  CUDA_TRY(cudaDeviceGetAttribute(&M->sm_count, cudaDevAttrMultiProcessorCount, M->device));
  return configure_fast(M);
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

// Event pairs around the stages of a call; `which` 0 opens a pair tagged `tag`, 1 closes it.
enum { T_ENCODE = 0, T_NORMALIZE = 1, T_MEMO_COLLECT = 2, T_MEMO_WORDS = 3, T_MEMO_EMIT = 4, T_COMPACT = 5, T_COUNT = 6 };
int kev_record(const spm_cuda_model *M, cudaStream_t st, int which, int tag = T_ENCODE) {
  while (M->kev.size() < 2 * (M->nk + 1)) {
    cudaEvent_t a;
    CUDA_TRY(cudaEventCreate(&a));
    M->kev.push_back(a);
  }
  if (M->kev_tag.size() < M->nk + 1) M->kev_tag.resize(M->nk + 1);
  CUDA_TRY(cudaEventRecord(M->kev[2 * M->nk + which], st));
  if (which == 0) M->kev_tag[M->nk] = tag;
  if (which == 1) ++M->nk;
  return kOk;
}

// After an exact-tier pass over `tn` listed sentences: when some met more than kCapc near-tied
// candidates at one position (status_or[1], published with the scalars), those run again with
// room for every candidate a position can hold (min(ring, 255)), so one pathological sentence
// neither fails the batch nor changes its ids.  d_scal[2] = status_or[0] | status_or[1] << 32.
int run_capc_retry(const spm_cuda_model *M, cudaStream_t st, const int32_t *list, int64_t tn, int64_t *d_scal,
                   int32_t *gends, double *launches) {
  if (M->hm.model_type != M_UNIGRAM || tn <= 0) return kOk;
  int rc;
  k_publish<<<1, 32, 0, st>>>(nullptr, d_scal, M->d_pin, -1);
  CUDA_TRY(cudaStreamSynchronize(st));
  if ((M->h_pin[kPubScal + 2] >> 32) == 0) return kOk;
  const int capc = std::min(M->hm.ring, 255);
  if ((rc = M->g_rbase.ensure((size_t)tn * 8 + 8))) return rc;
  unsigned long long *d_total = (unsigned long long *)M->g_rbase.p + tn;
  CUDA_TRY(cudaMemsetAsync(d_total, 0, 8, st));
  const unsigned grid = (unsigned)((tn + kThreads - 1) / kThreads);
  k_gen_retry_plan<<<grid, kThreads, 0, st>>>(M->mv, list, tn, (const int32_t *)M->g_nlen.p,
                                              (const uint32_t *)M->w_status.p, capc, (int64_t *)M->g_rbase.p, d_total);
  unsigned long long total = 0;
  CUDA_TRY(cudaMemcpyAsync(&total, d_total, 8, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  if ((rc = M->g_rscratch.ensure((size_t)total * 4 + 64))) return rc;
  uint32_t *d_stor = (uint32_t *)(d_scal + 2);
  CUDA_TRY(cudaMemsetAsync(d_stor + 1, 0, 4, st));
  k_gen_retry<<<grid, kThreads, 0, st>>>(M->mv, list, tn, (const int64_t *)M->g_noff.p, (const uint8_t *)M->g_norm.p,
                                         (const int32_t *)M->g_nlen.p, (const int64_t *)M->g_rbase.p, capc,
                                         (int32_t *)M->g_rscratch.p, (int32_t *)M->g_stage.p, (int32_t *)M->w_cnt.p,
                                         (uint32_t *)M->w_status.p, (int64_t *)M->w_srcoff.p, d_stor, gends);
  *launches += 2;
  CUDA_TRY(cudaGetLastError());
  return kOk;
}

// General tier over `gcount` sentences of the chunk (list: device array of
// chunk-relative indices, or nullptr for "all sentences 0..gcount-1").
// `timed`: this is the dominant kernel of a non-fast model (event pair recorded).
int run_general_pass(const spm_cuda_model *M, cudaStream_t st, const uint8_t *d_text, const int64_t *d_off, int64_t c0,
                     const int32_t *list, int64_t gcount, bool timed, double *launches) {
  if (gcount <= 0) return kOk;
  int rc;
  const int exp_ = M->hm.expansion;
  if ((rc = M->g_need.ensure(gcount * 4))) return rc;
  if ((rc = M->g_noff.ensure((gcount + 1) * 8))) return rc;
  if ((rc = M->g_nlen.ensure(gcount * 4))) return rc;
  if ((rc = M->g_words.ensure(gcount * 4))) return rc;
  if ((rc = M->g_woff.ensure((gcount + 1) * 8))) return rc;
  if ((rc = M->g_status.ensure(gcount * 4))) return rc;
  int64_t *d_tot = (int64_t *)M->w_scalars.p + 4;  // [4] norm bytes, [5] scratch words
  const unsigned grid = (unsigned)((gcount + kThreads - 1) / kThreads);
  k_gen_prepare<<<grid, kThreads, 0, st>>>(d_off, c0, list, nullptr, 0, gcount, exp_, (int32_t *)M->g_need.p);
  *launches += 1;
  if ((rc = run_scan(M, st, (const int32_t *)M->g_need.p, gcount, IdentityOp(), (int64_t *)M->g_noff.p, d_tot, 0,
                     launches)))
    return rc;
  k_publish<<<1, 32, 0, st>>>(nullptr, (const int64_t *)M->w_scalars.p, M->d_pin, -1);
  CUDA_TRY(cudaStreamSynchronize(st));
  const int64_t total_norm = M->h_pin[kPubScal + 4];
  if ((rc = M->g_norm.ensure((size_t)total_norm + 64))) return rc;
  if ((rc = M->g_stage.ensure((size_t)total_norm * 4 + 64))) return rc;
  k_gen_normalize<<<grid, kThreads, 0, st>>>(M->mv, d_text, d_off, c0, list, nullptr, 0, gcount,
                                             (const int32_t *)M->g_need.p, (const int64_t *)M->g_noff.p,
                                             (uint8_t *)M->g_norm.p, (int32_t *)M->g_nlen.p,
                                             (int32_t *)M->g_words.p, (uint32_t *)M->g_status.p);
  *launches += 1;
  if ((rc = run_scan(M, st, (const int32_t *)M->g_words.p, gcount, IdentityOp(), (int64_t *)M->g_woff.p, d_tot + 1,
                     0, launches)))
    return rc;
  k_publish<<<1, 32, 0, st>>>(nullptr, (const int64_t *)M->w_scalars.p, M->d_pin, -1);
  CUDA_TRY(cudaStreamSynchronize(st));
  const int64_t total_words = M->h_pin[kPubScal + 5];
  const int64_t kScratchBudgetWords = (int64_t)1 << 30;  // 4 GiB of Viterbi / agenda scratch per launch
  uint32_t *d_stor = (uint32_t *)((int64_t *)M->w_scalars.p + 2);  // OR of all status words of the call
  auto launch = [&](int64_t a, int64_t b, int64_t wbase) -> int {
    const int64_t tn = b - a;
    const unsigned g = (unsigned)((tn + kThreads - 1) / kThreads);
    if (timed && (rc = kev_record(M, st, 0))) return rc;
    k_gen_encode<<<g, kThreads, 0, st>>>(M->mv, list, nullptr, a, tn, (const int64_t *)M->g_noff.p + a,
                                         (const uint8_t *)M->g_norm.p, (const int32_t *)M->g_nlen.p + a,
                                         (const int64_t *)M->g_woff.p + a, wbase, (int32_t *)M->g_scratch.p,
                                         (int32_t *)M->g_stage.p, (const uint32_t *)M->g_status.p + a,
                                         (int32_t *)M->w_cnt.p, (uint32_t *)M->w_status.p,
                                         (int64_t *)M->w_srcoff.p, d_stor, nullptr);
    *launches += 1;
    if (timed && (rc = kev_record(M, st, 1))) return rc;
    CUDA_TRY(cudaGetLastError());
    return kOk;
  };
  if (total_words <= kScratchBudgetWords) {
    if ((rc = M->g_scratch.ensure((size_t)total_words * 4 + 64))) return rc;
    if (gcount > 16384) return launch(0, gcount, 0);
    if (timed && (rc = kev_record(M, st, 0))) return rc;
    k_gen_encode_block<<<(unsigned)gcount, 32, 0, st>>>(
        M->mv, list, gcount, (const int64_t *)M->g_noff.p, (const uint8_t *)M->g_norm.p,
        (const int32_t *)M->g_nlen.p, (const int32_t *)M->g_words.p, (const int64_t *)M->g_woff.p,
        (int32_t *)M->g_scratch.p, (int32_t *)M->g_stage.p, (const uint32_t *)M->g_status.p, (int32_t *)M->w_cnt.p,
        (uint32_t *)M->w_status.p, (int64_t *)M->w_srcoff.p, d_stor);
    *launches += 1;
    if (timed && (rc = kev_record(M, st, 1))) return rc;
    CUDA_TRY(cudaGetLastError());
    return kOk;
  }
  // rare: slice the list so each launch stays within the scratch budget
  M->h_woff.resize(gcount + 1);
  CUDA_TRY(cudaMemcpyAsync(M->h_woff.data(), M->g_woff.p, (gcount + 1) * 8, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  int64_t max_one = 0;
  for (int64_t t = 0; t < gcount; ++t) max_one = std::max(max_one, M->h_woff[t + 1] - M->h_woff[t]);
  if ((rc = M->g_scratch.ensure((size_t)std::max(max_one, kScratchBudgetWords) * 4 + 64))) return rc;
  for (int64_t a = 0; a < gcount;) {
    int64_t b = a + 1;
    while (b < gcount && M->h_woff[b + 1] - M->h_woff[a] <= kScratchBudgetWords) ++b;
    if ((rc = launch(a, b, M->h_woff[a]))) return rc;
    a = b;
  }
  return kOk;
}

int run_general(const spm_cuda_model *M, cudaStream_t st, const uint8_t *d_text, const int64_t *d_off, int64_t c0,
                const int32_t *list, int64_t gcount, bool timed, double *launches) {
  if (gcount <= 0) return kOk;
  int rc = run_general_pass(M, st, d_text, d_off, c0, list, gcount, timed, launches);
  if (rc) return rc;
  return run_capc_retry(M, st, list, gcount, (int64_t *)M->w_scalars.p, nullptr, launches);
}

}  // namespace

// ---------------------------------------------------------------- C ABI
extern "C" {

const char *spm_cuda_last_error(void) { return g_err.c_str(); }

}  // extern "C"
static int run_self_test(spm_cuda_model *M);   // spm_pieces.inc
extern "C" {

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
  {  // initial option values from the environment (A/B measurement scripts); read per load
    auto env = [](const char *k, int64_t *v) { const char *e = getenv(k); if (e && *e) *v = atoll(e); };
    if (getenv("SPM_CUDA_NO_MEMO")) M->opt.memo = 0;
    if (getenv("SPM_CUDA_NO_WARP")) M->opt.warp = 0;
    if (getenv("SPM_CUDA_SCAN_PER_SENTENCE")) M->opt.scan_per_sentence = 1;
    if (getenv("SPM_CUDA_PIECES_EXACT")) M->opt.pieces_exact = 1;
    env("SPM_CUDA_MEMO_BITS", &M->opt.memo_bits);
    env("SPM_CUDA_HOST_CHUNKS", &M->opt.host_chunks);
    env("SPM_CUDA_EMIT_BLOCKS", &M->opt.emit_blocks);
    env("SPM_CUDA_WARP_ROUTE", &M->opt.warp_route);
  }
  M->bos_id = piece_to_id(M->hm, M->hm.bos_piece);
  M->eos_id = piece_to_id(M->hm, M->hm.eos_piece);
  M->pad_id = piece_to_id(M->hm, M->hm.pad_piece);
  DeviceGuard guard(device);
  if (guard.err != cudaSuccess) { g_err = "cudaSetDevice failed"; delete M; return kInternal; }
  rc = upload_model(M);
  if (rc) { spm_cuda_model_free(M); return rc; }
  if (cudaHostAlloc((void **)&M->h_pin, (kPubTot + 4096) * 8, cudaHostAllocMapped) != cudaSuccess ||
      cudaHostGetDevicePointer((void **)&M->d_pin, M->h_pin, 0) != cudaSuccess) {
    g_err = "cudaHostAlloc failed";
    spm_cuda_model_free(M);
    return kInternal;
  }
  cudaEventCreate(&M->ev0);
  cudaEventCreate(&M->ev1);
  // self-test samples carried by the model (src/sentencepiece_processor.cc:134-154)
  rc = run_self_test(M);
  if (rc) { spm_cuda_model_free(M); return rc; }
  *out = M;
  return kOk;
}

void spm_cuda_model_free(spm_cuda_model *M) {
  if (!M) return;
  DeviceGuard guard(M->device);
  for (DevBuf *b : M->all_bufs()) b->release();
  if (M->s_in) cudaStreamDestroy(M->s_in);
  if (M->s_out) cudaStreamDestroy(M->s_out);
  for (cudaEvent_t e : M->ev_in) cudaEventDestroy(e);
  for (cudaEvent_t e : M->ev_done) cudaEventDestroy(e);
  if (M->ev_start) cudaEventDestroy(M->ev_start);
  if (M->ev_out) cudaEventDestroy(M->ev_out);
  if (M->ev_mid) cudaEventDestroy(M->ev_mid);
  if (M->h_pin) cudaFreeHost(M->h_pin);
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
  DEVICE_GUARD(M->device);
  CUDA_TRY(cudaMemcpy(M->d_leaf.p, M->hm.pieces.leaf.data(), M->hm.pieces.leaf.size() * 8, cudaMemcpyHostToDevice));
  M->mv.has_unused = M->hm.has_unused;
  return configure_fast(M);
}

static int64_t *option_slot(Options *o, const char *name) {
  const struct { const char *name; int64_t *p; } table[] = {
      {"memo", &o->memo}, {"memo_bits", &o->memo_bits}, {"host_chunks", &o->host_chunks},
      {"host_chunk_min", &o->host_chunk_min}, {"max_chunk", &o->max_chunk}, {"warp", &o->warp},
      {"scan_per_sentence", &o->scan_per_sentence}, {"warp_route", &o->warp_route}, {"fast_side", &o->fast_side}, {"mid_lines", &o->mid_lines},
      {"emit_blocks", &o->emit_blocks}, {"pieces_exact", &o->pieces_exact}, {"pieces_plain", &o->pieces_plain}, {"bpe_route", &o->bpe_route}};
  for (const auto &e : table)
    if (strcmp(e.name, name) == 0) return e.p;
  return nullptr;
}

int spm_cuda_model_set_option(spm_cuda_model *M, const char *name, int64_t value) {
  if (!M || !name) { g_err = "null argument"; return kInvalidArgument; }
  std::lock_guard<std::mutex> lk(M->mu);
  int64_t *slot = option_slot(&M->opt, name);
  if (!slot) { g_err = std::string("unknown option: ") + name; return kNotFound; }
  if (value < 0) { g_err = std::string("option value must be >= 0: ") + name; return kInvalidArgument; }
  *slot = value;
  return kOk;
}

int spm_cuda_model_get_option(const spm_cuda_model *M, const char *name, int64_t *value) {
  if (!M || !name || !value) { g_err = "null argument"; return kInvalidArgument; }
  std::lock_guard<std::mutex> lk(M->mu);
  const int64_t *slot = option_slot(const_cast<Options *>(&M->opt), name);
  if (!slot) { g_err = std::string("unknown option: ") + name; return kNotFound; }
  *value = *slot;
  return kOk;
}

// Frees the workspace of past calls (it grows to the largest batch seen and is otherwise kept
// for the next call); the model tables stay.
int spm_cuda_model_release_workspace(spm_cuda_model *M) {
  if (!M) { g_err = "model is null"; return kInvalidArgument; }
  std::lock_guard<std::mutex> lk(M->mu);
  DEVICE_GUARD(M->device);
  CUDA_TRY(cudaDeviceSynchronize());
  for (DevBuf *b : M->workspace_bufs()) b->release();
  M->lshist_zeroed = false;
  M->pt_ready = false;
  M->dec_ready = false;
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
  for (int i = 0; i < n && i < 16; ++i) stats[i] = M->stats[i];
  return kOk;
}

}  // extern "C"

// The body of spm_cuda_encode_batch; the caller holds M->mu (the pieces path runs it first
// for the ids and then adds the spans, spm_pieces.inc).
static int encode_batch_locked(const spm_cuda_model *M, const char *text, const int64_t *text_offsets, int64_t n,
                               uint32_t extra_options, int32_t *ids, int64_t ids_capacity, int64_t *id_offsets,
                               int memory_space, void *cuda_stream);

extern "C" {

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
  return encode_batch_locked(M, text, text_offsets, n, extra_options, ids, ids_capacity, id_offsets, memory_space,
                             cuda_stream);
}

}  // extern "C"

static int encode_batch_locked(const spm_cuda_model *M, const char *text, const int64_t *text_offsets, int64_t n,
                               uint32_t extra_options, int32_t *ids, int64_t ids_capacity, int64_t *id_offsets,
                               int memory_space, void *cuda_stream) {
  DEVICE_GUARD(M->device);
  cudaStream_t st = (cudaStream_t)cuda_stream;
  static const bool trace = getenv("SPM_CUDA_TRACE") != nullptr;   // host-side timeline on stderr
  const auto t_begin = std::chrono::steady_clock::now();
  auto tr = [&](const char *what, long k) {
    if (trace)
      fprintf(stderr, "[spm_cuda] %8.3f ms  %s %ld\n",
              std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count(), what, k);
  };
  const bool host = memory_space == SPM_CUDA_HOST;
  double launches = 0;
  int rc;

  // Offsets stay on the device; the host only ever reads a handful of them.
  const int64_t *d_off;
  const uint8_t *d_text;
  int64_t text_lo = 0, text_hi = 0;
  if (host) {
    text_lo = text_offsets[0]; text_hi = text_offsets[n];
  } else {
    CUDA_TRY(cudaMemcpyAsync(&text_lo, text_offsets, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(&text_hi, text_offsets + n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
  }
  if (text_hi < text_lo) { g_err = "text_offsets must be non-decreasing"; return kInvalidArgument; }
  if (host) {
    if ((rc = M->w_off.ensure((n + 1) * 8))) return rc;
    if ((rc = M->w_text.ensure((size_t)(text_hi - text_lo) + 16))) return rc;
    if ((rc = M->w_idoff.ensure((n + 1) * 8))) return rc;
    CUDA_TRY(cudaMemcpyAsync(M->w_off.p, text_offsets, (n + 1) * 8, cudaMemcpyHostToDevice, st));
    d_off = (const int64_t *)M->w_off.p;   // (the text follows chunk by chunk on the copy-in stream)
    d_text = (const uint8_t *)M->w_text.p - text_lo;  // indexed by absolute offsets
  } else {
    d_off = text_offsets;
    d_text = (const uint8_t *)text;
  }
  // one offset, read by the host (chunk planning only; a single-chunk call never needs it)
  auto read_off = [&](int64_t i, int64_t *v) -> int {
    if (i == 0) { *v = text_lo; return kOk; }
    if (i == n) { *v = text_hi; return kOk; }
    if (host) { *v = text_offsets[i]; return kOk; }
    CUDA_TRY(cudaMemcpyAsync(v, text_offsets + i, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return kOk;
  };
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
  if ((rc = M->w_counters.ensure((C_COUNT + 4) * 4))) return rc;   // + per-call counters (not reset per chunk)
  int64_t *d_carry = (int64_t *)M->w_scalars.p;        // [0] running CSR total
  uint32_t *d_counters = (uint32_t *)M->w_counters.p;
  CUDA_TRY(cudaMemsetAsync(M->w_scalars.p, 0, 64, st));
  CUDA_TRY(cudaMemsetAsync(d_counters + C_COUNT, 0, 16, st));
  int64_t max_len = 0, max_raw = 0, max_long = 0;
  if (n > 0) {
    const unsigned vgrid = (unsigned)std::min<int64_t>((n + 255) / 256, 4 * (int64_t)M->sm_count);
    k_validate<<<vgrid, 256, 0, st>>>(d_off, n, (unsigned long long *)M->w_scalars.p);
    launches += 1;
    k_publish<<<1, 32, 0, st>>>(nullptr, (const int64_t *)M->w_scalars.p, M->d_pin, -1);
    CUDA_TRY(cudaStreamSynchronize(st));
    const int64_t *hv = M->h_pin + kPubScal;
    tr("validated", n);
    if (hv[3]) { g_err = "text_offsets must be non-decreasing"; return kInvalidArgument; }
    max_len = hv[6]; max_raw = hv[7]; max_long = hv[1];
    if (max_len > ((int64_t)1 << 26)) { g_err = "sentence longer than 64 MiB"; return kOutOfRange; }
  }
  if (M->fast && M->w_hist.cap == 0) {
    if ((rc = M->w_hist.ensure((size_t)kFastMaxRaw * 8))) return rc;   // hist | bincur
    CUDA_TRY(cudaMemsetAsync(M->w_hist.p, 0, (size_t)kFastMaxRaw * 8, st));
  }
  uint32_t *d_hist = (uint32_t *)M->w_hist.p, *d_bincur = d_hist ? d_hist + kFastMaxRaw : nullptr;
  M->nk = 0;
  CUDA_TRY(cudaEventRecord(M->ev0, st));

  const ExtraPlan plan = make_extra_plan(extra_options, M->bos_id, M->eos_id);
  const int exp_ = M->hm.expansion;
  // word memo (spm_memo.cuh): one table per call, zeroed here -- nothing survives the call.
  // Option "memo" = 0 switches the tier off (the Viterbi / merge kernels alone).
  const bool memo = M->fast && M->hm.memo_ok && M->opt.memo != 0 && n > 0;
  MemoTable memo_tb;
  memset(&memo_tb, 0, sizeof(memo_tb));
  if (memo) {
    const uint64_t want = 2 * (uint64_t)n + (uint64_t)(text_hi - text_lo) / 16;
    uint64_t entries = 1 << 12;
    const int max_bits = (int)std::max<int64_t>(12, std::min<int64_t>(26, M->opt.memo_bits));
    while (entries < want && entries < (1ull << max_bits)) entries <<= 1;
    if ((rc = M->w_memo.ensure(entries * sizeof(MemoKey)))) return rc;
    if ((rc = M->w_mres.ensure(entries * sizeof(MemoRes)))) return rc;
    if ((rc = M->w_newlist.ensure(entries / 2 * 4))) return rc;
    // keys: empty.  The result records are not zeroed (256 MB per call): a slot is only ever read after it
    // was claimed in this call, and every claimed slot gets its meta word from the words kernel or -- beyond
    // max_new -- from the claimer itself (meta 0 = "not encoded", never served).
    CUDA_TRY(cudaMemsetAsync(M->w_memo.p, 0, entries * sizeof(MemoKey), st));
    memo_tb.keys = (MemoKey *)M->w_memo.p;
    memo_tb.res = (MemoRes *)M->w_mres.p;
    memo_tb.mask = (uint32_t)(entries - 1);
    memo_tb.new_list = (uint32_t *)M->w_newlist.p;
    memo_tb.max_new = (uint32_t)(entries / 2);
    if ((rc = M->w_movf.ensure((size_t)memo_tb.max_new * kMemoOverflowIds * 4))) return rc;
    memo_tb.overflow = (int32_t *)M->w_movf.p;
    memo_tb.n_ovf = d_counters + C_COUNT;
    if (M->fast_kind == 2) {
      if ((rc = M->w_wtext.ensure((size_t)memo_tb.max_new * 64))) return rc;
      if ((rc = M->w_wlen.ensure((size_t)memo_tb.max_new * 4))) return rc;
    }
    launches += 1;
  }
  // chunk = as many sentences as fit the byte budget (fast tier: the raw text; general
  // tier: the worst-case normalized text) -- one chunk for all but huge batches
  const int64_t kRawBudget = M->fast ? (2048ll << 20) : (256ll << 20) / exp_;
  const int64_t kMaxChunk = M->opt.max_chunk > 0 ? M->opt.max_chunk : (M->fast ? (1 << 24) : (1 << 20));
  uint32_t stor = 0;
  int64_t n_general = 0, n_viterbi = 0, n_memo_words = 0;
  struct Chunk { int64_t c0, c1, o0, o1; };
  std::vector<Chunk> chunks;
  {
    // host-memory calls of some size are cut into equal chunks so that copy-in, kernels and copy-out
    // pipeline.  The call is bound by PCIe (profiles/r1_v23_summary.md: the kernels are done long
    // before the last ids have left), so what counts is how early the copy-out can start and how
    // little is left when the copy-in ends -- against the fixed cost of a chunk (the leftover
    // tiers' latency, ~30 launches and a host round trip: ~0.8 ms; BPE, whose merge kernel takes
    // ~1.7 ms over a chunk's leftover sentences whatever the chunk holds: ~2.5 ms).  Measured per
    // 10 M sentences: unigram 8 / 12 / 16 / 24 chunks 28.0 / 26.9 / 26.7 / 29.1 ms (text with many
    // rule hits: 12: 27.7, 16: 31.0), chunks tapered at both ends 27.2; BPE 4 / 5 / 6 / 7: 33.2 /
    // 33.0 / 34.4 / 36.4.  Options "host_chunks" / "host_chunk_min" override count and threshold.
    int64_t want = kMaxChunk;
    if (host && n >= std::max<int64_t>(1, M->opt.host_chunk_min)) {
      const int64_t parts = std::max<int64_t>(
          1, std::min<int64_t>(64, M->opt.host_chunks > 0 ? M->opt.host_chunks : (M->fast_kind == 2 ? 8 : 12)));
      want = std::max<int64_t>(1, std::min(kMaxChunk, (n + parts - 1) / parts));
    }
    int64_t c0 = 0, off_c0 = text_lo;
    while (c0 < n) {
      int64_t c1 = std::min(n, c0 + want), off_c1 = 0;
      if ((rc = read_off(c1, &off_c1))) return rc;
      while (off_c1 - off_c0 > kRawBudget && c1 - c0 > 1) {
        c1 = c0 + (c1 - c0) / 2;
        if ((rc = read_off(c1, &off_c1))) return rc;
      }
      chunks.push_back(Chunk{c0, c1, off_c0, off_c1});
      c0 = c1;
      off_c0 = off_c1;
    }
  }
  const size_t nch = chunks.size();
  if (host && nch > 0) {
    if (!M->s_in) {
      CUDA_TRY(cudaStreamCreateWithFlags(&M->s_in, cudaStreamNonBlocking));
      CUDA_TRY(cudaStreamCreateWithFlags(&M->s_out, cudaStreamNonBlocking));
      CUDA_TRY(cudaEventCreateWithFlags(&M->ev_start, cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&M->ev_out, cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&M->ev_mid, cudaEventDisableTiming));
    }
    if (nch + 1 > 4096) { g_err = "too many chunks"; return kOutOfRange; }
    while (M->ev_in.size() < nch) {
      cudaEvent_t a, b2;
      CUDA_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&b2, cudaEventDisableTiming));
      M->ev_in.push_back(a);
      M->ev_done.push_back(b2);
    }
    // all text copies are queued now; chunk k's kernels wait for ev_in[k] only
    CUDA_TRY(cudaEventRecord(M->ev_start, st));
    CUDA_TRY(cudaStreamWaitEvent(M->s_in, M->ev_start, 0));
    CUDA_TRY(cudaStreamWaitEvent(M->s_out, M->ev_start, 0));
    for (size_t k = 0; k < nch; ++k) {
      const Chunk &ch = chunks[k];
      if (ch.o1 > ch.o0)
        CUDA_TRY(cudaMemcpyAsync((uint8_t *)M->w_text.p + (ch.o0 - text_lo), text + ch.o0, ch.o1 - ch.o0,
                                 cudaMemcpyHostToDevice, M->s_in));
      CUDA_TRY(cudaEventRecord(M->ev_in[k], M->s_in));
    }
  }
  // copy-out of finished chunks (host mode): ids and offsets of chunk k leave on s_out as
  // soon as its running total has reached the host, while later chunks compute
  size_t flushed = 0;
  int64_t ids_done = 0;
  if (host)
    for (size_t k = 0; k < nch; ++k) M->h_pin[kPubTot + k] = -1;   // "total of chunk k not published yet"
  auto flush_outputs = [&](size_t upto) -> int {   // chunks [flushed, upto) have their totals in h_pin
    for (; flushed < upto; ++flushed) {
      const Chunk &ch = chunks[flushed];
      const int64_t tot = M->h_pin[kPubTot + flushed];
      CUDA_TRY(cudaStreamWaitEvent(M->s_out, M->ev_done[flushed], 0));
      CUDA_TRY(cudaMemcpyAsync(id_offsets + ch.c0, d_idoff + ch.c0, (ch.c1 - ch.c0 + 1) * 8, cudaMemcpyDeviceToHost,
                               M->s_out));
      const int64_t hi = std::min(tot, ids_capacity);
      if (hi > ids_done)
        CUDA_TRY(cudaMemcpyAsync(ids + ids_done, d_ids + ids_done, (hi - ids_done) * 4, cudaMemcpyDeviceToHost,
                                 M->s_out));
      ids_done = std::max(ids_done, hi);
    }
    return kOk;
  };
  for (size_t kc = 0; kc < nch; ++kc) {
    const int64_t c0 = chunks[kc].c0, c1 = chunks[kc].c1, off_c0 = chunks[kc].o0, off_c1 = chunks[kc].o1;
    if (host) CUDA_TRY(cudaStreamWaitEvent(st, M->ev_in[kc], 0));
    const int64_t cn = c1 - c0, raw = off_c1 - off_c0;
    if ((rc = M->w_cnt.ensure(cn * 4))) return rc;
    if ((rc = M->w_status.ensure(cn * 4))) return rc;
    if ((rc = M->w_srcoff.ensure(cn * 8))) return rc;
    CUDA_TRY(cudaMemsetAsync(d_counters, 0, C_COUNT * 4, st));
    int32_t *d_vstage = nullptr;
    if (M->fast) {
      if ((rc = M->w_norm.ensure((size_t)(3 * raw + 160 * cn + 256)))) return rc;
      if ((rc = M->w_stage.ensure((size_t)(raw + kFastExtraChars * cn + 64) * 4))) return rc;
      // with the word memo in front, `stage` holds the packed ids of whole groups (k_memo_emit); the
      // sentences it leaves to the Viterbi / merge kernels stage into their own slots of a second buffer
      if (memo && (rc = M->w_vstage.ensure((size_t)(raw + kFastExtraChars * cn + 64) * 4))) return rc;
      if (memo && (rc = M->w_spos.ensure(cn * 4))) return rc;
      d_vstage = (int32_t *)(memo ? M->w_vstage.p : M->w_stage.p);
      if ((rc = M->w_nlen.ensure(cn * 4))) return rc;
      if ((rc = M->w_perm.ensure(cn * 4))) return rc;
      if ((rc = M->w_genlist.ensure(cn * 4))) return rc;
      // back-pointer rows: one per character (<= raw + kFastExtraChars) + BOS/EOS rows
      // (+ the side area of fast_unigram's near-tied candidates, when the model's ring fits its bit mask)
      const int side_words = (M->fast_kind == 1 && M->hm.max_piece_chars + 1 <= 31 && M->opt.fast_side != 0) ? kFastSideWords : 0;
      const int64_t back_rows = ((max_raw + kFastExtraChars + 8 + 3) & ~(int64_t)3) + side_words;
      // long lines (kFastMaxRaw <= bytes < kLongMaxRaw): same kernels over long_list, on as
      // many warps per SM as a 2 GiB back-pointer budget allows
      const bool has_long = max_len >= kFastMaxRaw && max_long >= kFastMaxRaw;
      const bool warp_ok = M->fast_kind == 1 && M->opt.warp != 0 && M->hm.ring <= kWarpRing &&
                           M->hm.max_piece_chars + 1 <= kWarpHits && M->hm.piece.size() < (1u << 24);
      const int64_t back_rows_long = (max_long + kFastExtraChars + 8 + 3) & ~(int64_t)3;
      int long_warps = M->fast_kind == 1 ? kFastThreads / 32 : kBpeThreads / 32;
      if (M->fast_kind == 1)
        long_warps = (int)std::max<int64_t>(
            1, std::min<int64_t>(long_warps, (2048ll << 20) / (back_rows_long * 4 * 32 * M->sm_count)));
      // k_encode_fast hands a list of up to warp_route sentences to k_encode_warp as a whole: a chunk
      // that small never needs the per-thread back-pointer rows (2.4 GB for one 4 KB line otherwise)
      const uint32_t warp_route = (uint32_t)std::max<int64_t>(0, std::min<int64_t>(M->opt.warp_route, 0x7FFFFFFF));
      const bool fast_may_run = !(warp_ok && warp_route != 0 && (uint64_t)cn <= warp_route);
      size_t back_bytes = fast_may_run ? (size_t)back_rows * M->sm_count * kFastThreads * 4 : 0;
      // Option "mid_lines": lines of 4..8 KiB normalized in pieces like the long ones, then a second pass of
      // k_encode_fast (one thread per line is ~6 x cheaper per byte than the warp kernel; needs the side area: a
      // fifth of these lines meets a near tie).  OFF: 2.2 M Japanese lines hold 58 k such lines -- 12 warps per SM,
      // every one a chain of 8 KB of dependent trie steps and scattered back-record stores: 33.8 ms for the
      // Viterbi stage against 28.2 ms with these lines in the warp kernel.
      const bool use_mid = warp_ok && side_words != 0 && fast_may_run && has_long && M->opt.mid_lines != 0;
      const int64_t back_rows_mid =
          ((std::min<int64_t>(max_long, kMidMaxRaw - 1) + kFastExtraChars + 8 + 3) & ~(int64_t)3) + side_words;
      if (use_mid) {
        back_bytes = std::max(back_bytes, (size_t)back_rows_mid * M->sm_count * kFastThreads * 4);
        if ((rc = M->w_midlist.ensure((size_t)std::min<int64_t>(cn, raw / kFastMaxRaw + 1) * 4))) return rc;
      }
      if (has_long && !warp_ok)
        back_bytes = std::max(back_bytes, (size_t)back_rows_long * M->sm_count * long_warps * 32 * 4);
      if (M->fast_kind == 1 && back_bytes > 0 && (rc = M->w_back.ensure(back_bytes))) return rc;
      // (with k_encode_warp the long-line list also takes the sentences k_encode_fast flags)
      const int64_t long_cap = warp_ok ? cn : std::min<int64_t>(cn, raw / kFastMaxRaw + 1);
      if ((has_long || warp_ok) && (rc = M->w_vitl.ensure((size_t)long_cap * 4))) return rc;
      if ((has_long || warp_ok) && (rc = M->w_longlist.ensure((size_t)long_cap * 4))) return rc;
      const int64_t max_n_long = has_long ? std::min<int64_t>(cn, raw / kFastMaxRaw + 1) : 0;
      const unsigned gsort = (unsigned)((cn + kSortTile - 1) / kSortTile);
      const unsigned gsent = (unsigned)((cn + 255) / 256), glong = (unsigned)((max_n_long + 255) / 256);
      static_assert(kMemoThreads == 256 && kScanWarps * 32 == 256, "grids assume 256 threads");
      // Word memo first (spm_memo.cuh): what it cannot serve is listed in w_vit and takes the
      // sort -> normalize -> Viterbi / merge path below, like every sentence does without the memo.
      const int32_t *todo = nullptr;     // nullptr: all sentences of the chunk
      if (memo) {
        // lower half: the words of 32 sentences side by side (k_memo_scan_words); upper half: one list per sentence
        const uint32_t wslot_half = (uint32_t)(raw / 2 + 4 * cn + 64);
        if ((rc = M->w_wslots.ensure((size_t)wslot_half * 2 * 4))) return rc;
        if ((rc = M->w_wbase.ensure(cn * 4))) return rc;
        if ((rc = M->w_nwords.ensure(cn * 4))) return rc;
        if ((rc = M->w_vit.ensure(cn * 4))) return rc;
        MemoTable tb = memo_tb;
        tb.n_new = d_counters + C_NNEW;
        const unsigned gnew = (tb.max_new + 255) / 256;
        if ((rc = kev_record(M, st, 0, T_MEMO_COLLECT))) return rc;
        if (M->opt.scan_per_sentence != 0)   // the older thread-per-sentence scan
          k_memo_scan_plain<<<gsent, kScanWarps * 32, 0, st>>>(
              M->mv, tb, d_text, d_off, c0, cn, d_text + text_lo, d_text + text_hi, (uint32_t *)M->w_wslots.p,
              (uint32_t *)M->w_wbase.p, wslot_half, (int32_t *)M->w_nwords.p, (int32_t *)M->w_cnt.p, (uint32_t *)M->w_status.p, (int32_t *)M->w_vit.p,
              d_counters);
        else
          k_memo_scan_words<<<(unsigned)((cn + kSwWarps * 32 - 1) / (kSwWarps * 32)), kSwWarps * 32, 0, st>>>(
              M->mv, tb, d_text, d_off, c0, cn, d_text + text_lo, d_text + text_hi, (uint32_t *)M->w_wslots.p,
              (uint32_t *)M->w_wbase.p, wslot_half, (int32_t *)M->w_nwords.p, (int32_t *)M->w_cnt.p, (uint32_t *)M->w_status.p, (int32_t *)M->w_vit.p,
              d_counters);
        // the listed sentences by descending length (w_perm is free until the leftover tier sorts into it)
        if ((rc = M->w_lshist.ensure(2 * kLsBins * 4))) return rc;
        if (!M->lshist_zeroed) { CUDA_TRY(cudaMemsetAsync(M->w_lshist.p, 0, 2 * kLsBins * 4, st)); M->lshist_zeroed = true; }
        const unsigned gls = (unsigned)std::min<int64_t>((cn + kLsThreads - 1) / kLsThreads, 2 * (int64_t)M->sm_count);
        k_ls_hist<<<gls, kLsThreads, 0, st>>>(d_off, c0, (const int32_t *)M->w_vit.p, d_counters, C_NSLOW,
                                             (uint32_t *)M->w_lshist.p);
        k_ls_scan<<<1, kLsBins, 0, st>>>((uint32_t *)M->w_lshist.p, (uint32_t *)M->w_lshist.p + kLsBins);
        k_ls_scatter<<<gls, kLsThreads, 0, st>>>(d_off, c0, (const int32_t *)M->w_vit.p, d_counters, C_NSLOW,
                                                (uint32_t *)M->w_lshist.p + kLsBins, (int32_t *)M->w_perm.p);
        launches += 3;
        k_memo_scan<<<(unsigned)((cn + kMemoScanThreads - 1) / kMemoScanThreads), kMemoScanThreads, 0, st>>>(
            M->mv, tb, d_text, d_off, c0, d_text + text_lo, d_text + text_hi, (const int32_t *)M->w_perm.p,
            (uint32_t *)M->w_wslots.p + wslot_half, (int32_t *)M->w_nwords.p, d_counters);
        launches += 1;
        if ((rc = kev_record(M, st, 1))) return rc;
        if ((rc = kev_record(M, st, 0, T_MEMO_WORDS))) return rc;
        if (M->fast_kind == 1) {
          k_memo_words_unigram<<<gnew, kMemoThreads, 0, st>>>(M->fp, M->hm.neg_only ? 1 : 0, tb);
          launches += 1;
        } else {
          k_memo_words_text<<<gnew, kMemoThreads, 0, st>>>(tb, (uint8_t *)M->w_wtext.p, (int32_t *)M->w_wlen.p);
          k_memo_words_bpe<<<(tb.max_new + kMemoBpeThreads - 1) / kMemoBpeThreads, kMemoBpeThreads, 0, st>>>(
              M->fp, tb, (const uint8_t *)M->w_wtext.p, (const int32_t *)M->w_wlen.p);
          launches += 2;
        }
        if ((rc = kev_record(M, st, 1))) return rc;
        if ((rc = kev_record(M, st, 0, T_MEMO_EMIT))) return rc;
        // grid-stride kernel: exactly as many CTAs as are resident at once (6 per SM at 40 registers), so
        // that the sentences are dealt out in ONE wave -- with 8 per SM the last quarter of the CTAs ran
        // alone at a third of the machine
        if (M->emit_blocks == 0) {
          int occ = 0;
          CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_memo_emit, kEmitWarps * 32, 0));
          M->emit_blocks = std::max(1, occ);
        }
        const int64_t emit_per_sm = M->opt.emit_blocks > 0 ? M->opt.emit_blocks : M->emit_blocks;
        const unsigned gemit = (unsigned)std::min<int64_t>((cn + 32 * kEmitWarps - 1) / (32 * kEmitWarps),
                                                           emit_per_sm * (int64_t)M->sm_count);
        k_memo_emit<<<gemit, kEmitWarps * 32, 0, st>>>(
            tb, M->fast_kind == 1, d_off, c0, cn, (const uint32_t *)M->w_wslots.p, (const uint32_t *)M->w_wbase.p,
            (const int32_t *)M->w_nwords.p,
            (int32_t *)M->w_stage.p, (int32_t *)M->w_cnt.p, (uint32_t *)M->w_status.p, (int32_t *)M->w_spos.p,
            (int32_t *)M->w_vit.p, d_counters, C_NVIT);
        if ((rc = kev_record(M, st, 1))) return rc;
        launches += 2;
        todo = (const int32_t *)M->w_vit.p;
      }
      k_hist<<<gsort, kSortThreads, 0, st>>>(d_off, c0, cn, todo, C_NVIT, (int32_t *)M->w_cnt.p,
                                   (uint32_t *)M->w_status.p, d_hist, (int32_t *)M->w_genlist.p,
                                   (int32_t *)M->w_longlist.p, use_mid ? (int32_t *)M->w_midlist.p : nullptr, d_counters);
      k_binscan<<<1, 1024, 0, st>>>(d_hist, d_bincur, d_counters);
      k_scatter<<<gsort, kSortThreads, 0, st>>>(d_off, c0, cn, todo, C_NVIT, d_bincur, (int32_t *)M->w_perm.p,
                                      d_counters);
      if ((rc = kev_record(M, st, 0, T_NORMALIZE))) return rc;
      k_normalize_fast<<<(unsigned)std::min<int64_t>((cn + kNormThreads - 1) / kNormThreads, (int64_t)M->sm_count * SPM_NORM_BLOCKS), kNormThreads, 0, st>>>(
          M->mv, d_text, d_off, c0, d_text + text_lo, d_text + text_hi, (const int32_t *)M->w_perm.p, C_NFAST,
          (uint8_t *)M->w_norm.p, (int32_t *)M->w_nlen.p, (int32_t *)M->w_genlist.p, d_counters);
      if (has_long) {   // one block per long line; the grid covers the most long lines the chunk's bytes can hold
        if ((rc = M->w_ntmp.ensure((size_t)(3 * raw + 88 * (raw / kNormChunk + 1) + 96 * cn + 256)))) return rc;
        k_normalize_long<<<(unsigned)max_n_long, kNormLongThreads, 0, st>>>(
            M->mv, d_text, d_off, c0, d_text + text_lo, d_text + text_hi, (const int32_t *)M->w_longlist.p, C_NLONG,
            (uint8_t *)M->w_ntmp.p, (uint8_t *)M->w_norm.p, (int32_t *)M->w_nlen.p, (int32_t *)M->w_genlist.p,
            d_counters);
        if (use_mid) {
          k_normalize_long<<<(unsigned)max_n_long, kNormLongThreads, 0, st>>>(
              M->mv, d_text, d_off, c0, d_text + text_lo, d_text + text_hi, (const int32_t *)M->w_midlist.p, C_NMID,
              (uint8_t *)M->w_ntmp.p, (uint8_t *)M->w_norm.p, (int32_t *)M->w_nlen.p, (int32_t *)M->w_genlist.p,
              d_counters);
          launches += 1;
        }
      }
      if ((rc = kev_record(M, st, 1))) return rc;
      launches += has_long ? 5 : 4;
      const int32_t *list_a = (const int32_t *)M->w_perm.p, *list_b = (const int32_t *)M->w_longlist.p;
      const int cnt_a = C_NFAST, work_a = C_WORK, cnt_b = C_NLONG, work_b = C_WORK_LONG;
      if ((rc = kev_record(M, st, 0, T_ENCODE))) return rc;
      if (M->fast_kind == 1) {
        // sentences the one-candidate kernel flags (near ties) are appended to the long-line list
        // when k_encode_warp follows: it keeps kWarpCapc candidates per position itself
        k_encode_fast<<<M->sm_count, kFastThreads, M->fast_smem, st>>>(
            M->fp, d_off, c0, (const uint8_t *)M->w_norm.p, (const int32_t *)M->w_nlen.p, list_a, cnt_a, work_a,
            kFastThreads / 32, (uint32_t *)M->w_back.p, back_rows, d_vstage, (int32_t *)M->w_cnt.p,
            (uint32_t *)M->w_status.p, (int32_t *)(warp_ok ? M->w_longlist.p : M->w_genlist.p),
            warp_ok ? C_NLONG : C_NGEN, d_counters, warp_ok ? warp_route : 0u, C_NFAST, side_words);
        if (use_mid) {
          k_encode_fast<<<M->sm_count, kFastThreads, M->fast_smem, st>>>(
              M->fp, d_off, c0, (const uint8_t *)M->w_norm.p, (const int32_t *)M->w_nlen.p,
              (const int32_t *)M->w_midlist.p, C_NMID, C_WORK_MID, kFastThreads / 32, (uint32_t *)M->w_back.p,
              back_rows_mid, d_vstage, (int32_t *)M->w_cnt.p, (uint32_t *)M->w_status.p, (int32_t *)M->w_longlist.p,
              C_NLONG, d_counters, warp_route, C_NFAST, side_words);
          launches += 1;
        }
        if (warp_ok) {
          // one warp per line, exact candidate lists (spm_warp.cuh); back records: kWarpCapc rows of
          // `wstride` per resident warp, at most 8 GiB in all (the kernel is a chain of dependent shared-memory steps per
          // character: what it needs is resident warps, and a 33 KB line costs each 1.6 MB of records); longest lines first
          const int64_t wstride = 3 * max_long + 136;
          int64_t wblocks = std::max<int64_t>(
              1, std::min<int64_t>(8 * (int64_t)M->sm_count, (8192ll << 20) / (wstride * 4 * kWarpCapc * kWarpWarps)));
          wblocks = std::min<int64_t>(wblocks, (cn + kWarpWarps - 1) / kWarpWarps);   // a warp per listed line at most
          if ((rc = M->w_wback.ensure((size_t)(wblocks * kWarpWarps * kWarpCapc * wstride * 4)))) return rc;
          k_sort_long<<<1, 1024, 0, st>>>(d_off, c0, (const int32_t *)M->w_longlist.p, C_NLONG,
                                          (int32_t *)M->w_vitl.p, d_counters);
          if (M->hm.max_piece_chars + 1 <= kWarpHitsSmall)
            k_encode_warp<kWarpHitsSmall><<<(unsigned)wblocks, kWarpWarps * 32, 0, st>>>(
                M->fp, d_off, c0, (const uint8_t *)M->w_norm.p, (const int32_t *)M->w_nlen.p,
                (const int32_t *)M->w_vitl.p, cnt_b, work_b, (uint32_t *)M->w_wback.p, wstride, d_vstage,
                (int32_t *)M->w_cnt.p, (int32_t *)M->w_genlist.p, d_counters);
          else
            k_encode_warp<kWarpHits><<<(unsigned)wblocks, kWarpWarps * 32, 0, st>>>(
                M->fp, d_off, c0, (const uint8_t *)M->w_norm.p, (const int32_t *)M->w_nlen.p,
                (const int32_t *)M->w_vitl.p, cnt_b, work_b, (uint32_t *)M->w_wback.p, wstride, d_vstage,
                (int32_t *)M->w_cnt.p, (int32_t *)M->w_genlist.p, d_counters);
          launches += 2;
        } else if (has_long) {
          k_encode_fast<<<M->sm_count, kFastThreads, M->fast_smem, st>>>(
              M->fp, d_off, c0, (const uint8_t *)M->w_norm.p, (const int32_t *)M->w_nlen.p, list_b, cnt_b, work_b,
              long_warps, (uint32_t *)M->w_back.p, back_rows_long, d_vstage, (int32_t *)M->w_cnt.p,
              (uint32_t *)M->w_status.p, (int32_t *)M->w_genlist.p, C_NGEN, d_counters, 0u, cnt_b, 0);
          launches += 1;
        }
      } else {
        k_encode_bpe_fast<<<M->sm_count, kBpeThreads, M->fast_smem, st>>>(
            M->fp, d_off, c0, (const uint8_t *)M->w_norm.p, (const int32_t *)M->w_nlen.p, list_a, cnt_a, work_a,
            kBpeThreads / 32, d_vstage, (int32_t *)M->w_cnt.p, (uint32_t *)M->w_status.p,
            (int32_t *)M->w_genlist.p, d_counters,
            (uint32_t)std::max<int64_t>(0, std::min<int64_t>(M->opt.bpe_route, 0x7FFFFFFF)));
        if (has_long)
          k_encode_bpe_fast<<<M->sm_count, kBpeThreads, M->fast_smem, st>>>(
              M->fp, d_off, c0, (const uint8_t *)M->w_norm.p, (const int32_t *)M->w_nlen.p, list_b, cnt_b, work_b,
              long_warps, d_vstage, (int32_t *)M->w_cnt.p, (uint32_t *)M->w_status.p,
              (int32_t *)M->w_genlist.p, d_counters, 0u);
        launches += has_long ? 1 : 0;
      }
      if ((rc = kev_record(M, st, 1))) return rc;
      launches += 1;
      CUDA_TRY(cudaGetLastError());
      k_publish<<<1, 32, 0, st>>>(d_counters, (const int64_t *)M->w_scalars.p, M->d_pin, -1);
      if (host) {
        // While this chunk waits for its text, the chunk before it finishes: its ids leave the
        // moment its total shows up in the mapped snapshot, not when this chunk is half done.
        CUDA_TRY(cudaEventRecord(M->ev_mid, st));
        for (;;) {
          const cudaError_t q = cudaEventQuery(M->ev_mid);
          if (q == cudaSuccess) break;
          if (q != cudaErrorNotReady) CUDA_TRY(q);
          if (flushed < kc && ((volatile int64_t *)M->h_pin)[kPubTot + flushed] >= 0 && (rc = flush_outputs(flushed + 1)))
            return rc;
        }
      } else {
        CUDA_TRY(cudaStreamSynchronize(st));
      }
      int64_t hc[C_COUNT];
      for (int k = 0; k < C_COUNT; ++k) hc[k] = M->h_pin[k];
      tr("fast kernel done, chunk", (long)kc);
      if (host && (rc = flush_outputs(kc))) return rc;   // earlier chunks are complete by now
      n_general += hc[C_NGEN];
      n_viterbi += hc[C_NFAST] + hc[C_NMID] + hc[C_NLONG] - hc[C_NRETRY];
      n_memo_words += hc[C_NNEW];
      if ((rc = run_general(M, st, d_text, d_off, c0, (const int32_t *)M->w_genlist.p, hc[C_NGEN], false,
                            &launches)))
        return rc;
    } else {
      n_general += cn;
      if ((rc = run_general(M, st, d_text, d_off, c0, nullptr, cn, true, &launches))) return rc;
    }
    if ((rc = kev_record(M, st, 0, T_COMPACT))) return rc;
    FinalCountOp fop;
    fop.extra = plan.npre + plan.nsuf;
    if ((rc = run_scan(M, st, (const int32_t *)M->w_cnt.p, cn, fop, d_idoff + c0, d_carry, 1, &launches))) return rc;
    const unsigned cgrid = (unsigned)((cn + 255) / 256);  // 8 warps x 32 sentences per block
    k_compact<<<cgrid, 256, 0, st>>>(plan, d_off, c0, cn, (const int32_t *)M->w_stage.p, d_vstage,
                                     (const int32_t *)M->g_stage.p, (const int64_t *)M->w_srcoff.p,
                                     (const uint32_t *)M->w_status.p, (const int32_t *)M->w_cnt.p,
                                     (const int32_t *)M->w_spos.p, d_idoff, d_ids, d_cap);
    if ((rc = kev_record(M, st, 1))) return rc;
    launches += 1;
    CUDA_TRY(cudaGetLastError());
    tr("chunk enqueued", (long)kc);
    if (host) {
      k_publish<<<1, 32, 0, st>>>(nullptr, (const int64_t *)M->w_scalars.p, M->d_pin, (int)kc);
      CUDA_TRY(cudaEventRecord(M->ev_done[kc], st));
    }
  }
  CUDA_TRY(cudaEventRecord(M->ev1, st));
  k_publish<<<1, 32, 0, st>>>(nullptr, (const int64_t *)M->w_scalars.p, M->d_pin, -1);
  CUDA_TRY(cudaStreamSynchronize(st));
  const int64_t hsc[3] = {M->h_pin[kPubScal], M->h_pin[kPubScal + 1], M->h_pin[kPubScal + 2]};
  const int64_t total = hsc[0];
  stor = (uint32_t)hsc[2];
  M->last_stor = stor;
  if (n == 0) {
    const int64_t z = 0;
    if (host) id_offsets[0] = 0;
    else CUDA_TRY(cudaMemcpyAsync(id_offsets, &z, 8, cudaMemcpyHostToDevice, st));
  }
  if (host && n > 0) {
    if ((rc = flush_outputs(nch))) return rc;      // st is synchronized: every total is on the host
    CUDA_TRY(cudaEventRecord(M->ev_out, M->s_out));
    CUDA_TRY(cudaStreamWaitEvent(st, M->ev_out, 0));
  }
  tr("kernels done", 0);
  CUDA_TRY(cudaStreamSynchronize(st));
  tr("all done", 0);
  float ms = 0;
  cudaEventElapsedTime(&ms, M->ev0, M->ev1);
  M->stats[0] = launches;
  M->stats[1] = (double)n_general;
  M->stats[2] = ms;
  double tms[T_COUNT] = {0, 0, 0, 0, 0, 0};
  int tn[T_COUNT] = {0, 0, 0, 0, 0, 0};
  for (size_t k = 0; k < M->nk; ++k) {
    float t = 0;
    cudaEventElapsedTime(&t, M->kev[2 * k], M->kev[2 * k + 1]);
    tms[M->kev_tag[k]] += t;
    ++tn[M->kev_tag[k]];
  }
  int dom = 0;
  for (int k = 1; k < T_COUNT; ++k) if (tms[k] > tms[dom]) dom = k;
  M->stats[4] = tms[dom];            // the dominant stage of this call
  M->stats[5] = (double)tn[dom];
  M->stats[8] = (double)dom;
  for (int k = 0; k < T_COUNT; ++k) M->stats[9 + k] = tms[k];
  M->stats[15] = (double)n_viterbi;
  (void)n_memo_words;
  M->stats[6] = (double)n;
  M->stats[7] = (double)total;
  size_t ws = 0;
  for (const DevBuf *b : M->all_bufs()) ws += b->cap;
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

extern "C" {

// ---------------------------------------------------------------- line splitting C ABI
int spm_cuda_split_lines(const spm_cuda_model *M, const char *buf, int64_t len, char *text, int64_t *offsets,
                         int64_t capacity, int64_t *n_lines, int memory_space, void *cuda_stream) {
  if (!M) { g_err = "model is not loaded"; return kInvalidArgument; }
  if (len < 0 || capacity < 0 || !offsets || !n_lines || (len > 0 && (!buf || !text))) {
    g_err = "invalid argument";
    return kInvalidArgument;
  }
  std::lock_guard<std::mutex> lk(M->mu);
  DEVICE_GUARD(M->device);
  cudaStream_t st = (cudaStream_t)cuda_stream;
  const bool host = memory_space == SPM_CUDA_HOST;
  int rc;
  double launches = 0;
  const int64_t ntiles = (len + kNlTile - 1) / kNlTile;
  const uint8_t *d_buf = (const uint8_t *)buf;
  uint8_t *d_text = (uint8_t *)text;
  int64_t *d_off = offsets;
  if (host) {
    if ((rc = M->w_text.ensure((size_t)len + 16))) return rc;
    if ((rc = M->w_norm.ensure((size_t)len + 16))) return rc;
    if ((rc = M->w_off.ensure((size_t)(capacity + 1) * 8))) return rc;
    if (len > 0) CUDA_TRY(cudaMemcpyAsync(M->w_text.p, buf, len, cudaMemcpyHostToDevice, st));
    d_buf = (const uint8_t *)M->w_text.p; d_text = (uint8_t *)M->w_norm.p; d_off = (int64_t *)M->w_off.p;
  }
  if ((rc = M->w_cnt.ensure((size_t)(ntiles + 1) * 4))) return rc;
  if ((rc = M->w_srcoff.ensure((size_t)(ntiles + 2) * 8))) return rc;
  if ((rc = M->w_scalars.ensure(64))) return rc;
  int64_t *d_total = (int64_t *)M->w_scalars.p;
  int64_t total_nl = 0;
  uint8_t last = '\n';
  if (len > 0) {
    k_nl_count<<<(unsigned)ntiles, kNlThreads, 0, st>>>(d_buf, len, (int32_t *)M->w_cnt.p);
    if ((rc = run_scan(M, st, (const int32_t *)M->w_cnt.p, ntiles, IdentityOp(), (int64_t *)M->w_srcoff.p, d_total, 0,
                       &launches)))
      return rc;
    k_nl_scatter<<<(unsigned)ntiles, kNlThreads, 0, st>>>(d_buf, len, (const int64_t *)M->w_srcoff.p, d_text, d_off,
                                                          capacity);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(&total_nl, d_total, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(&last, d_buf + len - 1, 1, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
  }
  const int64_t n = total_nl + ((len > 0 && last != '\n') ? 1 : 0);
  *n_lines = n;
  if (n > capacity) {
    g_err = "offsets capacity too small: " + std::to_string(n) + " lines";
    return kResourceExhausted;
  }
  const int64_t zero = 0, end = len - total_nl;
  if (len == 0) {
    if (host) offsets[0] = 0; else CUDA_TRY(cudaMemcpyAsync(d_off, &zero, 8, cudaMemcpyHostToDevice, st));
  } else {
    if (last != '\n') CUDA_TRY(cudaMemcpyAsync(d_off + n, &end, 8, cudaMemcpyHostToDevice, st));
    if (host) {
      CUDA_TRY(cudaMemcpyAsync(offsets, d_off, (size_t)(n + 1) * 8, cudaMemcpyDeviceToHost, st));
      if (end > 0) CUDA_TRY(cudaMemcpyAsync(text, d_text, (size_t)end, cudaMemcpyDeviceToHost, st));
    }
  }
  CUDA_TRY(cudaStreamSynchronize(st));
  return kOk;
}

// ---------------------------------------------------------------- line splitting without any allocation
// spm_cuda_split_lines_noalloc: the same two kernels (k_nl_count, k_nl_scatter) with their scratch --
// the newline count of every 4 KB tile, its exclusive scan, the total -- in a file-scope device
// array instead of the handle's cudaMalloc'ed workspace, for callers that cannot allocate (code
// under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that
// loads the library holds the array until exit; every call shares it (calls that overlap on one
// GPU race).  Nothing in it has to start at zero: each call overwrites the entries it reads.
constexpr int64_t kNoAllocMaxLen = 256ll << 20;                     // largest buffer the entry point takes
constexpr int kNoAllocMaxTiles = (int)(kNoAllocMaxLen / kNlTile);   // 65,536 tiles of 4 KB
struct NlScratch {
  int64_t base[kNoAllocMaxTiles + 1];   // newlines before tile k
  int32_t count[kNoAllocMaxTiles];      // newlines inside tile k
  int64_t result[2];                    // [0] newlines in the buffer, [1] its last byte
};
static __device__ __align__(256) unsigned char g_nl_scratch[sizeof(NlScratch)];   // 786,448 bytes

// One block: exclusive scan of count[0..ntiles) into base[], the totals, and the end offset of an
// unterminated last line (the scatter kernel writes offsets[k] at the k-th newline only).
__global__ void __launch_bounds__(kScanThreads)
k_nl_scan_static(NlScratch *sc, int64_t ntiles, const uint8_t *buf, int64_t len, int64_t *offsets, int64_t capacity) {
  const int64_t per = (ntiles + kScanThreads - 1) / kScanThreads;
  const int64_t a = (int64_t)threadIdx.x * per, b = a + per < ntiles ? a + per : ntiles;
  int64_t s = 0;
  for (int64_t k = a; k < b; ++k) s += sc->count[k];
  int64_t tot;
  int64_t run = block_exclusive_scan(s, &tot);
  for (int64_t k = a; k < b; ++k) { sc->base[k] = run; run += sc->count[k]; }
  if (threadIdx.x == 0) {
    const uint8_t last = buf[len - 1];
    sc->result[0] = tot;
    sc->result[1] = last;
    if (last != '\n' && tot + 1 <= capacity) offsets[tot + 1] = len - tot;
  }
}

namespace {
std::mutex g_nl_mu;
NlScratch *g_nl_addr[64] = {};   // per device: the array's DEVICE address (the symbol itself is a host-side stand-in)
int nl_scratch_of_current_device(NlScratch **out) {
  int dev = -1;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { g_err = "device ordinal out of range"; return kOutOfRange; }
  std::lock_guard<std::mutex> lk(g_nl_mu);
  if (!g_nl_addr[dev]) {   // once per device, with that device current
    void *p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(&p, g_nl_scratch));
    g_nl_addr[dev] = (NlScratch *)p;
  }
  *out = g_nl_addr[dev];
  return kOk;
}
}  // namespace

int spm_cuda_split_lines_noalloc(const spm_cuda_model *M, const char *buf, int64_t len, char *text, int64_t *offsets,
                                 int64_t capacity, int64_t *n_lines, void *cuda_stream) {
  if (len < 0 || capacity < 0 || !offsets || !n_lines || (len > 0 && (!buf || !text))) {
    g_err = "invalid argument";
    return kInvalidArgument;
  }
  if (len > kNoAllocMaxLen) {
    g_err = "spm_cuda_split_lines_noalloc: buffer of " + std::to_string(len) + " bytes exceeds the static scratch (" +
            std::to_string(kNoAllocMaxLen) + " bytes at most; use spm_cuda_split_lines)";
    return kOutOfRange;
  }
  int dev = -1;
  if (M) dev = M->device; else CUDA_TRY(cudaGetDevice(&dev));
  DEVICE_GUARD(dev);
  cudaStream_t st = (cudaStream_t)cuda_stream;
  if (len == 0) {
    CUDA_TRY(cudaMemsetAsync(offsets, 0, 8, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    *n_lines = 0;
    return kOk;
  }
  NlScratch *sc = nullptr;
  int rc = nl_scratch_of_current_device(&sc);
  if (rc) return rc;
  const int64_t ntiles = (len + kNlTile - 1) / kNlTile;
  k_nl_count<<<(unsigned)ntiles, kNlThreads, 0, st>>>((const uint8_t *)buf, len, sc->count);
  k_nl_scan_static<<<1, kScanThreads, 0, st>>>(sc, ntiles, (const uint8_t *)buf, len, offsets, capacity);
  k_nl_scatter<<<(unsigned)ntiles, kNlThreads, 0, st>>>((const uint8_t *)buf, len, sc->base, (uint8_t *)text, offsets,
                                                        capacity);
  CUDA_TRY(cudaGetLastError());
  int64_t res[2] = {0, '\n'};
  CUDA_TRY(cudaMemcpyAsync(res, sc->result, 16, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  const int64_t n = res[0] + (res[1] != '\n' ? 1 : 0);
  *n_lines = n;
  if (n > capacity) {
    g_err = "offsets capacity too small: " + std::to_string(n) + " lines";
    return kResourceExhausted;
  }
  return kOk;
}

// ---------------------------------------------------------------- E-step C ABI
struct spm_cuda_estep {
  int device = 0;
  int V = 0;
  int max_ctas = 1;        // resident CTAs of k_estep on the device (the grid of the persistent kernel)
  EstepView ev;
  DevBuf d_units, d_leaf, d_hot, d_lse;
  mutable std::mutex mu;
  mutable DevBuf w_words, w_off, w_freq, w_scratch, w_expected, w_out, w_logz, w_perm, w_hist;
};

int spm_cuda_estep_create(const char *pieces, const float *scores, int V, int device, spm_cuda_estep **out) {
  if (out == nullptr) { g_err = "output container is null"; return kInternal; }
  *out = nullptr;
  if (pieces == nullptr || scores == nullptr || V <= 0) { g_err = "no pieces are loaded."; return kInvalidArgument; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    g_err = "no CUDA device available (this library has no CPU path)";
    return kInternal;
  }
  if (device < 0 || device >= ndev) { g_err = "invalid device ordinal"; return kInvalidArgument; }
  std::vector<std::string> ps;
  std::vector<float> sc(scores, scores + V);
  const char *q = pieces;
  for (int i = 0; i < V; ++i) { ps.emplace_back(q); q += ps.back().size() + 1; }
  HostTrie trie;
  float min_score = 0;
  std::vector<int32_t> hot_ids;
  int rc = build_estep_trie(ps, sc, &trie, &min_score, &g_err, &hot_ids);
  if (rc) return rc;
  std::unique_ptr<spm_cuda_estep> E(new spm_cuda_estep);
  E->device = device;
  E->V = V;
  DEVICE_GUARD(device);
  if ((rc = upload(&E->d_units, trie.units.data(), trie.units.size() * 4))) return rc;
  if ((rc = upload(&E->d_leaf, trie.leaf.data(), trie.leaf.size() * 8))) return rc;
  {
    std::vector<double> lse;
    build_lse_table(&lse);
    if ((rc = upload(&E->d_lse, lse.data(), lse.size() * 8))) return rc;
    E->ev.lse = (const double *)E->d_lse.p;
  }
  E->ev.units = (const uint32_t *)E->d_units.p;
  E->ev.leaf = (const uint2 *)E->d_leaf.p;
  E->ev.units_hot = nullptr;   // (the kernel points it at its shared-memory copy)
  E->ev.k_units = (uint32_t)std::min<size_t>(trie.units.size(), kEstepHotUnits);
  hot_ids.resize(kEstepHot, -1);
  if ((rc = upload(&E->d_hot, hot_ids.data(), hot_ids.size() * 4))) return rc;
  E->ev.unk_score = min_score - 10.0f;   // kUnkPenalty, src/unigram_model.cc:550-551
  E->ev.id_mask = V <= (1 << 20) ? 0xFFFFFu : 0x0FFFFFFFu;
  E->ev.unk_id = 0;                      // TrainerModel never calls InitializePieces
  for (int k = 1; k < kEstepHot && V <= (1 << 20); ++k)
    if (hot_ids[k] == 0) E->ev.unk_id = (uint32_t)k << 20;
  CUDA_TRY(cudaFuncSetAttribute(k_estep<kEstepPos>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                estep_smem_bytes(kEstepPos)));
  {
    int occ = 0, sms = 1;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_estep<kEstepPos>, kEstepThreads,
                                                           estep_smem_bytes(kEstepPos)));
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    E->max_ctas = std::max(1, occ) * std::max(1, sms);
  }
  *out = E.release();
  return kOk;
}

void spm_cuda_estep_free(spm_cuda_estep *E) {
  if (!E) return;
  DeviceGuard guard(E->device);
  for (DevBuf *b : {&E->d_units, &E->d_leaf, &E->d_hot, &E->d_lse, &E->w_words, &E->w_off, &E->w_freq, &E->w_scratch,
                    &E->w_expected, &E->w_out, &E->w_logz, &E->w_perm, &E->w_hist})
    b->release();
  delete E;
}

int spm_cuda_estep_vocab_size(const spm_cuda_estep *E) { return E ? E->V : 0; }

int spm_cuda_estep_run(const spm_cuda_estep *E, const char *words, const int64_t *word_offsets, const int64_t *freq,
                       int64_t n, double all_freq, double *expected, double *obj_ntokens, float *logz,
                       int memory_space, void *cuda_stream) {
  if (!E) { g_err = "E-step handle is null"; return kInvalidArgument; }
  if (n < 0 || (n > 0 && (!words || !word_offsets || !freq)) || !expected || !obj_ntokens || !(all_freq > 0)) {
    g_err = "invalid argument";
    return kInvalidArgument;
  }
  std::lock_guard<std::mutex> lk(E->mu);
  DEVICE_GUARD(E->device);
  cudaStream_t st = (cudaStream_t)cuda_stream;
  const bool host = memory_space == SPM_CUDA_HOST;
  int rc;
  const uint8_t *d_words = (const uint8_t *)words;
  const int64_t *d_off = word_offsets, *d_freq = freq;
  double *d_exp = expected, *d_out = obj_ntokens;
  float *d_logz = logz;
  int64_t lo = 0, hi = 0;
  if (n > 0) {
    if (host) { lo = word_offsets[0]; hi = word_offsets[n]; }
    else {
      CUDA_TRY(cudaMemcpyAsync(&lo, word_offsets, 8, cudaMemcpyDeviceToHost, st));
      CUDA_TRY(cudaMemcpyAsync(&hi, word_offsets + n, 8, cudaMemcpyDeviceToHost, st));
      CUDA_TRY(cudaStreamSynchronize(st));
    }
    if (hi < lo) { g_err = "word_offsets must be non-decreasing"; return kInvalidArgument; }
  }
  if ((rc = E->w_out.ensure(32))) return rc;
  if (host) {
    if ((rc = E->w_words.ensure((size_t)(hi - lo) + 16))) return rc;
    if ((rc = E->w_off.ensure((size_t)(n + 1) * 8))) return rc;
    if ((rc = E->w_freq.ensure((size_t)(n + 1) * 8))) return rc;
    if ((rc = E->w_expected.ensure((size_t)E->V * 8))) return rc;
    if (logz && (rc = E->w_logz.ensure((size_t)(n + 1) * 4))) return rc;
    if (n > 0) {
      CUDA_TRY(cudaMemcpyAsync(E->w_words.p, words + lo, hi - lo, cudaMemcpyHostToDevice, st));
      CUDA_TRY(cudaMemcpyAsync(E->w_off.p, word_offsets, (n + 1) * 8, cudaMemcpyHostToDevice, st));
      CUDA_TRY(cudaMemcpyAsync(E->w_freq.p, freq, n * 8, cudaMemcpyHostToDevice, st));
    }
    CUDA_TRY(cudaMemcpyAsync(E->w_expected.p, expected, (size_t)E->V * 8, cudaMemcpyHostToDevice, st));
    d_words = (const uint8_t *)E->w_words.p - lo;
    d_off = (const int64_t *)E->w_off.p; d_freq = (const int64_t *)E->w_freq.p;
    d_exp = (double *)E->w_expected.p; d_out = (double *)E->w_out.p;
    d_logz = logz ? (float *)E->w_logz.p : nullptr;
    CUDA_TRY(cudaMemcpyAsync(d_out, obj_ntokens, 16, cudaMemcpyHostToDevice, st));
  }
  if (n > 0) {
    if ((rc = E->w_scratch.ensure((size_t)(4 * ((hi - lo) + n) + 16) * 4))) return rc;
    if ((rc = E->w_perm.ensure((size_t)n * 4))) return rc;
    if ((rc = E->w_hist.ensure(2 * kEstepBins * 4))) return rc;
    uint32_t *d_hist = (uint32_t *)E->w_hist.p, *d_cursor = d_hist + kEstepBins;
    CUDA_TRY(cudaMemsetAsync(d_hist, 0, kEstepBins * 4, st));
    const unsigned gs = (unsigned)std::min<int64_t>((n + 255) / 256, 1184);
    k_estep_hist<<<gs, 256, 0, st>>>(d_off, n, d_hist);
    k_estep_binscan<<<1, 32, 0, st>>>(d_hist, d_cursor);
    k_estep_scatter<<<gs, 256, 0, st>>>(d_off, n, d_cursor, (uint32_t *)E->w_perm.p);
    // (after the scatter cursor[kEstepBins - 1] = n: the end of the shortest bin)
    const unsigned ge = (unsigned)std::min<int64_t>((n + kEstepThreads - 1) / kEstepThreads, E->max_ctas);
    k_estep<kEstepPos><<<ge, kEstepThreads, estep_smem_bytes(kEstepPos), st>>>(
        E->ev, (const int32_t *)E->d_hot.p, d_words, d_off, d_freq, nullptr, d_cursor + (kEstepBins - 1),
        (const uint32_t *)E->w_perm.p, 1.0 / all_freq, (float *)E->w_scratch.p, d_exp, d_out, d_logz);
    CUDA_TRY(cudaGetLastError());
  }
  // Device-resident calls return with the work enqueued (no host sync between this kernel and a
  // collective the caller puts on the same stream); a word of 1 MiB or more makes obj NaN there.
  if (!host) return kOk;
  CUDA_TRY(cudaMemcpyAsync(expected, d_exp, (size_t)E->V * 8, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaMemcpyAsync(obj_ntokens, d_out, 16, cudaMemcpyDeviceToHost, st));
  if (logz && n > 0) CUDA_TRY(cudaMemcpyAsync(logz, d_logz, n * 4, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  const bool bad = obj_ntokens[0] != obj_ntokens[0];
  if (bad) { g_err = "word longer than 1 MiB"; return kOutOfRange; }
  return kOk;
}

}  // extern "C"

// pieces / SentencePieceText spans, batch Decode, dense / sparse layouts
#include "spm_pieces.inc"
