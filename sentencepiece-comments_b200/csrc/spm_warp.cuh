// spm_warp.cuh -- one WARP per sentence: exact unigram lattice + Viterbi for long lines.
//
// Same arithmetic and the same candidate rule as unigram_encode<CAPC> of spm_core.cuh
// (which restates src/unigram_model.cc:535-604 + :222-261): per pending end position the
// candidates that some right node can still select (within eps of the running maximum,
// exact duplicates dropped), in begin order; a right node picks the first strictly
// greater fl(bt_l + score).  What changes is the schedule:
//   phase A  the 32 lanes walk the trie from 32 consecutive BYTE positions at once
//            (lanes on continuation bytes idle) and park their hits (end, id, score)
//            in shared memory -- the walks do not depend on the DP;
//   phase B  the positions are visited in order; the hits of one position go to
//            distinct end positions, so lane h scores and pushes hit h without conflicts
//            (pick over <= CAPC candidates, push into the ring slot of its end).
// The candidate ring (ring <= 64 byte positions x CAPC) and the hit lists live in shared
// memory; back records (CAPC per byte position) in a global scratch slice of the warp.
// A thread-per-sentence kernel needs ~20 ms for one 33 KB line; this one ~3 ms, and it
// keeps near-tied candidates itself instead of flagging the line for the general tier.
// Device only (warp intrinsics); the candidate rule itself is covered on the CPU through
// unigram_encode<CAPC>, and the kernel by the GPU parity tests on long lines.
#pragma once
#include "spm_fast.cuh"

namespace spmb {

constexpr int kWarpCapc = 4;      // candidates kept per end position
constexpr int kWarpRing = 64;     // byte positions in flight (>= max piece bytes + 5)
constexpr int kWarpHits = 20;     // hits per start position (>= max piece chars + 1) ...
constexpr int kWarpHitsSmall = 12;  // ... or 12 for models whose pieces have at most 11 characters: 6.4 KB of shared
                                    // memory per warp instead of 9 KB -> 8 instead of 6 resident CTAs per SM

template <int HITS>
struct WarpScratch {              // shared memory of one warp
  uint32_t cnt[kWarpRing];
  float maxbt[kWarpRing];
  float bt[kWarpRing * kWarpCapc];
  uint32_t rec[kWarpRing * kWarpCapc];
  uint32_t hit_id[32 * HITS];
  float hit_sc[32 * HITS];
  uint16_t hit_end[32 * HITS];   // end offset relative to the round base
};

#if defined(__CUDACC__)
// back record: piece id (24 bits) | piece bytes (6 bits, < kWarpRing) | predecessor's candidate slot (2 bits)
SPM_HD uint32_t warp_pack_rec(uint32_t id, uint32_t len, uint32_t prevslot) {
  return id | (len << 24) | (prevslot << 30);
}

// s: normalized valid UTF-8 text, n bytes (zero padded).  back: kWarpCapc rows of `stride`
// records.  out: ids in REVERSE order (unknown runs merged); returns the count, or -1 when
// more than kWarpCapc near-tied candidates met at one position (caller: general tier).
template <int HITS>
__device__ inline int32_t warp_unigram(const FastView &fv, WarpScratch<HITS> &ws, const uint8_t *s, int32_t n,
                                       uint32_t *back, int64_t stride, int32_t *out) {
  const int lane = threadIdx.x & 31;
  const uint32_t mask = kWarpRing - 1;
  for (int k = lane; k < kWarpRing; k += 32) ws.cnt[k] = 0;
  __syncwarp();
  if (lane == 0) { ws.cnt[0] = 1; ws.maxbt[0] = 0.0f; ws.bt[0] = 0.0f; ws.rec[0] = warp_pack_rec(0, 0, 0); }   // BOS
  bool overflow = false;
  const uint32_t root = fv.units[0];
  for (int32_t P = 0; P < n; P += 32) {
    __syncwarp();
    // ---- phase A: trie walk from byte P + lane
    const int32_t p = P + lane;
    int nh = 0;
    bool start = false;
    if (p < n) {
      const uint32_t b0 = s[p];
      start = (b0 & 0xC0u) != 0x80u;
      if (start) {
        const int cl = one_char_len((uint8_t)b0);
        uint32_t u = root;
        int nchars = 0;
        bool single = false;
        for (int32_t i = p; i < n; ++i) {
          const uint32_t c = s[i];
          if (!(u & 0x200u) || c == 0) break;
          const uint32_t child = (u >> 10) ^ c;
          const uint32_t v = __ldg(fv.units + child);
          if ((v & 0xFFu) != c || (v & 0x300u) == 0) break;
          u = v;
          if ((c & 0xC0u) != 0x80u) ++nchars;
          if ((v & 0x100u) && nh < HITS) {
            ws.hit_id[lane * HITS + nh] = __ldg(fv.leaf_id + child);
            ws.hit_sc[lane * HITS + nh] = __ldg(fv.score + child);
            ws.hit_end[lane * HITS + nh] = (uint16_t)(i + 1 - P);
            ++nh;
            if (nchars == 1) single = true;
          }
        }
        if (!single && nh < HITS) {   // UNK node of one character (src/unigram_model.cc:598-602)
          ws.hit_id[lane * HITS + nh] = (uint32_t)fv.unk_id;
          ws.hit_sc[lane * HITS + nh] = fv.unk_score;
          ws.hit_end[lane * HITS + nh] = (uint16_t)(p + cl - P);
          ++nh;
        }
      }
    }
    __syncwarp();
    // ---- phase B: positions in order
    uint32_t todo = __ballot_sync(0xFFFFFFFFu, start);
    while (todo) {
      const int j = __ffs(todo) - 1;
      todo &= todo - 1;
      const int32_t pj = P + j;
      const uint32_t r = (uint32_t)pj & mask;
      const uint32_t cnt = ws.cnt[r];
      const int nhj = __shfl_sync(0xFFFFFFFFu, nh, j);
      if ((uint32_t)lane < cnt) back[(int64_t)lane * stride + pj] = ws.rec[r * kWarpCapc + lane];
      if (lane < nhj) {
        const float sr = ws.hit_sc[j * HITS + lane];
        const uint32_t id = ws.hit_id[j * HITS + lane];
        const int32_t e = P + (int32_t)ws.hit_end[j * HITS + lane];
        // best predecessor (src/unigram_model.cc:229-246): first strictly greater, in begin order
        float best = ws.bt[r * kWarpCapc] + sr;
        uint32_t bs = 0;
        for (uint32_t k = 1; k < cnt; ++k) {
          const float v = ws.bt[r * kWarpCapc + k] + sr;
          if (v > best) { best = v; bs = k; }
        }
        // push into the slot of e (only this lane touches it in this step)
        const uint32_t re = (uint32_t)e & mask;
        const uint32_t ce = ws.cnt[re];
        const uint32_t rec = warp_pack_rec(id, (uint32_t)(e - pj), bs);
        if (ce == 0) {
          ws.maxbt[re] = best; ws.bt[re * kWarpCapc] = best; ws.rec[re * kWarpCapc] = rec; ws.cnt[re] = 1;
        } else {
          const float M = ws.maxbt[re];
          const float am = fabsf(M), ab = fabsf(best);
          const float eps = ((am > ab ? am : ab) + fv.smax) * 2.384185791015625e-07f;
          if (best - M > eps) {
            ws.maxbt[re] = best; ws.bt[re * kWarpCapc] = best; ws.rec[re * kWarpCapc] = rec; ws.cnt[re] = 1;
          } else if (!(M - best > eps)) {
            bool dup = false;
            for (uint32_t k = 0; k < ce; ++k)
              if (ws.bt[re * kWarpCapc + k] == best) dup = true;   // loses the first-wins tie for every score
            if (!dup) {
              if (ce == kWarpCapc) {
                overflow = true;
              } else {
                ws.bt[re * kWarpCapc + ce] = best; ws.rec[re * kWarpCapc + ce] = rec; ws.cnt[re] = ce + 1;
                if (best > M) ws.maxbt[re] = best;
              }
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) ws.cnt[r] = 0;
      __syncwarp();
    }
    if (__any_sync(0xFFFFFFFFu, (start && nh >= HITS) || overflow)) return -1;
  }
  __syncwarp();
  // ---- EOS (score 0) picks among the candidates ending at n; backtrace by lane 0
  const uint32_t r = (uint32_t)n & mask;
  const uint32_t cnt = ws.cnt[r];
  if ((uint32_t)lane < cnt) back[(int64_t)lane * stride + n] = ws.rec[r * kWarpCapc + lane];
  __syncwarp();
  int32_t total = 0;
  if (lane == 0) {
    float best = ws.bt[r * kWarpCapc] + 0.0f;
    uint32_t slot = 0;
    for (uint32_t k = 1; k < cnt; ++k) {
      const float v = ws.bt[r * kWarpCapc + k] + 0.0f;
      if (v > best) { best = v; slot = k; }
    }
    __threadfence_block();
    int32_t pos = n;
    bool prev_unk = false;
    while (pos > 0) {
      const uint32_t rec = back[(int64_t)slot * stride + pos];
      const int32_t id = (int32_t)(rec & 0x00FFFFFFu);
      const bool is_unk = id == fv.unk_id;
      if (!(is_unk && prev_unk)) out[total++] = id;     // runs of unknown pieces collapse (processor.cc:525-529)
      prev_unk = is_unk;
      pos -= (int32_t)((rec >> 24) & 0x3Fu);
      slot = rec >> 30;
    }
  }
  return __shfl_sync(0xFFFFFFFFu, total, 0);
}
#endif

}  // namespace spmb
