/* Deterministic synthetic corpora for the bench and the full-size parity tests
 * (SURVEY.md section 8(d), configs 2/3/4/5).  Workload generator only -- not on
 * the product path.  Sentence i depends only on (seed, i), so any shard of the
 * corpus can be produced independently (multi-GPU ranks generate their own).
 *
 * English-like: words drawn Zipf(s) from a caller-supplied word list, Poisson
 * word count clipped to [3,60], first letter upper-cased, trailing '.'.
 * "dirty" mode adds NFKC-triggering code points, double spaces / tabs and
 * malformed UTF-8 at the given per-mille rates.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline uint64_t splitmix(uint64_t *s) {
  uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
static inline double u01(uint64_t *s) { return (splitmix(s) >> 11) * (1.0 / 9007199254740992.0); }

typedef struct {
  const char *words;       /* concatenated word bytes */
  const int32_t *woff;     /* nwords+1 */
  const double *cdf;       /* nwords, cumulative Zipf probabilities */
  int32_t nwords;
  double lambda;           /* Poisson mean of words per sentence */
  int32_t dirty;           /* 0 clean, 1 dirty */
} synth_cfg;

static const char *kDirty[] = {
    "\xef\xbc\xa1\xef\xbc\xa2\xef\xbc\xa3",  /* fullwidth ABC */
    "\xef\xac\x81",                          /* fi ligature */
    "\xe3\x8d\xbf",                          /* U+337F square kabushiki-gaisha */
    "e\xcc\x81",                             /* e + combining acute */
    "\xef\xbd\xb8\xef\xbe\x9e",              /* halfwidth katakana ku + voiced mark */
    "\xe2\x91\xa0",                          /* circled digit one */
    "\xc2\xbd",                              /* vulgar fraction one half */
    "\xe2\x80\xa6",                          /* horizontal ellipsis */
    "\xef\xbc\x91\xef\xbc\x92",              /* fullwidth 12 */
    "A\xcc\x8a",                             /* A + ring above */
};
static const char *kBroken[] = {"\x80", "\xe3\x81", "\xff", "\xc0\xaf", "\xf0\x9f\x98"};

static int pick_word(const synth_cfg *c, uint64_t *rng) {
  double u = u01(rng);
  int lo = 0, hi = c->nwords - 1;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (c->cdf[mid] < u) lo = mid + 1; else hi = mid;
  }
  return lo;
}

/* Generates sentence `idx`; writes into out when non-NULL; returns its length. */
static int64_t gen_sentence(const synth_cfg *c, uint64_t seed, int64_t idx, char *out) {
  uint64_t rng = seed * 0xD1342543DE82EF95ull + (uint64_t)idx * 0x2545F4914F6CDD1Dull + 1;
  splitmix(&rng);
  /* Poisson(lambda) by Knuth's product method */
  double L = exp(-c->lambda), p = 1.0;
  int k = 0;
  do { ++k; p *= u01(&rng); } while (p > L);
  int nw = k - 1;
  if (nw < 3) nw = 3;
  if (nw > 60) nw = 60;
  int64_t n = 0;
  for (int w = 0; w < nw; ++w) {
    if (w > 0) {
      int sep = 0;
      if (c->dirty) {
        uint64_t r = splitmix(&rng) % 1000;
        if (r < 3) sep = 1; else if (r < 5) sep = 2;
      }
      if (sep == 0) { if (out) out[n] = ' '; n += 1; }
      else if (sep == 1) { if (out) { out[n] = ' '; out[n + 1] = ' '; } n += 2; }
      else { if (out) out[n] = '\t'; n += 1; }
    }
    int special = 0;
    if (c->dirty) {
      uint64_t r = splitmix(&rng) % 1000;
      if (r < 10) special = 1; else if (r < 11) special = 2;
    }
    if (special == 1) {
      const char *s = kDirty[splitmix(&rng) % (sizeof(kDirty) / sizeof(kDirty[0]))];
      size_t l = strlen(s);
      if (out) memcpy(out + n, s, l);
      n += l;
    } else if (special == 2) {
      const char *s = kBroken[splitmix(&rng) % (sizeof(kBroken) / sizeof(kBroken[0]))];
      size_t l = strlen(s);
      if (out) memcpy(out + n, s, l);
      n += l;
    } else {
      int wi = pick_word(c, &rng);
      int l = c->woff[wi + 1] - c->woff[wi];
      if (out) {
        memcpy(out + n, c->words + c->woff[wi], l);
        if (w == 0 && out[n] >= 'a' && out[n] <= 'z') out[n] -= 32;
      }
      n += l;
    }
  }
  if (out) out[n] = '.';
  n += 1;
  return n;
}

/* Pass 1: off[0..n] (exclusive prefix sum of lengths, off[0]=0). */
int64_t synth_lengths(const synth_cfg *c, uint64_t seed, int64_t first, int64_t n, int64_t *off) {
  off[0] = 0;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) off[i + 1] = gen_sentence(c, seed, first + i, NULL);
  for (int64_t i = 0; i < n; ++i) off[i + 1] += off[i];
  return off[n];
}

/* Pass 2: write bytes at the offsets from pass 1. */
void synth_fill(const synth_cfg *c, uint64_t seed, int64_t first, int64_t n, const int64_t *off,
                char *out) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) gen_sentence(c, seed, first + i, out + off[i]);
}

/* The (word, freq) list of the clean corpus without materialising it: hist[3 * wi + v] counts
 * word type wi in variant v (0 as is, 1 first letter upper-cased = first word of a sentence,
 * 2 with the sentence's trailing '.').  What SplitSentencesByWhitespace would hand the unigram
 * trainer's E-step for sentences [first, first + n) (src/trainer_interface.cc:465-477). */
void synth_word_hist(const synth_cfg *c, uint64_t seed, int64_t first, int64_t n, int64_t *hist) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    uint64_t rng = seed * 0xD1342543DE82EF95ull + (uint64_t)(first + i) * 0x2545F4914F6CDD1Dull + 1;
    splitmix(&rng);
    double L = exp(-c->lambda), p = 1.0;
    int k = 0;
    do { ++k; p *= u01(&rng); } while (p > L);
    int nw = k - 1;
    if (nw < 3) nw = 3;
    if (nw > 60) nw = 60;
    for (int w = 0; w < nw; ++w) {
      const int wi = pick_word(c, &rng);
      int v = 0;
      if (w == 0 && c->words[c->woff[wi]] >= 'a' && c->words[c->woff[wi]] <= 'z') v = 1;
      if (w == nw - 1) v = 2;
#pragma omp atomic
      hist[3 * (int64_t)wi + v] += 1;
    }
  }
}

/* CJK-style corpus: re-sample whole lines of a base text (config 4). */
int64_t synth_resample_lengths(const int64_t *loff, int64_t nlines, uint64_t seed, int64_t first,
                               int64_t n, int64_t *off) {
  off[0] = 0;
  for (int64_t i = 0; i < n; ++i) {
    uint64_t rng = seed * 0x9E3779B97F4A7C15ull + (uint64_t)(first + i);
    int64_t li = (int64_t)(splitmix(&rng) % (uint64_t)nlines);
    off[i + 1] = off[i] + (loff[li + 1] - loff[li]);
  }
  return off[n];
}
void synth_resample_fill(const char *lines, const int64_t *loff, int64_t nlines, uint64_t seed,
                         int64_t first, int64_t n, const int64_t *off, char *out) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    uint64_t rng = seed * 0x9E3779B97F4A7C15ull + (uint64_t)(first + i);
    int64_t li = (int64_t)(splitmix(&rng) % (uint64_t)nlines);
    memcpy(out + off[i], lines + loff[li], loff[li + 1] - loff[li]);
  }
}
