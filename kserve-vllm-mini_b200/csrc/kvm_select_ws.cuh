// kvm_select_ws.cuh - workspace layout of the exact multi-rank select (csrc/kvm_select.cu), shared
// with the fused one-pass log kernel (csrc/kvm_fused.cu), which writes the digit parameters and
// the per-set histograms straight into it.  Everything lives in an anonymous namespace (one copy
// per translation unit).
#pragma once

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"

// (library-internal, defined in kvm_select.cu)
int kvm_select_plan_internal(const kvm_set_info* info, const kvm_query* queries, int32_t nq,
                             const kvm_hist_spec* hists, void* workspace, size_t workspace_bytes,
                             int64_t arena_capacity, int32_t refine_rounds, int64_t refine_min_keys,
                             int32_t preset, void* stream);

namespace {

constexpr int SEL_THREADS = 512;
constexpr int MAXQ = KVM_MAX_QUERIES;
constexpr int MAX_SRC = 16;
constexpr int STATS_SCRATCH = 256 + 8 * (KVM_MAX_SETS + 2) * 148 * 8;
constexpr int NH = KVM_MAX_HISTS;   // value histograms per problem
constexpr int NB_MAX = KVM_HIST_MAX_BINS;

struct ColParam {
  uint64_t kbase;
  int32_t shift;  // digit = (key - kbase) >> shift
  int32_t width;  // significant bits of (kmax - kmin); 0 => column holds one distinct key
};

struct Bucket {
  uint64_t base;  // keys of the bucket lie in [base, base + 2^shift)
  int32_t shift;
  int32_t col;
  uint32_t set_mask;
  uint32_t bin;
  int64_t count;  // global number of keys in the bucket
  int32_t run;    // buckets sharing (col, bin) starting at this one (valid in the first)
  int32_t pad;
};

struct QState {
  int64_t n;
  int64_t rank;
  int64_t resid;  // rank inside the bucket
  uint64_t key;   // valid when resolved
  int32_t status; // KVM_Q_*
  int32_t bucket; // -1 when resolved without candidates
  int32_t col;
  int32_t resolved;
};

struct SelWs {
  int32_t status;
  int32_t nq;
  int32_t nbuckets;
  int32_t refine;          // 1: the candidate buckets are still too large - one more refine round runs
  int64_t arena_capacity;
  int64_t arena_used;
  int32_t refine_budget;   // refine rounds the host will still launch
  int32_t level;           // 1 after resolve, +1 per refine round that ran
  int64_t refine_min;      // refine while the buckets hold more keys than this in total
  int32_t nparents;        // level-1 buckets (the lookup table's targets)
  int32_t preset_cols;     // 1: col[] came from a sample (fused pass): bins 0 and BINS-1 are catch-alls
  kvm_set_info info;
  kvm_query queries[MAXQ];
  ColParam col[KVM_MAX_COLS];
  QState q[MAXQ];
  Bucket bucket[MAXQ];           // the candidate buckets keys are appended to ("targets")
  int64_t bucket_offsets[MAXQ + 1];
  int64_t bucket_counts[MAXQ];  // local append cursors (compact)
  Bucket parent[MAXQ];           // the level-1 buckets the lookup table points at; after a refine
  uint8_t child_first[MAXQ];     // round each one owns the contiguous targets
  uint8_t child_n[MAXQ];         // [child_first, child_first + child_n) that subdivide it
  int32_t tparent[MAXQ];         // level-1 bucket of each target
  kvm_hist_spec hspec;
  int32_t active20[NH];
  int32_t col_hist[KVM_MAX_COLS];  // value histogram fed by each column, -1 = none
  uint64_t thr20[NH][NB_MAX];
  uint32_t flag20[NH][BINS / 32];
  unsigned long long edge20[NH][NB_MAX];
  uint8_t edge_code[KVM_MAX_COLS][BINS];  // 0: no histogram edge in this fine bin; b: exactly edge b;
                                          // 255: several edges (rank against all of them)
  uint16_t tab[KVM_MAX_COLS][BINS];       // low byte: first bucket with that (col, bin), 255 = none;
                                          // high byte: edge_code
  uint64_t hist[KVM_MAX_SETS][BINS];
  uint64_t hist2[MAXQ][BINS];  // refine round: per target, histogram of the next 11 key bits
  unsigned char stats_scratch[STATS_SCRATCH];
};

constexpr size_t WS_HEADER = (sizeof(SelWs) + 255) / 256 * 256;

struct QueryPack {
  kvm_query q[MAXQ];
  int32_t nq;
};

struct SrcPack {
  const uint64_t* arena[MAX_SRC];
  int32_t nsrc;
};

constexpr uint64_t DIGITS_NONE = 0x0FFFFFFFFFFFFFFFull;  // five 12-bit fields, all "not in a set"
constexpr int REC_COLD_BIT = 60;                           // bit 60 of a digit record: the row's KVM_F_COLD
__host__ __device__ __forceinline__ uint64_t rec_blank(uint32_t flags) {
  return DIGITS_NONE | ((flags & KVM_F_COLD) ? (1ull << REC_COLD_BIT) : 0ull);
}

struct SeriesArgs {
  const double* start_ms;
  const kvm_log_stats* stats;  // min_start_ms read on the device
  double window_ms;
  int64_t max_bins;
  unsigned long long* series;  // [3][max_bins]
  unsigned long long* dropped;
};

}  // namespace
