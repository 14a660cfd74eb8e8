// kernels.h — host-visible launchers of the sm_100a kernels (internal to libvsann.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "common.cuh"

namespace vsann {

constexpr uint32_t WORK_RING_SLOTS = 256;
constexpr uint32_t KNN_BATCH = 32;     // intra-batch exact neighbours merged into the candidates

struct DeviceInfo {
    int device = 0;
    int sm_count = 148;
    size_t smem_per_block = 227 * 1024;
    uint32_t hash_min = 1024;          // smallest visited hash of the staged kernels (VSANN_HASH_MIN)
    uint32_t hash_factor = 8;          // visited-hash entries per unit of ef (VSANN_HASH_FACTOR); measured: 8 beats 16 (occupancy) and 4 (re-evaluations)
    bool force_smem_list = false;      // VSANN_SMEM_LIST=1: shared-memory list kernel for every ef (A/B testing)
    uint32_t stage_bytes = 4096;       // per-warp bulk-copy staging area (VSANN_STAGE_BYTES; 0 = register-gather kernels)
    bool gather4 = true;               // TMA tile::gather4 for rows of at most 256 floats (VSANN_GATHER4=0: one bulk copy per row)
    bool select_smem = true;      // K4b keeps the selected neighbours' rows in shared memory (VSANN_SELECT_SMEM=0: gathers)
    uint32_t* work_ring = nullptr;     // WORK_RING_SLOTS device words: work counters of the persistent kernels
};

struct SearchPlan { uint32_t efp, hsize, warps_per_block, blocks_per_sm; size_t smem; };

// One insertion batch: slots [first_slot, first_slot + count). Level-l items are the slots
// divisible by M^l; their per-level scratch index is slot / M^l - first_row[l].
struct BuildBatch {
    uint32_t first_slot, count;
    int new_top;                               // top level once this batch has landed
    uint32_t first_row[MAX_LEVELS];
    uint32_t items[MAX_LEVELS];                // items on level l
    uint32_t item_off[MAX_LEVELS + 1];         // prefix sums over levels (k_select work list)
    uint32_t total_items;
    uint32_t* cand_id[MAX_LEVELS];             // [items_l][efp] graph candidates, ascending
    float* cand_d[MAX_LEVELS];
    uint32_t* cand_n[MAX_LEVELS];              // [items_l]
    uint32_t* knn_id[MAX_LEVELS];              // [items_l][kbp] intra-batch exact neighbours (may be null)
    float* knn_d[MAX_LEVELS];
};

// Reverse-edge inbox machinery (K5). Rows of all levels share one index space:
// global row = row_base[l] + slot / M^l.
struct LinkTables {
    uint32_t row_base[MAX_LEVELS];
    uint32_t* inbox_cnt;      // [total_rows]  edges registered this batch (0 between batches)
    uint32_t* inbox_idx;      // [total_rows]  -> touched index
    uint32_t* n_edges;        // device scalars: n_edges, n_touched, inbox cursor (16 bytes, zeroed before every batch)
    uint32_t* n_touched;
    uint32_t* inbox_cursor;
    uint32_t* edge_row;       // [edge_cap]
    uint32_t* edge_src;       // [edge_cap]
    uint32_t* touched_row;    // [edge_cap]
    uint32_t* touched_slot;   // [edge_cap]
    uint8_t* touched_level;   // [edge_cap]
    uint32_t* inbox_fill;     // [edge_cap]   per touched row: entries written so far
    uint32_t* inbox_off;      // [edge_cap]   per touched row: start of its segment in `inbox` (every edge has a place: no edge is dropped)
    uint32_t* inbox;          // [edge_cap]   reverse-edge sources, one segment per touched row
};

// Exact (brute-force) search over a strided set of rows.
struct ExactCorpus {
    const float* vecs;        // first row
    const float* sqn;         // squared norm of row i at sqn[i * row_step]
    size_t row_stride;        // floats between consecutive corpus rows (dpad * row_step)
    uint32_t row_step;        // slot step between consecutive rows
    uint32_t first_slot;      // slot id of row 0
    uint32_t n;               // rows
    const uint32_t* deleted;  // tombstones by slot (may be null)
    const uint64_t* slot_key; // tie-break keys by slot (null => tie-break by slot)
    uint32_t n_slots;
};
struct ExactQueries {
    const float* q;           // row i at q + i * stride
    size_t stride;
    uint32_t nq;
    uint32_t dims, dpad;      // floats to read per row / padded length of corpus rows
    uint32_t self_first;      // query i is slot self_first + i * self_step (excluded); self_step == 0 => no exclusion
    uint32_t self_step;
};
struct ExactScratch {         // sized by exact_scratch_bytes
    void* ptr; size_t bytes;
};

size_t exact_scratch_bytes(uint32_t nq, uint32_t k, int sm_count);
size_t exact_scratch_bytes_upto(uint32_t max_nq, uint32_t k, int sm_count);
// Outputs: ascending top-k per query. out_slots (u32, NONE padded) and/or out_keys (u64, UINT64_MAX padded).
cudaError_t launch_exact(const ExactCorpus& c, const ExactQueries& q, int metric, uint32_t k, uint32_t out_ld,
                         uint32_t* out_slots, uint64_t* out_keys, float* out_dist,
                         const ExactScratch& scratch, const DeviceInfo& dev, cudaStream_t st);

// K1s (exact_kernels.cu): true when launch_exact answers this call with the HBM-streaming scan (few queries, contiguous corpus).
bool exact_stream_supported(const ExactCorpus& c, const ExactQueries& q, uint32_t k, const DeviceInfo& dev);

struct TcStats {                    // host-side counters of the tensor-core exact path
    uint64_t queries = 0, fp32_fallback = 0, second_scan = 0, scans = 0;
    double scan_ms = 0;             // device time of the full (emit) scans, measured with events when time_scan is set
    bool time_scan = false;
    unsigned long long* dev = nullptr;   // device counters {second-scan queries, fallback queries}, accumulated by the exact path itself (it never syncs to read them)
};
// Tensor-core (tcgen05 TF32 + exact FP32 re-rank) form of launch_exact for contiguous corpora; same results.
bool exact_tc_supported(const ExactCorpus& c, const ExactQueries& q, uint32_t k);
size_t exact_tc_scratch_bytes(uint32_t nq, uint32_t n, uint32_t k, uint32_t dpad, int sm_count);
cudaError_t launch_exact_tc(const ExactCorpus& c, const ExactQueries& q, int metric, uint32_t k, uint32_t out_ld,
                            uint32_t* out_slots, uint64_t* out_keys, float* out_dist,
                            const ExactScratch& scratch, const ExactScratch& fb_scratch, const DeviceInfo& dev, cudaStream_t st, TcStats* stats);

// up to 8 more copies of every result list (peer memory in the sharded search); plain data, passed to the kernels by value
struct OutFan { uint32_t n = 0; uint64_t* keys[8] = {}; float* dist[8] = {}; };
cudaError_t launch_search(const GraphView& g, const float* d_queries, uint32_t nq, uint32_t k, uint32_t ef, const DeviceInfo& dev,
                          uint64_t* d_keys, float* d_dist, unsigned long long* counters, cudaStream_t st, const OutFan* fan = nullptr);
uint32_t build_efp(uint32_t ef);
cudaError_t launch_build_search(const GraphView& g, const BuildBatch& b, uint32_t ef, const DeviceInfo& dev,
                                unsigned long long* counters, cudaStream_t st);
cudaError_t launch_link(const GraphView& g, const BuildBatch& b, const LinkTables& t, uint32_t efp, uint32_t kbp, const DeviceInfo& dev,
                        unsigned long long* counters, cudaStream_t st);
cudaError_t launch_row_prepare(const float* d_src, float* vecs, float* sqn, uint32_t first, uint32_t n, uint32_t dims, uint32_t dpad, cudaStream_t st);
cudaError_t launch_tombstone(uint32_t* deleted, const uint32_t* d_slots, uint32_t n, cudaStream_t st);
cudaError_t launch_merge_topk_ptrs(uint32_t parts, uint32_t nq, uint32_t k, const uint64_t* const* part_keys, const float* const* part_dist,
                                   uint64_t* keys_out, float* dist_out, cudaStream_t st);
cudaError_t launch_merge_topk_flags(uint32_t parts, uint32_t nq, uint32_t k, const uint64_t* const* part_keys, const float* const* part_dist,
                                    uint32_t* const* flags, uint32_t my_part, uint32_t epoch, uint64_t* keys_out, float* dist_out, cudaStream_t st);
cudaError_t launch_count_found(const uint64_t* keys, uint32_t nq, uint32_t k, uint32_t* found, cudaStream_t st);
cudaError_t launch_merge_topk(uint32_t parts, uint32_t nq, uint32_t k, const uint64_t* keys_in, const float* dist_in,
                              uint64_t* keys_out, float* dist_out, cudaStream_t st);

}  // namespace vsann
