// step.cuh -- program descriptor of the persistent per-token decode kernel (step.cu)
#pragma once
#include "gemv.cuh"

namespace b2 {

enum { OP_EMBED = 0, OP_GEMV = 1, OP_ATTN = 2, OP_ARGMAX = 3, OP_XCHG = 4 };

// One stage of a decode step.  Stages run in order with a grid-wide barrier between them.
struct StepOp {
    int kind;
    GemvArgs g;   // OP_GEMV; OP_XCHG (tensor-parallel shard: wait for the ranks' partial rows and add them into x) uses x, K and the x_* / xg_* fields
    // OP_EMBED: x = dequantise(table[*token])
    int e_type, e_d_model, e_vocab;
    const uint8_t* e_table;
    int64_t e_row_bytes;
    const int32_t* e_token;
    float* e_x;
    // OP_ATTN: single-query attention over the fp16 cache at positions [0, *pos]
    const float* a_q;
    const __half *a_k, *a_v;
    float* a_out;
    int a_n_head, a_n_kv, a_hd, a_n_ctx, a_max_chunks;
    int a_min_chunk;         // shortest chunk of tokens per work item (multiple of 32, <= STEP_ATTN_MAX_CHUNK)
    int a_cluster_merge;     // 1: merge the chunks of a kv head inside a cluster when they fit (B2_ATTN_NO_CLUSTER=1 clears it)
    const int* a_pos;
    float* a_partials;       // [n_head][max_chunks][hd + 2]
    unsigned int* a_counters;  // [n_kv], zero between steps
    // OP_ARGMAX (+ the device-resident loop's bookkeeping)
    const float* s_logits;
    int64_t s_n;
    int32_t *s_token, *s_pos, *s_step, *s_out;
    int s_cap;
    float* s_val;            // [grid] partial maxima
    int* s_idx;              // [grid]
    unsigned int* s_counter;
    unsigned int* s_epoch;   // tensor-parallel shard: launches so far (the exchange stages' sequence numbers), else NULL
};

struct StepProgram {
    const StepOp* ops;          // device memory
    int n_ops;
    int max_k;                  // largest mat-vec K (sizes the activation buffers)
    // grid barrier: its own 128-byte line (STEP_SYNC_BYTES), zeroed before every launch; word 0 counts arrivals
    unsigned int* sync_counter;
    // optional profiling trace (B2_STEP_TRACE=1): [n_ops][4] globaltimer stamps written by CTA 0:
    // 0 = compute warps passed the grid barrier, 1 = activation vector built, 2 = last tile handed to the
    // epilogue warp, 3 = this CTA arrived on the stage's grid barrier
    unsigned long long* trace;
    // step2.cu: ring steps the L2 look-ahead warp runs in front of the producer (0 = no look-ahead; B2_STEP_L2_AHEAD)
    int l2_ahead;
    // B2_STEP_FINE=<stage index>: per ring step of that stage, CTA 0 stamps {wait begin, data there, step done, copy issued}
    // into trace[n_ops * STEP_TRACE_SLOTS + 4 k ..] (STEP_FINE_STEPS steps at most); -1 = off
    int fine_op;
    // ranks of the tensor-parallel group this program exchanges with (0: one GPU; selects the kernel instantiation)
    int tp;
};

// tokens one CTA handles per attention work item: chosen per step from the sequence length so that the
// items fill the grid once (multiples of 32 between these bounds)
constexpr int STEP_ATTN_MIN_CHUNK = 32, STEP_ATTN_MAX_CHUNK = 256;
constexpr int STEP_TRACE_SLOTS = 12;  // time stamps per stage in StepProgram::trace (4 .. 7: finer marks inside a stage, 8 .. 10: tile-loop cycles)
constexpr int STEP_FINE_STEPS = 512;
constexpr int STEP_MAX_GRID = 1024;   // CTAs (argmax scratch)
constexpr size_t STEP_SYNC_BYTES = 128;

// scratch sizes for a program's attention stages (grid = step_grid_size())
size_t step_attn_partials_bytes(int n_head, int head_dim, int n_ctx, int grid);
int step_attn_max_chunks(int n_ctx, int grid);
// number of CTAs the persistent kernel runs with for this activation format and K (co-resident clusters x 8)
int step_grid_size(int act, int max_k);
int launch_step(const StepProgram& p, int act, cudaStream_t st);
// second design (step2.cu, the default; B2_STEP_V=1 keeps the first): needs every mat-vec K to be a multiple of 256
bool step2_supports(int act, int K);
int step2_grid_size(int act, int max_k, bool tp);
int launch_step2(const StepProgram& p, int act, cudaStream_t st);

}  // namespace b2
