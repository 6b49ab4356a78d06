// qe_codegen.h -- turns a bound pipeline into CUDA source: a "pipeline program"
// struct P with straight-line, fully typed code for every expression tree, plus
// the __global__ entry point that plugs it into the hand-written kernel body of
// qe_device.cuh.  NVRTC compiles the result for sm_100a (qe_jit.cpp).
//
// This is the B200 counterpart of the reference's JVM bytecode back end
// (M/evaluator/BytecodeCompiler.kt:134-175 `compile`, :37-132 `compileProjection`):
// specialise once per query so the per-row cost has no dispatch, no boxing and
// no operand stack -- the whole tree lives in registers.
#ifndef QE_CODEGEN_H
#define QE_CODEGEN_H

#include <string>
#include <vector>

#include "qe_program.h"

namespace qe {

enum KernelMode { KM_FILTER = 0, KM_PROJECT = 1, KM_AGG = 2, KM_GROUPBY_PRIVATE = 3, KM_GROUPBY_ATOMIC = 4, KM_GROUPBY_SHARED = 5,
                  KM_GROUPBY_REG = 6 };

// Accumulator words of one group (qe_device.cuh, "group-by"): word 0 = first row << 32 | rows of the group,
// then per aggregate a value word (SUM / AVG / MIN / MAX) and, when its input can be null, its own count of
// non-null inputs (otherwise the group's row count is that count: cnt_word = -1).
struct GroupPlan {
    int nkeys = 0, nwords = 0;
    std::vector<int> kind;                 // QE_GK_* of every word
    std::vector<unsigned long long> init;  // initial bits of every word
    std::vector<int> val_word, cnt_word;   // per aggregate; val_word = -1 for COUNT
    int threads = 256;                     // private variant: threads per CTA (512 when enough groups still fit)
    int lmax = 0, cslots = 0;              // private variant: groups per CTA that fit shared memory, key-cache slots
    size_t private_smem() const { return (size_t)nwords * (size_t)lmax * (size_t)threads * 8u; }
    // shared-memory variants with 2^i copies of the words per CTA, i = 0..5 (qe_groupby_copies_body): groups per CTA and
    // key-cache slots; 0 groups = the variant does not fit
    int clmax[6] = {0, 0, 0, 0, 0, 0}, ccslots[6] = {64, 64, 64, 64, 64, 64};
    static int copies_index(int copies) { int i = 0; while ((1 << i) < copies && i < 5) i++; return i; }
    size_t copies_smem(int copies) const;  // accumulators + key cache, all dynamic
    int slmaxw = 0, scslotsw = 0;          // ... and with one copy per WARP, updated without atomics (threads / 32 copies)
    size_t shared_smemw() const;
    // two-pass variant: a record is rec_words 8-byte words -- [0] row | key null flags << 32 | input validity << 40, then
    // the key words, then one word per DISTINCT aggregate input (rec_word[j] of aggregate j; SUM(x), COUNT(x) share one)
    int rec_words = 0;
    std::vector<int> rec_word;
    static void scatter_shape(int rec_words, int &npart, int &pcap);   // partitions and staging records that fit 200 KB
    static size_t scatter_smem(int rec_words, int npart, int pcap) { return (size_t)npart * ((size_t)rec_words * (size_t)pcap * 8u + 4u); }
    int rmax = 0;                          // register variant: groups per CTA whose accumulators every thread keeps in registers (0: not applicable)
    int rmax_regs = 0;                     // ... by the register budget alone, without the cost rule
};
GroupPlan plan_groupby(const BoundPipeline &bp, int threads = 256, int lane_acc_kb = 0);

struct TileShape {
    int steps = 2;          // steps per warp per tile
    int vec = 2;            // adjacent rows per lane per step (2: 128-bit loads, 4: 256-bit loads)
    int min_blocks = 1;     // __launch_bounds__ second argument
    bool staged = false;    // filter kernel: columns arrive through cp.async.bulk into shared-memory stages (vec must be 2)
    int nstages = 3;
    bool pipelined = true;  // filter kernel (register-sourced): look-back and scatter deferred by one tile
    // filter kernel, ring variant (qe_filter_ring_body): persistent warp-specialised CTA around a ring of
    // shared-memory stages filled by cp.async.bulk; steps / vec are ignored
    bool ring = false;
    int tile_rows = 4096;   // rows per tile = rows per stage; a multiple of 32 * ncw and of 128
    int ncw = 16;           // consumer warps
    int nlb = 1;            // look-back warps
    int defer = 3;          // tiles between a tile's evaluation (phase A) and its scatter (phase C)
    int loads = 0;          // bulk-copy tiles in flight at most (0 = as many as there are free stages)
    int list_cap = 0;       // sparse tiles: a consumer warp that kept at most this many rows of a tile leaves their row numbers
                            // as a packed list in shared memory (0 = masks only; 32, 64, 96, 128) ...
    int nsw = 0;            // ... and `nsw` sweeper warps store the listed rows of every consumer warp once the tile's prefix is known
                            // (the consumers then never wait for a prefix); 0: every consumer warp stores its own list in phase C
    int gthreads = 256;     // group-by, private variant: threads per CTA (256 or 512)
    // group-by, register variant (qe_groupby_reg_body): the same ring of bulk-copy stages (tile_rows, ncw, nstages),
    // one producer warp + ncw consumer warps, no look-back
    bool greg = false;
    // dense projection through the same ring of stages (qe_project_ring_body): one producer warp + ncw consumer warps
    bool pring = false;
    int lane_acc_kb = 0;    // group-by, lane copies: cap on the KB of the copies (0 = none, see plan_groupby)
    int scopies = 1;        // group-by, CTA-shared variant: copies of the accumulator words per CTA (1, 8 = one per lane & 7, 32 = one
                            // per lane, or -1 = one per warp, updated with plain loads and stores)
    int rmax_hint = 0;      // register variant: compile for this many groups instead of the plan's maximum (0: the maximum)
    // group-by in two passes for more groups than a CTA's shared memory holds (qe_groupby_scatter_body / _gather_body):
    // gpass 1 = rows -> records, partitioned by key hash into npart partitions (pcap records of staging each);
    // gpass 2 = one CTA per partition aggregates its records with `scopies` copies.  0 = neither.
    int gpass = 0, npart = 0, pcap = 0;

    int rows_per_tile() const { return (ring || greg || pring) ? tile_rows : (gthreads / 32) * steps * 32 * vec; }
    int block_threads() const { return (greg || pring) ? 32 * (1 + ncw) : ring ? 32 * (2 + nlb + (list_cap ? nsw : 0) + ncw) : 256; }
};

// outputs whose validity bitmap is all ones because the filter keeps no row where they are null (ring filter kernel)
std::vector<bool> implied_valid_outputs(const BoundPipeline &bp);

// bytes of dynamic shared memory the staged filter kernel needs (0 when not staged)
size_t staged_smem_bytes(const BoundPipeline &bp, const TileShape &shape);

// name of the generated entry point: one per kernel body, so that a profiler's launch list says which
// kernel of DESIGN.md section 4 a launch was (qe_filter_ring_kernel, qe_project_kernel, qe_agg_kernel, ...)
const char *jit_entry_name(KernelMode mode, const TileShape &shape);

std::string generate_source(const BoundPipeline &bp, KernelMode mode, const TileShape &shape);

}  // namespace qe
#endif
