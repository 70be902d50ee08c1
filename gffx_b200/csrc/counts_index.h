// counts_index.h — index-time layout for depth / coverage (K6 / K7), derived once from the line table.
//
// Reference semantics (src/commands/depth.rs:121-293,
// src/commands/coverage.rs:208-450): a region is ROUTED to every root whose tree interval it overlaps
// (deduplicated by root fid); inside the root's block every line carrying an ID is tested against the
// routed regions; results are keyed by the ID STRING and merged across blocks.
//
// Layout:
//   unit    = (block, ID) — the lines of one ID inside one block, sorted by start, plus their merged
//             segments (touching lines merge; the gaps between segments are what a region must not
//             fall into to be counted once).
//   space   = a set of regions + a coordinate set.  Space c < n_vseqs is "every valid region on
//             virtual sequence c" (VSeqMap, host_index.h) and holds the units of REGULAR blocks (exactly one tree interval, every
//             line nested inside it — then overlapping a line implies being routed).  Every other
//             block with usable lines gets a private space fed only by the regions routed to it.
//   X       = per space the sorted distinct line coordinates; slot k of a space = "rank k in X".
//             depth: histograms of region starts / ends over the slots (prefix sums S, E) give the
//             number of regions overlapping any [X_i, X_j); coverage: a bitmap over the space's
//             coordinate hull gives the covered bases.
#pragma once
#include "host_index.h"
#include "winjoin_build.h"

namespace gffx {

struct RankSeqHost {  // same 32-byte layout as RankSeq (filter.cuh)
    uint32_t x0, cnt, dir_base, nb, shift, slot0, canon, pad;
};
struct CovSpaceHost {  // 16 B
    uint32_t lo, nbits;   // bitmap covers coordinates [lo, lo+nbits)
    uint64_t word0;       // first 64-bit word of the space; the space owns nbits/64+1 words
};

// gap_bits marks the coordinates that start a gap SHORTER than this; a region of at least this length looks for gaps regardless
constexpr uint32_t GAP_SHORT = 1024;

struct CountsIndexHost {
    uint32_t n_spaces = 0, n_slots = 0, n_units = 0, n_ids = 0, n_segs = 0, n_clines = 0;
    uint64_t n_words = 0;
    std::vector<RankSeqHost> rank;
    std::vector<CovSpaceHost> cov;
    std::vector<uint32_t> X, dir;
    // units
    std::vector<uint32_t> u_space, u_chrom, u_blk, u_line0, u_seg0;   // u_line0/u_seg0: n_units+1
    std::vector<uint32_t> u_base;   // what the unit's coordinates below carry on top of the sequence's own (VSeqMap base; 0 in a private space)
    // lines grouped by unit, sorted by start: coordinates and global slots of both ends
    std::vector<uint32_t> l_fs, l_fe, l_sfs, l_sfe;
    // merged segments grouped by unit
    std::vector<uint32_t> g_s, g_e, g_ss, g_se;
    // per ID: its units in ascending block (= file offset) order
    std::vector<uint32_t> id_unit0, id_units;
    // gaps between consecutive segments of a unit, [a, b] = [seg.e, next.s], sorted by (space, a), one (0xFFFFFFFF) sentinel after
    // each space's gaps, space s owning [gsp0[s], gsp0[s+1]); gap_bits: one bit per rank slot, set when the slot's coordinate is
    // the `a` of a gap shorter than GAP_SHORT — a read shorter than that only looks for the gaps inside it when it steps over such
    // a coordinate (counts.cuh)
    // gs_first[slot] = the first gap of the slot's space whose a is >= the slot's coordinate: where that look starts
    std::vector<uint32_t> gap_a, gap_b, gap_seg, gsp0, gap_bits, gs_first;
    // regular blocks: the block's one tree interval per unit (NONE32 for a unit of an irregular block): coverage derives
    // "some read was routed to this block" from the covered bases inside it
    std::vector<uint32_t> u_rs, u_re;
    // tree intervals of irregular blocks: seq = sequence number, payload = private space
    // the tree intervals of irregular blocks by (sequence, start): sequence v owns [irr_seq0[v], irr_seq0[v+1]); irr_pm = running
    // maximum of the ends inside the sequence (the walk back from the last start below a region's end stops when it falls to the
    // region's start); irr_space = the block's private space
    std::vector<uint32_t> irr_seq0, irr_s, irr_e, irr_pm, irr_space;
    std::vector<uint32_t> irr_grp, irr_pos;          // per irr entry: duplicate-fid group (NONE32 = single) and rank inside it
    std::vector<uint32_t> irr_base;                  // per irr entry: VSeqMap base of its sequence (regions are moved back by it)
    std::vector<uint32_t> irr_blk;                   // per irr entry: its block
    std::vector<uint32_t> grp_off, grp_seq, grp_s, grp_e;  // group members in rank order
    std::vector<uint32_t> blk_space;                 // per block: private space of an irregular block, else NONE32
    uint64_t n_irregular = 0;
};

void build_counts_index(const HostIndex& H, CountsIndexHost* out);

}  // namespace gffx
