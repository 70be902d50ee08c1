// reads.h — alignment files (SAM text, BAM, bgzip-ed / gzip-ed SAM) as a source of read intervals for depth / coverage:
// what process_bam (src/commands/depth.rs:297-423) and collect_by_root_from_bam (src/commands/coverage.rs:125-205) take
// from rust-htslib.  See reads.cpp for the record rules.
#pragma once
#include <functional>
#include <mutex>
#include <string>
#include <unordered_map>
#include "host_index.h"

namespace gffx {

struct ReadStats {
    uint64_t records = 0;       // alignment records seen
    uint64_t unmapped = 0;      // skipped: unmapped flag or no target (depth.rs:337, :345)
    uint64_t unknown_ref = 0;   // skipped: the index has no sequence of the target's name (depth.rs:349)
    uint64_t skipped = 0;       // skipped: pos < 0 or end <= pos (depth.rs:351, :356)
    int format = 0;             // 0 SAM text, 1 BGZF (BAM or bgzip-ed SAM), 2 gzip-ed SAM
};

// Every kept record appends (sequence number, pos, reference_end) to `out`, in file order.  Throws Err(E_FORMAT) where htslib
// would fail the read (the reference then aborts: `let rec = r?;`), Err(E_ARG) for CRAM, Err(E_IO) when the file cannot be opened.
void read_alignments(const std::string& path, const std::unordered_map<std::string, uint32_t>& seq_num, int threads, Regions* out, ReadStats* st);

}  // namespace gffx
