// libgffx_synth.so: the synthetic-workload TEXT generator on its own (synth.cpp, no engine, no CUDA), so that the CPU arm of
// bench.py (--impl reference) can write the benchmark's GFF3 without mapping the engine's library.
#include <cstdint>
#include <exception>
#include <string>

#include <vector>

#include "engine.h"
#include "synth_common.h"

extern "C" int gffx_synth_gff_file(const char* path, uint64_t seed, uint32_t n_genes, int quirks, uint32_t n_scaffolds) {
    try {
        gffx::synth_gff_file(path ? std::string(path) : std::string(), seed, n_genes, quirks != 0, n_scaffolds);
        return 0;
    } catch (const std::exception&) {
        return -1;
    }
}

// regions [first, first+n) of the counter-based stream `seed` placed on sequences of the given extents (extent[c] = largest
// root end of sequence c): the same function of (seed, index) as gffx_synth_regions_host / _device (synth_common.h)
extern "C" int gffx_synth_regions_ext(const uint32_t* extent, uint32_t n_seqs, uint64_t seed, uint64_t first, uint64_t n, uint32_t length,
                                      uint32_t* chr, uint32_t* start, uint32_t* end) {
    if (!extent || n_seqs == 0) return -1;
    std::vector<uint64_t> cum(n_seqs + 1, 0);
    for (uint32_t i = 0; i < n_seqs; i++) cum[i + 1] = cum[i] + extent[i];
    if (cum.back() == 0) return -1;
    for (uint64_t i = 0; i < n; i++) gffx::synth_region(seed, first + i, length, cum.data(), extent, n_seqs, chr + i, start + i, end + i);
    return 0;
}
