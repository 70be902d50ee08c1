// cli_main.cpp — `gffx-b200 <intersect|depth|coverage>`: the reference's CLI surface for the hot path
// (src/main.rs:17-25; flags of src/commands/intersect.rs:43-70 + src/utils/common.rs:17-52,
// src/commands/depth.rs:40-74, src/commands/coverage.rs:37-57) over the C ABI.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/gffx_b200.h"

static int usage() {
    fprintf(stderr,
            "Usage: gffx-b200 <COMMAND>\n\nCommands:\n"
            "  intersect  Extract models by a region or regions from a BED file\n"
            "  coverage   Compute coverage breadth across genomic feature.\n"
            "  depth      Compute coverage depth across genomic features\n\n"
            "intersect: -i FILE (-r chr:start-end | -b BED) [-c | -C | -O] [-I] [-e] [-T TYPES] [-o FILE] [-t NUM] [-v]\n"
            "depth:     -i FILE -s SOURCE.bed [-o FILE] [--bin-shift K] [-t NUM] [-v]\n"
            "coverage:  -i FILE -s SOURCE.bed [-o FILE] [-t NUM] [-v]\n"
            "(index/extract/search/sample are not part of this engine: use the reference `gffx`)\n"
            "Environment: GFFX_DEVICE=<ordinal> selects the GPU.\n");
    return 2;
}

int main(int argc, char** argv) {
    if (argc < 2) return usage();
    std::string cmd = argv[1];
    std::vector<std::string> a(argv + 2, argv + argc);
    int device = getenv("GFFX_DEVICE") ? atoi(getenv("GFFX_DEVICE")) : 0;
    auto val = [&](size_t& i) -> const char* {
        if (i + 1 >= a.size()) {
            fprintf(stderr, "error: a value is required for '%s' but none was supplied\n", a[i].c_str());
            exit(2);
        }
        return a[++i].c_str();
    };
    int rc;
    if (cmd == "intersect") {
        gffx_intersect_args x;
        memset(&x, 0, sizeof x);
        x.mode = GFFX_MODE_OVERLAP;
        x.device = device;
        int nmode = 0;
        for (size_t i = 0; i < a.size(); i++) {
            const std::string& f = a[i];
            if (f == "-i" || f == "--input") x.input = val(i);
            else if (f == "-o" || f == "--output") x.output = val(i);
            else if (f == "-e" || f == "--entire_group") x.entire_group = 1;
            else if (f == "-T" || f == "--types") x.types = val(i);
            else if (f == "-t" || f == "--threads") val(i);
            else if (f == "-v" || f == "--verbose") x.verbose = 1;
            else if (f == "-r" || f == "--region") x.region = val(i);
            else if (f == "-b" || f == "--bed") x.bed = val(i);
            else if (f == "-c" || f == "--contained") { x.mode = GFFX_MODE_CONTAINED; nmode++; }
            else if (f == "-C" || f == "--contains-region") { x.mode = GFFX_MODE_CONTAINS_REGION; nmode++; }
            else if (f == "-O" || f == "--overlap") { x.mode = GFFX_MODE_OVERLAP; nmode++; }
            else if (f == "-I" || f == "--invert") x.invert = 1;
            else { fprintf(stderr, "error: unexpected argument '%s' found\n", f.c_str()); return 2; }
        }
        if (!x.input) { fprintf(stderr, "error: the following required arguments were not provided: --input <FILE>\n"); return 2; }
        if (nmode > 1) { fprintf(stderr, "error: --contained, --contains-region and --overlap cannot be used together\n"); return 2; }
        if ((x.bed != nullptr) == (x.region != nullptr)) { fprintf(stderr, "error: exactly one of --region <REGION> / --bed <BED> is required\n"); return 2; }
        rc = gffx_intersect_run(&x);
    } else if (cmd == "depth" || cmd == "coverage") {
        gffx_counts_args x;
        memset(&x, 0, sizeof x);
        x.bin_shift = 12;
        x.threads = 12;
        x.device = device;
        for (size_t i = 0; i < a.size(); i++) {
            const std::string& f = a[i];
            if (f == "-i" || f == "--input") x.input = val(i);
            else if (f == "-s" || f == "--source") x.source = val(i);
            else if (f == "-o" || f == "--output") x.output = val(i);
            else if (f == "-t" || f == "--threads") x.threads = atoi(val(i));
            else if (f == "--bin-shift" && cmd == "depth") x.bin_shift = (uint32_t)atoi(val(i));
            else if (f == "-v" || f == "--verbose") x.verbose = 1;
            else { fprintf(stderr, "error: unexpected argument '%s' found\n", f.c_str()); return 2; }
        }
        if (!x.input || !x.source) { fprintf(stderr, "error: --input <FILE> and --source <SOURCE> are required\n"); return 2; }
        rc = cmd == "depth" ? gffx_depth_run(&x) : gffx_coverage_run(&x);
    } else {
        return usage();
    }
    if (rc != 0) {
        fprintf(stderr, "Error: %s\n", gffx_last_error());
        return 1;
    }
    return 0;
}
