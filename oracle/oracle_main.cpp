// oracle_main.cpp — `gffx_oracle <index|intersect|depth|coverage> …`: the CPU restatement behind the
// reference's CLI flags (src/main.rs:17-25, src/commands/intersect.rs:43-70, src/commands/depth.rs:40-74,
// src/commands/coverage.rs:37-57, src/utils/common.rs:17-52).  TEST INFRASTRUCTURE / CPU BASELINE ONLY.
#include "gffx_oracle.hpp"
#include <omp.h>

using namespace gffx_oracle;

static bool ends_with_ci(const std::string& s, const char* ext) {
    size_t d = s.rfind('.');
    if (d == std::string::npos) return false;
    std::string e = s.substr(d + 1);
    for (auto& c : e) c = (char)tolower(c);
    return e == ext;
}

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: gffx_oracle <index|intersect|depth|coverage> [flags]\n"); return 2; }
    std::string cmd = argv[1];
    try {
        std::vector<std::string> a(argv + 2, argv + argc);
        auto val = [&](size_t& i) -> std::string {
            if (i + 1 >= a.size()) throw Error("missing value for " + a[i]);
            return a[++i];
        };
        if (cmd == "index") {
            std::string in, skip = "remark,note,comment,region,gap,assembly_gap,contig,scaffold,source";
            for (size_t i = 0; i < a.size(); i++) {
                if (a[i] == "-i" || a[i] == "--input") in = val(i);
                else if (a[i] == "-s" || a[i] == "--skip-types") skip = val(i);
                else if (a[i] == "-a" || a[i] == "--attribute") val(i);
                else if (a[i] == "-v" || a[i] == "--verbose") {}
                else throw Error("unexpected argument " + a[i]);
            }
            build_index(in, skip);
            return 0;
        }
        if (cmd == "intersect") {
            IntersectArgs ia;
            for (size_t i = 0; i < a.size(); i++) {
                const std::string& f = a[i];
                if (f == "-i" || f == "--input") ia.input = val(i);
                else if (f == "-o" || f == "--output") ia.output = val(i);
                else if (f == "-e" || f == "--entire_group") ia.entire_group = true;
                else if (f == "-T" || f == "--types") { ia.has_types = true; ia.types = val(i); }
                else if (f == "-t" || f == "--threads") val(i);  // intersect never configures the pool
                else if (f == "-v" || f == "--verbose") ia.verbose = true;
                else if (f == "-r" || f == "--region") ia.region = val(i);
                else if (f == "-b" || f == "--bed") ia.bed = val(i);
                else if (f == "-c" || f == "--contained") ia.contained = true;
                else if (f == "-C" || f == "--contains-region") ia.contains_region = true;
                else if (f == "-O" || f == "--overlap") ia.overlap = true;
                else if (f == "-I" || f == "--invert") ia.invert = true;
                else throw Error("unexpected argument " + f);
            }
            if ((int)ia.contained + (int)ia.contains_region + (int)ia.overlap > 1) throw Error("-c, -C and -O are mutually exclusive");
            if (ia.bed.empty() == ia.region.empty()) throw Error("exactly one of --region / --bed is required");
            intersect_run(ia);
            return 0;
        }
        if (cmd == "depth" || cmd == "coverage") {
            std::string in, src, out;
            uint32_t bin_shift = 12;
            int threads = 12;
            for (size_t i = 0; i < a.size(); i++) {
                const std::string& f = a[i];
                if (f == "-i" || f == "--input") in = val(i);
                else if (f == "-s" || f == "--source") src = val(i);
                else if (f == "-o" || f == "--output") out = val(i);
                else if (f == "-t" || f == "--threads") threads = atoi(val(i).c_str());
                else if (f == "--bin-shift" && cmd == "depth") bin_shift = (uint32_t)atoi(val(i).c_str());
                else if (f == "-v" || f == "--verbose") {}
                else throw Error("unexpected argument " + f);
            }
            if (threads == 0) threads = omp_get_num_procs();
            omp_set_num_threads(threads);
            if (!ends_with_ci(src, "bed")) {
                if (ends_with_ci(src, "bam") || ends_with_ci(src, "sam") || ends_with_ci(src, "cram"))
                    throw Error("BAM/SAM/CRAM sources need htslib, which is not available in this build");
                throw Error("Unsupported file type: \"" + src + "\". Expected .bam/.sam/.cram or .bed");
            }
            GofMap gof = load_gof(in);
            MappedFile gff(in);
            TreeIndexData idx = TreeIndexData::load_tree_index(in);
            FILE* o = out.empty() ? stdout : fopen(out.c_str(), "wb");
            if (!o) throw Error("cannot create output " + out);
            if (cmd == "depth") {
                IdMap m = depth_process_bed(src, idx, gof, gff.data, gff.len, bin_shift);
                write_depth_results(m, o);
            } else {
                ByRoot by_root = collect_by_root_from_bed(src, idx);
                IdMap m = finalize_compute_breadth(by_root, gof, gff.data, gff.len);
                write_breadth_results(m, o);
            }
            if (!out.empty()) fclose(o); else fflush(o);
            return 0;
        }
        throw Error("unknown subcommand " + cmd);
    } catch (const std::exception& e) {
        fprintf(stderr, "Error: %s\n", e.what());
        return 1;
    }
}
