// gm_search.cpp — `gm_search -c <collection dir> -p <pattern file> [-t 0|1] [-b 0|1]`
// (benchmark/gapped-matching/src/gm_search.cpp:35-184) for the B200 index: loads
// <dir>/index/index-<name()>-<hash>.sdsl, runs every pattern (info -> prepare -> search, each search timed),
// prints "TIMING = <us>" per pattern and the "# key = value" summary with the same keys (num_results, checksum,
// *_time_mus, prep_*_time_mus), so outputs diff against the reference harness.
// -b 1 (extension) answers the whole pattern file in ONE device pass (index_b200::search_batch); the per-pattern
// TIMING lines then carry the batch time divided by the number of patterns.
#include <getopt.h>

#include "gm_b200.hpp"

#ifndef INDEX_TYPE
#define INDEX_TYPE gm_b200::index_b200
#endif

using namespace std::chrono;

static void print_summary(const char* prefix, gm_b200::timing_results& t)
{
    auto ts = t.summary();
    std::cout << "# " << prefix << "total_time_mus = " << duration_cast<microseconds>(ts.total).count() << std::endl;
    std::cout << "# " << prefix << "min_time_mus = " << duration_cast<microseconds>(ts.min).count() << std::endl;
    std::cout << "# " << prefix << "qrt_1st_time_mus = " << duration_cast<microseconds>(ts.qrt_1st).count() << std::endl;
    std::cout << "# " << prefix << "mean_time_mus = " << duration_cast<microseconds>(ts.mean).count() << std::endl;
    std::cout << "# " << prefix << "median_time_mus = " << duration_cast<microseconds>(ts.median).count() << std::endl;
    std::cout << "# " << prefix << "qrt_3rd_time_mus = " << duration_cast<microseconds>(ts.qrt_3rd).count() << std::endl;
    std::cout << "# " << prefix << "max_time_mus = " << duration_cast<microseconds>(ts.max).count() << std::endl;
}

int main(int argc, char* const argv[])
{
    std::string dir, pattern_file;
    bool string_patterns = true, batch = false;
    int op;
    while ((op = getopt(argc, argv, "c:p:t:b:")) != -1) {
        switch (op) {
            case 'c': dir = optarg; break;
            case 'p': pattern_file = optarg; break;
            case 't': string_patterns = std::string(optarg) == "1"; break;
            case 'b': batch = std::string(optarg) == "1"; break;
        }
    }
    if (dir.empty() || pattern_file.empty()) {
        fprintf(stdout, "%s -c <collection dir> -p <pattern file> [-t <string patterns>] [-b <one batch>]\n", argv[0]);
        return EXIT_FAILURE;
    }
    try {
        gm_b200::collection col(dir);
        auto patterns = gm_b200::parse_pattern_file(pattern_file, string_patterns);
        INDEX_TYPE idx;
        const std::string input_file = col.path + "/index/index-" + idx.name() + "-" + idx.class_hash() + ".sdsl";
        std::ifstream ifs(input_file.c_str(), std::ios::binary);
        if (!ifs.is_open()) {
            std::cerr << "cannot read index from file : " << input_file << std::endl;
            return EXIT_FAILURE;
        }
        idx.load(ifs);

        size_t num_results = 0, checksum = 0;
        gm_b200::timing_results t_prep, t;
        std::string total_info;
        if (batch) {
            for (const auto& pat : patterns) total_info += "," + idx.info(pat);
            auto t0 = high_resolution_clock::now();
            auto all = idx.search_batch(patterns);
            auto dur = high_resolution_clock::now() - t0;
            for (size_t i = 0; i < patterns.size(); i++) {
                t_prep.add_timing(nanoseconds(0));
                t.add_timing(duration_cast<nanoseconds>(dur) / patterns.size());
                std::cout << "TIMING = " << duration_cast<microseconds>(dur).count() / patterns.size() << std::endl;
                for (uint64_t pos : all[i].positions) {
                    checksum += pos;
                    num_results++;
                }
            }
        } else {
            for (const auto& pat : patterns) {
                total_info += "," + idx.info(pat);
                auto p0 = high_resolution_clock::now();
                idx.prepare(pat);
                t_prep.add_timing(duration_cast<nanoseconds>(high_resolution_clock::now() - p0));
                auto t0 = high_resolution_clock::now();
                auto res = idx.search(pat);
                auto dur = high_resolution_clock::now() - t0;
                t.add_timing(duration_cast<nanoseconds>(dur));
                std::cout << "TIMING = " << duration_cast<microseconds>(dur).count() << std::endl;
                for (uint64_t pos : res.positions) {
                    checksum += pos;
                    num_results++;
                }
            }
        }
        std::cout << "# info =" << total_info << std::endl;
        std::cout << "# num_results = " << num_results << std::endl;
        std::cout << "# checksum = " << checksum << std::endl;
        print_summary("", t);
        print_summary("prep_", t_prep);
    } catch (const std::exception& e) {
        std::cerr << "gm_search: " << e.what() << std::endl;
        return EXIT_FAILURE;
    }
    return 0;
}
