// gm_search.cpp — `gm_search -c <collection dir> -p <pattern file> [-t 0|1] [-b 0|1]`
// (benchmark/gapped-matching/src/gm_search.cpp:35-184) for the B200 index: loads
// <dir>/index/index-<name()>-<hash>.sdsl, runs every pattern (info -> prepare -> search, each search timed),
// prints "TIMING = <us>" per pattern and the "# key = value" summary with the same keys (num_results, checksum,
// *_time_mus, prep_*_time_mus), so outputs diff against the reference harness.
// -b 1 (extension) answers the whole pattern file in ONE device pass (index_b200::search_batch); the per-pattern
// TIMING lines then carry the batch time divided by the number of patterns.
// -g N / --gpus N (extension, implies -b 1, byte texts): the text of the collection is cut into N overlapping shards,
// one per GPU, one host thread per GPU; the cursor exchange runs over NCCL inside the library (vlgb_search_sharded).
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
    int op, gpus = 1;
    static const option long_opts[] = {{"gpus", required_argument, nullptr, 'g'}, {"wcsearch3-gaps", no_argument, nullptr, 'w'}, {nullptr, 0, nullptr, 0}};
    while ((op = getopt_long(argc, argv, "c:p:t:b:g:", long_opts, nullptr)) != -1) {
        switch (op) {
            case 'c': dir = optarg; break;
            case 'p': pattern_file = optarg; break;
            case 't': string_patterns = std::string(optarg) == "1"; break;
            case 'b': batch = std::string(optarg) == "1"; break;
            case 'g': gpus = atoi(optarg); break;
            case 'w': gm_b200::set_gap_semantics(gm_b200::GAPS_WCSEARCH3); break;  // index_wcsearch3.hpp:412-413: gaps[0] + |P_1| for every level
        }
    }
    if (dir.empty() || pattern_file.empty()) {
        fprintf(stdout, "%s -c <collection dir> -p <pattern file> [-t <string patterns>] [-b <one batch>] [--gpus <n>] [--wcsearch3-gaps]\n", argv[0]);
        return EXIT_FAILURE;
    }
    try {
        gm_b200::collection col(dir);
        auto patterns = gm_b200::parse_pattern_file(pattern_file, string_patterns);
        if (gpus > 1) {  // sharded over several GPUs: built from the collection's text, one shard per device
            const std::vector<uint64_t> sym = vlg_b200::read_sdsl_int_vector(col.file_map["TEXT"]);
            std::vector<uint8_t> text(sym.size());
            for (size_t i = 0; i < sym.size(); i++) {
                if (sym[i] == 0 || sym[i] > 255) throw std::runtime_error("--gpus needs a byte text");
                text[i] = (uint8_t)sym[i];
            }
            gm_b200::multi_gpu_run run = gm_b200::search_batch_multi_gpu(text, patterns, gpus);
            size_t num_results = 0, checksum = 0;
            gm_b200::timing_results t_prep, t;
            for (size_t i = 0; i < patterns.size(); i++) {
                t_prep.add_timing(nanoseconds(0));
                t.add_timing(nanoseconds((long long)(run.seconds * 1e9 / patterns.size())));
                std::cout << "TIMING = " << (long long)(run.seconds * 1e6 / patterns.size()) << std::endl;
                for (uint64_t pos : run.results[i].positions) {
                    checksum += pos;
                    num_results++;
                }
            }
            std::cout << "# gpus = " << gpus << std::endl;
            std::cout << "# exchange_rounds = " << run.stats.rounds << std::endl;
            std::cout << "# shards_reentered = " << run.stats.reentered_shards << std::endl;
            std::cout << "# num_results = " << num_results << std::endl;
            std::cout << "# checksum = " << checksum << std::endl;
            print_summary("", t);
            print_summary("prep_", t_prep);
            return 0;
        }
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
