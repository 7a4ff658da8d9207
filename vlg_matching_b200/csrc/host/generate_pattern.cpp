// generate_pattern.cpp — `generate-pattern.x file count length [human readable] [charset]`
// (examples/generate-pattern.cpp:1-120) on the B200 index: prints the `count` most frequent phrases of `length`
// symbols, one per line, in the reference's order.  The reference builds a compressed suffix tree of the file and
// caches it as <file>.sdsl; here the suffix array is sorted on the device in a fraction of a second, so nothing is
// cached.  The patterns of the experiment grid are made from this list (benchmark/gapped-matching/Makefile:90-106).
#include <cstdlib>
#include <iostream>
#include <string>
#include <vector>

#include "../../../include/vlgb200.h"

int main(int argc, char** argv)
{
    if (argc <= 3) {
        std::cout << "Usage: " << argv[0] << " file count length [charset]" << std::endl;
        std::cout << " - file:       The data to extract the patterns from" << std::endl;
        std::cout << " - count:      Number of patterns to extract" << std::endl;
        std::cout << " - length:     Minimum length of patterns to extract" << std::endl;
        std::cout << " - human readable text: 0 or 1 (default)" << std::endl;
        std::cout << " - charset (only applies if human readable):" << std::endl;
        std::cout << "   - 0 = alpha-numerical only (default)" << std::endl;
        std::cout << "   - 1 = alpha-numerical with whitespace" << std::endl;
        std::cout << std::endl;
        std::cout << " Sorts the suffixes on the GPU to extract most common patterns/phrases matching provided criteria." << std::endl;
        std::cout << std::endl;
        return 1;
    }
    const std::string file(argv[1]);
    const uint32_t count = (uint32_t)atoi(argv[2]), length = (uint32_t)atoi(argv[3]);
    const bool human_readable = argc <= 4 ? true : (atoi(argv[4]) == 1);
    const uint32_t charset = !human_readable ? 2u : (argc <= 5 ? 0u : (uint32_t)atoi(argv[5]));
    vlgb_index* idx = nullptr;
    int rc = vlgb_index_build_file(file.c_str(), &idx);
    if (rc == VLGB_OK) rc = vlgb_index_attach_rank(idx, VLGB_RANK_KEEP_SA);
    std::vector<uint8_t> out((size_t)count * length + 1);
    std::vector<uint64_t> occ(count + 1);
    uint32_t found = 0;
    if (rc == VLGB_OK) rc = vlgb_index_frequent_phrases(idx, length, count, charset, out.data(), occ.data(), &found);
    if (rc != VLGB_OK) {
        std::cerr << "generate-pattern: " << vlgb_last_error() << std::endl;
        vlgb_index_free(idx);
        return 1;
    }
    if (found) std::cerr << "Found " << found << " phrases with min. #occurrences of " << occ[0] << std::endl;
    for (uint32_t i = 0; i < count; i++) {
        if (human_readable) {
            std::cout << std::string(reinterpret_cast<const char*>(out.data()) + (size_t)i * length, length) << std::endl;
        } else {
            for (uint32_t j = 0; j < length; j++) std::cout << (j ? " " : "") << (unsigned)out[(size_t)i * length + j];
            std::cout << std::endl;
        }
    }
    vlgb_index_free(idx);
    return 0;
}
