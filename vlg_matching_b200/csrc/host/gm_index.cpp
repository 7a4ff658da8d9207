// gm_index.cpp — `gm_index -c <collection dir>` (benchmark/gapped-matching/src/gm_index.cpp:18-81) for the
// B200 index: reads <dir>/text.TEXT, builds INDEX_TYPE(col) (the text goes to HBM), writes
// <dir>/index/index-<name()>-<hash>.sdsl and the .html space report next to it.
#include <getopt.h>

#include "gm_b200.hpp"

#ifndef INDEX_TYPE
#define INDEX_TYPE gm_b200::index_b200
#endif

int main(int argc, char* const argv[])
{
    std::string dir;
    int op;
    while ((op = getopt(argc, argv, "c:")) != -1)
        if (op == 'c') dir = optarg;
    if (dir.empty()) {
        fprintf(stdout, "%s -c <collection dir>\nwhere\n  -c <collection dir>  : the collection dir.\n", argv[0]);
        return EXIT_FAILURE;
    }
    try {
        gm_b200::collection col(dir);
        auto start = std::chrono::high_resolution_clock::now();
        INDEX_TYPE idx(col);
        auto stop = std::chrono::high_resolution_clock::now();
        std::cerr << "index construction in (s): "
                  << std::chrono::duration_cast<std::chrono::milliseconds>(stop - start).count() / 1000.0f << std::endl;
        const std::string output_file = col.path + "/index/index-" + idx.name() + "-" + idx.class_hash() + ".sdsl";
        std::ofstream ofs(output_file.c_str(), std::ios::binary);
        if (!ofs.is_open()) {
            std::cerr << "cannot write index to file : " << output_file << std::endl;
            return EXIT_FAILURE;
        }
        std::cerr << "writing index to file : " << output_file << std::endl;
        auto bytes = idx.serialize(ofs);
        std::cerr << "index size : " << bytes / (1024 * 1024) << " MiB" << std::endl;
        std::ofstream vofs((output_file + ".html").c_str());
        vofs << "<html><body><h1>" << idx.name() << "</h1><p>text shard in HBM: " << bytes << " bytes serialized</p></body></html>\n";
    } catch (const std::exception& e) {
        std::cerr << "gm_index: " << e.what() << std::endl;
        return EXIT_FAILURE;
    }
    return 0;
}
