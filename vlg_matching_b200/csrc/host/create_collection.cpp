// create_collection.cpp — `create_collection -i <input file> -c <collection dir> [-n <num bytes>]`
// (benchmark/gapped-matching/src/create_collection.cpp:26-115): makes <dir>/{tmp,index,patterns,results} and
// writes <dir>/text.TEXT as a bit-compressed serialized int_vector<0>.  Byte input (-n 1, default): the bytes
// \0 \n \r \f become ' ' (:87-90).  Integer input (-n 2/4/8): little-endian integers, cut at the first 0 (:99-101).
// Host-only tool (no GPU work).
#include <getopt.h>

#include "gm_b200.hpp"

int main(int argc, char* const argv[])
{
    std::string input, dir;
    int num_bytes = 1, op;
    while ((op = getopt(argc, argv, "i:c:n:")) != -1) {
        switch (op) {
            case 'i': input = optarg; break;
            case 'c': dir = optarg; break;
            case 'n': num_bytes = atoi(optarg); break;
        }
    }
    if (input.empty() || dir.empty() || num_bytes < 1 || num_bytes > 8) {
        fprintf(stdout, "%s -i <input file> -c <col dir> [-n <num bytes>]\n", argv[0]);
        return EXIT_FAILURE;
    }
    try {
        gm_b200::create_directory(dir);
        for (const char* sub : {"/tmp", "/index", "/patterns", "/results"}) gm_b200::create_directory(dir + sub);
        std::ifstream ifs(input.c_str(), std::ios::binary);
        if (!ifs) {
            std::cerr << "Error opening input file '" << input << "'" << std::endl;
            return EXIT_FAILURE;
        }
        std::vector<char> raw((std::istreambuf_iterator<char>(ifs)), std::istreambuf_iterator<char>());
        std::vector<uint64_t> sym;
        if (num_bytes == 1) {
            sym.resize(raw.size());
            for (size_t i = 0; i < raw.size(); i++) {
                uint8_t b = (uint8_t)raw[i];
                if (b == 0 || b == '\n' || b == '\r' || b == '\f') b = ' ';
                sym[i] = b;
            }
        } else {
            sym.resize(raw.size() / num_bytes);
            for (size_t i = 0; i < sym.size(); i++) {
                uint64_t v = 0;
                for (int b = 0; b < num_bytes; b++) v |= (uint64_t)(uint8_t)raw[i * num_bytes + b] << (8 * b);
                if (v == 0) {
                    sym.resize(i);
                    break;
                }
                sym[i] = v;
            }
        }
        uint64_t mx = 0;
        for (uint64_t v : sym) mx = std::max(mx, v);
        uint8_t width = 1;  // util::bit_compress: hi(max) + 1 bits
        while (width < 64 && (mx >> width)) width++;
        vlg_b200::write_sdsl_int_vector(dir + "/text.TEXT", sym, width);
        std::cerr << "num_syms = " << sym.size() << "\nlog2(sigma) = " << (int)width << std::endl;
    } catch (const std::exception& e) {
        std::cerr << "create_collection: " << e.what() << std::endl;
        return EXIT_FAILURE;
    }
    return EXIT_SUCCESS;
}
