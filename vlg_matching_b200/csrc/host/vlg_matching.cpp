// vlg_matching.cpp — the entry point of /root/reference/examples/vlg_matching.cpp served by the B200 matcher:
// same arguments, same stdout.  With no argument it indexes "abracadabrasimsalabim" as bytes and as integers
// and answers the reference's six queries (examples/vlg_matching.cpp:40-42,58-60); with a file argument it
// indexes the file as bytes, writes <file>.info.html, then (like the reference, :52-53) tries to read the same
// file as a serialized int_vector<> for the int-alphabet block.
#include "vlg_b200.hpp"

using namespace vlg_b200;
using namespace std;

template <typename index>
void dump_query_results(index idx, string qry)
{
    cout << endl;
    cout << "count(" << qry << ")=" << count(idx, qry) << endl;
    cout << "locate(" << qry << ")=" << endl;
    auto res = locate(idx, qry);
    size_t occ = 1;
    for (auto it = res.begin(); it != res.end(); ++it, ++occ) {
        cout << "  " << occ << ". occ starting at position " << *it << endl;
        cout << "     Subpattern positions:";
        for (size_t i = 0; i < it.size(); ++i) cout << " " << it[i];
        cout << endl;
    }
}

int main(int argc, char* argv[])
{
    try {
        {  // byte alphabet
            vlg_index<> idx;
            if (argc <= 1) {
                construct_im(idx, "abracadabrasimsalabim", 1);
            } else {
                string filename(argv[1]);
                construct(idx, filename, 1);
                ofstream out(filename + ".info.html");
                write_structure<HTML_FORMAT>(idx, out);
            }
            dump_query_results(idx, "ac.{2,5}?a.{4,8}?b");
            dump_query_results(idx, "a.{0,10}?a.{0,10}?a");
            dump_query_results(idx, "foo.{0,10}?bar");
        }
        {  // int alphabet
            vlg_index<int_alphabet_tag> idx;
            if (argc <= 1) {
                construct_im(idx, vector<uint64_t>({97, 98, 114, 97, 99, 97, 100, 97, 98, 114, 97, 115, 105, 109, 115, 97, 108, 97, 98, 105, 109}));
            } else {
                string filename(argv[1]);
                construct(idx, filename, 0);
                ofstream out(filename + ".info.html");
                write_structure<HTML_FORMAT>(idx, out);
            }
            dump_query_results(idx, "97 99 .{2,5}? 97 .{4,8}? 98");
            dump_query_results(idx, "97 .{0,10}? 97 .{0,10}? 97");
            dump_query_results(idx, "1337 .{0,10}? 42");
        }
    } catch (const std::exception& e) {
        cerr << "vlg_matching: " << e.what() << endl;  // no GPU, bad file, bad query: fail loudly
        return 1;
    }
    return 0;
}
