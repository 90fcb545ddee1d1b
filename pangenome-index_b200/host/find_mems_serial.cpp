// find_mems_serial — the reference's own main loop (src/find_mems.cpp:13-147: one read at a time, find_all_mems,
// print, per-MEM TagArray query that prints) written against the facade header instead of the reference's headers.
// It exists to show that reference-style code compiles and produces the same stdout through the C ABI; it issues one
// small GPU launch per call and is therefore slow — the batched driver is find_mems.cpp.
//   find_mems_serial <r_index.ri> <tags> <reads.txt> <min_mem_length> <min_occ> [compact|bytecode]
#include <fstream>
#include <iostream>
#include <string>
#include "panindexer.hpp"

using namespace panindexer;

int main(int argc, char **argv) {
    if (argc < 6) { std::cerr << "usage: find_mems_serial <ri> <tags> <reads> <min_len> <min_occ> [compact|bytecode]" << std::endl; return EXIT_FAILURE; }
    std::string r_index_file = argv[1], tag_array_index = argv[2], reads_file = argv[3];
    size_t mem_length = std::stoi(argv[4]), min_occ = std::stoi(argv[5]);
    const bool bytecode = argc > 6 && std::string(argv[6]) == "bytecode";
    try {
        FastLocate r_index;
        {
            std::ifstream rin(r_index_file, std::ios::binary);
            if (!rin) { std::cerr << "Cannot open r-index: " << r_index_file << std::endl; std::exit(EXIT_FAILURE); }
            r_index.load_encoded(rin);
        }
        TagArray tag_array;
        std::ifstream in_ds(tag_array_index, std::ios::binary);
        if (bytecode) tag_array.load_compressed_tags(in_ds); else tag_array.load_compressed_tags_compact(in_ds);
        std::ifstream reads(reads_file);
        if (!reads) { std::cerr << "Cannot open reads file: " << reads_file << std::endl; std::exit(EXIT_FAILURE); }
        std::string read; int i = 0;
        while (std::getline(reads, read)) {
            if (read.empty()) continue;
            i++;
            auto mems = find_all_mems(read, mem_length, min_occ, r_index);
            std::cout << "Seq: " << i << std::endl;
            for (const auto &mem : mems) {
                std::cout << "MEM START: " << mem.start << ", MEM END: " << mem.end << " BWT START: " << mem.bwt_start << " SIZE: " << mem.size << std::endl;
                size_t tag_nums = 0;
                if (bytecode) tag_array.query_compressed(mem.bwt_start, mem.bwt_start + mem.size - 1, tag_nums);
                else tag_array.query_compressed_compact(mem.bwt_start, mem.bwt_start + mem.size - 1, tag_nums);
            }
            std::cout << std::endl;
        }
    } catch (const std::exception &e) { std::cerr << "find_mems_serial: " << e.what() << std::endl; return EXIT_FAILURE; }
    return 0;
}
