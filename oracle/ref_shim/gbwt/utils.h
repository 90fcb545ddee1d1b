// ORACLE -- TEST INFRASTRUCTURE ONLY (oracle/ref_shim): stand-in for the few GBWT declarations the reference's
// src/r-index.cpp and src/tag_arrays.cpp use (gbwt is neither vendored nor installed; see ../sdsl_shim.hpp).
// Restated published semantics only: size_type / byte_type / range_type, ByteCode (7 data bits per byte, low group
// first, bit 7 = continue), Verbosity, readTimer, parallelQuickSort, Version::R_INDEX_VERSION = 1 (the shipped .ri
// files carry version 1).
#pragma once
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <iostream>
#include <string>
#include <utility>
#include <vector>
#include <omp.h>
#include "../sdsl_shim.hpp"

namespace gbwt {

typedef std::uint64_t size_type;
typedef std::uint8_t byte_type;
typedef std::uint32_t node_type;
typedef std::pair<size_type, size_type> range_type;

struct Version { constexpr static std::uint32_t R_INDEX_VERSION = 1; };

struct Verbosity {
    static size_type level;
    constexpr static size_type SILENT = 0, BASIC = 1, EXTENDED = 2, FULL = 3, DEFAULT = SILENT;
    static void set(size_type l) { level = l; }
};

inline double readTimer() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
inline double inMegabytes(size_type bytes) { return bytes / (1024.0 * 1024.0); }
inline double inGigabytes(size_type bytes) { return bytes / (1024.0 * 1024.0 * 1024.0); }
inline void printHeader(const std::string &header, size_type indent = 18) {
    std::string padding; if (header.length() + 1 < indent) padding = std::string(indent - 1 - header.length(), ' ');
    std::cout << header << ":" << padding;
}

template <class Iterator, class Comparator> inline void parallelQuickSort(Iterator first, Iterator last, const Comparator &comp) { std::stable_sort(first, last, comp); }
template <class Iterator> inline void parallelQuickSort(Iterator first, Iterator last) { std::stable_sort(first, last); }
template <class Iterator, class Comparator> inline void sequentialSort(Iterator first, Iterator last, const Comparator &comp) { std::stable_sort(first, last, comp); }
template <class Iterator> inline void sequentialSort(Iterator first, Iterator last) { std::stable_sort(first, last); }

struct ByteCode {
    typedef gbwt::size_type value_type;
    typedef gbwt::byte_type code_type;
    constexpr static size_type DATA_BITS = 7;
    constexpr static code_type DATA_MASK = 0x7F, NEXT_BYTE = 0x80;
    // Bytes past the end of the array read as 0 (undefined in GBWT; defined like oracle/).
    template <class ByteArray> static value_type read(ByteArray &array, size_type &i) {
        value_type res = 0; size_type offset = 0;
        for (;;) {
            const code_type b = i < (size_type) array.size() ? (code_type) array[i] : (code_type) 0; ++i;
            if (offset < 64) res |= ((value_type) (b & DATA_MASK)) << offset;
            offset += DATA_BITS;
            if (!(b & NEXT_BYTE)) return res;
        }
    }
    template <class ByteArray> static void write(ByteArray &array, value_type value) {
        while (value > DATA_MASK) { array.push_back((value & DATA_MASK) | NEXT_BYTE); value >>= DATA_BITS; }
        array.push_back(value);
    }
};

} // namespace gbwt
