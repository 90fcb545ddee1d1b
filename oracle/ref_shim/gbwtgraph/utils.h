// ORACLE -- TEST INFRASTRUCTURE ONLY (oracle/ref_shim): the GBWTGraph declarations the reference's tag arrays use.
// Position::encode is THE one mapping on the find_mems path that no file or test of the reference pins (gbwtgraph is
// not vendored): value = id << 11 | is_rev << 10 | (offset & 0x3FF), the published gbwtgraph layout, which the
// reference's own encode_run_length_compact mirrors (src/tag_arrays.cpp:38-45). It STAYS A SHIM.
#pragma once
#include <cstdint>
#include <string>
#include <cassert>
#include <map>
#include <set>
#include <unordered_map>
#include <unordered_set>
#include <functional>
#include "../gbwt/utils.h"
#include "../handlegraph/types.hpp"
namespace gbwtgraph {
using handlegraph::nid_t; using handlegraph::offset_t; using handlegraph::pos_t;
using handlegraph::make_pos_t; using handlegraph::id; using handlegraph::is_rev; using handlegraph::offset;
struct Position {
    typedef std::uint64_t code_type;
    code_type value;
    constexpr static std::size_t OFFSET_BITS = 10, REV_OFFSET = 10, ID_OFFSET = 11;
    constexpr static code_type OFF_MASK = 0x3FF, REV_MASK = 0x400;
    constexpr static code_type NO_VALUE = 0;
    static Position no_value() { return Position{NO_VALUE}; }
    static Position encode(pos_t pos) {
        return Position{((code_type) handlegraph::id(pos) << ID_OFFSET) | ((code_type) handlegraph::is_rev(pos) << REV_OFFSET) | ((code_type) handlegraph::offset(pos) & OFF_MASK)};
    }
    std::uint64_t id() const { return value >> ID_OFFSET; }
    bool is_rev() const { return value & REV_MASK; }
    std::uint64_t offset() const { return value & OFF_MASK; }
    pos_t decode() const { return make_pos_t((nid_t) id(), is_rev(), (offset_t) offset()); }
    bool operator==(Position o) const { return value == o.value; }
    bool operator!=(Position o) const { return value != o.value; }
    bool operator<(Position o) const { return value < o.value; }
};
} // namespace gbwtgraph
