// ORACLE -- TEST INFRASTRUCTURE ONLY (oracle/ref_shim): libhandlegraph's position types, restated
// (pos_t = (node id, is_reverse, offset) tuple with make_pos_t / id / is_rev / offset accessors).
#pragma once
#include <cstddef>
#include <cstdint>
#include <tuple>
namespace handlegraph {
typedef long long int nid_t;
typedef std::size_t offset_t;
typedef std::tuple<nid_t, bool, offset_t> pos_t;
inline pos_t make_pos_t(const nid_t &id, bool is_rev, const offset_t &off) { return std::make_tuple(id, is_rev, off); }
inline nid_t id(const pos_t &pos) { return std::get<0>(pos); }
inline bool is_rev(const pos_t &pos) { return std::get<1>(pos); }
inline offset_t offset(const pos_t &pos) { return std::get<2>(pos); }
} // namespace handlegraph
