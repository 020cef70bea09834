// libbdsg stand-in: declarations only, every method throws.  TEST INFRASTRUCTURE (see ../README.md).
// Snarl decomposition from the graph is outside the path; the harness hands the analyzers ready-made Snarl_data_t.
#pragma once
#include <functional>
#include <set>
#include <map>
#include <unordered_set>
#include <climits>
#include <cassert>
#include <stdexcept>
#include <string>
#include <handlegraph/handle_graph.hpp>
#include <handlegraph/path_handle_graph.hpp>
#include <handlegraph/path_position_handle_graph.hpp>
#include <handlegraph/snarl_decomposition.hpp>
#include <handlegraph/util.hpp>

namespace bdsg {
using handlegraph::handle_t;
using handlegraph::net_handle_t;
using handlegraph::nid_t;
using handlegraph::path_handle_t;
using handlegraph::step_handle_t;

[[noreturn]] inline void ref_stub_called() { throw std::logic_error("oracle/_ref: libbdsg stand-in called"); }

class SnarlDistanceIndex {
public:
    enum endpoint_t { START, END, TIP };
    void deserialize(const std::string&) { ref_stub_called(); }
    size_t get_depth(const net_handle_t&) const { ref_stub_called(); }
    net_handle_t get_bound(const net_handle_t&, bool, bool) const { ref_stub_called(); }
    net_handle_t get_node_from_sentinel(const net_handle_t&) const { ref_stub_called(); }
    net_handle_t get_parent(const net_handle_t&) const { ref_stub_called(); }
    net_handle_t get_root() const { ref_stub_called(); }
    nid_t node_id(const net_handle_t&) const { ref_stub_called(); }
    size_t node_length(const net_handle_t&) const { ref_stub_called(); }
    size_t minimum_length(const net_handle_t&) const { ref_stub_called(); }
    size_t maximum_length(const net_handle_t&) const { ref_stub_called(); }
    endpoint_t ends_at(const net_handle_t&) const { ref_stub_called(); }
    bool starts_at_start(const net_handle_t&) const { ref_stub_called(); }
    bool is_node(const net_handle_t&) const { ref_stub_called(); }
    bool is_root(const net_handle_t&) const { ref_stub_called(); }
    bool is_snarl(const net_handle_t&) const { ref_stub_called(); }
    bool is_chain(const net_handle_t&) const { ref_stub_called(); }
    bool is_trivial_chain(const net_handle_t&) const { ref_stub_called(); }
    bool is_sentinel(const net_handle_t&) const { ref_stub_called(); }
    std::string net_handle_as_string(const net_handle_t&) const { ref_stub_called(); }
    handle_t get_handle(const net_handle_t&, const handlegraph::HandleGraph*) const { ref_stub_called(); }
    template <class F> bool for_each_child(const net_handle_t&, const F&) const { ref_stub_called(); }
    template <class F> bool follow_net_edges(const net_handle_t&, const handlegraph::HandleGraph*, bool, const F&) const {
        ref_stub_called();
    }
};

class PositionOverlay {
public:
    PositionOverlay() = default;
    explicit PositionOverlay(const handlegraph::PathHandleGraph*) { ref_stub_called(); }
    size_t get_position_of_step(const step_handle_t&) const { ref_stub_called(); }
};
class PackedPositionOverlay : public PositionOverlay {};

}  // namespace bdsg
