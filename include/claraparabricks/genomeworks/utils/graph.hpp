// B200 cudapoa engine -- graph containers returned by Batch::get_graphs().
// API-compatible with GenomeWorks common/base/include/claraparabricks/genomeworks/utils/graph.hpp:28-277
// (Graph, DirectedGraph, UndirectedGraph; same member names, same DOT text).
#pragma once

#include <cstdint>
#include <functional>
#include <sstream>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

namespace claraparabricks
{
namespace genomeworks
{

/// Hash of a pair as the XOR of the two member hashes.
struct PairHash
{
    template <class A, class B>
    size_t operator()(const std::pair<A, B>& p) const
    {
        return std::hash<A>()(p.first) ^ std::hash<B>()(p.second);
    }
};

/// Adjacency + weighted edge set + optional node labels.
class Graph
{
public:
    using node_id_t     = int32_t;
    using edge_weight_t = int32_t;
    using edge_t        = std::pair<node_id_t, node_id_t>;

    /// Nodes reachable over one edge from `node` (empty if it has none).
    const std::vector<node_id_t>& get_adjacent_nodes(node_id_t node) const
    {
        auto it = adjacent_nodes_.find(node);
        return it == adjacent_nodes_.end() ? empty_ : it->second;
    }

    /// Ids of all nodes that have at least one adjacent node.
    const std::vector<node_id_t> get_node_ids() const
    {
        std::vector<node_id_t> ids;
        ids.reserve(adjacent_nodes_.size());
        for (const auto& kv : adjacent_nodes_)
        {
            ids.push_back(kv.first);
        }
        return ids;
    }

    /// All edges with their weights.
    const std::vector<std::pair<edge_t, edge_weight_t>> get_edges() const
    {
        return std::vector<std::pair<edge_t, edge_weight_t>>(edges_.begin(), edges_.end());
    }

    /// Attach a label to a node id (first label wins).
    void set_node_label(node_id_t node, const std::string& label) { node_labels_.insert({node, label}); }

    /// Label of a node, or "" when none was set.
    std::string get_node_label(node_id_t node) const
    {
        auto it = node_labels_.find(node);
        return it == node_labels_.end() ? std::string() : it->second;
    }

protected:
    bool directed_edge_exists(edge_t edge) { return edges_.count(edge) != 0; }

    void update_adject_nodes(edge_t edge) { adjacent_nodes_[edge.first].push_back(edge.second); }

    void node_labels_to_dot(std::ostringstream& dot_str) const
    {
        for (const auto& kv : node_labels_)
        {
            dot_str << kv.first << " [label=\"" << kv.second << "\"];\n";
        }
    }

    void edges_to_dot(std::ostringstream& dot_str, const std::string& node_separator) const
    {
        for (const auto& kv : edges_)
        {
            dot_str << kv.first.first << " " << node_separator << " " << kv.first.second;
            dot_str << " [label=\"" << kv.second << "\"];\n";
        }
    }

    std::unordered_map<node_id_t, std::vector<node_id_t>> adjacent_nodes_;
    std::unordered_map<edge_t, edge_weight_t, PairHash> edges_;
    std::unordered_map<node_id_t, std::string> node_labels_;
    const std::vector<node_id_t> empty_;
};

/// Directed graph; duplicate edges are ignored.
class DirectedGraph : public Graph
{
public:
    using Graph::edge_t;
    using Graph::edge_weight_t;
    using Graph::node_id_t;

    void add_edge(node_id_t node_id_from, node_id_t node_id_to, edge_weight_t weight = 0)
    {
        edge_t e(node_id_from, node_id_to);
        if (directed_edge_exists(e))
        {
            return;
        }
        edges_.insert({e, weight});
        update_adject_nodes(e);
    }

    std::string serialize_to_dot() const
    {
        std::ostringstream dot_str;
        dot_str << "digraph g {\n";
        node_labels_to_dot(dot_str);
        edges_to_dot(dot_str, "->");
        dot_str << "}\n";
        return dot_str.str();
    }
};

/// Undirected graph; an edge and its reverse count as the same edge.
class UndirectedGraph : public Graph
{
public:
    using Graph::edge_t;
    using Graph::edge_weight_t;
    using Graph::node_id_t;

    void add_edge(node_id_t node_id_from, node_id_t node_id_to, edge_weight_t weight = 0)
    {
        edge_t e(node_id_from, node_id_to);
        edge_t r(node_id_to, node_id_from);
        if (directed_edge_exists(e) || directed_edge_exists(r))
        {
            return;
        }
        edges_.insert({e, weight});
        update_adject_nodes(e);
        update_adject_nodes(r);
    }

    std::string serialize_to_dot() const
    {
        std::ostringstream dot_str;
        dot_str << "graph g {\n";
        node_labels_to_dot(dot_str);
        edges_to_dot(dot_str, "--");
        dot_str << "}\n";
        return dot_str.str();
    }
};

} // namespace genomeworks
} // namespace claraparabricks
