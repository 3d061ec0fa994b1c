// TEST INFRASTRUCTURE ONLY (oracle). i2l shim — tree. PARITY UNPINNED (i2l absent).
//
// Only what place.cpp touches: get_node_count() (place.cpp:92,99,166) and
// get_by_postorder_id(id) -> optional-like of pointer-like with get_branch_length()
// (place.cpp:103,110,429,435). Newick parsing lives in the real i2l and is not restated here: the
// oracle's tree is built from per-node arrays indexed by post-order id.
#ifndef I2L_SHIM_PHYLO_TREE_H
#define I2L_SHIM_PHYLO_TREE_H

#include <optional>
#include <i2l/phylo_kmer.h>

namespace i2l
{
    class phylo_node
    {
    public:
        using id_type = int;
        using branch_length_type = double;

        phylo_node(id_type postorder_id, branch_length_type branch_length)
            : _postorder_id(postorder_id), _branch_length(branch_length) {}

        id_type get_postorder_id() const noexcept { return _postorder_id; }
        branch_length_type get_branch_length() const noexcept { return _branch_length; }

    private:
        id_type _postorder_id;
        branch_length_type _branch_length;
    };

    class phylo_tree
    {
    public:
        explicit phylo_tree(const std::vector<double>& branch_lengths)
        {
            _nodes.reserve(branch_lengths.size());
            for (size_t i = 0; i < branch_lengths.size(); ++i)
            {
                _nodes.emplace_back(static_cast<phylo_node::id_type>(i), branch_lengths[i]);
            }
        }

        size_t get_node_count() const noexcept { return _nodes.size(); }

        std::optional<const phylo_node*> get_by_postorder_id(phylo_node::id_type postorder_id) const noexcept
        {
            if (postorder_id < 0 || static_cast<size_t>(postorder_id) >= _nodes.size())
            {
                return std::nullopt;
            }
            return &_nodes[static_cast<size_t>(postorder_id)];
        }

    private:
        std::vector<phylo_node> _nodes;
    };
}

#endif
