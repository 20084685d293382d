// Minimal YAML reader for the Cantera-YAML dialect used by the reference's mechanisms/
// directory (block maps/sequences, flow maps/sequences that may span lines, '#' comments,
// block scalars, quoted scalars).  Replaces the YAML front end of Cantera::newSolution that
// the reference reaches through CanteraMixture (src/dfCanteraMixture/CanteraMixture.C:65).
#pragma once
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace dfchem {
namespace yaml {

struct Node;
using NodePtr = std::shared_ptr<Node>;

struct Node {
    enum Kind { Null, Scalar, Seq, Map } kind = Null;
    std::string scalar;
    std::vector<NodePtr> seq;
    std::vector<std::pair<std::string, NodePtr>> map;  // insertion-ordered

    bool has(const std::string& key) const;
    const Node& at(const std::string& key) const;  // throws if missing
    const Node& at(size_t i) const;
    size_t size() const { return kind == Seq ? seq.size() : kind == Map ? map.size() : 0; }
    double as_double() const;
    const std::string& as_string() const;
    bool is_scalar() const { return kind == Scalar; }
    bool is_seq() const { return kind == Seq; }
    bool is_map() const { return kind == Map; }
};

NodePtr parse(const std::string& text);
NodePtr parse_file(const std::string& path);

}  // namespace yaml
}  // namespace dfchem
